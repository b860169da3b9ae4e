// assemble_args.h — argument block of the assembly kernels (include/b2f/assemble.cuh).  Plain struct, no device code: it is
// filled by the library from a handle (b2f_user_assemble_begin in b2f.h) and passed by value to the kernels, both by the
// library's own operators and by user-written ones (include/b2f/user_operator.cuh).
#pragma once
#include <cstdint>

namespace b2f {

struct AssembleArgs
{
  const int32_t* cell_node;
  const double* coords;
  const uint8_t* node_is_own; // may be null
  const uint8_t* cell_is_own; // may be null; 0 = ghost cell (pattern only, not assembled)
  long long nb_cell;
  long long nb_node;
  const int32_t* row_index; // nb_node + 1
  const int32_t* columns;
  double* values;
  const int32_t* nc_ptr; // node -> cell list offsets, nb_node + 1
  const uint32_t* nc_list; // entries: cell * 4 + local index of the node in the cell
  const uint32_t* nc_slots; // per node->cell entry: column slots of the cell's other vertices, 10 bits each (see k_node_gather_slots)
  int slots_ok; // nc_slots is usable (every row has at most 1023 blocks)
  const int32_t* fan_ptr; // node -> fan-walk step offsets, nb_node + 1
  const uint32_t* fan_steps; // one word per step (see k_node_fan)
  const uint32_t* fan_cell; // cell id of every step (read by operators with per-cell data only)
  int fan_ok; // the fan plan is usable
  int tile_extents[6]; // largest tile of the plan: {blocks, steps or pairs} for 128-, 64- and 32-row tiles
  unsigned long long* counters; // [0] columns not found, [1] gather tiles on the global fallback path
  int accumulate;
  int fan_lockstep = 0; // the step words are in the lockstep form (tets)
  long long first_tile = 0; // fan kernel only: the launch covers the tiles [first_tile, first_tile + nb_tile) (host pipeline)
  long long nb_tile = -1; // -1: every tile
  double* rhs = nullptr; // fused constant source (fan kernel): rhs[row * BS] += rhs_factor * (sum of 6*volume / 2*area of the incident cells)
  double rhs_factor = 0.0;
  int diag_last = 0; // rows are stored neighbours first (ascending), diagonal last (b2f_pattern_layout)
  // Reorder mode of the fan kernel (b2f_options.reorder): the kernel walks tiles of INTERNAL rows (a space-filling order of the
  // nodes); row_index / columns / coords / fan_* above are then the internal copies (rows permuted, column ids and coordinates in
  // the internal numbering) and the values are written to the caller's layout through row_perm / out_row_index.
  // filled from the handle (null without reordering); every kernel but the fan kernel ignores them:
  const int32_t* r_perm = nullptr; // internal row -> caller row
  const int32_t* r_row_index = nullptr; // internal copies: rows permuted, column ids and coordinates in the internal numbering
  const int32_t* r_columns = nullptr;
  const double* r_coords = nullptr;
  const uint16_t* r_slot_map = nullptr; // per internal block: its slot in the caller's row
  // set by the fan kernel's launcher, which swaps the internal copies into row_index / columns / coords:
  const int32_t* row_perm = nullptr; // internal row -> caller row; null = rows are walked in the caller's order
  const int32_t* out_row_index = nullptr; // caller row_index: where a row's values go
  const int32_t* caller_columns = nullptr; // caller-numbered columns / coordinates for the tiles that take the global fallback path
  const double* caller_coords = nullptr;
  const uint16_t* slot_map = nullptr;
  // 8192 when the step words carry first-visit marks AND this launch may use them (overwrite mode, every cell assembled): the
  // first load of a slot then starts from zero and the slab is not cleared; 0 otherwise
  uint32_t fan_first_mask = 0;
};

enum { VARIANT_CELL_ATOMIC = 1, VARIANT_CELL_ATOMIC_WARPAGG = 2, VARIANT_NODE_GATHER = 3, VARIANT_NODE_GATHER_SLOTS = 4 };

} // namespace b2f
