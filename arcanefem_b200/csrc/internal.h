// internal.h — handle layout and helpers shared by the translation units of libb200fem.so.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "b2f.h"

namespace b2f {

void set_error(const char* fmt, ...);
int fail(int status, const char* fmt, ...);
int fail_cuda(cudaError_t e, const char* what);

#define B2F_CUDA(expr)                                   \
  do {                                                   \
    cudaError_t b2f_e__ = (expr);                        \
    if (b2f_e__ != cudaSuccess)                          \
      return ::b2f::fail_cuda(b2f_e__, #expr);           \
  } while (0)

// Owning device buffer (plain cudaMalloc: buffers live as long as the pattern; temporaries are freed
// explicitly by b2f_release_workspace).
struct DevBuf
{
  void* p = nullptr;
  size_t bytes = 0;
  cudaError_t alloc(size_t n)
  {
    if (p && bytes >= n && n * 2 >= bytes)
      return cudaSuccess;
    release();
    if (n == 0)
      n = 16;
    cudaError_t e = cudaMalloc(&p, n);
    if (e == cudaSuccess)
      bytes = n;
    else
      p = nullptr;
    return e;
  }
  void release()
  {
    if (p)
      cudaFree(p);
    p = nullptr;
    bytes = 0;
  }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct GhostPlan
{
  bool ready = false;
  int rank = 0, world = 1;
  // per peer: rows / blocks this rank sends (its ghost rows owned by the peer) and receives (its own rows that are ghosts on the peer)
  std::vector<int64_t> send_rows_n, recv_rows_n, send_blocks, recv_blocks;
  std::vector<int64_t> send_off, recv_off; // offsets (in blocks) of every peer's part of the packed buffers
  DevBuf send_index, recv_index, send_vals, recv_vals; // value index per exchanged scalar; packed values
  DevBuf recv_dst, recv_src; // recv_index in a stable sorted order + the position of each entry in recv_vals (fixed summation order)
  int64_t total_send = 0, total_recv = 0; // scalars
  int64_t send_row_lo = 0, send_row_hi = 0; // span of the local rows that are sent (ghost rows an assembled cell touches)
  int64_t recv_val_lo = 0, recv_val_hi = -1; // span (inclusive) of the value indices that receive partial sums
};

} // namespace b2f

struct b2f_context_s
{
  b2f_mesh_view mesh{};
  b2f_options opt{};
  int bs = 1, nv = 4, variant = B2F_ASM_NODE_GATHER;
  bool per_block = true;
  int diag_last = 0; // b2f_pattern_layout: rows are stored neighbours first, diagonal last
  int device = 0;

  bool has_pattern = false;
  bool pattern_by_lists = false; // the pattern came from the row-wise build (no global sort)
  bool has_csr = false;
  int64_t nb_col = 0, nb_nz = 0, nb_edge = 0, nb_dof = 0;

  b2f::DevBuf row_index, columns, nb_nz_per_row, values; // BSRMatrix arrays
  b2f::DevBuf nc_ptr, nc_list, nc_slots; // node -> cell lists and the column-slot plan of the gather kernel
  bool slots_ok = false;
  b2f::DevBuf fan_ptr, fan_steps, fan_cell; // fan-walk plan of the default gather kernel (one word per step, cell id per step)
  // b2f_reorder: the plan above is stored in the library's own node order; internal -> caller row, the internal copies of
  // row_index / columns (column ids in the internal numbering) and of the coordinates (refreshed by every assembly)
  bool reordered = false;
  b2f::DevBuf perm_rows, r_row_index, r_columns, r_coords, r_slot_map;
  bool inner = false; // the shadow context reorder_plan builds the internal pattern + plan in: no value array, no system state
  bool fan_ok = false;
  bool fan_lockstep = false; // step words converted to the lockstep form (sparsity.cu k_fan_lockstep)
  double fan_lockstep_bodies = 0.0; // step bodies a warp would execute per warp step in the lockstep form (1 = rows in step)
  int64_t nb_fan_step = 0;
  int64_t nb_pair = 0; // node/cell incidences of the assembled cells
  int tile_extents[6] = { 0, 0, 0, 0, 0, 0 }; // largest gather tile: {blocks, adj} for 128-row tiles, then 64-row, then 32-row tiles
  b2f::DevBuf csr_rows, csr_nbcol, csr_cols; // BS > 1 CSR expansion
  b2f::DevBuf ij_flag, ij_off, ij_ncols, ij_rows, ij_row_local, ij_cols, ij_values, ij_rhs; // own-row arrays of b2f_to_hypre_ij
  b2f::DevBuf rhs, forced_info, forced_value, elim_info, elim_value; // linear-system state
  b2f::DevBuf counters; // 12 x u64 ([0] not found, [1] fallback tiles, [2] bad cells, [3] flags, [4..6] tile extents, [8..9] lockstep cost)
  b2f::DevBuf ws_keys_a, ws_keys_b, ws_tmp, ws_i32a, ws_i64; // workspace
  b2f::DevBuf coords_stage; // b2f_assemble_host
  // b2f_assemble_host pipeline: row chunks, the highest node id each chunk reads, value offsets at the chunk boundaries
  int64_t range_first_row = 0, range_rows = -1; // >= 0 while a chunk of the pipeline is being launched
  bool fan_first = false; // the step words carry the first-visit marks (k_fan_first_flags)
  bool range_keep_counters = false; // a later range of the same assembly: the counters are not cleared again
  cudaStream_t comm_stream = nullptr; // b2f_assemble_exchange: ghost rows travel here while the other rows are assembled
  cudaEvent_t comm_events[2] = { nullptr, nullptr };
  std::vector<int64_t> chunk_row, chunk_value_off, chunk_node_hi;
  cudaStream_t pipe_h2d = nullptr, pipe_d2h = nullptr;
  std::vector<cudaEvent_t> pipe_events;
  b2f::GhostPlan ghost;
  int64_t last_not_found = 0;
  int64_t last_fallback_tiles = 0;
};

namespace b2f {
// sparsity.cu
int build_pattern(b2f_context_s* h, cudaStream_t stream);
// sparsity.cu
int reorder_plan(b2f_context_s* h, cudaStream_t stream);
int permute_coords(b2f_context_s* h, cudaStream_t stream);
// assemble.cu
struct AssembleArgs;
void fill_assemble_args(const b2f_context_s* h, AssembleArgs* A); // pointers, plan and counters of the handle (accumulate = 0, whole matrix)
int assemble(b2f_context_s* h, const b2f_operator* op, cudaStream_t stream, const double* qdot = nullptr); // qdot: fused constant source
// boundary.cu
int rhs_fill(b2f_context_s* h, double v, cudaStream_t s);
int clear_values(b2f_context_s* h, cudaStream_t s);
int rhs_constant_source(b2f_context_s* h, double qdot, cudaStream_t s);
int rhs_neumann(b2f_context_s* h, const b2f_neumann* bc, cudaStream_t s);
int rhs_nodal_source(b2f_context_s* h, const double* node_field, double scale, const uint8_t* node_skip, cudaStream_t s);
int add_face_mass(b2f_context_s* h, int64_t nb_face, const int32_t* face_node, double coefficient, cudaStream_t s);
int dirichlet_flag_penalty(b2f_context_s* h, const int32_t* ids, int64_t n, int dof_index, double value, double penalty, int weak, cudaStream_t s);
int dirichlet_flag_elimination(b2f_context_s* h, const int32_t* ids, int64_t n, int dof_index, double value, int mode, cudaStream_t s);
int dirichlet_apply(b2f_context_s* h, cudaStream_t s);
// convert.cu
int to_csr(b2f_context_s* h, cudaStream_t s);
int csr_global_numbering(b2f_context_s* h, const int32_t* dof_number, int32_t* rows_global, int32_t* cols_global, cudaStream_t s);
int to_hypre_ij(b2f_context_s* h, const int32_t* dof_number, b2f_ij_view* out, cudaStream_t s);
int pipeline_plan(b2f_context_s* h, int nchunk, cudaStream_t s);
int spmv(b2f_context_s* h, const double* x, double* y, cudaStream_t s);
// ghost.cu
int ghost_plan(b2f_context_s* h, void* comm, int rank, int world, const int32_t* node_owner, const int64_t* node_gid, cudaStream_t s);
int ghost_exchange(b2f_context_s* h, void* comm, cudaStream_t s);
int ghost_send_recv(b2f_context_s* h, void* comm, cudaStream_t s);
int ghost_add(b2f_context_s* h, cudaStream_t s);
} // namespace b2f
