// assemble.cu — dispatch of b2f_assemble onto the kernel templates of include/b2f/assemble.cuh.
// Replaces BSRFormat::assembleBilinear and its four kernels (femutils/BSRFormat.h:786-1111) together with the
// shipped element callbacks (modules/poisson/FemModule.h:139-186, modules/elasticity/FemModule.h:112-152,
// modules/heat/FemModule.cc:678-770).
#include "internal.h"
#include "b2f/assemble_args.h"

namespace b2f {

void fill_assemble_args(const b2f_context_s* h, AssembleArgs* out)
{
  AssembleArgs A{};
  A.cell_node = h->mesh.cell_node;
  A.coords = h->mesh.coords;
  A.node_is_own = h->mesh.node_is_own;
  A.cell_is_own = h->mesh.cell_is_own;
  A.nb_cell = h->mesh.nb_cell;
  A.nb_node = h->mesh.nb_node;
  A.row_index = h->row_index.as<int32_t>();
  A.columns = h->columns.as<int32_t>();
  A.values = h->values.as<double>();
  A.nc_ptr = h->nc_ptr.as<int32_t>();
  A.nc_list = h->nc_list.as<uint32_t>();
  A.nc_slots = h->nc_slots.as<uint32_t>();
  A.slots_ok = h->slots_ok ? 1 : 0;
  A.fan_ptr = h->fan_ptr.as<int32_t>();
  A.fan_steps = h->fan_steps.as<uint32_t>();
  A.fan_cell = h->fan_cell.as<uint32_t>();
  A.fan_ok = h->fan_ok ? 1 : 0;
  A.fan_lockstep = h->fan_lockstep ? 1 : 0;
  for (int i = 0; i < 6; ++i)
    A.tile_extents[i] = h->tile_extents[i];
  A.counters = h->counters.as<unsigned long long>();
  A.accumulate = 0;
  A.diag_last = h->diag_last;
  // reorder mode: the fan kernel's launcher swaps the internal copies in (launch_assemble_layout); every other kernel keeps
  // the caller's arrays above
  if (h->reordered) {
    A.r_perm = h->perm_rows.as<int32_t>();
    A.r_row_index = h->r_row_index.as<int32_t>();
    A.r_columns = h->r_columns.as<int32_t>();
    A.r_coords = h->r_coords.as<double>();
    A.r_slot_map = h->r_slot_map.as<uint16_t>();
  }
  *out = A;
}

// one translation unit per operator family (parallel compilation of the kernel instantiations)
int run_laplace(b2f_context_s* h, int accumulate, cudaStream_t stream, const double* qdot);
int run_elasticity(b2f_context_s* h, double lambda, double mu2, int accumulate, cudaStream_t stream, const double* qdot);
int run_heat(b2f_context_s* h, double lambda, double inv_dt, const double* cell_lambda, int accumulate, cudaStream_t stream, const double* qdot);

int assemble(b2f_context_s* h, const b2f_operator* op, cudaStream_t stream, const double* qdot)
{
  const int nv = h->nv, bs = h->bs;
  if (h->range_rows < 0 ? true : (h->range_first_row == 0 && !h->range_keep_counters))
    B2F_CUDA(cudaMemsetAsync(h->counters.p, 0, sizeof(unsigned long long) * 2, stream));
  switch (op->kind) {
  case B2F_OP_LAPLACE:
    if (bs != 1)
      return fail(B2F_ERR_INVALID_ARGUMENT, "B2F_OP_LAPLACE needs block_size 1 (got %d)", bs);
    return run_laplace(h, op->accumulate, stream, qdot);
  case B2F_OP_ELASTICITY:
    if ((nv == 3 && bs == 2) || (nv == 4 && bs == 3))
      return run_elasticity(h, op->lambda, op->mu2, op->accumulate, stream, qdot);
    return fail(B2F_ERR_INVALID_ARGUMENT, "B2F_OP_ELASTICITY needs block_size == dim (dim=%d, block_size=%d)", h->mesh.dim, bs);
  case B2F_OP_HEAT:
    if (bs != 1)
      return fail(B2F_ERR_INVALID_ARGUMENT, "B2F_OP_HEAT needs block_size 1 (got %d)", bs);
    return run_heat(h, op->lambda, op->inv_dt, op->cell_lambda, op->accumulate, stream, qdot);
  default:
    return fail(B2F_ERR_NOT_IMPLEMENTED, "unknown operator kind %d", op->kind);
  }
}

} // namespace b2f
