// assemble_run.cuh — fills AssembleArgs from the handle and launches the kernels of include/b2f/assemble.cuh for one
// operator type.  Included by one translation unit per operator family (assemble_laplace.cu, assemble_elasticity.cu,
// assemble_heat.cu) so that the kernel instantiations compile in parallel.
#pragma once
#include "internal.h"
#include "b2f/assemble.cuh"

namespace b2f {
template <class Op>
inline int run(b2f_context_s* h, const Op& op, int accumulate, cudaStream_t stream, const double* qdot)
  {
    AssembleArgs A;
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
    A.accumulate = accumulate;
    int variant = h->variant;
    if (variant == B2F_ASM_DEFAULT)
      variant = B2F_ASM_NODE_GATHER;
    // constant source fused into the walk (value per cell and node = qdot * measure / NV, ArcaneFemFunctionsGpu.h:402-429):
    // the fan kernel sums 6*volume / 2*area, hence the factor 1/(6*4) / 1/(2*3)
    if (h->range_rows >= 0) { // host pipeline: rows [range_first_row, range_first_row + range_rows), multiples of the tile height
      A.first_tile = h->range_first_row / FanCfg<Op>::ROWS;
      A.nb_tile = (h->range_rows + FanCfg<Op>::ROWS - 1) / FanCfg<Op>::ROWS;
    }
    const bool fused = qdot != nullptr && variant == B2F_ASM_NODE_GATHER && h->fan_ok && h->mesh.cell_is_own == nullptr;
    if (fused) {
      A.rhs = h->rhs.as<double>();
      A.rhs_factor = *qdot / (Op::NV == 4 ? 24.0 : 6.0);
    }
    B2F_CUDA(launch_assemble<Op>(A, op, variant, h->per_block, h->nb_nz, stream));
    if (qdot != nullptr && !fused)
      return rhs_constant_source(h, *qdot, stream);
    return B2F_OK;
  }

} // namespace b2f
