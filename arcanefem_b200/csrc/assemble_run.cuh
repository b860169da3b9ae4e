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
    fill_assemble_args(h, &A);
    A.accumulate = accumulate;
    // first-visit marks of the plan: usable when the slab starts from zero (overwrite mode) and every cell around an own row is
    // assembled (in the cell_is_own exchange mode a slot that only ghost cells touch is never stored: the slab is cleared instead)
#ifndef B2F_FAN_FIRST
#define B2F_FAN_FIRST 1 /* 0: ignore the marks (A/B builds) */
#endif
    A.fan_first_mask = (B2F_FAN_FIRST && h->fan_first && !h->fan_lockstep && !accumulate && h->mesh.cell_is_own == nullptr) ? 8192u : 0u;
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
    if (h->reordered && variant == B2F_ASM_NODE_GATHER && h->fan_ok && uses_fan<Op>::value) {
      const int rc = permute_coords(h, stream); // coords_internal[i] = coords[perm[i]]: the caller may have moved the nodes
      if (rc != B2F_OK)
        return rc;
    }
    B2F_CUDA(launch_assemble<Op>(A, op, variant, h->per_block, h->nb_nz, stream));
    if (qdot != nullptr && !fused)
      return rhs_constant_source(h, *qdot, stream);
    return B2F_OK;
  }

} // namespace b2f
