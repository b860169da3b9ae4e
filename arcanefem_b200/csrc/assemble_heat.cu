// Heat P1, mass + stiffness (modules/heat/FemModule.cc:678-770) and its tetrahedral twin: kernel instantiations of
// include/b2f/assemble.cuh.
#include "assemble_run.cuh"

namespace b2f {
int run_heat(b2f_context_s* h, double lambda, double inv_dt, const double* cell_lambda, int accumulate, cudaStream_t stream, const double* qdot)
{
  if (cell_lambda == nullptr) // one conductivity for the mesh: the instantiation without per-cell data (no cell ids in the walk)
    return h->nv == 3 ? run(h, HeatTriUniform{ lambda, inv_dt, nullptr }, accumulate, stream, qdot) : run(h, HeatTetUniform{ lambda, inv_dt, nullptr }, accumulate, stream, qdot);
  return h->nv == 3 ? run(h, HeatTri{ lambda, inv_dt, cell_lambda }, accumulate, stream, qdot) : run(h, HeatTet{ lambda, inv_dt, cell_lambda }, accumulate, stream, qdot);
}
} // namespace b2f
