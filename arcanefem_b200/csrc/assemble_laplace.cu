// Laplace P1 (modules/poisson/FemModule.h:139-186): kernel instantiations of include/b2f/assemble.cuh.
#include "assemble_run.cuh"

namespace b2f {
int run_laplace(b2f_context_s* h, int accumulate, cudaStream_t stream, const double* qdot)
{
  return h->nv == 3 ? run(h, LaplaceTri{}, accumulate, stream, qdot) : run(h, LaplaceTet{}, accumulate, stream, qdot);
}
} // namespace b2f
