// Linear elasticity P1, BS = 2 triangles (modules/elasticity/FemModule.h:112-152) and BS = 3 tetrahedra: kernel
// instantiations of include/b2f/assemble.cuh.
#include "assemble_run.cuh"

namespace b2f {
int run_elasticity(b2f_context_s* h, double lambda, double mu2, int accumulate, cudaStream_t stream, const double* qdot)
{
  return h->nv == 3 ? run(h, ElasticityTri{ lambda, mu2 }, accumulate, stream, qdot) : run(h, ElasticityTet{ lambda, mu2 }, accumulate, stream, qdot);
}
} // namespace b2f
