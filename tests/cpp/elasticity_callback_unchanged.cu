// elasticity_callback_unchanged.cu — the reference's 2D elasticity element callback (modules/elasticity/FemModule.h:112-152,
// text in tests/cpp/elasticity_callback_body.inc) passed UNCHANGED, inside the lambda of modules/elasticity/FemModule.cc:448-451,
// to Arcane::FemUtils::BSRFormat<2>::assembleBilinear.  Exercises FixedMatrix's initializer list, matrixTranspose,
// FixedMatrix<N,1> ^ FixedMatrix<1,N>, *, /, + on the device.  Both linear-system kinds: CSR consumer (per-row values) and the
// per-block branch of toLinearSystem (femutils/BSRFormat.h:261-280,416-427).  Checked against the library's own
// Operator::Elasticity (fan kernel) on the same handle: "bad=0".
#include <cmath>
#include <cstdio>
#include <vector>
#include <arcane/IMesh.h>
#include <arcane/accelerator/VariableViews.h>
#include <arcanefem_b200/arcane/ArcaneFemFunctionsGpu.h>
#include <arcanefem_b200/arcane/BSRFormat.h>
#include <arcanefem_b200/arcane/FemUtils.h>

using namespace Arcane;
using namespace Arcane::FemUtils;
namespace ax = Arcane::Accelerator;

// ======================================================= reference text: begin
#include "elasticity_callback_body.inc"
// ======================================================= reference text: end

struct Module // the call site of modules/elasticity/FemModule.cc:441-451 (public member: an extended lambda needs that)
{
  IMesh* m_mesh;
  ISubDomain* m_sd;
  VariableNodeReal3 m_node_coord;
  Real lambda, mu2;
  IMesh* mesh() const { return m_mesh; }
  ISubDomain* subDomain() const { return m_sd; }
  void assemble(BSRFormat<2>& m_bsr_format)
  {
    UnstructuredMeshConnectivityView m_connectivity_view(mesh());
    auto cn_cv = m_connectivity_view.cellNode();
    auto queue = subDomain()->acceleratorMng()->defaultQueue();
    auto command = makeCommand(queue);
    auto in_node_coord = ax::viewIn(command, m_node_coord);
    auto lambda_copy = lambda;
    auto mu2_copy = mu2;

    m_bsr_format.assembleBilinear([=] ARCCORE_HOST_DEVICE(CellLocalId cell_lid) { return computeElementMatrixTRIA3Gpu(cell_lid, cn_cv, in_node_coord, lambda_copy, mu2_copy); });
  }
};

static int run(bool use_csr)
{
  const int n = 11, n1 = n + 1;
  std::vector<Real3> coords;
  std::vector<Int32> cells;
  for (int j = 0; j < n1; ++j)
    for (int i = 0; i < n1; ++i) {
      const bool in = i > 0 && i < n && j > 0 && j < n;
      coords.push_back(Real3{ (i + (in ? 0.2 * std::sin(1.3 * i + 2.9 * j) : 0.0)) / n, (j + (in ? 0.2 * std::cos(2.1 * i - 1.7 * j) : 0.0)) / n, 0.0 });
    }
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < n; ++i) {
      const Int32 a = i + n1 * j;
      const Int32 t[2][3] = { { a, a + 1, a + n1 + 1 }, { a, a + n1 + 1, a + n1 } };
      for (auto& tri : t)
        cells.insert(cells.end(), tri, tri + 3);
    }
  ISubDomain sd;
  IMesh mesh(2, cells, coords, {}, sd.traceMng());
  FemDoFsOnNodes dofs(sd.traceMng());
  dofs.initialize(&mesh, 2);
  BSRFormat<2> bsr(sd.traceMng(), *sd.acceleratorMng()->defaultQueue(), dofs);
  bsr.initialize(&mesh, use_csr, false);
  bsr.computeSparsity();
  const Real E = 21.0e5, nu = 0.28; // modules/elasticity/inputs/bar.2D.traction.bsr.hypre.arc:24-25
  Module mod{ &mesh, &sd, mesh.nodesCoordinates(), E * nu / ((1 + nu) * (1 - 2 * nu)), (E / (2 * (1 + nu))) * 2 };
  mod.assemble(bsr);
  B200HypreLinearSystemFactory factory;
  DoFLinearSystem ls;
  ls.setLinearSystemFactory(&factory);
  ls.initialize(&sd, sd.acceleratorMng()->defaultRunner(), dofs.dofFamily(), "Solver");
  bsr.toLinearSystem(ls); // use_csr = false: no CSR view, the values stay attached (per-block branch)
  int bad = 0;
  if (bsr.matrix().orderValuePerBlock() != !use_csr)
    ++bad;
  if (use_csr && ls.getCSRValues().nbValue() != bsr.matrix().nbNz())
    ++bad;
  ls.rhsVariable().fill(0.0);
  ls.getForcedInfo().fill(false);
  cudaDeviceSynchronize();
  const Int64 nnz = bsr.matrix().nbNz();
  std::vector<double> cb((size_t)nnz), builtin((size_t)nnz);
  cudaMemcpy(cb.data(), bsr.matrix().values().to1DSpan().data(), sizeof(double) * (size_t)nnz, cudaMemcpyDeviceToHost);
  bsr.assembleBilinear(ArcaneFemB200::Operator::Elasticity(E, nu));
  cudaDeviceSynchronize();
  cudaMemcpy(builtin.data(), bsr.matrix().values().to1DSpan().data(), sizeof(double) * (size_t)nnz, cudaMemcpyDeviceToHost);
  double scale = 0.0, worst = 0.0;
  for (double v : builtin)
    scale = std::max(scale, std::fabs(v));
  for (Int64 q = 0; q < nnz; ++q)
    worst = std::max(worst, std::fabs(cb[(size_t)q] - builtin[(size_t)q]) / scale);
  if (!(worst <= 1e-12) || scale == 0.0)
    ++bad;
  // BSRMatrix::getValue through DoFLocalIds agrees with the raw array (findValueIndex, femutils/BSRFormat.h:145-171)
  const double d00 = bsr.matrix().getValue(DoFLocalId(2 * 40 + 1), DoFLocalId(2 * 40 + 1));
  if (!(d00 > 0.0))
    ++bad;
  std::printf("elasticity_callback_unchanged use_csr=%d: nnz=%lld, callback vs built-in operator %.2e (scale %.3e) bad=%d\n", (int)use_csr, (long long)nnz, worst, scale, bad);
  return bad;
}

int main()
{
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    std::printf("elasticity_callback_unchanged: no CUDA device (the library has no CPU fallback)\n");
    return 3;
  }
  int bad = 0;
  try {
    bad += run(true);
    bad += run(false);
  }
  catch (const std::exception& e) {
    std::printf("elasticity_callback_unchanged: exception: %s\n", e.what());
    return 4;
  }
  std::printf("elasticity_callback_unchanged: bad=%d\n", bad);
  return bad == 0 ? 0 : 2;
}
