// C++ driver over the shim headers, mirroring the call sequence of the reference's Poisson module
// (modules/poisson/FemModule.cc:24-117,152-191,255-286): initialize -> computeSparsity -> assembleBilinear(Laplace)
// -> toLinearSystem -> source + Dirichlet penalty -> solve() prelude.  Checks invariants on the host and exits 0.
// Without a CUDA device, initialize() must throw (no CPU fallback) and the driver exits 3.
#include <cmath>
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include "arcanefem_b200/BSRFormat.h"
#include "arcanefem_b200/BoundaryConditions.h"

using namespace ArcaneFemB200;

template <class T> T* to_device(const std::vector<T>& v)
{
  T* d = nullptr;
  cudaMalloc(&d, sizeof(T) * v.size());
  cudaMemcpy(d, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice);
  return d;
}
template <class T> std::vector<T> to_host(const T* d, Int64 n)
{
  std::vector<T> v((size_t)n);
  cudaMemcpy(v.data(), d, sizeof(T) * (size_t)n, cudaMemcpyDeviceToHost);
  return v;
}

int main()
{
  const int N = 6, n1 = N + 1;
  std::vector<double> coords;
  for (int k = 0; k < n1; ++k) for (int j = 0; j < n1; ++j) for (int i = 0; i < n1; ++i) {
    coords.push_back(i / double(N) + 0.01 * std::sin(3.0 * j + k));
    coords.push_back(j / double(N));
    coords.push_back(k / double(N) + 0.01 * std::cos(2.0 * i));
  }
  // 6 Kuhn tets per hex (paths 000 -> 111), corner id = dx + 2 dy + 4 dz
  const int kuhn[6][4] = { {0,1,3,7}, {0,3,2,7}, {0,2,6,7}, {0,6,4,7}, {0,4,5,7}, {0,5,1,7} };
  std::vector<Int32> cells;
  for (int k = 0; k < N; ++k) for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i)
    for (auto& t : kuhn) for (int a = 0; a < 4; ++a) {
      const int c = t[a];
      cells.push_back((i + (c & 1)) + n1 * ((j + ((c >> 1) & 1)) + n1 * (k + ((c >> 2) & 1))));
    }
  const Int64 nn = (Int64)coords.size() / 3, nc = (Int64)cells.size() / 4;
  b2f_mesh_view mesh{};
  mesh.dim = 3; mesh.nodes_per_cell = 4; mesh.nb_cell = nc; mesh.nb_node = nn;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    try {
      BSRFormat<1> bsr;
      std::vector<Int32> dummy(4, 0);
      mesh.cell_node = dummy.data();
      bsr.initialize(mesh, true);
    }
    catch (const std::exception& e) {
      std::printf("no device: initialize threw as expected: %s\n", e.what());
      return 3;
    }
    return 1;
  }
  mesh.cell_node = to_device(cells);
  mesh.coords = to_device(coords);

  BSRFormat<1> bsr;
  bsr.initialize(mesh, /*does_linear_system_use_csr=*/true);
  bsr.computeSparsity();
  bsr.assembleBilinear(Operator::Laplace());
  DoFLinearSystem ls;
  bsr.toLinearSystem(ls);
  cudaDeviceSynchronize();

  auto& M = bsr.matrix();
  const CSRFormatView& csr = ls.getCSRValues();
  auto rows = to_host(csr.rows().data(), csr.nbRow());
  auto cols = to_host(csr.columns().data(), csr.nbColumn());
  auto vals = to_host(csr.values().data(), csr.nbValue());
  rows.push_back((Int32)csr.nbValue());
  int bad = 0;
  double amax = 0;
  for (double v : vals) amax = std::fmax(amax, std::fabs(v));
  for (Int64 r = 0; r < nn; ++r) {
    double s = 0;
    if (cols[rows[r]] != r) ++bad; // diagonal first
    for (Int32 q = rows[r]; q < rows[r + 1]; ++q) {
      s += vals[q];
      if (q > rows[r] + 1 && cols[q] <= cols[q - 1]) ++bad; // ascending neighbours
    }
    if (std::fabs(s) > 1e-12 * amax) ++bad; // constants are in the kernel of the Laplacian
  }
  // symmetry through getValue
  for (Int64 r = 0; r < nn; r += 37)
    for (Int32 q = rows[r]; q < rows[r + 1]; ++q)
      if (std::fabs(M.getValue(r, cols[q]) - M.getValue(cols[q], r)) > 1e-12 * amax) ++bad;
  if (M.nbRow() != nn || M.nbCol() != (Int32)cols.size() || M.orderValuePerBlock()) ++bad;

  // RHS + Dirichlet by penalty on the plane z index 0 (nodes 0 .. n1*n1-1), then the solve() prelude
  std::vector<Int32> group;
  for (Int32 i = 0; i < n1 * n1; ++i) group.push_back(i);
  Gpu::NodeGroup g{ to_device(group), (Int64)group.size() };
  ls.clearValues();
  check(b2f_rhs_fill(ls.handle(), 0.0, nullptr));
  Gpu::BoundaryConditions3D::applyConstantSourceToRhs(5.5, ls);
  Gpu::BoundaryConditions::applyDirichletViaPenalty(ls, g, 0.5, 1.0e30);
  ls.applyDirichlet();
  cudaDeviceSynchronize();
  auto rhs = to_host(ls.rhsVariable().data(), nn);
  if (M.getValue(3, 3) != 1.0e30 || rhs[3] != 0.5 * 1.0e30) ++bad;
  double total = 0;
  for (Int64 r = n1 * n1; r < nn; ++r) total += rhs[r];
  // missing value must throw like BSRMatrix::findValueIndex (BSRFormat.h:170)
  bool threw = false;
  try { M.getValue(0, nn - 1); } catch (const std::runtime_error&) { threw = true; }
  if (!threw) ++bad;
  // The per-block branch of toLinearSystem (femutils/BSRFormat.h:261-280,416-427): BSRFormat<3> built for a linear system that
  // does not take CSR - no CSR view is made, the system is attached all the same and the Dirichlet entries work on it.
  {
    BSRFormat<3> el;
    el.initialize(mesh, /*does_linear_system_use_csr=*/false);
    el.computeSparsity();
    el.assembleBilinear(Operator::Elasticity(21.0e5, 0.28));
    DoFLinearSystem els;
    el.toLinearSystem(els); // must not throw (round 1: B2F_ERR_STATE from toCsr)
    if (!el.matrix().orderValuePerBlock() || els.getCSRValues().nbValue() != 0) ++bad;
    els.clearValues();
    check(b2f_rhs_fill(els.handle(), 0.0, nullptr));
    Gpu::BoundaryConditions::applyDirichletViaRowEliminationVectorial(els, g, 1, 0.25);
    els.applyDirichlet();
    cudaDeviceSynchronize();
    auto erhs = to_host(els.rhsVariable().data(), 3 * nn);
    if (el.matrix().getValue(3 * 5 + 1, 3 * 5 + 1) != 1.0 || el.matrix().getValue(3 * 5 + 1, 3 * 5 + 0) != 0.0 || erhs[3 * 5 + 1] != 0.25) ++bad;
    if (el.matrix().getValue(3 * 5 + 0, 3 * 5 + 0) == 1.0) ++bad; // the other components keep their rows
  }
  std::printf("poisson_driver: nn=%lld nc=%lld nnz=%d rhs_sum=%.12g bad=%d\n", (long long)nn, (long long)nc, M.nbNz(), total, bad);
  return bad == 0 ? 0 : 2;
}
