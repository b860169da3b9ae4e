// C++ driver over the shim headers for the transient heat module (modules/heat/FemModule.cc:113-131,386-537,775-840): pattern
// once, then per time step conduction + mass / dt, the convection boundary term, the (T_old / dt) source, the convection
// load and the module's "Penalty" Dirichlet (a_ii += P).  Checks conservation identities of the assembled system on the host:
//   sum of all matrix entries = area / dt + h * L_conv   (the stiffness sums to zero, the mass to the area, the face mass to h L)
//   sum of the right-hand side = (sum of T_old * lumped mass) / dt + h * Text * L_conv
// Without a CUDA device initialize() throws (no CPU fallback) and the driver exits 3.
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
  const int N = 12, n1 = N + 1;
  const double Lx = 2.0, Ly = 1.0, lambda = 1.75, dt = 0.4, h = 1.5, Text = 20.0, Tfix = 10.0, P = 1.0e30;
  std::vector<double> coords;
  for (int j = 0; j < n1; ++j)
    for (int i = 0; i < n1; ++i) {
      const bool inner = i > 0 && i < N && j > 0 && j < N;
      coords.push_back(Lx * i / N + (inner ? 0.02 * std::sin(3.0 * j + i) : 0.0));
      coords.push_back(Ly * j / N + (inner ? 0.01 * std::cos(2.0 * i + j) : 0.0));
      coords.push_back(0.0);
    }
  std::vector<Int32> cells;
  for (int j = 0; j < N; ++j)
    for (int i = 0; i < N; ++i) {
      const Int32 a = i + n1 * j, b = a + 1, c = a + n1 + 1, d = a + n1;
      cells.insert(cells.end(), { a, b, c, a, c, d });
    }
  std::vector<Int32> conv_faces, fixed_nodes; // convection on x = Lx, Dirichlet on x = 0
  for (int j = 0; j < N; ++j) {
    conv_faces.push_back(N + n1 * j);
    conv_faces.push_back(N + n1 * (j + 1));
  }
  for (int j = 0; j < n1; ++j)
    fixed_nodes.push_back(n1 * j);
  const Int64 nn = (Int64)coords.size() / 3, nc = (Int64)cells.size() / 3;
  std::vector<uint8_t> fixed_mask((size_t)nn, 0);
  for (Int32 n : fixed_nodes)
    fixed_mask[(size_t)n] = 1;
  std::vector<double> T_old((size_t)nn);
  for (Int64 n = 0; n < nn; ++n)
    T_old[(size_t)n] = fixed_mask[(size_t)n] ? Tfix : 30.0 + 0.1 * (double)(n % 7);

  b2f_mesh_view mesh{};
  mesh.dim = 2; mesh.nodes_per_cell = 3; mesh.nb_cell = nc; mesh.nb_node = nn;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    try {
      BSRFormat<1> bsr;
      mesh.cell_node = cells.data();
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
  Gpu::FaceGroup conv{ to_device(conv_faces), (Int64)conv_faces.size() / 2, nullptr };
  Gpu::NodeGroup fixed{ to_device(fixed_nodes), (Int64)fixed_nodes.size() };
  const double* T_old_d = to_device(T_old);
  const uint8_t* mask_d = to_device(fixed_mask);

  BSRFormat<1> bsr;
  bsr.initialize(mesh, true);
  bsr.computeSparsity(); // once: every step reuses the pattern and the plan
  DoFLinearSystem ls;
  bsr.toLinearSystem(ls);
  namespace BC = Gpu::BoundaryConditions2D;
  int bad = 0;
  double last_sum = 0;
  for (int step = 0; step < 2; ++step) {
    bsr.assembleBilinear(Operator::Heat(lambda, dt));
    BC::applyConvectionToMatrix(ls, conv, h);
    ls.clearValues();
    BC::applyNodalSourceToRhs(ls, T_old_d, 1.0 / dt, mask_d);
    BC::applyConvectionToRhs(ls, conv, h, Text);
    cudaDeviceSynchronize();
    // identities before the Dirichlet treatment
    const CSRFormatView& csr = ls.getCSRValues();
    const auto vals = to_host(csr.values().data(), csr.nbValue());
    const auto rhs = to_host(ls.rhsVariable().data(), nn);
    double msum = 0, rsum = 0;
    for (double v : vals) msum += v;
    for (double v : rhs) rsum += v;
    // lumped mass per node = sum of area / 3 of the incident cells (signed 2D area, positive here)
    std::vector<double> lumped((size_t)nn, 0.0);
    double area = 0;
    for (Int64 c = 0; c < nc; ++c) {
      const double* p0 = &coords[3 * (size_t)cells[3 * c]];
      const double* p1 = &coords[3 * (size_t)cells[3 * c + 1]];
      const double* p2 = &coords[3 * (size_t)cells[3 * c + 2]];
      const double a = 0.5 * ((p1[0] - p0[0]) * (p2[1] - p0[1]) - (p2[0] - p0[0]) * (p1[1] - p0[1]));
      area += a;
      for (int k = 0; k < 3; ++k)
        lumped[(size_t)cells[3 * c + k]] += a / 3.0;
    }
    double src = 0;
    for (Int64 n = 0; n < nn; ++n)
      if (!fixed_mask[(size_t)n])
        src += T_old[(size_t)n] * lumped[(size_t)n] / dt;
    const double want_m = area / dt + h * Ly, want_r = src + h * Text * Ly;
    if (std::fabs(msum - want_m) > 1e-9 * want_m || std::fabs(rsum - want_r) > 1e-9 * want_r)
      ++bad;
    if (step == 1 && std::fabs(msum - last_sum) > 1e-12 * std::fabs(msum))
      ++bad; // re-assembly on the same pattern reproduces the system (the face term uses atomics: equal up to their order)
    last_sum = msum;
    // the module's "Penalty": a_ii += P, b_i = P * T (FemModule.cc:386-406) = WeakPenalty here, flagged last
    Gpu::BoundaryConditions::applyDirichletViaWeakPenalty(ls, fixed, Tfix, P);
    ls.applyDirichlet();
    cudaDeviceSynchronize();
    if (bsr.matrix().getValue(fixed_nodes[3], fixed_nodes[3]) != P || to_host(ls.rhsVariable().data(), nn)[(size_t)fixed_nodes[3]] != P * Tfix)
      ++bad;
    std::printf("heat_driver step %d: matrix sum %.12g (want %.12g) rhs sum %.12g (want %.12g)\n", step, msum, want_m, rsum, want_r);
  }
  std::printf("heat_driver: nn=%lld nc=%lld bad=%d\n", (long long)nn, (long long)nc, bad);
  return bad == 0 ? 0 : 2;
}
