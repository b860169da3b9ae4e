// User-written element callbacks through include/b2f/user_operator.cuh — the replacement of the reference's
// `template <class Function> assembleBilinear(Function compute_element_matrix)` (femutils/BSRFormat.h:786-787,1105-1111)
// for callbacks the library does not ship.  Two callbacks in the reference's shape (full element matrix per cell):
//   * StiffnessFromNormalisedGradients (BS = 1): V * (grad N_a . grad N_b) with the gradients divided by 6V first, the
//     way the reference's Poisson callback is written (modules/poisson/FemModule.h:177-186) - a different order of
//     operations from the built-in LaplaceTet, so the comparison with b2f_assemble(B2F_OP_LAPLACE) is a real cross-check;
//   * DensityMass (BS = 3): rho[cell] * V/20 * (1 + delta_ab) * I3 - per-cell data reached through the cell id.
// Each is assembled on the device for the node-gather and the cell-atomic variants and both value layouts and compared
// with a plain host loop over the same callback (1e-12 of the row scale); accumulate mode is checked too.
// Without a CUDA device the driver exits 3 (no CPU fallback).
#include <cmath>
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include "b2f/user_operator.cuh"
#include "arcanefem_b200/BSRFormat.h"
#include "arcanefem_b200/FemFunctionsGpu.cuh"

template <class T> T* to_device(const std::vector<T>& v)
{
  T* d = nullptr;
  cudaMalloc(&d, sizeof(T) * v.size());
  cudaMemcpy(d, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice);
  return d;
}
template <class T> std::vector<T> to_host(const T* d, long long n)
{
  std::vector<T> v((size_t)n);
  cudaMemcpy(v.data(), d, sizeof(T) * (size_t)n, cudaMemcpyDeviceToHost);
  return v;
}

__host__ __device__ inline void cross(const double* u, const double* v, double* w)
{
  w[0] = u[1] * v[2] - u[2] * v[1];
  w[1] = u[2] * v[0] - u[0] * v[2];
  w[2] = u[0] * v[1] - u[1] * v[0];
}

__host__ __device__ inline double tet_gradients(const double (&X)[4][3], double (&g)[4][3])
{
  double e[3][3];
  for (int a = 0; a < 3; ++a)
    for (int d = 0; d < 3; ++d)
      e[a][d] = X[a + 1][d] - X[0][d];
  cross(e[1], e[2], g[1]);
  cross(e[2], e[0], g[2]);
  cross(e[0], e[1], g[3]);
  const double det = e[0][0] * g[1][0] + e[0][1] * g[1][1] + e[0][2] * g[1][2];
  for (int d = 0; d < 3; ++d) {
    g[1][d] /= det;
    g[2][d] /= det;
    g[3][d] /= det;
    g[0][d] = -(g[1][d] + g[2][d] + g[3][d]);
  }
  return fabs(det) / 6.0;
}

struct StiffnessFromNormalisedGradients
{
  __host__ __device__ void operator()(long long, const double (&X)[4][3], double (&Ke)[4][4]) const
  {
    double g[4][3];
    const double vol = tet_gradients(X, g);
    for (int a = 0; a < 4; ++a)
      for (int b = 0; b < 4; ++b)
        Ke[a][b] = vol * (g[a][0] * g[b][0]) + vol * (g[a][1] * g[b][1]) + vol * (g[a][2] * g[b][2]);
  }
};

struct DensityMass
{
  const double* rho; // [nb_cell], device pointer on the device, host pointer in the host loop
  __host__ __device__ void operator()(long long cell, const double (&X)[4][3], double (&Ke)[12][12]) const
  {
    double g[4][3];
    const double m = rho[cell] * tet_gradients(X, g) / 20.0;
    for (int r = 0; r < 12; ++r)
      for (int c = 0; c < 12; ++c)
        Ke[r][c] = (r % 3 == c % 3) ? m * ((r / 3 == c / 3) ? 2.0 : 1.0) : 0.0;
  }
};

// Callbacks written with the reference's element-kernel vocabulary (include/arcanefem_b200/FemFunctionsGpu.cuh): the body of a
// module's callback (modules/poisson/FemModule.h:139-147,177-186) with the gathered vertices in place of (cell_lid, cn_cv, coords).
struct PoissonTetra4Callback
{
  __host__ __device__ ArcaneFemB200::FixedMatrix<4, 4> operator()(long long, const double (&X)[4][3]) const
  {
    using namespace ArcaneFemB200;
    const Real volume = Gpu::MeshOperation::computeVolumeTetra4(X);
    const Real4 dxU = Gpu::FeOperation3D::computeGradientXTetra4(X);
    const Real4 dyU = Gpu::FeOperation3D::computeGradientYTetra4(X);
    const Real4 dzU = Gpu::FeOperation3D::computeGradientZTetra4(X);
    return volume * (dxU ^ dxU) + volume * (dyU ^ dyU) + volume * (dzU ^ dzU);
  }
};

// Host check of the vocabulary itself (runs without a device): the callback above against the normalised-gradient
// stiffness, a triangle against its closed form, FixedMatrix arithmetic.
static int host_check_vocabulary()
{
  using namespace ArcaneFemB200;
  int bad = 0;
  const double T[2][4][3] = { { { 0.1, 0.0, 0.2 }, { 1.2, 0.1, 0.0 }, { 0.3, 0.9, 0.1 }, { 0.2, 0.3, 1.1 } },
                              { { 0.1, 0.0, 0.2 }, { 0.3, 0.9, 0.1 }, { 1.2, 0.1, 0.0 }, { 0.2, 0.3, 1.1 } } }; // second: negatively oriented
  for (auto& X : T) {
    double ref[4][4];
    StiffnessFromNormalisedGradients{}(0, X, ref);
    const FixedMatrix<4, 4> K = PoissonTetra4Callback{}(0, X);
    for (int a = 0; a < 4; ++a)
      for (int b = 0; b < 4; ++b)
        if (std::fabs(K(a, b) - ref[a][b]) > 1e-13 * std::fabs(ref[0][0]))
          ++bad;
  }
  const double P[3][3] = { { 0.0, 0.0, 0.0 }, { 2.0, 0.0, 0.0 }, { 0.0, 3.0, 0.0 } };
  if (std::fabs(Gpu::MeshOperation::computeAreaTria3(P) - 3.0) > 1e-15)
    ++bad;
  const Real3 dx = Gpu::FeOperation2D::computeGradientXTria3(P), dy = Gpu::FeOperation2D::computeGradientYTria3(P);
  const double gx[3] = { -0.5, 0.5, 0.0 }, gy[3] = { -1.0 / 3, 0.0, 1.0 / 3 };
  for (int a = 0; a < 3; ++a)
    if (std::fabs(dx[a] - gx[a]) > 1e-15 || std::fabs(dy[a] - gy[a]) > 1e-15)
      ++bad;
  const FixedMatrix<3, 3> L = 3.0 * (dx ^ dx) + (dy ^ dy) * 3.0; // area * (grad . grad)
  if (std::fabs(L(0, 0) - 3.0 * (0.25 + 1.0 / 9)) > 1e-15 || std::fabs(L(1, 2)) > 1e-15 || std::fabs((L / 3.0 - L / 3.0)(1, 1)) != 0.0)
    ++bad;
  return bad;
}

// Host loop with the reference's scatter rule and value layouts (femutils/BSRFormat.h:145-171,807-836,862-897).
template <int BS, class F>
std::vector<double> host_assemble(const F& f, const std::vector<int32_t>& cells, const std::vector<double>& coords, const std::vector<int32_t>& ri,
                                  const std::vector<int32_t>& co, bool per_block)
{
  std::vector<double> v((size_t)co.size() * BS * BS, 0.0);
  const long long nc = (long long)cells.size() / 4;
  for (long long c = 0; c < nc; ++c) {
    double X[4][3], Ke[4 * BS][4 * BS];
    for (int a = 0; a < 4; ++a)
      for (int d = 0; d < 3; ++d)
        X[a][d] = coords[3 * (size_t)cells[4 * c + a] + d];
    f(c, X, Ke);
    for (int a = 0; a < 4; ++a) {
      const int32_t r = cells[4 * c + a], rb = ri[r], k = ri[r + 1] - rb;
      for (int b = 0; b < 4; ++b) {
        int x = -1;
        for (int q = 0; q < k; ++q)
          if (co[rb + q] == cells[4 * c + b])
            x = q;
        for (int i = 0; i < BS; ++i)
          for (int j = 0; j < BS; ++j) {
            const size_t idx = per_block ? ((size_t)(rb + x) * BS * BS + i * BS + j) : ((size_t)rb * BS * BS + (size_t)i * BS * k + x * BS + j);
            v[idx] += Ke[BS * a + i][BS * b + j];
          }
      }
    }
  }
  return v;
}

static int compare(const std::vector<double>& got, const std::vector<double>& ref, const char* what)
{
  double amax = 0;
  for (double x : ref)
    amax = std::fmax(amax, std::fabs(x));
  int bad = 0;
  for (size_t i = 0; i < ref.size(); ++i)
    if (!(std::fabs(got[i] - ref[i]) <= 1e-12 * amax))
      ++bad;
  if (bad)
    std::printf("  %s: %d of %zu entries differ\n", what, bad, ref.size());
  return bad;
}

#define CHECK_OK(expr)                                                                   \
  do {                                                                                   \
    const int st__ = (expr);                                                             \
    if (st__ != B2F_OK) {                                                                \
      std::printf("%s -> %s: %s\n", #expr, b2f_status_string(st__), b2f_last_error_string()); \
      return 1000;                                                                       \
    }                                                                                    \
  } while (0)

template <int BS, class DevF, class HostF, class OtherF>
int run_case(const b2f_mesh_view& mesh, const std::vector<int32_t>& cells, const std::vector<double>& coords, const DevF& fdev, const HostF& fhost,
             const OtherF& other_block_size, int variant, bool use_csr, const char* name)
{
  b2f_options opt{};
  opt.block_size = BS;
  opt.use_csr = use_csr ? 1 : 0;
  opt.assembly_variant = variant;
  b2f_handle h = nullptr;
  CHECK_OK(b2f_create(&mesh, &opt, &h));
  CHECK_OK(b2f_compute_sparsity(h, nullptr));
  b2f_bsr_view v{};
  CHECK_OK(b2f_get_bsr(h, &v));
  const auto ri = to_host(v.row_index, v.nb_row + 1);
  const auto co = to_host(v.columns, v.nb_col);
  const auto ref = host_assemble<BS>(fhost, cells, coords, ri, co, !use_csr);
  int bad = 0;
  CHECK_OK((b2f::assemble_with_element_matrix<4, BS>(h, fdev, false, nullptr)));
  cudaDeviceSynchronize();
  bad += compare(to_host(v.values, v.nb_nz), ref, name);
  b2f_stats st{};
  CHECK_OK(b2f_get_stats(h, &st));
  if (st.columns_not_found != 0)
    ++bad;
  // accumulate mode: a second pass on top of the first gives twice the values
  CHECK_OK((b2f::assemble_with_element_matrix<4, BS>(h, fdev, true, nullptr)));
  cudaDeviceSynchronize();
  std::vector<double> twice(ref);
  for (double& x : twice)
    x *= 2.0;
  bad += compare(to_host(v.values, v.nb_nz), twice, "accumulate");
  if (BS == 1) {
    // the normalised-gradient callback against the built-in operator (different order of operations, same matrix)
    b2f_operator op{};
    op.kind = B2F_OP_LAPLACE;
    CHECK_OK(b2f_assemble(h, &op, nullptr));
    cudaDeviceSynchronize();
    bad += compare(to_host(v.values, v.nb_nz), ref, "built-in Laplace vs callback");
  }
  // a callback for another block size is refused (in the reference a FixedMatrix of the wrong size would not compile)
  if (b2f::assemble_with_element_matrix<4, (BS == 1 ? 3 : 1)>(h, other_block_size, false, nullptr) != B2F_ERR_INVALID_ARGUMENT)
    ++bad;
  CHECK_OK(b2f_destroy(h));
  std::printf("  %-28s variant=%d layout=%s bad=%d\n", name, variant, use_csr ? "per-row" : "per-block", bad);
  return bad;
}

int main()
{
  const int vocab_bad = host_check_vocabulary();
  std::printf("element-kernel vocabulary (host): bad=%d\n", vocab_bad);
  if (vocab_bad)
    return 4;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    std::printf("no CUDA device: user operators need the GPU (no CPU fallback)\n");
    return 3;
  }
  const int N = 5, n1 = N + 1;
  std::vector<double> coords;
  for (int k = 0; k < n1; ++k)
    for (int j = 0; j < n1; ++j)
      for (int i = 0; i < n1; ++i) {
        coords.push_back(i / double(N) + 0.02 * std::sin(3.0 * j + k));
        coords.push_back(j / double(N) + 0.02 * std::sin(2.0 * k + i));
        coords.push_back(k / double(N) + 0.02 * std::cos(2.0 * i + j));
      }
  const int kuhn[6][4] = { { 0, 1, 3, 7 }, { 0, 3, 2, 7 }, { 0, 2, 6, 7 }, { 0, 6, 4, 7 }, { 0, 4, 5, 7 }, { 0, 5, 1, 7 } };
  std::vector<int32_t> cells;
  for (int k = 0; k < N; ++k)
    for (int j = 0; j < N; ++j)
      for (int i = 0; i < N; ++i)
        for (auto& t : kuhn)
          for (int a = 0; a < 4; ++a) {
            const int c = t[a];
            cells.push_back((i + (c & 1)) + n1 * ((j + ((c >> 1) & 1)) + n1 * (k + ((c >> 2) & 1))));
          }
  const long long nn = (long long)coords.size() / 3, nc = (long long)cells.size() / 4;
  std::vector<double> rho((size_t)nc);
  for (long long c = 0; c < nc; ++c)
    rho[(size_t)c] = 1.0 + 0.5 * std::sin(0.37 * (double)c);
  b2f_mesh_view mesh{};
  mesh.dim = 3;
  mesh.nodes_per_cell = 4;
  mesh.nb_cell = nc;
  mesh.nb_node = nn;
  mesh.cell_node = to_device(cells);
  mesh.coords = to_device(coords);
  const double* rho_dev = to_device(rho);

  int bad = 0;
  const int variants[3] = { B2F_ASM_DEFAULT, B2F_ASM_CELL_ATOMIC, B2F_ASM_NODE_GATHER_SLOTS };
  for (int variant : variants)
    for (int use_csr = 0; use_csr < 2; ++use_csr) {
      bad += run_case<1>(mesh, cells, coords, StiffnessFromNormalisedGradients{}, StiffnessFromNormalisedGradients{}, DensityMass{ rho_dev }, variant, use_csr != 0,
                         "stiffness (BS=1)");
      bad += run_case<3>(mesh, cells, coords, DensityMass{ rho_dev }, DensityMass{ rho.data() }, StiffnessFromNormalisedGradients{}, variant, use_csr != 0,
                         "density mass (BS=3)");
    }
  {
    // the same through the C++ shim: BSRFormat<1>::assembleBilinear<NV>(callback) beside assembleBilinear(Operator)
    ArcaneFemB200::BSRFormat<1> bsr;
    bsr.initialize(mesh, true);
    bsr.computeSparsity();
    bsr.assembleBilinear<4>(StiffnessFromNormalisedGradients{});
    cudaDeviceSynchronize();
    const auto user = to_host(bsr.matrix().values().data(), bsr.matrix().values().size());
    bsr.assembleBilinear(ArcaneFemB200::Operator::Laplace());
    cudaDeviceSynchronize();
    const int b = compare(to_host(bsr.matrix().values().data(), bsr.matrix().values().size()), user, "shim: built-in Laplace vs callback");
    std::printf("  shim BSRFormat<1>::assembleBilinear<4>(callback) bad=%d\n", b);
    bad += b;
    // a callback that returns a FixedMatrix, written with the reference's helpers, against the built-in operator
    if (ArcaneFemB200::assemble_with_fixed_matrix<4, 1>(bsr.handle(), PoissonTetra4Callback{}, false, nullptr) != B2F_OK)
      ++bad;
    cudaDeviceSynchronize();
    const int b2 = compare(to_host(bsr.matrix().values().data(), bsr.matrix().values().size()), user, "reference-style FixedMatrix callback vs built-in Laplace");
    std::printf("  assemble_with_fixed_matrix<4, 1>(PoissonTetra4Callback) bad=%d\n", b2);
    bad += b2;
  }
  std::printf("user_operator_driver: nn=%lld nc=%lld bad=%d\n", nn, nc, bad);
  return bad == 0 ? 0 : 2;
}
