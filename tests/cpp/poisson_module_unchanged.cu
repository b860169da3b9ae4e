// poisson_module_unchanged.cu — the reference's Poisson module, BSR + Hypre path, compiled UNCHANGED against the drop-in
// headers (include/arcanefem_b200/arcane/*.h) and run on the device.  Only the #include block differs from
// modules/poisson/FemModule.h:17-44 (INTEGRATION.md shows the diff): the femutils headers on the hot path are replaced by
// the library's, Arcane itself and the femutils pieces off the path (FemDoFsOnNodes, IArcaneFemBC, the .axl-generated base
// class) come from the model under tests/cpp/arcane_stub because Arcane is not installed here.  Everything between the two
// markers below is the reference's text (tests/cpp/poisson_module_body.inc, checked against /root/reference by
// tests/test_cpp_shim.py); the functions after it are this driver's stand-ins for the module's solver / output stages.
//
// Checks (exit code 0 and "bad=0"): the matrix the unchanged callback path assembles equals the library's own Laplace operator
// (fan kernel) to 1e-12 and a host evaluation of the reference formula, in 2D and 3D; rhs = source + Neumann + penalty rows;
// solve() leaves P on the Dirichlet diagonals and produces the HYPRE_IJ argument arrays.
#include <cmath>
#include <cstdio>
#include <functional>
#include <vector>

// ---- replaces the #include block of modules/poisson/FemModule.h:17-44
#include <arcane/utils/CommandLineArguments.h>
#include <arcane/utils/PlatformUtils.h>
#include <arcane/utils/ArcaneGlobal.h>
#include <arcane/utils/StringList.h>
#include <arcane/utils/NumArray.h>
#include <arcane/ITimeLoopMng.h>
#include <arcane/IItemFamily.h>
#include <arcane/ItemGroup.h>
#include <arcane/ICaseMng.h>
#include <arcane/IMesh.h>
#include <arccore/base/NotImplementedException.h>
#include <arcane/accelerator/core/IAcceleratorMng.h>
#include <arcane/accelerator/VariableViews.h>
#include <arcanefem_b200/arcane/DoFLinearSystem.h> // "IDoFLinearSystemFactory.h" + "DoFLinearSystem.h"
#include <arcanefem_b200/arcane/ArcaneFemFunctionsGpu.h> // "ArcaneFemFunctionsGpu.h"
#include "FemDoFsOnNodes.h"
#include "IArcaneFemBC.h"
#include <arcanefem_b200/arcane/BSRFormat.h> // "BSRFormat.h"
#include <arcanefem_b200/arcane/FemUtils.h> // "FemUtils.h"
#include "Fem_axl.h"
namespace Arcane::FemUtils { struct ArcaneFemFunctions { struct MeshOperation {}; }; } // "ArcaneFemFunctions.h": CPU helpers, legacy path only

// ======================================================= reference text: begin
#define private public // the checks below read m_linear_system / m_bsr_format of the unchanged class
#include "poisson_module_body.inc"
#undef private
// ======================================================= reference text: end

// --- stand-ins for the module stages that are not on the assembly path
void FemModule::_solve() { m_linear_system.solve(); }
void FemModule::_updateVariables() {}
void FemModule::_validateResults() {}
void FemModule::_assembleLinearOperator() { ARCANE_THROW(NotImplementedException, "CPU rhs assembly: legacy path"); }
void FemModule::_printArcaneFemTime(const String label, const Real value) { info() << label << " " << value; }
FixedMatrix<3, 3> FemModule::_computeElementMatrixTria3(Cell) { ARCANE_THROW(NotImplementedException, "legacy CPU path"); }
FixedMatrix<4, 4> FemModule::_computeElementMatrixTetra4(Cell) { ARCANE_THROW(NotImplementedException, "legacy CPU path"); }
template <int N>
void FemModule::_assembleBilinear(const std::function<FixedMatrix<N, N>(const Cell&)>&) { ARCANE_THROW(NotImplementedException, "legacy CPU path"); }

// --- a small case: structured mesh, source f, Dirichlet on one side (penalty), Neumann on the opposite side
namespace {
struct Dirichlet : BC::IDirichletBoundaryCondition
{
  FaceGroup g;
  Real v, p;
  Dirichlet(const FaceGroup& gg, Real vv, Real pp) : g(gg), v(vv), p(pp) {}
  FaceGroup getSurface() override { return g; }
  Real getValue() override { return v; }
  Real getPenalty() override { return p; }
  String getEnforceDirichletMethod() override { return "Penalty"; }
};
struct Neumann : BC::INeumannBoundaryCondition
{
  FaceGroup g;
  Real v;
  Neumann(const FaceGroup& gg, Real vv) : g(gg), v(vv) {}
  FaceGroup getSurface() override { return g; }
  Real getValue() override { return v; }
  Real getValueX() override { return 0; }
  Real getValueY() override { return 0; }
  Real getValueZ() override { return 0; }
  bool hasValue() const override { return true; }
  bool hasValueX() const override { return false; }
  bool hasValueY() const override { return false; }
  bool hasValueZ() const override { return false; }
};
struct Conditions : BC::IArcaneFemBC
{
  UniqueArray<BC::IDirichletPointCondition*> pts;
  UniqueArray<BC::IDirichletBoundaryCondition*> dir;
  UniqueArray<BC::INeumannBoundaryCondition*> neu;
  ConstArrayView<BC::IDirichletPointCondition*> dirichletPointConditions() override { return pts; }
  ConstArrayView<BC::IDirichletBoundaryCondition*> dirichletBoundaryConditions() override { return dir; }
  ConstArrayView<BC::INeumannBoundaryCondition*> neumannBoundaryConditions() override { return neu; }
};

struct HostMesh
{
  int dim;
  std::vector<Int32> cells, faces_lo, faces_hi; // boundary faces on the planes x_last = 0 and x_last = 1
  std::vector<Real3> coords;
  std::vector<Int32> nodes_lo;
};

HostMesh make_mesh(int dim, int n)
{
  HostMesh m;
  m.dim = dim;
  const int n1 = n + 1;
  auto jig = [](int i, int j, int k) { return 0.13 * std::sin(1.7 * i + 2.3 * j + 3.1 * k); };
  if (dim == 2) {
    for (int j = 0; j < n1; ++j)
      for (int i = 0; i < n1; ++i) {
        const bool in = i > 0 && i < n && j > 0 && j < n;
        m.coords.push_back(Real3{ (i + (in ? jig(i, j, 0) : 0.0)) / n, (j + (in ? jig(j, i, 1) : 0.0)) / n, 0.0 });
      }
    for (int j = 0; j < n; ++j)
      for (int i = 0; i < n; ++i) {
        const Int32 a = i + n1 * j;
        const Int32 t[2][3] = { { a, a + 1, a + n1 + 1 }, { a, a + n1 + 1, a + n1 } };
        for (auto& tri : t)
          m.cells.insert(m.cells.end(), tri, tri + 3);
      }
    for (int i = 0; i < n; ++i) {
      m.faces_lo.insert(m.faces_lo.end(), { i, i + 1 });
      m.faces_hi.insert(m.faces_hi.end(), { i + n1 * n, i + 1 + n1 * n });
    }
    for (int i = 0; i < n1; ++i)
      m.nodes_lo.push_back(i);
  }
  else {
    for (int k = 0; k < n1; ++k)
      for (int j = 0; j < n1; ++j)
        for (int i = 0; i < n1; ++i) {
          const bool in = i > 0 && i < n && j > 0 && j < n && k > 0 && k < n;
          m.coords.push_back(Real3{ (i + (in ? jig(i, j, k) : 0.0)) / n, (j + (in ? jig(j, k, i) : 0.0)) / n, (k + (in ? jig(k, i, j) : 0.0)) / n });
        }
    const int kuhn[6][4] = { { 0, 1, 3, 7 }, { 0, 3, 2, 7 }, { 0, 2, 6, 7 }, { 0, 6, 4, 7 }, { 0, 4, 5, 7 }, { 0, 5, 1, 7 } };
    for (int k = 0; k < n; ++k)
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i) {
          Int32 c[8];
          for (int q = 0; q < 8; ++q)
            c[q] = (i + (q & 1)) + n1 * ((j + ((q >> 1) & 1)) + n1 * (k + ((q >> 2) & 1)));
          for (auto& t : kuhn)
            for (int a = 0; a < 4; ++a)
              m.cells.push_back(c[t[a]]);
        }
    for (int j = 0; j < n; ++j)
      for (int i = 0; i < n; ++i) {
        const Int32 a = i + n1 * j, top = n1 * n1 * n;
        m.faces_lo.insert(m.faces_lo.end(), { a, a + 1, a + n1 + 1 });
        m.faces_lo.insert(m.faces_lo.end(), { a, a + n1 + 1, a + n1 });
        m.faces_hi.insert(m.faces_hi.end(), { top + a, top + a + 1, top + a + n1 + 1 });
        m.faces_hi.insert(m.faces_hi.end(), { top + a, top + a + n1 + 1, top + a + n1 });
      }
    for (int q = 0; q < n1 * n1; ++q)
      m.nodes_lo.push_back(q);
  }
  return m;
}

// the reference formula on the host: Ke = V * (dx ^ dx + dy ^ dy [+ dz ^ dz]), femutils/ArcaneFemFunctions.h:59-93,535-569,701-818
void host_matrix(const HostMesh& m, std::vector<std::vector<std::pair<Int32, double>>>& rows, std::vector<double>& measure)
{
  const int nv = m.dim + 1;
  const size_t nc = m.cells.size() / nv;
  rows.assign(m.coords.size(), {});
  measure.assign(nc, 0.0);
  for (size_t c = 0; c < nc; ++c) {
    const Int32* cn = &m.cells[c * nv];
    double X[4][3];
    for (int a = 0; a < nv; ++a) {
      X[a][0] = m.coords[cn[a]].x;
      X[a][1] = m.coords[cn[a]].y;
      X[a][2] = m.coords[cn[a]].z;
    }
    double K[4][4];
    if (nv == 3) {
      const double X3[3][3] = { { X[0][0], X[0][1], X[0][2] }, { X[1][0], X[1][1], X[1][2] }, { X[2][0], X[2][1], X[2][2] } };
      const double area = ArcaneFemB200::Gpu::MeshOperation::computeAreaTria3(X3);
      const auto gx = ArcaneFemB200::Gpu::FeOperation2D::computeGradientXTria3(X3), gy = ArcaneFemB200::Gpu::FeOperation2D::computeGradientYTria3(X3);
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b)
          K[a][b] = area * gx[a] * gx[b] + area * gy[a] * gy[b];
      measure[c] = area;
    }
    else {
      const double vol = ArcaneFemB200::Gpu::MeshOperation::computeVolumeTetra4(X);
      const auto gx = ArcaneFemB200::Gpu::FeOperation3D::computeGradientXTetra4(X), gy = ArcaneFemB200::Gpu::FeOperation3D::computeGradientYTetra4(X),
                 gz = ArcaneFemB200::Gpu::FeOperation3D::computeGradientZTetra4(X);
      for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 4; ++b)
          K[a][b] = vol * gx[a] * gx[b] + vol * gy[a] * gy[b] + vol * gz[a] * gz[b];
      measure[c] = vol;
    }
    for (int a = 0; a < nv; ++a)
      for (int b = 0; b < nv; ++b) {
        auto& row = rows[cn[a]];
        bool hit = false;
        for (auto& e : row)
          if (e.first == cn[b]) {
            e.second += K[a][b];
            hit = true;
          }
        if (!hit)
          row.push_back({ cn[b], K[a][b] });
      }
  }
}

int run_case(int dim, int n)
{
  int bad = 0;
  HostMesh hm = make_mesh(dim, n);
  ISubDomain sd;
  std::vector<Int32> all_faces = hm.faces_lo;
  all_faces.insert(all_faces.end(), hm.faces_hi.begin(), hm.faces_hi.end());
  IMesh mesh(dim, hm.cells, hm.coords, all_faces, sd.traceMng());
  const Int32 nf_lo = (Int32)(hm.faces_lo.size() / dim), nf_hi = (Int32)(hm.faces_hi.size() / dim);
  std::vector<Int32> lo_ids, hi_ids;
  for (Int32 i = 0; i < nf_lo; ++i)
    lo_ids.push_back(i);
  for (Int32 i = 0; i < nf_hi; ++i)
    hi_ids.push_back(nf_lo + i);
  NodeGroup lo_nodes("lo_nodes", hm.nodes_lo);
  FaceGroup lo("lo", lo_ids, lo_nodes), hi("hi", hi_ids, NodeGroup("hi_nodes", {}));
  const Real P = 1.0e31, g = 0.5, q = 1.25, f = 5.5;
  Dirichlet dbc(lo, g, P);
  Neumann nbc(hi, q);
  Conditions bc;
  bc.dir.add(&dbc);
  bc.neu.add(&nbc);
  B200HypreLinearSystemFactory factory;

  ModuleBuildInfo mbi(&sd, &mesh);
  FemModule module(mbi);
  module.options()->bsr.value = true;
  module.options()->bsrAtomicFree.value = false;
  module.options()->f.value = f;
  module.options()->linearSystem.factory = &factory;
  module.options()->bc = &bc;
  module.startInit(); // the reference's text from here on
  module.compute();
  cudaDeviceSynchronize();

  // --- what the unchanged module left on the device
  auto* impl = dynamic_cast<B200HypreLinearSystemImpl*>(module.m_linear_system.implementation());
  if (!impl)
    return 100;
  b2f_handle h = impl->handle();
  b2f_bsr_view v{};
  ArcaneFemB200::check(b2f_get_bsr(h, &v));
  std::vector<Int32> ri((size_t)v.nb_row + 1), co((size_t)v.nb_col);
  std::vector<double> val((size_t)v.nb_nz), rhs((size_t)v.nb_row);
  cudaMemcpy(ri.data(), v.row_index, sizeof(Int32) * ri.size(), cudaMemcpyDeviceToHost);
  cudaMemcpy(co.data(), v.columns, sizeof(Int32) * co.size(), cudaMemcpyDeviceToHost);
  cudaMemcpy(val.data(), v.values, sizeof(double) * val.size(), cudaMemcpyDeviceToHost);
  b2f_system_view sv{};
  ArcaneFemB200::check(b2f_get_system(h, &sv));
  cudaMemcpy(rhs.data(), sv.rhs, sizeof(double) * rhs.size(), cudaMemcpyDeviceToHost);

  std::vector<std::vector<std::pair<Int32, double>>> rows;
  std::vector<double> measure;
  host_matrix(hm, rows, measure);
  std::vector<char> is_dir(hm.coords.size(), 0);
  for (Int32 nd : hm.nodes_lo)
    is_dir[nd] = 1;
  double worst = 0.0;
  for (size_t r = 0; r < rows.size(); ++r) {
    if ((size_t)(ri[r + 1] - ri[r]) != rows[r].size())
      ++bad;
    double scale = 0.0;
    for (auto& e : rows[r])
      scale = std::max(scale, std::fabs(e.second));
    for (Int32 x = ri[r]; x < ri[r + 1]; ++x) {
      double ref = 0.0;
      bool hit = false;
      for (auto& e : rows[r])
        if (e.first == co[x]) {
          ref = e.second;
          hit = true;
        }
      if (!hit)
        ++bad;
      if (is_dir[r] && co[x] == (Int32)r)
        ref = P; // _applyForcedValuesToLhs: diagonal <- P (femutils/HypreDoFLinearSystem.cc:356-382)
      const double d = std::fabs(val[x] - ref) / std::max(std::fabs(ref), scale);
      worst = std::max(worst, d);
    }
  }
  if (!(worst <= 1e-12))
    ++bad;
  // rhs: source f * |cell| / Nv, Neumann q * |face| / Nvf, penalty rows P * g
  std::vector<double> rref(hm.coords.size(), 0.0);
  const int nv = dim + 1;
  for (size_t c = 0; c < measure.size(); ++c)
    for (int a = 0; a < nv; ++a)
      rref[hm.cells[c * nv + a]] += f * measure[c] / nv;
  for (Int32 i = 0; i < nf_hi; ++i) {
    const Int32* fn = &hm.faces_hi[(size_t)i * dim];
    double meas;
    if (dim == 2) {
      const Real3 a = hm.coords[fn[0]], b = hm.coords[fn[1]];
      meas = std::sqrt((b.x - a.x) * (b.x - a.x) + (b.y - a.y) * (b.y - a.y));
    }
    else {
      const Real3 a = hm.coords[fn[0]], b = hm.coords[fn[1]], c = hm.coords[fn[2]];
      const Real3 n = math::cross(b - a, c - a);
      meas = 0.5 * std::sqrt(math::dot(n, n));
    }
    for (int a = 0; a < dim; ++a)
      rref[fn[a]] += q * meas / dim;
  }
  for (Int32 nd : hm.nodes_lo)
    rref[nd] = P * g;
  double worst_rhs = 0.0, rscale = 0.0;
  for (size_t i = 0; i < rref.size(); ++i)
    if (!is_dir[i])
      rscale = std::max(rscale, std::fabs(rref[i]));
  for (size_t i = 0; i < rref.size(); ++i)
    worst_rhs = std::max(worst_rhs, std::fabs(rhs[i] - rref[i]) / std::max(std::fabs(rref[i]), rscale));
  if (!(worst_rhs <= 1e-12))
    ++bad;
  // solve() produced the HYPRE_IJMatrixSetValues / HYPRE_IJVectorSetValues arguments for every (own) row
  const b2f_ij_view& ij = impl->lastHypreArguments();
  if (ij.nb_row != v.nb_row || ij.nb_value != v.nb_nz || !ij.ncols || !ij.rows || !ij.cols || !ij.values || !ij.rhs)
    ++bad;

  // --- the same assembly through the library's own operator (fan kernel) on the same handle: the fast path of INTEGRATION.md
  std::vector<double> fast((size_t)v.nb_nz);
  module.m_bsr_format.assembleBilinear(ArcaneFemB200::Operator::Laplace());
  cudaDeviceSynchronize();
  cudaMemcpy(fast.data(), v.values, sizeof(double) * fast.size(), cudaMemcpyDeviceToHost);
  double worst_fast = 0.0;
  for (size_t r = 0; r < rows.size(); ++r) {
    double scale = 0.0;
    for (Int32 x = ri[r]; x < ri[r + 1]; ++x)
      scale = std::max(scale, std::fabs(fast[x]));
    for (Int32 x = ri[r]; x < ri[r + 1]; ++x) {
      if (is_dir[r] && co[x] == (Int32)r)
        continue; // val holds P there
      worst_fast = std::max(worst_fast, std::fabs(val[x] - fast[x]) / std::max(std::fabs(fast[x]), scale));
    }
  }
  if (!(worst_fast <= 1e-12))
    ++bad;
  std::printf("poisson_module_unchanged %dD n=%d: rows=%lld nnz=%lld  matrix vs reference formula %.2e, rhs %.2e, callback path vs built-in operator %.2e  bad=%d\n", dim, n,
              (long long)v.nb_row, (long long)v.nb_nz, worst, worst_rhs, worst_fast, bad);
  return bad;
}
} // namespace

int main()
{
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    std::printf("poisson_module_unchanged: no CUDA device (the library has no CPU fallback)\n");
    return 3;
  }
  int bad = 0;
  try {
    bad += run_case(2, 9);
    bad += run_case(3, 6);
  }
  catch (const std::exception& e) {
    std::printf("poisson_module_unchanged: exception: %s\n", e.what());
    return 4;
  }
  std::printf("poisson_module_unchanged: bad=%d\n", bad);
  return bad == 0 ? 0 : 2;
}
