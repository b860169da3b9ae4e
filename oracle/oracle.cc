// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the shipped product.
//
// Sequential CPU restatement of the ArcaneFEM bilinear-form assembly hot path
// (reference: toutane/arcanefem, paths below are relative to /root/reference).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// `--impl reference` legs may load this library.  The product (libb200fem.so)
// never links, loads or calls it.
//
// The reference itself cannot be compiled here: every translation unit needs
// the Arcane framework >= 3.14.14 (CMakeLists.txt:10), which is absent.  This
// file restates the algorithm function by function; each function cites the
// reference lines it follows.  Arcane primitives used by the reference
// (GenericSorter, Scanner::exclusiveSum, doAtomic<Add>, math::cross/dot/normL2)
// are restated with their textbook semantics executed sequentially, which is
// what Arcane's sequential runner does with the very same lambdas.
//
// Parity pinning: the reference holds no matrix-level golden vectors.  The
// oracle is pinned through the reference's solution-level fixtures
// (modules/poisson/check/*.txt, modules/testlab/tests/*.txt,
// modules/elasticity/check/*.txt) replayed in tests/test_oracle_golden.py.
//
// Build: g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC (see oracle/Makefile).
// -ffp-contract=off pins the arithmetic to plain IEEE mul/add like the
// reference's host build.

#include <algorithm>
#if defined(_OPENMP)
#include <parallel/algorithm>
#endif
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#if defined(_OPENMP)
#include <omp.h>
#endif

namespace {

struct R3 { double x, y, z; };

inline R3 sub(R3 a, R3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
// Arcane math::cross / math::dot / Real3::normL2 (textbook definitions).
inline R3 cross(R3 u, R3 v) { return { u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x }; }
inline double dot(R3 u, R3 v) { return u.x * v.x + u.y * v.y + u.z * v.z; }
inline double normL2(R3 u) { return std::sqrt(u.x * u.x + u.y * u.y + u.z * u.z); }

inline R3 vtx(const double* coords, int32_t n) { return { coords[3 * (int64_t)n], coords[3 * (int64_t)n + 1], coords[3 * (int64_t)n + 2] }; }

// femutils/ArcaneFemFunctionsGpu.h:75-83 (CPU twin ArcaneFemFunctions.h:83-93)
inline double area_tria3(const int32_t* cn, const double* coords)
{
  R3 v0 = vtx(coords, cn[0]), v1 = vtx(coords, cn[1]), v2 = vtx(coords, cn[2]);
  R3 v = cross(sub(v1, v0), sub(v2, v0));
  return normL2(v) / 2.0;
}

// femutils/ArcaneFemFunctionsGpu.h:110-122 (CPU twin ArcaneFemFunctions.h:59-72)
inline double volume_tetra4(const int32_t* cn, const double* coords)
{
  R3 p0 = vtx(coords, cn[0]), p1 = vtx(coords, cn[1]), p2 = vtx(coords, cn[2]), p3 = vtx(coords, cn[3]);
  R3 v0 = sub(p1, p0), v1 = sub(p2, p0), v2 = sub(p3, p0);
  return std::fabs(dot(v0, cross(v1, v2))) / 6.0;
}

// femutils/ArcaneFemFunctionsGpu.h:218-227
inline void grad_x_tria3(const int32_t* cn, const double* coords, double* dx)
{
  R3 v0 = vtx(coords, cn[0]), v1 = vtx(coords, cn[1]), v2 = vtx(coords, cn[2]);
  double A2 = ((v1.x - v0.x) * (v2.y - v0.y) - (v2.x - v0.x) * (v1.y - v0.y));
  dx[0] = (v1.y - v2.y) / A2;
  dx[1] = (v2.y - v0.y) / A2;
  dx[2] = (v0.y - v1.y) / A2;
}

// femutils/ArcaneFemFunctionsGpu.h:243-252
inline void grad_y_tria3(const int32_t* cn, const double* coords, double* dy)
{
  R3 v0 = vtx(coords, cn[0]), v1 = vtx(coords, cn[1]), v2 = vtx(coords, cn[2]);
  double A2 = ((v1.x - v0.x) * (v2.y - v0.y) - (v2.x - v0.x) * (v1.y - v0.y));
  dy[0] = (v2.x - v1.x) / A2;
  dy[1] = (v0.x - v2.x) / A2;
  dy[2] = (v1.x - v0.x) / A2;
}

inline double v6_tetra4(R3 p0, R3 p1, R3 p2, R3 p3)
{
  R3 v0 = sub(p1, p0), v1 = sub(p2, p0), v2 = sub(p3, p0);
  return std::fabs(dot(v0, cross(v1, v2)));
}

// femutils/ArcaneFemFunctionsGpu.h:280-302
inline void grad_x_tetra4(const int32_t* cn, const double* coords, double* dx)
{
  R3 p0 = vtx(coords, cn[0]), p1 = vtx(coords, cn[1]), p2 = vtx(coords, cn[2]), p3 = vtx(coords, cn[3]);
  double V6 = v6_tetra4(p0, p1, p2, p3);
  dx[0] = (p1.y * (p3.z - p2.z) + p2.y * (p1.z - p3.z) + p3.y * (p2.z - p1.z)) / V6;
  dx[1] = (p0.y * (p2.z - p3.z) + p2.y * (p3.z - p0.z) + p3.y * (p0.z - p2.z)) / V6;
  dx[2] = (p0.y * (p3.z - p1.z) + p1.y * (p0.z - p3.z) + p3.y * (p1.z - p0.z)) / V6;
  dx[3] = (p0.y * (p1.z - p2.z) + p1.y * (p2.z - p0.z) + p2.y * (p0.z - p1.z)) / V6;
}

// femutils/ArcaneFemFunctionsGpu.h:325-347
inline void grad_y_tetra4(const int32_t* cn, const double* coords, double* dy)
{
  R3 p0 = vtx(coords, cn[0]), p1 = vtx(coords, cn[1]), p2 = vtx(coords, cn[2]), p3 = vtx(coords, cn[3]);
  double V6 = v6_tetra4(p0, p1, p2, p3);
  dy[0] = (p1.z * (p3.x - p2.x) + p2.z * (p1.x - p3.x) + p3.z * (p2.x - p1.x)) / V6;
  dy[1] = (p0.z * (p2.x - p3.x) + p2.z * (p3.x - p0.x) + p3.z * (p0.x - p2.x)) / V6;
  dy[2] = (p0.z * (p3.x - p1.x) + p1.z * (p0.x - p3.x) + p3.z * (p1.x - p0.x)) / V6;
  dy[3] = (p0.z * (p1.x - p2.x) + p1.z * (p2.x - p0.x) + p2.z * (p0.x - p1.x)) / V6;
}

// femutils/ArcaneFemFunctionsGpu.h:370-392
inline void grad_z_tetra4(const int32_t* cn, const double* coords, double* dz)
{
  R3 p0 = vtx(coords, cn[0]), p1 = vtx(coords, cn[1]), p2 = vtx(coords, cn[2]), p3 = vtx(coords, cn[3]);
  double V6 = v6_tetra4(p0, p1, p2, p3);
  dz[0] = (p1.x * (p3.y - p2.y) + p2.x * (p1.y - p3.y) + p3.x * (p2.y - p1.y)) / V6;
  dz[1] = (p0.x * (p2.y - p3.y) + p2.x * (p3.y - p0.y) + p3.x * (p0.y - p2.y)) / V6;
  dz[2] = (p0.x * (p3.y - p1.y) + p1.x * (p0.y - p3.y) + p3.x * (p1.y - p0.y)) / V6;
  dz[3] = (p0.x * (p1.y - p2.y) + p1.x * (p2.y - p0.y) + p2.x * (p0.y - p1.y)) / V6;
}

// Operator ids shared with include/b2f.h (kept numerically identical; the
// oracle does not include the product header on purpose).
enum { OP_LAPLACE = 0, OP_ELASTICITY = 1, OP_HEAT = 2 };

struct OpParams
{
  int32_t op;
  double lambda; // elasticity: Lame lambda; heat: default conductivity
  double mu2; // elasticity: 2*mu
  double inv_dt; // heat: 1/dt
  const double* cell_lambda; // heat: optional per-cell conductivity
};

// Element matrix, row-major [(NV*BS) x (NV*BS)], entry (BS*a+i, BS*b+j)
// (contract of the callback: femutils/BSRFormat.h:823,881).
void element_matrix(int nv, int bs, const OpParams& p, int64_t cell, const int32_t* cn, const double* coords, double* Ke)
{
  const int n = nv * bs;
  if (p.op == OP_LAPLACE && nv == 3 && bs == 1) {
    // modules/poisson/FemModule.h:139-147: area*(dx^dx) + area*(dy^dy)
    double area = area_tria3(cn, coords);
    double dx[3], dy[3];
    grad_x_tria3(cn, coords, dx);
    grad_y_tria3(cn, coords, dy);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        Ke[i * 3 + j] = area * (dx[i] * dx[j]) + area * (dy[i] * dy[j]);
    return;
  }
  if (p.op == OP_LAPLACE && nv == 4 && bs == 1) {
    // modules/poisson/FemModule.h:177-186: V*(dx^dx) + V*(dy^dy) + V*(dz^dz), left to right
    double volume = volume_tetra4(cn, coords);
    double dx[4], dy[4], dz[4];
    grad_x_tetra4(cn, coords, dx);
    grad_y_tetra4(cn, coords, dy);
    grad_z_tetra4(cn, coords, dz);
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j)
        Ke[i * 4 + j] = (volume * (dx[i] * dx[j]) + volume * (dy[i] * dy[j])) + volume * (dz[i] * dz[j]);
    return;
  }
  if (p.op == OP_ELASTICITY && nv == 3 && bs == 2) {
    // modules/elasticity/FemModule.h:112-152 (computeElementMatrixTRIA3Base), evaluated entry by
    // entry with the reference's association:
    //   step1 = (((a1 + a2) + a3) + a4) * lambda / (4*area)
    //   step2 = (((b1 + b2) + (((c1 + c2) + c3) + c4) * 0.5) * mu2) / (4*area)
    R3 m0 = vtx(coords, cn[0]), m1 = vtx(coords, cn[1]), m2 = vtx(coords, cn[2]);
    double area = area_tria3(cn, coords);
    double dPx[3] = { m1.y - m2.y, m2.y - m0.y, m0.y - m1.y };
    double dPy[3] = { m2.x - m1.x, m0.x - m2.x, m1.x - m0.x };
    double dxU1[6], dyU1[6], dxU2[6], dyU2[6];
    for (int a = 0; a < 3; ++a) {
      dxU1[2 * a] = dPx[a]; dxU1[2 * a + 1] = 0.;
      dyU1[2 * a] = dPy[a]; dyU1[2 * a + 1] = 0.;
      dxU2[2 * a] = 0.; dxU2[2 * a + 1] = dPx[a];
      dyU2[2 * a] = 0.; dyU2[2 * a + 1] = dPy[a];
    }
    const double four_area = 4 * area;
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) {
        // (dxV1 ^ dxU1)(i,j) = dxU1[i]*dxU1[j] etc. (V = transpose of U)
        double s1 = (((dxU1[i] * dxU1[j] + dxU1[i] * dyU2[j]) + dyU2[i] * dxU1[j]) + dyU2[i] * dyU2[j]);
        double step1 = (s1 * p.lambda) / four_area;
        double s2a = dxU1[i] * dxU1[j] + dyU2[i] * dyU2[j];
        double s2b = (((dyU1[i] * dyU1[j] + dyU1[i] * dxU2[j]) + dxU2[i] * dyU1[j]) + dxU2[i] * dxU2[j]);
        double step2 = ((s2a + s2b * 0.5) * p.mu2) / four_area;
        Ke[i * 6 + j] = step1 + step2;
      }
    return;
  }
  if (p.op == OP_ELASTICITY && nv == 4 && bs == 3) {
    // NOT in the reference (elasticity there is 2D only).  Tetrahedral extension of
    // modules/elasticity/FemModule.h:112-140:
    //   Ke[(a,i),(b,j)] = V*(lambda*g_a^i*g_b^j + mu*g_a^j*g_b^i + mu*delta_ij*(g_a.g_b))
    // with g from the 3D gradient helpers (ArcaneFemFunctionsGpu.h:280-392), mu = mu2/2.
    double volume = volume_tetra4(cn, coords);
    double g[3][4];
    grad_x_tetra4(cn, coords, g[0]);
    grad_y_tetra4(cn, coords, g[1]);
    grad_z_tetra4(cn, coords, g[2]);
    const double mu = p.mu2 * 0.5;
    for (int a = 0; a < 4; ++a)
      for (int b = 0; b < 4; ++b) {
        double gab = (g[0][a] * g[0][b] + g[1][a] * g[1][b]) + g[2][a] * g[2][b];
        for (int i = 0; i < 3; ++i)
          for (int j = 0; j < 3; ++j) {
            double v = p.lambda * (g[i][a] * g[j][b]) + mu * (g[j][a] * g[i][b]);
            if (i == j)
              v += mu * gab;
            Ke[(3 * a + i) * n + (3 * b + j)] = volume * v;
          }
      }
    return;
  }
  if (p.op == OP_HEAT && nv == 3 && bs == 1) {
    // modules/heat/FemModule.cc:678-770 (_computeElementMatrixTRIA3), signed area :606-612.
    R3 m0 = vtx(coords, cn[0]), m1 = vtx(coords, cn[1]), m2 = vtx(coords, cn[2]);
    double area = 0.5 * ((m1.x - m0.x) * (m2.y - m0.y) - (m2.x - m0.x) * (m1.y - m0.y));
    double lambda = p.cell_lambda ? p.cell_lambda[cell] : p.lambda;
    double dPx[3] = { m1.y - m2.y, m2.y - m0.y, m0.y - m1.y };
    double dPy[3] = { m2.x - m1.x, m0.x - m2.x, m1.x - m0.x };
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        // bT(i) = dPhi_i * 0.5 ; b(j) = (dPhi_j / area) * 0.5 ; matrixMultiplication: 0.0 + a*b
        double kxx = 0.0 + (dPx[i] * 0.5) * ((dPx[j] / area) * 0.5);
        double kyy = 0.0 + (dPy[i] * 0.5) * ((dPy[j] / area) * 0.5);
        double k = ((0. + kxx) + kyy) * lambda;
        double uv = 0.0 + ((area / 3.) * 0.5) * (1. * 0.5);
        if (i == j)
          uv *= 2.;
        uv *= p.inv_dt; // reference: multInPlace(1./dt)
        Ke[i * 3 + j] = k + uv;
      }
    return;
  }
  if (p.op == OP_HEAT && nv == 4 && bs == 1) {
    // NOT in the reference (heat there is 2D only).  3D twin: lambda*V*(grad.grad) + (V/20)(1+delta)/dt
    // (P1 consistent mass on a tetrahedron), gradients from ArcaneFemFunctionsGpu.h:280-392.
    double volume = volume_tetra4(cn, coords);
    double lambda = p.cell_lambda ? p.cell_lambda[cell] : p.lambda;
    double dx[4], dy[4], dz[4];
    grad_x_tetra4(cn, coords, dx);
    grad_y_tetra4(cn, coords, dy);
    grad_z_tetra4(cn, coords, dz);
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j) {
        double k = ((volume * (dx[i] * dx[j]) + volume * (dy[i] * dy[j])) + volume * (dz[i] * dz[j])) * lambda;
        double uv = (volume / 20.) * (i == j ? 2. : 1.) * p.inv_dt;
        Ke[i * 4 + j] = k + uv;
      }
    return;
  }
  for (int i = 0; i < n * n; ++i)
    Ke[i] = std::nan("");
}

// femutils/BSRFormat.h:583-588
inline uint64_t pack(int32_t n0, int32_t n1)
{
  int32_t mn = n0 > n1 ? n1 : n0;
  int32_t mx = n0 > n1 ? n0 : n1;
  return ((uint64_t)mn << 32) | (uint64_t)(uint32_t)mx;
}
// femutils/BSRFormat.h:593-597
inline void unpack(uint64_t e, int32_t& n0, int32_t& n1)
{
  n0 = (int32_t)(e >> 32);
  n1 = (int32_t)(e & 0xFFFFFFFF);
}

// Row end without a sentinel (femutils/BSRFormat.h:151,817): the oracle carries
// row_index with nb_node+1 entries, entry nb_node == nb_col, which is equivalent.

// Value index of scalar (row dof, col dof): femutils/BSRFormat.h:145-171 (findValueIndex)
inline int64_t value_index(int bs, bool per_block, const int32_t* row_index, int64_t block_row, int64_t x /*block position in row*/, int i, int j)
{
  int64_t start = row_index[block_row];
  int64_t k = row_index[block_row + 1] - start;
  if (!per_block)
    return start * bs * bs + (int64_t)bs * x + (int64_t)i * bs * k + j;
  return (start + x) * bs * bs + (i * bs + j);
}

} // namespace

extern "C" {

int orc_num_threads()
{
#if defined(_OPENMP)
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void orc_free(void* p) { std::free(p); }

// ---------------------------------------------------------------------------
// Sparsity, sorted-edge ("atomic") variant executed sequentially.
// femutils/BSRFormat.h:602-651 (computeSortedEdges), :656-688 (computeNeighbors,
// computeRowIndex), :693-744 (registerEdgeInColumns, computeColumns), :749-770.
// Output: row_index[nb_node+1] (caller-allocated), *columns_out malloc'd [nnzb].
// Returns nnzb (= 2E + nb_node, BSRFormat.h:401), or -1 on error.
// ---------------------------------------------------------------------------
int64_t orc_build_pattern(int nv, int64_t nb_cell, int64_t nb_node, const int32_t* cell_node, int32_t* row_index, int32_t** columns_out)
{
  const int epe = (nv == 3) ? 3 : 6;
  const int64_t nb_edge_total = nb_cell * epe;
  std::vector<uint64_t> edges((size_t)nb_edge_total);
  for (int64_t c = 0; c < nb_cell; ++c) {
    const int32_t* cn = cell_node + c * nv;
    uint64_t* e = edges.data() + c * epe;
    if (nv == 3) {
      e[0] = pack(cn[0], cn[1]);
      e[1] = pack(cn[0], cn[2]);
      e[2] = pack(cn[1], cn[2]);
    }
    else {
      e[0] = pack(cn[0], cn[1]);
      e[1] = pack(cn[0], cn[2]);
      e[2] = pack(cn[0], cn[3]);
      e[3] = pack(cn[1], cn[2]);
      e[4] = pack(cn[1], cn[3]);
      e[5] = pack(cn[2], cn[3]);
    }
  }
  // GenericSorter::apply, BSRFormat.h:648-650 (a sort of integers: the result does not depend on the algorithm)
#if defined(_OPENMP)
  __gnu_parallel::sort(edges.begin(), edges.end());
#else
  std::sort(edges.begin(), edges.end());
#endif

  std::vector<int32_t> neighbors((size_t)nb_node, 1); // fill(1): the diagonal, :682
  for (int64_t t = 0; t < nb_edge_total; ++t) {
    uint64_t cur = edges[t];
    if (t == nb_edge_total - 1 || cur != edges[t + 1]) {
      int32_t n0, n1;
      unpack(cur, n0, n1);
      neighbors[n0] += 1;
      neighbors[n1] += 1;
    }
  }
  // exclusive scan (:685-687); entry nb_node closes the last row (== nb_col)
  int64_t acc = 0;
  for (int64_t n = 0; n < nb_node; ++n) {
    row_index[n] = (int32_t)acc;
    acc += neighbors[n];
  }
  row_index[nb_node] = (int32_t)acc;
  const int64_t nnzb = acc;

  int32_t* columns = (int32_t*)std::malloc(sizeof(int32_t) * (size_t)std::max<int64_t>(nnzb, 1));
  if (!columns)
    return -1;
  for (int64_t n = 0; n < nb_node; ++n)
    columns[row_index[n]] = (int32_t)n; // diagonal first, :712-717
  std::vector<int32_t> offsets((size_t)nb_node, 1); // :724
  for (int64_t t = 0; t < nb_edge_total; ++t) {
    uint64_t cur = edges[t];
    if (t == nb_edge_total - 1 || cur != edges[t + 1]) {
      int32_t n0, n1;
      unpack(cur, n0, n1);
      columns[row_index[n0] + offsets[n0]++] = n1; // registerEdgeInColumns(n0, n1), :739
      columns[row_index[n1] + offsets[n1]++] = n0; // registerEdgeInColumns(n1, n0), :740
    }
  }
  *columns_out = columns;
  return nnzb;
}

// nb_nz_per_row: femutils/BSRFormat.h:435-474
void orc_nz_per_row(int64_t nb_node, const int32_t* row_index, int32_t* nb_nz_per_row)
{
  for (int64_t i = 0; i < nb_node; ++i)
    nb_nz_per_row[i] = row_index[i + 1] - row_index[i];
}

// Element matrix for one cell (exposed so tests can cross-check Ke against independent derivations).
void orc_element_matrix(int nv, int bs, int op, double lambda, double mu2, double inv_dt, const double* cell_lambda, int64_t cell, const int32_t* cell_node, const double* coords, double* Ke)
{
  OpParams p{ op, lambda, mu2, inv_dt, cell_lambda };
  element_matrix(nv, bs, p, cell, cell_node + cell * nv, coords, Ke);
}

// ---------------------------------------------------------------------------
// Scatter-add, cell-wise (atomic variant executed sequentially).
// per-block: femutils/BSRFormat.h:786-837 ; per-row (CSR-compatible): :842-898.
// Deviation (documented in DESIGN.md): for the per-row layout the row offset is
// i*(BS*k) as in findValueIndex (:162) instead of the `if (i != 0)` single step
// (:883-884) which is only correct for BS <= 2 — identical results for BS <= 2.
// node_is_own may be NULL (all own).  Returns the number of (row,col) pairs whose
// column was not found in the pattern (the reference skips them silently).
// ---------------------------------------------------------------------------
int64_t orc_assemble_cellwise(int nv, int bs, int per_block, int op, double lambda, double mu2, double inv_dt, const double* cell_lambda,
                              int64_t nb_cell, int64_t nb_node, const int32_t* cell_node, const double* coords, const uint8_t* node_is_own,
                              const int32_t* row_index, const int32_t* columns, double* values)
{
  OpParams p{ op, lambda, mu2, inv_dt, cell_lambda };
  const int n = nv * bs;
  std::vector<double> Ke((size_t)n * n);
  int64_t not_found = 0;
  for (int64_t cell = 0; cell < nb_cell; ++cell) {
    const int32_t* cn = cell_node + cell * nv;
    element_matrix(nv, bs, p, cell, cn, coords, Ke.data());
    for (int a = 0; a < nv; ++a) {
      int32_t row = cn[a];
      if (node_is_own && !node_is_own[row])
        continue;
      for (int b = 0; b < nv; ++b) {
        int32_t col = cn[b];
        int32_t begin = row_index[row], end = row_index[row + 1];
        bool found = false;
        for (int32_t x = begin; x < end; ++x) {
          if (columns[x] == col) {
            for (int i = 0; i < bs; ++i)
              for (int j = 0; j < bs; ++j)
                values[value_index(bs, per_block != 0, row_index, row, x - begin, i, j)] += Ke[(bs * a + i) * n + (bs * b + j)];
            found = true;
            break;
          }
        }
        if (!found)
          ++not_found;
      }
    }
  }
  (void)nb_node;
  return not_found;
}

// ---------------------------------------------------------------------------
// Scatter-add, node-wise ("atomic-free") variant: femutils/BSRFormat.h:937-998
// (per-block) and :1003-1068 (per-row).  Restated with the two indexing defects
// fixed (local-index search limited to i<=3 on triangles :962,1029; column index
// incremented inside the isOwn branch :1063).  node->cell lists are built in
// ascending cell order (Arcane's nodeCell() order is not reproducible here; the
// order only changes the fp64 summation order).
// ---------------------------------------------------------------------------
int64_t orc_assemble_nodewise(int nv, int bs, int per_block, int op, double lambda, double mu2, double inv_dt, const double* cell_lambda,
                              int64_t nb_cell, int64_t nb_node, const int32_t* cell_node, const double* coords, const uint8_t* node_is_own,
                              const int32_t* row_index, const int32_t* columns, double* values)
{
  OpParams p{ op, lambda, mu2, inv_dt, cell_lambda };
  const int n = nv * bs;
  std::vector<int64_t> ptr((size_t)nb_node + 1, 0);
  for (int64_t c = 0; c < nb_cell * nv; ++c)
    ptr[cell_node[c] + 1]++;
  for (int64_t i = 0; i < nb_node; ++i)
    ptr[i + 1] += ptr[i];
  std::vector<int64_t> cur(ptr.begin(), ptr.end() - 1);
  std::vector<int64_t> lst((size_t)(nb_cell * nv));
  for (int64_t c = 0; c < nb_cell; ++c)
    for (int a = 0; a < nv; ++a)
      lst[cur[cell_node[c * nv + a]]++] = c * nv + a;
  std::vector<double> Ke((size_t)n * n);
  int64_t not_found = 0;
  for (int64_t row = 0; row < nb_node; ++row) {
    if (node_is_own && !node_is_own[row])
      continue;
    for (int64_t q = ptr[row]; q < ptr[row + 1]; ++q) {
      int64_t cell = lst[q] / nv;
      int a = (int)(lst[q] % nv);
      const int32_t* cn = cell_node + cell * nv;
      element_matrix(nv, bs, p, cell, cn, coords, Ke.data());
      for (int b = 0; b < nv; ++b) {
        int32_t col = cn[b];
        int32_t begin = row_index[row], end = row_index[row + 1];
        bool found = false;
        for (int32_t x = begin; x < end; ++x) {
          if (columns[x] == col) {
            for (int i = 0; i < bs; ++i)
              for (int j = 0; j < bs; ++j)
                values[value_index(bs, per_block != 0, row_index, row, x - begin, i, j)] += Ke[(bs * a + i) * n + (bs * b + j)];
            found = true;
            break;
          }
        }
        if (!found)
          ++not_found;
      }
    }
  }
  return not_found;
}

// ---------------------------------------------------------------------------
// BSR -> CSR: femutils/BSRFormat.h:194-256 (toCsr) for the per-row value layout
// (values are shared as is).  rows has BS*nb_node+1 entries here (reference: BS*nb_node,
// and for BS=3 one entry is left unset, :213-250 — fixed).
// ---------------------------------------------------------------------------
void orc_to_csr(int bs, int64_t nb_node, const int32_t* row_index, const int32_t* columns, int32_t* csr_rows, int32_t* csr_rows_nb_column, int32_t* csr_columns)
{
  int64_t off = 0;
  int64_t r = 0;
  for (int64_t n = 0; n < nb_node; ++n) {
    int32_t start = row_index[n], end = row_index[n + 1];
    for (int i = 0; i < bs; ++i) {
      csr_rows[r] = (int32_t)off;
      csr_rows_nb_column[r] = (end - start) * bs;
      for (int32_t x = start; x < end; ++x)
        for (int k = 0; k < bs; ++k)
          csr_columns[off++] = columns[x] * bs + k;
      ++r;
    }
  }
  csr_rows[r] = (int32_t)off;
}

// Dense-free SpMV on the CSR arrays, y = A x (check helper; sequential, left-to-right row sums).
void orc_csr_spmv(int64_t nb_row, const int32_t* rows, const int32_t* columns, const double* values, const double* x, double* y)
{
  for (int64_t r = 0; r < nb_row; ++r) {
    double s = 0.;
    for (int32_t q = rows[r]; q < rows[r + 1]; ++q)
      s += values[q] * x[columns[q]];
    y[r] = s;
  }
}

// ---------------------------------------------------------------------------
// RHS: constant source.  femutils/ArcaneFemFunctionsGpu.h:402-429
// (applyConstantSourceToRhsBase), :605-608 (2D: area), :695-698 (3D: volume).
// rhs[dof(node,0)] += qdot*|cell|/nbNode(cell) for own nodes; dof(node,i) = node*bs+i
// (femutils/FemDoFsOnNodes.cc:79-109).
// ---------------------------------------------------------------------------
void orc_rhs_constant_source(int nv, int bs, double qdot, int64_t nb_cell, const int32_t* cell_node, const double* coords, const uint8_t* node_is_own, double* rhs)
{
  for (int64_t cell = 0; cell < nb_cell; ++cell) {
    const int32_t* cn = cell_node + cell * nv;
    double domain = (nv == 3) ? area_tria3(cn, coords) : volume_tetra4(cn, coords);
    for (int a = 0; a < nv; ++a) {
      if (node_is_own && !node_is_own[cn[a]])
        continue;
      double value = qdot * domain / nv;
      rhs[(int64_t)cn[a] * bs] += value;
    }
  }
}

// ---------------------------------------------------------------------------
// RHS: nodal-field source of the transient heat module, modules/heat/FemModule.cc:480-495:
// rhs[dof(node,0)] += (T_old[node] / dt) * area / 3 over every cell, for own nodes whose temperature is not fixed.
// The area is the module's own signed 2D formula (_computeAreaTriangle3, :605-612); nv = 4 is the 3D twin (|det| / 6).
// scale = 1 / dt, node_skip = m_node_is_temperature_fixed (may be null).
// ---------------------------------------------------------------------------
void orc_rhs_nodal_source(int nv, int bs, double scale, const double* node_field, const uint8_t* node_skip, int64_t nb_cell, const int32_t* cell_node,
                          const double* coords, const uint8_t* node_is_own, double* rhs)
{
  for (int64_t cell = 0; cell < nb_cell; ++cell) {
    const int32_t* cn = cell_node + cell * nv;
    double domain;
    if (nv == 3) {
      R3 m0 = vtx(coords, cn[0]), m1 = vtx(coords, cn[1]), m2 = vtx(coords, cn[2]);
      domain = 0.5 * ((m1.x - m0.x) * (m2.y - m0.y) - (m2.x - m0.x) * (m1.y - m0.y));
    }
    else
      domain = volume_tetra4(cn, coords);
    for (int a = 0; a < nv; ++a) {
      if ((node_is_own && !node_is_own[cn[a]]) || (node_skip && node_skip[cn[a]]))
        continue;
      rhs[(int64_t)cn[a] * bs] += (node_field[cn[a]] * scale) * domain / nv;
    }
  }
}

// ---------------------------------------------------------------------------
// RHS: Neumann.  2D femutils/ArcaneFemFunctionsGpu.h:621-677 with computeLengthFace :132-138 and
// computeNormalFace :149-165 ; 3D :711-768 with computeAreaTria :88-96 and computeNormalTriangle
// :176-197.  face_outside[f] restates FaceInfoListView::isSubDomainBoundaryOutside (NULL = all true).
// nvf = 2 (edges of a 2D mesh) or 3 (triangles of a 3D mesh).
// ---------------------------------------------------------------------------
void orc_rhs_neumann(int nvf, int bs, int has_value, double value, double vx, double vy, double vz, int64_t nb_face, const int32_t* face_node,
                     const uint8_t* face_outside, const double* coords, const uint8_t* node_is_own, double* rhs)
{
  for (int64_t f = 0; f < nb_face; ++f) {
    const int32_t* fn = face_node + f * nvf;
    double contrib;
    if (nvf == 2) {
      R3 m0 = vtx(coords, fn[0]), m1 = vtx(coords, fn[1]);
      double length = std::sqrt((m1.x - m0.x) * (m1.x - m0.x) + (m1.y - m0.y) * (m1.y - m0.y));
      if (has_value)
        contrib = value * length / 2.0;
      else {
        if (face_outside && !face_outside[f])
          std::swap(m0, m1);
        double norm_N = std::sqrt((m1.y - m0.y) * (m1.y - m0.y) + (m1.x - m0.x) * (m1.x - m0.x));
        double Nx = (m1.y - m0.y) / norm_N;
        double Ny = (m0.x - m1.x) / norm_N;
        contrib = (Nx * vx + Ny * vy) * length / 2.0;
      }
    }
    else {
      R3 p0 = vtx(coords, fn[0]), p1 = vtx(coords, fn[1]), p2 = vtx(coords, fn[2]);
      double area = normL2(cross(sub(p1, p0), sub(p2, p0))) / 2.0;
      if (has_value)
        contrib = value * area / 3.0;
      else {
        if (face_outside && !face_outside[f])
          std::swap(p0, p1);
        R3 e1 = { p1.x - p0.x, p1.y - p0.y, p1.z - p0.z };
        R3 e2 = { p2.x - p0.x, p2.y - p0.y, p2.z - p0.z };
        R3 nrm = { e1.y * e2.z - e1.z * e2.y, e1.z * e2.x - e1.x * e2.z, e1.x * e2.y - e1.y * e2.x };
        double norm = std::sqrt(nrm.x * nrm.x + nrm.y * nrm.y + nrm.z * nrm.z);
        double nx = nrm.x / norm, ny = nrm.y / norm, nz = nrm.z / norm;
        contrib = (nx * vx + ny * vy + nz * vz) * area / 3.0;
      }
    }
    for (int a = 0; a < nvf; ++a) {
      if (node_is_own && !node_is_own[fn[a]])
        continue;
      rhs[(int64_t)fn[a] * bs] += contrib;
    }
  }
}

// ---------------------------------------------------------------------------
// Convection (Robin) boundary term of the heat module: modules/heat/FemModule.cc:628-672 (_computeElementMatrixEDGE2:
// bT = L/3, b = 1/2, diagonal doubled, times h  =>  h * L/6 * [[2,1],[1,2]], L from x and y only, :617-623) scattered by
// _assembleBilinearOperatorEDGE2 :813-840 into own rows.  nvf = 3 is the 3D twin (not in the reference):
// h * A/12 * (1 + delta_ab) over the triangle's nodes, A by the 3D cross product.  BS = 1.
// Returns the number of (row, col) pairs missing from the pattern.
// ---------------------------------------------------------------------------
int64_t orc_add_face_mass(int nvf, double h, int64_t nb_face, const int32_t* face_node, const double* coords, const uint8_t* node_is_own,
                          const int32_t* row_index, const int32_t* columns, double* values)
{
  int64_t not_found = 0;
  for (int64_t f = 0; f < nb_face; ++f) {
    const int32_t* fn = face_node + f * nvf;
    double measure;
    if (nvf == 2) {
      R3 m0 = vtx(coords, fn[0]), m1 = vtx(coords, fn[1]);
      measure = std::sqrt((m1.x - m0.x) * (m1.x - m0.x) + (m1.y - m0.y) * (m1.y - m0.y));
    }
    else {
      R3 p0 = vtx(coords, fn[0]), p1 = vtx(coords, fn[1]), p2 = vtx(coords, fn[2]);
      measure = normL2(cross(sub(p1, p0), sub(p2, p0))) / 2.0;
    }
    const double off = (nvf == 2) ? (measure / 3.0) * 0.5 * h : (measure / 12.0) * h;
    for (int a = 0; a < nvf; ++a) {
      if (node_is_own && !node_is_own[fn[a]])
        continue;
      for (int b = 0; b < nvf; ++b) {
        const double v = (a == b) ? off * 2.0 : off;
        int64_t idx = -1;
        for (int32_t q = row_index[fn[a]]; q < row_index[fn[a] + 1]; ++q)
          if (columns[q] == fn[b])
            idx = q;
        if (idx < 0)
          ++not_found;
        else
          values[idx] += v;
      }
    }
  }
  return not_found;
}

// ---------------------------------------------------------------------------
// Dirichlet flagging.  Penalty: femutils/ArcaneFemFunctionsGpu.h:434-456.
// Row elimination flag: :461-482 (mode 1 = ELIMINATE_ROW; 2 = ELIMINATE_ROW_COLUMN,
// femutils/AlephDoFLinearSystem.cc semantics).
// ---------------------------------------------------------------------------
void orc_dirichlet_flag_penalty(int bs, int dof_index, double value, double penalty, int weak, int64_t n, const int32_t* node_ids, const uint8_t* node_is_own,
                                uint8_t* forced_info, double* forced_value, double* rhs)
{
  for (int64_t k = 0; k < n; ++k) {
    int32_t node = node_ids[k];
    if (node_is_own && !node_is_own[node])
      continue;
    int64_t dof = (int64_t)node * bs + dof_index;
    forced_info[dof] = weak ? 2 : 1; // weak: modules/testlab/FemModule.cc:687-717 (a_ii += P)
    forced_value[dof] = penalty;
    rhs[dof] = penalty * value;
  }
}

void orc_dirichlet_flag_elimination(int bs, int dof_index, double value, int mode, int64_t n, const int32_t* node_ids, const uint8_t* node_is_own,
                                    uint8_t* elimination_info, double* elimination_value)
{
  for (int64_t k = 0; k < n; ++k) {
    int32_t node = node_ids[k];
    if (node_is_own && !node_is_own[node])
      continue;
    int64_t dof = (int64_t)node * bs + dof_index;
    elimination_info[dof] = (uint8_t)mode;
    elimination_value[dof] = value;
  }
}

// ---------------------------------------------------------------------------
// Dirichlet application on the CSR view, in the order of HypreDoFLinearSystemImpl::solve()
// (femutils/HypreDoFLinearSystem.cc:387-391): _applyRowElimination (:319-351) then
// _applyForcedValuesToLhs (:356-382, Gpu::Csr::findIndex ArcaneFemFunctionsGpu.h:53-59).
// Row+column elimination (mode 2) follows femutils/AlephDoFLinearSystem.cc:501-583:
//   phase 1: for every stored (r,c), r != c, r flagged ROW_COLUMN: rhs[c] -= A[r,c]*g_r
//            (visited in (r,c) lexicographic order, the std::map order);
//   then A[r,:] = A[:,r] = 0, A[r,r] = 1, rhs[r] = g_r.
// ---------------------------------------------------------------------------
void orc_dirichlet_apply(int64_t nb_dof, const int32_t* rows, const int32_t* columns, double* values, const uint8_t* elimination_info, const double* elimination_value,
                         const uint8_t* forced_info, const double* forced_value, double* rhs)
{
  // row+column, phase 1 (needs a column-sorted visit of each row to mimic the map order)
  bool any_rc = false;
  for (int64_t d = 0; d < nb_dof; ++d)
    if (elimination_info[d] == 2) {
      any_rc = true;
      break;
    }
  if (any_rc) {
    std::vector<std::pair<int32_t, int32_t>> ord;
    for (int64_t r = 0; r < nb_dof; ++r) {
      if (elimination_info[r] != 2)
        continue;
      ord.clear();
      for (int32_t q = rows[r]; q < rows[r + 1]; ++q)
        ord.push_back({ columns[q], q });
      std::sort(ord.begin(), ord.end());
      for (auto& cq : ord) {
        if (cq.first == r)
          continue;
        rhs[cq.first] = rhs[cq.first] - values[cq.second] * elimination_value[r];
      }
    }
    for (int64_t r = 0; r < nb_dof; ++r)
      for (int32_t q = rows[r]; q < rows[r + 1]; ++q) {
        int32_t c = columns[q];
        if (elimination_info[r] == 2 || elimination_info[c] == 2)
          values[q] = (r == c) ? 1.0 : 0.0;
      }
    for (int64_t r = 0; r < nb_dof; ++r)
      if (elimination_info[r] == 2)
        rhs[r] = elimination_value[r];
  }
  // _applyRowElimination
  for (int64_t d = 0; d < nb_dof; ++d) {
    if (elimination_info[d] == 1) {
      for (int32_t q = rows[d]; q < rows[d + 1]; ++q)
        values[q] = (columns[q] == d) ? 1 : 0;
      rhs[d] = elimination_value[d];
    }
  }
  // _applyForcedValuesToLhs
  for (int64_t d = 0; d < nb_dof; ++d) {
    if (forced_info[d]) {
      for (int32_t q = rows[d]; q < rows[d + 1]; ++q)
        if (columns[q] == d) {
          values[q] = (forced_info[d] == 2) ? values[q] + forced_value[d] : forced_value[d]; // 2: WeakPenalty
          break;
        }
    }
  }
}

// ---------------------------------------------------------------------------
// CPU baseline (timed by bench.py): the reference's legacy CPU assembly loop,
// modules/poisson/FemModule.cc:300-323 (cell loop -> Ke -> matrixAddValue per entry) with the
// CSR add of femutils/CsrFormatMatrix.h:58-87 (skip exact zeros, linear column search).
// threads <= 1: sequential (Arcane sequential runner).  threads > 1: OpenMP over cells with
// atomic adds — the analogue of running the same lambdas with Arcane's `-A,T=n` thread runner.
// Returns wall seconds of the loop.
// ---------------------------------------------------------------------------
double orc_legacy_assemble_csr(int nv, int op, double lambda, double mu2, double inv_dt, int64_t nb_cell, const int32_t* cell_node, const double* coords,
                               const int32_t* row_index, const int32_t* columns, double* values, int threads)
{
  OpParams p{ op, lambda, mu2, inv_dt, nullptr };
#if defined(_OPENMP)
  double t0 = omp_get_wtime();
#else
  double t0 = 0;
#endif
  if (threads <= 1) {
    double Ke[16];
    for (int64_t cell = 0; cell < nb_cell; ++cell) {
      const int32_t* cn = cell_node + cell * nv;
      element_matrix(nv, 1, p, cell, cn, coords, Ke);
      for (int a = 0; a < nv; ++a)
        for (int b = 0; b < nv; ++b) {
          double v = Ke[a * nv + b];
          if (v == 0.0)
            continue;
          int32_t row = cn[a], col = cn[b];
          for (int32_t x = row_index[row]; x < row_index[row + 1]; ++x)
            if (columns[x] == col) {
              values[x] += v;
              break;
            }
        }
    }
  }
  else {
#if defined(_OPENMP)
#pragma omp parallel for schedule(static) num_threads(threads)
#endif
    for (int64_t cell = 0; cell < nb_cell; ++cell) {
      double Ke[16];
      const int32_t* cn = cell_node + cell * nv;
      element_matrix(nv, 1, p, cell, cn, coords, Ke);
      for (int a = 0; a < nv; ++a)
        for (int b = 0; b < nv; ++b) {
          double v = Ke[a * nv + b];
          if (v == 0.0)
            continue;
          int32_t row = cn[a], col = cn[b];
          for (int32_t x = row_index[row]; x < row_index[row + 1]; ++x)
            if (columns[x] == col) {
#if defined(_OPENMP)
#pragma omp atomic
#endif
              values[x] += v;
              break;
            }
        }
    }
  }
#if defined(_OPENMP)
  return omp_get_wtime() - t0;
#else
  return 0;
#endif
}

} // extern "C"
