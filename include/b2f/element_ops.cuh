// element_ops.cuh — P1 element operators evaluated in registers (sm_100a device code).
//
// Each operator replaces one shipped instance of the reference's element callback
// `FixedMatrix<BS*NV,BS*NV> compute_element_matrix(CellLocalId)` (femutils/BSRFormat.h:786-787,1105;
// entry (BS*a+i, BS*b+j) couples dof i of local node a with dof j of local node b, :823,881):
//   LaplaceTri / LaplaceTet      modules/poisson/FemModule.h:139-147 / :177-186
//                                (helpers femutils/ArcaneFemFunctionsGpu.h:75-83,110-122,218-252,280-392)
//   ElasticityTri (BS=2)         modules/elasticity/FemModule.h:112-152
//   ElasticityTet (BS=3)         tetrahedral extension (not in the reference)
//   HeatTri                      modules/heat/FemModule.cc:678-770 ; HeatTet: 3D twin (not in the reference)
//
// B200-first formulation: the reference gathers the cell's vertices four times and divides every
// cofactor by V6; here the vertices are gathered once, gradients are kept as un-normalised cofactors of
// the edge vectors (no absolute-coordinate cancellation), and a single reciprocal scales the result:
//   tet:  g1 = e2 x e3, g2 = e3 x e1, g3 = e1 x e2, g0 = -(g1+g2+g3), det = e1.g1
//         V * (g_a/det).(g_b/det) = (g_a.g_b) / (6|det|)
//   tri:  n_a = (y_{a+1}-y_{a+2}, x_{a+2}-x_{a+1}), A2 = signed 2*area
//         A * (n_a/A2).(n_b/A2) = (n_a.n_b) / (2|A2|)
// Results agree with the reference formulas to rounding (tests: 1e-12 relative, row-scale floor).
//
// Operator concept used by the assembly kernels (include/b2f/assemble.cuh):
//   static constexpr int NV, BS;
//   static constexpr int FAN_WARPS_PER_SM;               // resident warps per SM the fan kernel is compiled for
//   static constexpr bool ROW_SUM_ZERO;                  // rows of every element matrix sum to zero (BS = 1 only)
//   static constexpr bool NEEDS_CELL;                    // prepare() reads per-cell data (the cell id must be fetched)
//   struct Geom;                                         // per-cell prepared geometry, registers only
//   __device__ Geom prepare(const double (&X)[NV][3], long long cell) const;
//   __device__ void block(const Geom&, int a, int b, double (&out)[BS*BS]) const;  // a, b compile-time after unrolling
//   __device__ double cell_value(long long cell) const;  // the per-cell datum (prefetched by the fan kernel), 0 if none
//   tets:      __device__ Geom finish(const TetGeom&, double cell_value, bool odd) const;   // geometry already built (fan kernel)
//   triangles: __device__ Geom prepare_oriented(const double (&X)[3][3], double cell_value, bool odd) const;
//              odd = the vertices are listed in an odd permutation of the cell's own order
// Operators are invariant under even permutations of the cell's vertices, which the node-gather kernel
// uses to bring the row node to local index 0.
#pragma once
#include <cuda_runtime.h>

#ifndef B2F_FAN_WARPS
#define B2F_FAN_WARPS 24
#endif
#ifndef B2F_FAN_WARPS_ELTET
#define B2F_FAN_WARPS_ELTET 4
#endif

namespace b2f {

// 1/x to full double precision for normal x > 0: hardware seed (MUFU.RCP64H, from the upper 32 bits of x: relative error
// about 2^-20) and ONE cubic correction y (1 + e + e^2), e = 1 - x y: the error becomes e^3, below half an ulp, with a chain
// of three DFMAs where two Newton steps (B2F_RCP_CUBIC == 0) need four.  No range checks and no slow path like the
// compiler's division has (cells with a measure outside the normal range are degenerate anyway).
// tools/experiments/rcp_probe.cu measures both forms against 1.0 / x.
// Heat rows: diagonal from the row itself (off-diagonal sum + mass sum) instead of computing it per step.  Parity-exact to
// 1e-12 but measured slower on B200 (heat3d 1.58 vs 1.51 ms, heat2d 1.50 vs 1.39 ms: the mass term per step costs more
// than the diagonal it removes), so it is off; -DB2F_HEAT_ROW_SUM=1 builds it for A/B runs.
// HeatTet: with the cell ids of the steps staged in shared memory (B2F_FAN_STAGE_CELLS) 4 CTAs of 128 rows fit an SM, so the
// kernel is compiled for 16 warps (up to 128 registers: 107 used, no spill)
#ifndef B2F_FAN_WARPS_HEATTET
#define B2F_FAN_WARPS_HEATTET 16
#endif
#ifndef B2F_HEAT_ROW_SUM
#define B2F_HEAT_ROW_SUM 0
#endif
#ifndef B2F_RCP_CUBIC
#define B2F_RCP_CUBIC 1
#endif
__device__ __forceinline__ double fast_rcp(double x)
{
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double e = fma(-x, y, 1.0);
#if B2F_RCP_CUBIC
  return fma(y, fma(e, e, e), y);
#else
  y = fma(y, e, y);
  e = fma(-x, y, 1.0);
  return fma(y, e, y);
#endif
}

struct TetGeom
{
  double g[4][3]; // un-normalised gradients (cofactors)
  double s; // 1 / (6 |det|)
  double vol6; // |det| = 6 * volume
};

__device__ __forceinline__ TetGeom tet_geometry(const double (&X)[4][3])
{
  TetGeom G;
  const double e1x = X[1][0] - X[0][0], e1y = X[1][1] - X[0][1], e1z = X[1][2] - X[0][2];
  const double e2x = X[2][0] - X[0][0], e2y = X[2][1] - X[0][1], e2z = X[2][2] - X[0][2];
  const double e3x = X[3][0] - X[0][0], e3y = X[3][1] - X[0][1], e3z = X[3][2] - X[0][2];
  G.g[1][0] = e2y * e3z - e2z * e3y;
  G.g[1][1] = e2z * e3x - e2x * e3z;
  G.g[1][2] = e2x * e3y - e2y * e3x;
  G.g[2][0] = e3y * e1z - e3z * e1y;
  G.g[2][1] = e3z * e1x - e3x * e1z;
  G.g[2][2] = e3x * e1y - e3y * e1x;
  G.g[3][0] = e1y * e2z - e1z * e2y;
  G.g[3][1] = e1z * e2x - e1x * e2z;
  G.g[3][2] = e1x * e2y - e1y * e2x;
  G.g[0][0] = -(G.g[1][0] + G.g[2][0] + G.g[3][0]);
  G.g[0][1] = -(G.g[1][1] + G.g[2][1] + G.g[3][1]);
  G.g[0][2] = -(G.g[1][2] + G.g[2][2] + G.g[3][2]);
  const double det = e1x * G.g[1][0] + e1y * G.g[1][1] + e1z * G.g[1][2];
  G.vol6 = fabs(det);
  G.s = fast_rcp(6.0 * G.vol6);
  return G;
}

struct TriGeom
{
  double n[3][2]; // un-normalised gradients: n_a = (y_{a+1}-y_{a+2}, x_{a+2}-x_{a+1})
  double a2; // signed 2*area
};

__device__ __forceinline__ TriGeom tri_geometry(const double (&X)[3][3])
{
  TriGeom G;
  G.n[0][0] = X[1][1] - X[2][1];
  G.n[0][1] = X[2][0] - X[1][0];
  G.n[1][0] = X[2][1] - X[0][1];
  G.n[1][1] = X[0][0] - X[2][0];
  G.n[2][0] = X[0][1] - X[1][1];
  G.n[2][1] = X[1][0] - X[0][0];
  G.a2 = (X[1][0] - X[0][0]) * (X[2][1] - X[0][1]) - (X[2][0] - X[0][0]) * (X[1][1] - X[0][1]);
  return G;
}

// ---------------------------------------------------------------------------------------------------
struct LaplaceTet
{
  static constexpr int NV = 4, BS = 1;
  static constexpr int FAN_WARPS_PER_SM = B2F_FAN_WARPS; // occupancy the fan kernel is compiled for (register budget)
  static constexpr bool ROW_SUM_ZERO = true; // element matrices have zero row sums (constants in the kernel)
  static constexpr bool NEEDS_CELL = false;
  static constexpr bool LINEAR_POST = false;
  using Geom = TetGeom;
  __device__ __forceinline__ Geom prepare(const double (&X)[4][3], long long) const { return tet_geometry(X); }
  __device__ __forceinline__ double cell_value(long long) const { return 0.0; }
  __device__ __forceinline__ Geom finish(const TetGeom& T, double, bool) const { return T; }
  __device__ __forceinline__ void block(const Geom& G, int a, int b, double (&out)[1]) const
  {
    out[0] = (G.g[a][0] * G.g[b][0] + G.g[a][1] * G.g[b][1] + G.g[a][2] * G.g[b][2]) * G.s;
  }
};

struct LaplaceTri
{
  static constexpr int NV = 3, BS = 1;
  static constexpr int FAN_WARPS_PER_SM = B2F_FAN_WARPS; // occupancy the fan kernel is compiled for (register budget)
  static constexpr bool ROW_SUM_ZERO = true; // element matrices have zero row sums (constants in the kernel)
  static constexpr bool NEEDS_CELL = false;
  static constexpr bool LINEAR_POST = false;
  struct Geom
  {
    TriGeom t;
    double s; // 1 / (2 |A2|)
  };
  __device__ __forceinline__ Geom prepare(const double (&X)[3][3], long long) const
  {
    Geom G;
    G.t = tri_geometry(X);
    G.s = 1.0 / (2.0 * fabs(G.t.a2));
    return G;
  }
  __device__ __forceinline__ double cell_value(long long) const { return 0.0; }
  __device__ __forceinline__ Geom prepare_oriented(const double (&X)[3][3], double, bool) const { return prepare(X, 0); }
  __device__ __forceinline__ void block(const Geom& G, int a, int b, double (&out)[1]) const
  {
    out[0] = (G.t.n[a][0] * G.t.n[b][0] + G.t.n[a][1] * G.t.n[b][1]) * G.s;
  }
};

// modules/elasticity/FemModule.h:112-152: Ke = [lambda*(...) + mu2*(...)] / (4*area), area = |A2|/2.
// Block (a,b), with p = dPhi (un-normalised gradient n):
//   [ l*px_a*px_b + m2*(px_a*px_b + .5*py_a*py_b)   l*px_a*py_b + m2*.5*py_a*px_b          ]
//   [ l*py_a*px_b + m2*.5*px_a*py_b                  l*py_a*py_b + m2*(py_a*py_b + .5*px_a*px_b) ]
struct ElasticityTri
{
  static constexpr int NV = 3, BS = 2;
  static constexpr int FAN_WARPS_PER_SM = 16; // occupancy the fan kernel is compiled for (register budget)
  static constexpr bool ROW_SUM_ZERO = false; // element matrices have zero row sums (constants in the kernel)
  static constexpr bool NEEDS_CELL = false;
  static constexpr bool LINEAR_POST = true; // K = lambda*A + mu*A^T + mu*tr(A)*I with A = sum of s * n_a (x) n_b: see raw_add / post
  double lambda, mu2;
  struct Geom
  {
    TriGeom t;
    double s; // 1 / (4*area) = 1 / (2 |A2|)
  };
  __device__ __forceinline__ Geom prepare(const double (&X)[3][3], long long) const
  {
    Geom G;
    G.t = tri_geometry(X);
    G.s = 1.0 / (2.0 * fabs(G.t.a2));
    return G;
  }
  __device__ __forceinline__ double cell_value(long long) const { return 0.0; }
  __device__ __forceinline__ Geom prepare_oriented(const double (&X)[3][3], double, bool) const { return prepare(X, 0); }
  // The node-gather kernel accumulates the plain outer products A_(0,b) += (s*n_0) (x) n_b along its walk (4 FMAs per
  // block instead of the full entries) and applies the material law once per stored block at the end: the law is linear
  // in A.  mu = mu2 / 2: entry (0,0) = l*xx + mu2*xx + .5*mu2*yy = l*A00 + mu*A00 + mu*(A00 + A11).
  __device__ __forceinline__ void raw_begin(const Geom& G, double (&u)[2]) const
  {
    u[0] = G.s * G.t.n[0][0];
    u[1] = G.s * G.t.n[0][1];
  }
  __device__ __forceinline__ void raw_add(const Geom& G, const double (&u)[2], int b, double (&acc)[4]) const
  {
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j)
        acc[i * 2 + j] = fma(u[i], G.t.n[b][j], acc[i * 2 + j]);
  }
  __device__ __forceinline__ void post(double (&A)[4]) const
  {
    const double mu = 0.5 * mu2, tr = mu * (A[0] + A[3]);
    const double k00 = lambda * A[0] + mu * A[0] + tr, k01 = lambda * A[1] + mu * A[2];
    const double k10 = lambda * A[2] + mu * A[1], k11 = lambda * A[3] + mu * A[3] + tr;
    A[0] = k00;
    A[1] = k01;
    A[2] = k10;
    A[3] = k11;
  }
  __device__ __forceinline__ void block(const Geom& G, int a, int b, double (&out)[4]) const
  {
    const double pxa = G.t.n[a][0], pya = G.t.n[a][1], pxb = G.t.n[b][0], pyb = G.t.n[b][1];
    const double xx = pxa * pxb, yy = pya * pyb, xy = pxa * pyb, yx = pya * pxb;
    const double hm = 0.5 * mu2;
    out[0] = (lambda * xx + mu2 * xx + hm * yy) * G.s;
    out[1] = (lambda * xy + hm * yx) * G.s;
    out[2] = (lambda * yx + hm * xy) * G.s;
    out[3] = (lambda * yy + mu2 * yy + hm * xx) * G.s;
  }
};

// Tetrahedral extension: Ke[(a,i),(b,j)] = V*(lambda*G_a^i*G_b^j + mu*G_a^j*G_b^i + mu*delta_ij*G_a.G_b), G = g/det.
struct ElasticityTet
{
  static constexpr int NV = 4, BS = 3;
  static constexpr int FAN_WARPS_PER_SM = B2F_FAN_WARPS_ELTET; // occupancy the fan kernel is compiled for (register budget)
  static constexpr bool ROW_SUM_ZERO = false; // element matrices have zero row sums (constants in the kernel)
  static constexpr bool NEEDS_CELL = false;
  static constexpr bool LINEAR_POST = true; // K = lambda*A + mu*A^T + mu*tr(A)*I with A = sum of s * g_a (x) g_b: see raw_add / post
  double lambda, mu2;
  using Geom = TetGeom;
  __device__ __forceinline__ Geom prepare(const double (&X)[4][3], long long) const { return tet_geometry(X); }
  __device__ __forceinline__ double cell_value(long long) const { return 0.0; }
  __device__ __forceinline__ Geom finish(const TetGeom& T, double, bool) const { return T; }
  // node-gather kernel: A_(0,b) += (s*g_0) (x) g_b along the walk (9 FMAs per block), material law once per stored block
  __device__ __forceinline__ void raw_begin(const Geom& G, double (&u)[3]) const
  {
#pragma unroll
    for (int i = 0; i < 3; ++i)
      u[i] = G.s * G.g[0][i];
  }
  __device__ __forceinline__ void raw_add(const Geom& G, const double (&u)[3], int b, double (&acc)[9]) const
  {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j)
        acc[i * 3 + j] = fma(u[i], G.g[b][j], acc[i * 3 + j]);
  }
  __device__ __forceinline__ void post(double (&A)[9]) const
  {
    const double mu = 0.5 * mu2, tr = mu * (A[0] + A[4] + A[8]);
    double K[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        double v = lambda * A[i * 3 + j] + mu * A[j * 3 + i];
        if (i == j)
          v += tr;
        K[i * 3 + j] = v;
      }
#pragma unroll
    for (int q = 0; q < 9; ++q)
      A[q] = K[q];
  }
  __device__ __forceinline__ void block(const Geom& G, int a, int b, double (&out)[9]) const
  {
    const double mu = 0.5 * mu2;
    const double ls = lambda * G.s, ms = mu * G.s;
    const double gab = (G.g[a][0] * G.g[b][0] + G.g[a][1] * G.g[b][1] + G.g[a][2] * G.g[b][2]) * ms;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        double v = ls * (G.g[a][i] * G.g[b][j]) + ms * (G.g[a][j] * G.g[b][i]);
        if (i == j)
          v += gab;
        out[i * 3 + j] = v;
      }
  }
};

// modules/heat/FemModule.cc:678-770: lambda_cell * (dPhi_a.dPhi_b) * 0.25/area + (area/12)*(1+delta_ab)/dt, signed area (:606-612).
// CELL: the conductivity is a per-cell array (the reference's m_cell_lambda); false: one value for the mesh - that instantiation
// needs no cell id, so the walk reads neither the plan's cell stream nor the array.
template <bool CELL>
struct HeatTriT
{
  static constexpr int NV = 3, BS = 1;
  static constexpr int FAN_WARPS_PER_SM = 24; // occupancy the fan kernel is compiled for (register budget)
  static constexpr bool ROW_SUM_ZERO = false; // element matrices have zero row sums (constants in the kernel)
  static constexpr bool NEEDS_CELL = CELL; // per-cell conductivity
  static constexpr bool LINEAR_POST = false;
  static constexpr bool ROW_SUM_MASS = B2F_HEAT_ROW_SUM != 0; // row 0 = zero-row-sum stiffness + m (off the diagonal) / + 2 m (on it): see assemble.cuh
  double lambda, inv_dt;
  const double* cell_lambda;
  struct Geom
  {
    TriGeom t;
    double ks; // lambda * 0.25 / area
    double m; // area / 12 / dt
  };
  __device__ __forceinline__ double mass(const Geom& G) const { return G.m; }
  // unconditional load in the per-cell form: a select on the pointer would wait for the load where it is issued and undo the prefetch
  __device__ __forceinline__ double cell_value(long long cell) const
  {
    if constexpr (CELL)
      return __ldg(cell_lambda + cell);
    else
      return lambda;
  }
  __device__ __forceinline__ Geom prepare(const double (&X)[3][3], long long cell) const { return prepare_oriented(X, cell_value(cell), false); }
  // odd: X lists the cell's vertices in an odd permutation of the cell's own order (the signed area flips back)
  __device__ __forceinline__ Geom prepare_oriented(const double (&X)[3][3], double lam, bool odd) const
  {
    Geom G;
    G.t = tri_geometry(X);
    const double area = odd ? -0.5 * G.t.a2 : 0.5 * G.t.a2;
    G.ks = (CELL ? lam : lambda) * 0.25 / area;
    G.m = area * (1.0 / 12.0) * inv_dt;
    return G;
  }
  __device__ __forceinline__ void block(const Geom& G, int a, int b, double (&out)[1]) const
  {
    const double k = (G.t.n[a][0] * G.t.n[b][0] + G.t.n[a][1] * G.t.n[b][1]) * G.ks;
    out[0] = k + (a == b ? 2.0 * G.m : G.m);
  }
};

// 3D twin of the heat operator: lambda*V*(grad.grad) + (V/20)*(1+delta_ab)/dt.
template <bool CELL>
struct HeatTetT
{
  static constexpr int NV = 4, BS = 1;
  static constexpr int FAN_WARPS_PER_SM = CELL ? B2F_FAN_WARPS_HEATTET : 20; // occupancy the fan kernel is compiled for (register budget)
  static constexpr bool ROW_SUM_ZERO = false; // element matrices have zero row sums (constants in the kernel)
  static constexpr bool NEEDS_CELL = CELL;
  static constexpr bool LINEAR_POST = false;
  static constexpr bool ROW_SUM_MASS = B2F_HEAT_ROW_SUM != 0;
  double lambda, inv_dt;
  const double* cell_lambda;
  struct Geom
  {
    TetGeom t;
    double ks, m;
  };
  __device__ __forceinline__ double mass(const Geom& G) const { return G.m; }
  // unconditional load in the per-cell form: a select on the pointer would wait for the load where it is issued and undo the prefetch
  __device__ __forceinline__ double cell_value(long long cell) const
  {
    if constexpr (CELL)
      return __ldg(cell_lambda + cell);
    else
      return lambda;
  }
  __device__ __forceinline__ Geom prepare(const double (&X)[4][3], long long cell) const { return finish(tet_geometry(X), cell_value(cell), false); }
  __device__ __forceinline__ Geom finish(const TetGeom& T, double lam, bool) const
  {
    Geom G;
    G.t = T;
    G.ks = (CELL ? lam : lambda) * G.t.s;
    G.m = G.t.vol6 * (1.0 / 120.0) * inv_dt;
    return G;
  }
  __device__ __forceinline__ void block(const Geom& G, int a, int b, double (&out)[1]) const
  {
    const double k = (G.t.g[a][0] * G.t.g[b][0] + G.t.g[a][1] * G.t.g[b][1] + G.t.g[a][2] * G.t.g[b][2]) * G.ks;
    out[0] = k + (a == b ? 2.0 * G.m : G.m);
  }
};

using HeatTri = HeatTriT<true>;
using HeatTet = HeatTetT<true>;
using HeatTriUniform = HeatTriT<false>;
using HeatTetUniform = HeatTetT<false>;

} // namespace b2f
