// arcanefem_b200/FemFunctionsGpu.cuh — the element-kernel vocabulary of the reference for callbacks written by the caller:
//   femutils/FemUtils.h:34-40      Real4
//   femutils/FemUtils.h:53-186     FixedMatrix<N, M> (row-major, operator()(i, j), +, -, scalar * and /)
//   femutils/FemUtils.h:191-226    operator^ : outer product of two Real3 / Real4
//   femutils/ArcaneFemFunctionsGpu.h:75-83,110-122     FemUtils::Gpu::MeshOperation::computeAreaTria3 / computeVolumeTetra4
//   femutils/ArcaneFemFunctionsGpu.h:218-252           FemUtils::Gpu::FeOperation2D::computeGradient{X,Y}Tria3
//   femutils/ArcaneFemFunctionsGpu.h:280-392           FemUtils::Gpu::FeOperation3D::computeGradient{X,Y,Z}Tetra4
// so that an element callback of a module (e.g. modules/poisson/FemModule.h:139-147,177-186) keeps its body.  One change of
// signature: the reference's helpers take (cell_lid, cn_cv, in_node_coord) and gather the cell's vertices themselves - each
// of the four helpers of a Tetra4 callback gathers all four again; here the library gathers once and the helpers take the
// vertex array `const Real (&X)[NV][3]` the callback receives.  Device code (nvcc); the functions are also callable on the host.
//
// Same results as the reference's formulas up to rounding: gradients are cofactors of the edge vectors divided by
// A2 (signed, triangles) or V6 = |det| (tetrahedra, like the reference: the sign drops out of every quadratic form).
#pragma once
#include <cmath>
#include <initializer_list>
#include "../b2f/user_operator.cuh"

#define B2F_HD __host__ __device__ __forceinline__

namespace ArcaneFemB200 {

using Real = double;

struct Real3
{
  Real x, y, z;
  B2F_HD Real operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

struct Real4 // femutils/FemUtils.h:34-40
{
  Real data[4];
  B2F_HD Real& operator[](int i) { return data[i]; }
  B2F_HD const Real& operator[](int i) const { return data[i]; }
};

template <int N, int M>
struct FixedMatrix // femutils/FemUtils.h:53-186
{
  Real m_values[N * M];
  static constexpr int totalNbElement() { return N * M; }
  static constexpr int rows = N, cols = M;
  B2F_HD FixedMatrix()
  {
    for (int q = 0; q < N * M; ++q)
      m_values[q] = 0.0;
  }
  // femutils/FemUtils.h:62: row-major list; entries the list does not reach stay zero
  B2F_HD FixedMatrix(std::initializer_list<Real> init_list)
  {
    int q = 0;
    for (auto it = init_list.begin(); it != init_list.end() && q < N * M; ++it)
      m_values[q++] = *it;
    for (; q < N * M; ++q)
      m_values[q] = 0.0;
  }
  B2F_HD Real& operator()(int i, int j) { return m_values[i * M + j]; }
  B2F_HD Real operator()(int i, int j) const { return m_values[i * M + j]; }
  B2F_HD void multInPlace(Real f)
  {
    for (int q = 0; q < N * M; ++q)
      m_values[q] *= f;
  }
};

template <int N, int M>
B2F_HD FixedMatrix<N, M> operator+(const FixedMatrix<N, M>& a, const FixedMatrix<N, M>& b)
{
  FixedMatrix<N, M> r;
  for (int q = 0; q < N * M; ++q)
    r.m_values[q] = a.m_values[q] + b.m_values[q];
  return r;
}
template <int N, int M>
B2F_HD FixedMatrix<N, M> operator-(const FixedMatrix<N, M>& a, const FixedMatrix<N, M>& b)
{
  FixedMatrix<N, M> r;
  for (int q = 0; q < N * M; ++q)
    r.m_values[q] = a.m_values[q] - b.m_values[q];
  return r;
}
template <int N, int M>
B2F_HD FixedMatrix<N, M> operator*(Real f, const FixedMatrix<N, M>& a)
{
  FixedMatrix<N, M> r;
  for (int q = 0; q < N * M; ++q)
    r.m_values[q] = f * a.m_values[q];
  return r;
}
template <int N, int M>
B2F_HD FixedMatrix<N, M> operator*(const FixedMatrix<N, M>& a, Real f)
{
  return f * a;
}
template <int N, int M>
B2F_HD FixedMatrix<N, M> operator/(const FixedMatrix<N, M>& a, Real f)
{
  FixedMatrix<N, M> r;
  for (int q = 0; q < N * M; ++q)
    r.m_values[q] = a.m_values[q] / f;
  return r;
}
template <int N, int M>
B2F_HD FixedMatrix<N, M> operator-(const FixedMatrix<N, M>& a) // femutils/FemUtils.h:126-135
{
  FixedMatrix<N, M> r;
  for (int q = 0; q < N * M; ++q)
    r.m_values[q] = -a.m_values[q];
  return r;
}
// femutils/FemUtils.h:217-226: column ^ row = outer product
template <int N>
B2F_HD FixedMatrix<N, N> operator^(const FixedMatrix<N, 1>& lhs, const FixedMatrix<1, N>& rhs)
{
  FixedMatrix<N, N> r;
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j)
      r(i, j) = lhs(i, 0) * rhs(0, j);
  return r;
}
// femutils/FemUtils.h:266-276, :281-297, :302-313
template <int N, int M>
B2F_HD FixedMatrix<N, N> matrixAddition(const FixedMatrix<N, M>& a, const FixedMatrix<M, N>& b)
{
  FixedMatrix<N, N> r;
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j)
      r(i, j) = a(i, j) + b(i, j);
  return r;
}
template <int N, int M>
B2F_HD FixedMatrix<N, N> matrixMultiplication(const FixedMatrix<N, M>& a, const FixedMatrix<M, N>& b)
{
  FixedMatrix<N, N> r;
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) {
      Real x = 0.0;
      for (int k = 0; k < M; ++k)
        x += a(i, k) * b(k, j);
      r(i, j) += x;
    }
  return r;
}
template <int N, int M>
B2F_HD FixedMatrix<M, N> matrixTranspose(const FixedMatrix<N, M>& a)
{
  FixedMatrix<M, N> t;
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < M; ++j)
      t(j, i) = a(i, j);
  return t;
}
// femutils/FemUtils.h:191-226: u ^ v = u v^T
B2F_HD FixedMatrix<3, 3> operator^(const Real3& u, const Real3& v)
{
  FixedMatrix<3, 3> r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      r(i, j) = u[i] * v[j];
  return r;
}
B2F_HD FixedMatrix<4, 4> operator^(const Real4& u, const Real4& v)
{
  FixedMatrix<4, 4> r;
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      r(i, j) = u[i] * v[j];
  return r;
}

namespace Gpu {

namespace detail {
  // cofactors of the edge vectors e_k = X[k] - X[0]: c_1 = e_2 x e_3, c_2 = e_3 x e_1, c_3 = e_1 x e_2, c_0 = -(c_1 + c_2 + c_3),
  // det = e_1 . c_1; grad N_a = c_a / det.  Returns |det| (= 6 V).
  B2F_HD Real tetra_cofactors(const Real (&X)[4][3], Real (&c)[4][3])
  {
    Real e[3][3];
    for (int k = 0; k < 3; ++k)
      for (int d = 0; d < 3; ++d)
        e[k][d] = X[k + 1][d] - X[0][d];
    for (int k = 0; k < 3; ++k) {
      const Real* u = e[(k + 1) % 3];
      const Real* v = e[(k + 2) % 3];
      c[k + 1][0] = u[1] * v[2] - u[2] * v[1];
      c[k + 1][1] = u[2] * v[0] - u[0] * v[2];
      c[k + 1][2] = u[0] * v[1] - u[1] * v[0];
    }
    const Real det = e[0][0] * c[1][0] + e[0][1] * c[1][1] + e[0][2] * c[1][2];
    for (int d = 0; d < 3; ++d)
      c[0][d] = -(c[1][d] + c[2][d] + c[3][d]);
    return fabs(det); // the reference divides by V6 = |det|: its gradients carry the cell's orientation sign
  }
  B2F_HD Real4 tetra_gradient(const Real (&X)[4][3], int d)
  {
    Real c[4][3];
    const Real v6 = tetra_cofactors(X, c);
    Real4 g;
    for (int a = 0; a < 4; ++a)
      g[a] = c[a][d] / v6;
    return g;
  }
} // namespace detail

namespace MeshOperation {
  // ArcaneFemFunctionsGpu.h:75-83: norm of the 3D cross product / 2
  B2F_HD Real computeAreaTria3(const Real (&X)[3][3])
  {
    const Real ax = X[1][0] - X[0][0], ay = X[1][1] - X[0][1], az = X[1][2] - X[0][2];
    const Real bx = X[2][0] - X[0][0], by = X[2][1] - X[0][1], bz = X[2][2] - X[0][2];
    const Real cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
    return sqrt(cx * cx + cy * cy + cz * cz) / 2.0;
  }
  // ArcaneFemFunctionsGpu.h:110-122: |v0 . (v1 x v2)| / 6
  B2F_HD Real computeVolumeTetra4(const Real (&X)[4][3])
  {
    Real c[4][3];
    return detail::tetra_cofactors(X, c) / 6.0;
  }
} // namespace MeshOperation

namespace FeOperation2D {
  // ArcaneFemFunctionsGpu.h:218-227 / :243-252: (y1 - y2, y2 - y0, y0 - y1) / A2 and (x2 - x1, x0 - x2, x1 - x0) / A2, A2 signed
  B2F_HD Real signedArea2(const Real (&X)[3][3]) { return (X[1][0] - X[0][0]) * (X[2][1] - X[0][1]) - (X[2][0] - X[0][0]) * (X[1][1] - X[0][1]); }
  B2F_HD Real3 computeGradientXTria3(const Real (&X)[3][3])
  {
    const Real a2 = signedArea2(X);
    return Real3{ (X[1][1] - X[2][1]) / a2, (X[2][1] - X[0][1]) / a2, (X[0][1] - X[1][1]) / a2 };
  }
  B2F_HD Real3 computeGradientYTria3(const Real (&X)[3][3])
  {
    const Real a2 = signedArea2(X);
    return Real3{ (X[2][0] - X[1][0]) / a2, (X[0][0] - X[2][0]) / a2, (X[1][0] - X[0][0]) / a2 };
  }
} // namespace FeOperation2D

namespace FeOperation3D {
  // ArcaneFemFunctionsGpu.h:280-392: d N_a / dx, dy, dz as cofactors / V6
  B2F_HD Real4 computeGradientXTetra4(const Real (&X)[4][3]) { return detail::tetra_gradient(X, 0); }
  B2F_HD Real4 computeGradientYTetra4(const Real (&X)[4][3]) { return detail::tetra_gradient(X, 1); }
  B2F_HD Real4 computeGradientZTetra4(const Real (&X)[4][3]) { return detail::tetra_gradient(X, 2); }
} // namespace FeOperation3D

} // namespace Gpu

// Callback in the reference's return convention: FixedMatrix<BS*NV, BS*NV> f(cell, X)  ->  the array form of user_operator.cuh.
template <int NV, int BS, class F>
struct FixedMatrixCallback
{
  F f;
  __device__ __forceinline__ void operator()(long long cell, const Real (&X)[NV][3], Real (&Ke)[NV * BS][NV * BS]) const
  {
    const FixedMatrix<NV * BS, NV * BS> m = f(cell, X);
    for (int i = 0; i < NV * BS; ++i)
      for (int j = 0; j < NV * BS; ++j)
        Ke[i][j] = m(i, j);
  }
};

// b2f::assemble_with_element_matrix for a callback that returns a FixedMatrix (b2f_status).
template <int NV, int BS, class F>
inline int assemble_with_fixed_matrix(b2f_handle h, const F& f, bool accumulate, cudaStream_t stream)
{
  return b2f::assemble_with_element_matrix<NV, BS>(h, FixedMatrixCallback<NV, BS, F>{ f }, accumulate, stream);
}

} // namespace ArcaneFemB200
