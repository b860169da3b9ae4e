// user_operator.cuh — the reference's `template <class Function> assembleBilinear(Function compute_element_matrix)`
// (femutils/BSRFormat.h:786-787,1105-1111) for element callbacks the library does not ship.
//
// A C ABI cannot carry a device lambda, so the built-in operators travel as an enum (b2f_operator).  A module with its own
// element matrix includes this header in a .cu file compiled with
//     nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -I<repo>/include ... -lb200fem
// and calls
//     b2f::assemble_with(handle, op, accumulate, stream)                 // op: any type with the operator concept of element_ops.cuh
//     b2f::assemble_with_element_matrix<NV, BS>(handle, f, accumulate, stream)   // f: the reference's callback shape
// Both launch the same kernels as b2f_assemble (variant and value layout of the handle) with the handle's pattern and
// plan; only the element arithmetic is the caller's.
//
// Callback shape of assemble_with_element_matrix (mirrors `FixedMatrix<BS*NV, BS*NV> f(CellLocalId)`, whose entry
// (BS*a+i, BS*b+j) couples dof i of local node a with dof j of local node b, BSRFormat.h:823,881):
//     __device__ void operator()(long long cell, const double (&X)[NV][3], double (&Ke)[BS*NV][BS*NV]) const;
// X holds the cell's vertex coordinates (VariableNodeReal3 rows), already gathered.  The node-gather kernels hand the
// vertices over in an even permutation of the cell's own order (the row node first) and read row 0 of Ke, so the callback
// must be a function of the vertices it is given (every geometric element matrix is); per-cell data is reached
// through `cell`.  The fan walk needs the operator in its incremental form (finish / prepare_oriented), which a plain
// callback cannot provide: such operators run on the slot-word gather (or the cell-wise atomic kernel, as the handle says).
#pragma once
#include "../b2f.h"
#include "assemble.cuh"

namespace b2f {

template <int NV_, int BS_, class F>
struct ElementMatrixOp
{
  static constexpr int NV = NV_, BS = BS_;
  static constexpr int FAN_WARPS_PER_SM = 8; // not used: NO_FAN
  static constexpr bool ROW_SUM_ZERO = false, NEEDS_CELL = true, LINEAR_POST = false, NO_FAN = true;
  struct Geom
  {
    double ke[NV_ * BS_][NV_ * BS_];
  };
  F f;
  __device__ __forceinline__ Geom prepare(const double (&X)[NV_][3], long long cell) const
  {
    Geom G;
    f(cell, X, G.ke);
    return G;
  }
  __device__ __forceinline__ double cell_value(long long) const { return 0.0; }
  __device__ __forceinline__ void block(const Geom& G, int a, int b, double (&out)[BS_ * BS_]) const
  {
#pragma unroll
    for (int i = 0; i < BS_; ++i)
#pragma unroll
      for (int j = 0; j < BS_; ++j)
        out[i * BS_ + j] = G.ke[BS_ * a + i][BS_ * b + j];
  }
};

// Returns a b2f_status; b2f_last_error_string() describes a failure of the C-ABI part, cudaGetErrorString the launch.
template <class Op>
inline int assemble_with(b2f_handle h, const Op& op, bool accumulate, cudaStream_t stream)
{
  AssembleArgs A;
  int32_t variant = 0, per_block = 0;
  int64_t nb_values = 0;
  const int st = b2f_user_assemble_begin(h, Op::NV, Op::BS, &A, sizeof(A), &variant, &per_block, &nb_values, stream);
  if (st != B2F_OK)
    return st;
  A.accumulate = accumulate ? 1 : 0;
  return launch_assemble<Op>(A, op, variant, per_block != 0, nb_values, stream) == cudaSuccess ? B2F_OK : B2F_ERR_CUDA;
}

template <int NV, int BS, class F>
inline int assemble_with_element_matrix(b2f_handle h, const F& f, bool accumulate, cudaStream_t stream)
{
  return assemble_with(h, ElementMatrixOp<NV, BS, F>{ f }, accumulate, stream);
}

} // namespace b2f
