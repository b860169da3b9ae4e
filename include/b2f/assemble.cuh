// assemble.cuh — scatter-add of P1 element matrices into BSR values (sm_100a).
//
// Template entry point behind b2f_assemble(); also the escape hatch that keeps the reference's
// `template <class Function> assembleBilinear(Function)` (femutils/BSRFormat.h:1105-1111) usable with a
// user-written operator: any type satisfying the concept in element_ops.cuh can be passed to
// b2f::launch_assemble<Op>() from a .cu file of the caller.
//
// Two kernels, mirroring the reference's two variants:
//   k_cell_atomic   cell-wise (BSRFormat.h:786-898): one thread per cell, Ke blocks in registers, column slot by
//                   binary search in the sorted row, fp64 RED per scalar; optional warp pre-reduction of
//                   same-address adds (__match_any_sync) so a warp issues one RED per distinct block.
//   k_node_gather   node-wise, atomic-free (BSRFormat.h:937-1068): a CTA owns ROWS consecutive block rows, whose
//                   values are one contiguous slab.  The slab's accumulators, its column indices and the rows'
//                   node->cell lists are staged in shared memory; one thread per row walks its cells, permutes each
//                   cell so the row node is local vertex 0 (even permutation), computes only row 0 of Ke and adds
//                   it into its private accumulators; the slab is then stored with coalesced vector stores — every
//                   value is written exactly once and no zero-fill pass is needed.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <type_traits>
#include <utility>
#include <cuda_runtime.h>
#include "assemble_args.h"
#include "element_ops.cuh"

namespace b2f {


// Scalar index of entry (i,j) of the x-th block of a block row that starts at block rb and holds k blocks
// (femutils/BSRFormat.h:145-171).
template <int BS, bool PER_BLOCK>
__device__ __forceinline__ long long value_index(long long rb, int k, int x, int i, int j)
{
  if (PER_BLOCK)
    return (rb + x) * (BS * BS) + (i * BS + j);
  return rb * (BS * BS) + (long long)i * (BS * k) + x * BS + j;
}

// Slot of the off-diagonal column c in a row of k entries stored "diagonal first, then ascending" (dl = 0) or "ascending,
// diagonal last" (dl = 1, b2f_pattern_layout).  Returns -1 if absent.
template <class ColPtr>
__device__ __forceinline__ int find_slot(ColPtr cols, int k, int32_t c, int dl = 0)
{
  int lo = 1 - dl, hi = k - dl;
  const int end = hi;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (cols[mid] < c)
      lo = mid + 1;
    else
      hi = mid;
  }
  return (lo < end && cols[lo] == c) ? lo : -1;
}
__device__ __forceinline__ int diag_slot(int k, int dl) { return dl ? k - 1 : 0; }

template <int NV>
__device__ __forceinline__ void load_cell_nodes(const int32_t* __restrict__ cell_node, long long cell, int32_t (&n)[NV])
{
  if (NV == 4) {
    const int4 v = __ldg(reinterpret_cast<const int4*>(cell_node) + cell);
    n[0] = v.x;
    n[1] = v.y;
    n[2] = v.z;
    n[NV - 1] = v.w;
  }
  else {
#pragma unroll
    for (int k = 0; k < NV; ++k)
      n[k] = __ldg(cell_node + cell * NV + k);
  }
}

#ifndef B2F_CHECKED
#define B2F_CHECKED 0
#endif
__device__ __forceinline__ void load_xyz(const double* __restrict__ coords, int32_t node, double (&x)[3])
{
  const double* p = coords + 3ll * node;
  x[0] = __ldg(p);
  x[1] = __ldg(p + 1);
  x[2] = __ldg(p + 2);
}

// Coordinates as the fan kernel reads them: the plain array, plus - in the range-checked build (B2F_CHECKED, below) - the node
// count and the violation counter, so that every gathered node id is checked.
struct CoordView
{
  const double* p;
#if B2F_CHECKED
  long long nb_node;
  unsigned long long* bad;
#endif
};
__device__ __forceinline__ void load_xyz(const CoordView& c, int32_t node, double (&x)[3])
{
#if B2F_CHECKED
  if (node < 0 || node >= c.nb_node) {
    atomicAdd(c.bad, 1ull);
    x[0] = x[1] = x[2] = 0.0;
    return;
  }
#endif
  load_xyz(c.p, node, x);
}

// Even permutation bringing local vertex a to position 0: tets m[k] = n[a ^ k] (double transpositions),
// triangles m[k] = n[(a + k) % 3] (rotations).  Orientation, hence every signed quantity, is preserved.
template <int NV>
__device__ __forceinline__ void bring_to_front(int32_t (&n)[NV], int a)
{
  if (NV == 4) {
    if (a & 1) {
      int32_t t = n[0];
      n[0] = n[1];
      n[1] = t;
      t = n[2];
      n[2] = n[NV - 1];
      n[NV - 1] = t;
    }
    if (a & 2) {
      int32_t t = n[0];
      n[0] = n[2];
      n[2] = t;
      t = n[1];
      n[1] = n[NV - 1];
      n[NV - 1] = t;
    }
  }
  else {
    const int32_t n0 = n[0], n1 = n[1], n2 = n[2];
    n[0] = (a == 0) ? n0 : (a == 1 ? n1 : n2);
    n[1] = (a == 0) ? n1 : (a == 1 ? n2 : n0);
    n[2] = (a == 0) ? n2 : (a == 1 ? n0 : n1);
  }
}

// ---------------------------------------------------------------------------------------------------
// Cell-wise atomic variant
// ---------------------------------------------------------------------------------------------------
template <class Op, bool PER_BLOCK, bool WARPAGG>
__global__ void __launch_bounds__(256) k_cell_atomic(AssembleArgs A, Op op)
{
  constexpr int NV = Op::NV, BS = Op::BS, BS2 = BS * BS;
  const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = cell < A.nb_cell && (A.cell_is_own == nullptr || A.cell_is_own[cell] != 0);
  if (!WARPAGG && !valid)
    return;
  const unsigned lane = threadIdx.x & 31u;

  int32_t n[NV];
  double X[NV][3];
  typename Op::Geom G;
  if (valid) {
    load_cell_nodes<NV>(A.cell_node, cell, n);
#pragma unroll
    for (int k = 0; k < NV; ++k)
      load_xyz(A.coords, n[k], X[k]);
    G = op.prepare(X, cell);
  }

#pragma unroll
  for (int a = 0; a < NV; ++a) {
    bool own = valid;
    int32_t rb = 0;
    int k = 0;
    if (valid) {
      const int32_t row = n[a];
      own = (A.node_is_own == nullptr) || (A.node_is_own[row] != 0);
      rb = __ldg(A.row_index + row);
      k = __ldg(A.row_index + row + 1) - rb;
    }
#pragma unroll
    for (int b = 0; b < NV; ++b) {
      int x = -1;
      if (own) {
        x = (a == b) ? diag_slot(k, A.diag_last) : find_slot(A.columns + rb, k, n[b], A.diag_last);
        if (x < 0)
          atomicAdd(A.counters, 1ull);
      }
      double out[BS2];
      if (own && x >= 0)
        op.block(G, a, b, out);
      if (!WARPAGG) {
        if (own && x >= 0) {
#pragma unroll
          for (int i = 0; i < BS; ++i)
#pragma unroll
            for (int j = 0; j < BS; ++j)
              atomicAdd(A.values + value_index<BS, PER_BLOCK>(rb, k, x, i, j), out[i * BS + j]);
        }
      }
      else {
        // Pre-reduce adds that target the same block inside the warp: one RED per distinct block.
        const bool act = own && x >= 0;
        const long long key = act ? ((long long)rb + x) : (-1ll - (long long)lane);
        const unsigned grp = __match_any_sync(0xffffffffu, key);
        const int leader = __ffs(grp) - 1;
        const int cnt = __reduce_max_sync(0xffffffffu, __popc(grp));
#pragma unroll
        for (int e = 0; e < BS2; ++e) {
          const double v = act ? out[e] : 0.0;
          double sum = 0.0;
          unsigned m = grp;
          for (int it = 0; it < cnt; ++it) {
            const int src = m ? (__ffs(m) - 1) : (int)lane;
            const double t = __shfl_sync(0xffffffffu, v, src);
            if (m) {
              sum += t;
              m &= m - 1;
            }
          }
          if (act && (int)lane == leader)
            atomicAdd(A.values + value_index<BS, PER_BLOCK>(rb, k, x, e / BS, e % BS), sum);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Node-wise gather variant
// ---------------------------------------------------------------------------------------------------
template <class Op>
struct GatherCfg
{
  static constexpr int BS2 = Op::BS * Op::BS;
  static constexpr int ROWS = (Op::BS == 1) ? 128 : 64; // block rows (= threads) per CTA
  static constexpr int ACC_BLOCKS = ROWS * ((Op::BS == 3) ? 18 : 24); // accumulator capacity in blocks
  static constexpr int ADJ_CAP = ROWS * 40; // staged node->cell entries
  static constexpr size_t SMEM_BYTES = (size_t)ACC_BLOCKS * BS2 * sizeof(double) + (size_t)ACC_BLOCKS * sizeof(int32_t) + (size_t)ADJ_CAP * sizeof(uint32_t);
};

// Sum over the cells incident to `row` of 6*volume (tets) / 2*area (triangles, 3D cross product as computeAreaTria3):
// the constant-source term of the rows that take the global fallback path of the fan kernel.
template <int NV>
__device__ __forceinline__ double row_measure(const AssembleArgs& A, long long row)
{
  double sum = 0.0;
  for (int32_t q = __ldg(A.nc_ptr + row), qe = __ldg(A.nc_ptr + row + 1); q < qe; ++q) {
    const long long cell = __ldg(A.nc_list + q) >> 2;
    int32_t n[NV];
    load_cell_nodes<NV>(A.cell_node, cell, n);
    double X[NV][3];
#pragma unroll
    for (int b = 0; b < NV; ++b)
      load_xyz(A.coords, n[b], X[b]);
    const double ax = X[1][0] - X[0][0], ay = X[1][1] - X[0][1], az = X[1][2] - X[0][2];
    const double bx = X[2][0] - X[0][0], by = X[2][1] - X[0][1], bz = X[2][2] - X[0][2];
    const double cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
    if constexpr (NV == 3)
      sum += sqrt(cx * cx + cy * cy + cz * cz);
    else {
      const double dx = X[NV - 1][0] - X[0][0], dy = X[NV - 1][1] - X[0][1], dz = X[NV - 1][2] - X[0][2];
      sum += fabs(cx * dx + cy * dy + cz * dz);
    }
  }
  return sum;
}

// row_map (SMEM == false only): the rows handled are row_map[row0 .. row0 + nrows) instead of row0 .. row0 + nrows (reorder mode).
template <class Op, bool PER_BLOCK, bool SMEM>
__device__ __forceinline__ void gather_rows(const AssembleArgs& A, const Op& op, long long row0, int nrows, int32_t vbeg, int32_t vend, int32_t abeg, int32_t aend,
                                            double* acc_s, int32_t* cols_s, uint32_t* adj_s, const int32_t* row_map = nullptr)
{
  constexpr int NV = Op::NV, BS = Op::BS, BS2 = BS * BS;
  const int t = threadIdx.x;
  if (SMEM) {
    const int nblk = vend - vbeg;
    for (int q = t; q < nblk; q += blockDim.x)
      cols_s[q] = __ldg(A.columns + vbeg + q);
    for (int q = t; q < aend - abeg; q += blockDim.x)
      adj_s[q] = __ldg(A.nc_list + abeg + q);
    if (A.accumulate) {
      for (int q = t; q < nblk * BS2; q += blockDim.x)
        acc_s[q] = A.values[(long long)vbeg * BS2 + q];
    }
    else {
      for (int q = t; q < nblk * BS2; q += blockDim.x)
        acc_s[q] = 0.0;
    }
    __syncthreads();
  }

  if (t < nrows) {
    const long long row = (!SMEM && row_map != nullptr) ? (long long)__ldg(row_map + row0 + t) : row0 + t;
    const int32_t rb = __ldg(A.row_index + row);
    const int k = __ldg(A.row_index + row + 1) - rb;
    const bool own = (A.node_is_own == nullptr) || (A.node_is_own[row] != 0);
    double* acc = SMEM ? (acc_s + (size_t)(rb - vbeg) * BS2) : (A.values + (long long)rb * BS2);
    const int32_t* cols = SMEM ? (cols_s + (rb - vbeg)) : (A.columns + rb);
    if (!SMEM && !A.accumulate) {
      for (int q = 0; q < k * BS2; ++q)
        acc[q] = 0.0;
    }
    if (own) {
      const int32_t qb = __ldg(A.nc_ptr + row), qe = __ldg(A.nc_ptr + row + 1);
      double X[NV][3];
      load_xyz(A.coords, (int32_t)row, X[0]);
      for (int32_t q = qb; q < qe; ++q) {
        const uint32_t e = SMEM ? adj_s[q - abeg] : __ldg(A.nc_list + q);
        const long long cell = e >> 2;
        int32_t n[NV];
        load_cell_nodes<NV>(A.cell_node, cell, n);
        bring_to_front<NV>(n, (int)(e & 3u));
#pragma unroll
        for (int b = 1; b < NV; ++b)
          load_xyz(A.coords, n[b], X[b]);
        const typename Op::Geom G = op.prepare(X, cell);
#pragma unroll
        for (int b = 0; b < NV; ++b) {
          const int x = (b == 0) ? diag_slot(k, A.diag_last) : find_slot(cols, k, n[b], A.diag_last);
          if (x < 0) {
            atomicAdd(A.counters, 1ull);
            continue;
          }
          double out[BS2];
          op.block(G, 0, b, out);
#pragma unroll
          for (int i = 0; i < BS; ++i)
#pragma unroll
            for (int j = 0; j < BS; ++j)
              acc[value_index<BS, PER_BLOCK>(0, k, x, i, j)] += out[i * BS + j];
        }
      }
    }
  }

  if (SMEM) {
    __syncthreads();
    const long long base = (long long)vbeg * BS2;
    const int nval = (vend - vbeg) * BS2;
    // coalesced store of the slab; 16-byte vector stores on the aligned middle part
    const int head = (int)((2 - (base & 1)) & 1); // values is 16B-aligned at element 0
    if (t < head && t < nval)
      A.values[base + t] = acc_s[t];
    const int nvec = (nval - head) >> 1;
    double2* dst = reinterpret_cast<double2*>(A.values + base + head);
    for (int q = t; q < nvec; q += blockDim.x)
      dst[q] = make_double2(acc_s[head + 2 * q], acc_s[head + 2 * q + 1]);
    const int tail = head + 2 * nvec;
    if (t == 0 && tail < nval)
      A.values[base + tail] = acc_s[tail];
  }
}

template <class Op, bool PER_BLOCK>
__global__ void __launch_bounds__(GatherCfg<Op>::ROWS) k_node_gather(AssembleArgs A, Op op)
{
  using C = GatherCfg<Op>;
  extern __shared__ __align__(16) unsigned char b2f_smem[];
  double* acc_s = reinterpret_cast<double*>(b2f_smem);
  int32_t* cols_s = reinterpret_cast<int32_t*>(acc_s + (size_t)C::ACC_BLOCKS * C::BS2);
  uint32_t* adj_s = reinterpret_cast<uint32_t*>(cols_s + C::ACC_BLOCKS);

  const long long row0 = (long long)blockIdx.x * C::ROWS;
  const long long left = A.nb_node - row0;
  const int nrows = left < C::ROWS ? (int)left : C::ROWS;
  const int32_t vbeg = __ldg(A.row_index + row0), vend = __ldg(A.row_index + row0 + nrows);
  const int32_t abeg = __ldg(A.nc_ptr + row0), aend = __ldg(A.nc_ptr + row0 + nrows);
  const bool fits = (vend - vbeg <= C::ACC_BLOCKS) && (aend - abeg <= C::ADJ_CAP);
  if (fits)
    gather_rows<Op, PER_BLOCK, true>(A, op, row0, nrows, vbeg, vend, abeg, aend, acc_s, cols_s, adj_s);
  else {
    if (threadIdx.x == 0)
      atomicAdd(A.counters + 1, 1ull);
    gather_rows<Op, PER_BLOCK, false>(A, op, row0, nrows, vbeg, vend, abeg, aend, acc_s, cols_s, adj_s);
  }
}


// ---------------------------------------------------------------------------------------------------
// Node-wise gather, plan-driven ("slots") kernel — the default.
//
// The sparsity build also emits, for every (row node, incident cell) pair, one 32-bit word holding the column slots
// (10 bits each) of the cell's other NV-1 vertices inside the row, listed in the even-permuted order that puts the
// row node at local vertex 0.  With it the hot loop needs neither the cell's connectivity nor a column search: the
// neighbour ids are columns[row_begin + slot], so one staged word per pair drives NV-1 coordinate gathers, the row-0
// entries of Ke and NV accumulator updates.  The tile's column indices and slot words are contiguous in global memory
// and are brought into shared memory by two TMA bulk copies (cp.async.bulk -> mbarrier) while the accumulators are
// cleared; coordinates of the next pair are prefetched into registers while the current pair is evaluated.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_global, uint32_t bytes, unsigned long long* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(__cvta_generic_to_global(src_global)), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// Cache hints of the fan kernel's streams (B2F_FAN_STREAM): 1 = the slab write-out uses streaming stores (st.global.cs), 2 = also the
// bulk copies of columns and step words carry an L2 evict_first policy - both are touched once per launch, while the gathered
// coordinates are re-read ~14 times from L2 (hit rate 54 %, profiles/r02_fan_ncu.txt).
#ifndef B2F_FAN_STREAM
#define B2F_FAN_STREAM 0
#endif
__device__ __forceinline__ void tma_load_1d_stream(void* dst_smem, const void* src_global, uint32_t bytes, unsigned long long* bar)
{
#if B2F_FAN_STREAM >= 2
  unsigned long long policy;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(dst_smem)),
               "l"(__cvta_generic_to_global(src_global)), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
               : "memory");
#else
  tma_load_1d(dst_smem, src_global, bytes, bar);
#endif
}
// 1: the finished slab of a tile leaves shared memory as ONE bulk copy of the async proxy (cp.async.bulk.global.shared::cta) instead of
// LDS + STG.128 of every thread: the write-out then costs no wavefronts of the L1 data pipe (the unit this kernel loads most)
#ifndef B2F_FAN_BULK_STORE
#define B2F_FAN_BULK_STORE 1
#endif
__device__ __forceinline__ void tma_store_1d(void* dst_global, const void* src_smem, uint32_t bytes)
{
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(__cvta_generic_to_global(dst_global)), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read()
{
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
#ifndef B2F_FAN_PREFETCH
#define B2F_FAN_PREFETCH 0
#endif
__device__ __forceinline__ void prefetch_l2(const void* p)
{
#if B2F_FAN_PREFETCH == 2
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#elif B2F_FAN_PREFETCH == 1
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
}
// L2 prefetch of a contiguous range (no destination, no completion tracking); size in bytes, multiple of 16
__device__ __forceinline__ void bulk_prefetch_l2(const void* p, uint32_t bytes)
{
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(__cvta_generic_to_global(p)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity)
{
  uint32_t done;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!done);
}

template <class Op>
struct SlotsCfg
{
  static constexpr int BS2 = Op::BS * Op::BS;
  static constexpr int ROWS = (Op::BS == 1) ? 128 : 64; // block rows (= threads) per CTA
  static constexpr int MAX_BLOCKS_PER_ROW = (Op::BS == 3) ? 18 : 24; // caps of the shared-memory tile (per row, on average)
  static constexpr int MAX_ADJ_PER_ROW = 40;
  static constexpr int PAD = 8; // alignment slack of the bulk copies (source is aligned down to 16 bytes)
  // capacities are rounded to multiples of 4 so that every region stays 16-byte aligned
  static size_t smem_bytes(int acc_blocks, int adj_cap)
  {
    return 16 + (size_t)acc_blocks * BS2 * sizeof(double) + (size_t)(acc_blocks + PAD) * sizeof(int32_t) + (size_t)(adj_cap + PAD) * sizeof(uint32_t);
  }
};

// One (row node, cell) pair: vertices X (X[0] = the row node), slot word w -> row-0 entries of Ke into acc.
template <class Op, bool PER_BLOCK>
__device__ __forceinline__ void gather_pair(const Op& op, const double (&X)[Op::NV][3], uint32_t w, long long cell, int k, double* __restrict__ acc)
{
  constexpr int NV = Op::NV, BS = Op::BS, BS2 = BS * BS;
  const typename Op::Geom G = op.prepare(X, cell);
#pragma unroll
  for (int b = 0; b < NV; ++b) {
    const int x = (b == 0) ? 0 : (int)((w >> (10 * (b - 1))) & 1023u);
    double out[BS2];
    op.block(G, 0, b, out);
#pragma unroll
    for (int i = 0; i < BS; ++i)
#pragma unroll
      for (int j = 0; j < BS; ++j)
        acc[value_index<BS, PER_BLOCK>(0, k, x, i, j)] += out[i * BS + j];
  }
}

template <class Op>
__device__ __forceinline__ void fetch_pair(const double* __restrict__ coords, const int32_t* cols, uint32_t w, double (&X)[Op::NV][3])
{
#pragma unroll
  for (int b = 1; b < Op::NV; ++b)
    load_xyz(coords, cols[(w >> (10 * (b - 1))) & 1023u], X[b]);
}

template <class Op, bool PER_BLOCK>
__global__ void __launch_bounds__(SlotsCfg<Op>::ROWS) k_node_gather_slots(AssembleArgs A, Op op, int acc_blocks, int adj_cap)
{
  using C = SlotsCfg<Op>;
  constexpr int NV = Op::NV, BS = Op::BS, BS2 = BS * BS;
  extern __shared__ __align__(16) unsigned char b2f_smem[];
  unsigned long long* bar = reinterpret_cast<unsigned long long*>(b2f_smem);
  double* acc_s = reinterpret_cast<double*>(b2f_smem + 16);
  int32_t* cols_s = reinterpret_cast<int32_t*>(acc_s + (size_t)acc_blocks * BS2);
  uint32_t* slots_s = reinterpret_cast<uint32_t*>(cols_s + acc_blocks + C::PAD);

  const int t = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * C::ROWS;
  const long long left = A.nb_node - row0;
  const int nrows = left < C::ROWS ? (int)left : C::ROWS;
  const int32_t vbeg = __ldg(A.row_index + row0), vend = __ldg(A.row_index + row0 + nrows);
  const int32_t abeg = __ldg(A.nc_ptr + row0), aend = __ldg(A.nc_ptr + row0 + nrows);
  const int32_t vbeg_al = vbeg & ~3, abeg_al = abeg & ~3; // 16-byte aligned sources
  const int ncols_cp = ((vend - vbeg_al) + 3) & ~3, nadj_cp = ((aend - abeg_al) + 3) & ~3;
  const bool fits = (vend - vbeg <= acc_blocks) && (ncols_cp <= acc_blocks + C::PAD) && (nadj_cp <= adj_cap + C::PAD);
  if (!fits) { // oversized tile (very long rows): search-based path on global memory
    if (t == 0)
      atomicAdd(A.counters + 1, 1ull);
    gather_rows<Op, PER_BLOCK, false>(A, op, row0, nrows, vbeg, vend, abeg, aend, nullptr, nullptr, nullptr);
    return;
  }

  if (t == 0)
    mbar_init(bar, 1);
  __syncthreads();
  if (t == 0) {
    mbar_expect_tx(bar, (uint32_t)(ncols_cp + nadj_cp) * 4u);
    if (ncols_cp > 0)
      tma_load_1d(cols_s, A.columns + vbeg_al, (uint32_t)ncols_cp * 4u, bar);
    if (nadj_cp > 0)
      tma_load_1d(slots_s, A.nc_slots + abeg_al, (uint32_t)nadj_cp * 4u, bar);
  }
  const int nval = (vend - vbeg) * BS2;
  if (A.accumulate) {
    for (int q = t; q < nval; q += C::ROWS)
      acc_s[q] = A.values[(long long)vbeg * BS2 + q];
  }
  else {
    for (int q = t; q < nval; q += C::ROWS)
      acc_s[q] = 0.0;
  }
  // per-row setup overlaps the bulk copies
  long long row = row0 + t;
  int32_t rb = 0, qb = 0;
  int k = 0, npair = 0;
  bool own = false;
  double XA[NV][3], XB[NV][3];
  if (t < nrows) {
    rb = __ldg(A.row_index + row);
    k = __ldg(A.row_index + row + 1) - rb;
    own = (A.node_is_own == nullptr) || (A.node_is_own[row] != 0);
    qb = __ldg(A.nc_ptr + row);
    npair = own ? (__ldg(A.nc_ptr + row + 1) - qb) : 0;
    load_xyz(A.coords, (int32_t)row, XA[0]);
#pragma unroll
    for (int d = 0; d < 3; ++d)
      XB[0][d] = XA[0][d];
  }
  mbar_wait(bar, 0);
  __syncthreads();

  if (npair > 0) {
    double* acc = acc_s + (size_t)(rb - vbeg) * BS2;
    const int32_t* cols = cols_s + (rb - vbeg_al);
    const uint32_t* sl = slots_s + (qb - abeg_al);
    const uint32_t* cl = A.nc_list + qb;
    // two register sets used alternately: the vertices of pair p+1 are in flight while pair p is evaluated
    uint32_t wA = sl[0], wB = 0;
    fetch_pair<Op>(A.coords, cols, wA, XA);
    int p = 0;
    while (true) {
      if (p + 1 < npair) {
        wB = sl[p + 1];
        fetch_pair<Op>(A.coords, cols, wB, XB);
      }
      gather_pair<Op, PER_BLOCK>(op, XA, wA, Op::NEEDS_CELL ? (long long)(__ldg(cl + p) >> 2) : 0ll, k, acc);
      if (++p >= npair)
        break;
      if (p + 1 < npair) {
        wA = sl[p + 1];
        fetch_pair<Op>(A.coords, cols, wA, XA);
      }
      gather_pair<Op, PER_BLOCK>(op, XB, wB, Op::NEEDS_CELL ? (long long)(__ldg(cl + p) >> 2) : 0ll, k, acc);
      if (++p >= npair)
        break;
    }
  }

  __syncthreads();
  const long long base = (long long)vbeg * BS2;
  const int head = (int)(base & 1); // values is 16-byte aligned at element 0
  if (t < head && t < nval)
    A.values[base + t] = acc_s[t];
  const int nvec = (nval - head) >> 1;
  double2* dst = reinterpret_cast<double2*>(A.values + base + head);
  for (int q = t; q < nvec; q += C::ROWS)
    dst[q] = make_double2(acc_s[head + 2 * q], acc_s[head + 2 * q + 1]);
  const int tail = head + 2 * nvec;
  if (t == 0 && tail < nval)
    A.values[base + tail] = acc_s[tail];
}

// ---------------------------------------------------------------------------------------------------
// Node-wise gather, fan-walk kernel — the default.
//
// ncu of the slots kernel on the 50M-tet cube (profiles/r01_gather_slots.txt) shows the L1 data pipe saturated
// (l1tex__data_pipe_lsu_wavefronts 96.6 %): 9 scattered coordinate gathers + 8 shared-memory accumulator accesses per
// (row node, cell) pair.  The fan plan (sparsity.cu, k_build_fan) orders the cells around a node so that consecutive
// cells share a face through the node.  A thread then keeps a window of the shared vertices (tets: A, B with
// n = eA x eB; triangles: A) in registers together with their accumulators, and every step
//   * gathers ONE new vertex C (coordinates prefetched one step ahead) and loads its accumulator,
//   * builds the cell's un-normalised gradients from the carried cross product: gA = eB x eC, gB = eC x eA, gC = n,
//   * adds row 0 of Ke into the three register accumulators + the register diagonal,
//   * flushes the accumulator of the vertex that leaves the window (plain store: it was loaded on entry) and replaces
//     that vertex by C in place (n <- -gA or -gB).
// Per cell: 3 gathers + 1 LDS + 1 STS of accumulators instead of 9 + 4 + 4, and 12 fewer fp64 operations.
// ---------------------------------------------------------------------------------------------------
constexpr unsigned FAN_START = 0x80000000u;

// Operators whose row-0 entries are "zero-row-sum stiffness + m off the diagonal, + 2 m on it" with a per-cell mass term
// m = op.mass(G) (heat: lambda * stiffness + mass / dt) declare `static constexpr bool ROW_SUM_MASS = true`: over the cells around
// a row, diag = -(sum of the row's off-diagonal entries) + (NV + 1) * (sum of m) - one add per step instead of the diagonal's
// dot product (BS = 1, overwrite mode).
template <class Op, class = void>
struct row_sum_mass : std::false_type
{};
template <class Op>
struct row_sum_mass<Op, std::enable_if_t<Op::ROW_SUM_MASS>> : std::true_type
{};

// Operator entry points of the fan kernel, selected by arity (only the matching one is instantiated).
template <class Op>
__device__ __forceinline__ auto tet_finish(const Op& op, const TetGeom& T, double cell, bool odd) -> decltype(op.finish(T, cell, odd))
{
  return op.finish(T, cell, odd);
}
template <class Op>
__device__ __forceinline__ auto tri_prepare(const Op& op, const double (&X0)[3], const double (&XA)[3], const double (&XC)[3], double cell, bool odd)
  -> decltype(op.prepare_oriented(std::declval<const double (&)[3][3]>(), cell, odd))
{
  const double X[3][3] = { { X0[0], X0[1], X0[2] }, { XA[0], XA[1], XA[2] }, { XC[0], XC[1], XC[2] } };
  return op.prepare_oriented(X, cell, odd);
}

// 1: the per-row loads of the prologue are issued with the tile extents (one DRAM round trip instead of three)
#ifndef B2F_FAN_EARLY_ROW
#define B2F_FAN_EARLY_ROW 1
#endif
// 1: operators with per-cell data get the cell ids of the tile's steps by a third bulk copy into shared memory (+4 B per step)
#ifndef B2F_FAN_STAGE_CELLS
#define B2F_FAN_STAGE_CELLS 1
#endif
// 1: HeatTet gathers its coordinates two steps ahead (see D2 in k_node_fan)
#ifndef B2F_FAN_DEPTH2
#define B2F_FAN_DEPTH2 1
#endif
// 1: the per-cell value of the next step is requested at the top of a step, with the coordinate gathers, not at its end
#ifndef B2F_FAN_CELL_EARLY
#define B2F_FAN_CELL_EARLY 1
#endif
// 1: the outer-product operators on tets (elasticity 3D: 5 warps per SM, registers to spare) gather two steps ahead too
#ifndef B2F_FAN_DEPTH2_POST
#define B2F_FAN_DEPTH2_POST 1
#endif
// 1: HeatTet keeps the per-cell value two steps ahead (see CD2 in k_node_fan)
#ifndef B2F_FAN_CELL_DEPTH2
#define B2F_FAN_CELL_DEPTH2 1
#endif
#ifndef B2F_FAN_AA_CELL
#define B2F_FAN_AA_CELL 1
#endif
#ifndef B2F_FAN_ADDR_AHEAD
#define B2F_FAN_ADDR_AHEAD 1
#endif
#ifndef B2F_FAN_ASM_SMEM
#define B2F_FAN_ASM_SMEM 1
#endif
// Shared-memory accessors on 32-bit shared-window addresses (B2F_FAN_ASM_SMEM): the walk keeps three byte addresses in
// registers instead of re-deriving 64-bit generic pointers from the dynamic-smem base at every step.  The read-only arrays
// (columns, step words) use plain asm (free to schedule); the accumulators use volatile asm, which keeps their loads and
// stores in program order with respect to each other.
// B2F_CHECKED (tools/build_variant.sh checked -DB2F_CHECKED=1): every shared-memory access of the walk is checked against the
// calling thread's OWN range (its row's slots of the slab, its row's column ids, its row's step words) and every gathered node id
// against [0, nb_node); violations are counted in counters[0] (b2f_stats.columns_not_found, which every parity test asserts to be
// zero) and the access is skipped.  Stands in for compute-sanitizer, which is closed on the measurement pool: a walk that stays
// inside its own ranges cannot race with another thread's, and cannot leave the tile.
#ifndef B2F_CHECKED
#define B2F_CHECKED 0
#endif
#if B2F_CHECKED
#define B2F_RANGE_MEMBERS \
  int lo, hi;             \
  unsigned long long* bad;
#define B2F_RANGE_TEST(i, fail_value)   \
  if ((i) < lo || (i) >= hi) {          \
    atomicAdd(bad, 1ull);               \
    return fail_value;                  \
  }
#else
#define B2F_RANGE_MEMBERS
#define B2F_RANGE_TEST(i, fail_value)
#endif
struct SmemI32
{
  uint32_t a;
  B2F_RANGE_MEMBERS
  __device__ __forceinline__ int32_t operator[](int i) const
  {
    B2F_RANGE_TEST(i, 0)
    int32_t v;
    asm("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a + 4u * (uint32_t)i));
    return v;
  }
};
struct SmemU32
{
  uint32_t a;
  B2F_RANGE_MEMBERS
  __device__ __forceinline__ uint32_t operator[](int i) const
  {
    B2F_RANGE_TEST(i, 0x80000000u)
    uint32_t v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a + 4u * (uint32_t)i));
    return v;
  }
};
struct SmemF64
{
  uint32_t a;
  B2F_RANGE_MEMBERS
  __device__ __forceinline__ double ld(int i) const
  {
    B2F_RANGE_TEST(i, 0.0)
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a + 8u * (uint32_t)i));
    return v;
  }
  __device__ __forceinline__ double operator[](int i) const { return ld(i); }
  __device__ __forceinline__ void st(int i, double v) const
  {
    B2F_RANGE_TEST(i, )
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(a + 8u * (uint32_t)i), "d"(v));
  }
};
template <int BS, bool PER_BLOCK>
__device__ __forceinline__ void load_block(SmemF64 acc, int k, int x, double (&out)[BS * BS])
{
#pragma unroll
  for (int i = 0; i < BS; ++i)
#pragma unroll
    for (int j = 0; j < BS; ++j)
      out[i * BS + j] = acc.ld((int)value_index<BS, PER_BLOCK>(0, k, x, i, j));
}
template <int BS, bool PER_BLOCK>
__device__ __forceinline__ void store_block(SmemF64 acc, int k, int x, const double (&v)[BS * BS])
{
#pragma unroll
  for (int i = 0; i < BS; ++i)
#pragma unroll
    for (int j = 0; j < BS; ++j)
      acc.st((int)value_index<BS, PER_BLOCK>(0, k, x, i, j), v[i * BS + j]);
}

template <int BS, bool PER_BLOCK>
__device__ __forceinline__ void load_block(const double* __restrict__ acc, int k, int x, double (&out)[BS * BS])
{
#pragma unroll
  for (int i = 0; i < BS; ++i)
#pragma unroll
    for (int j = 0; j < BS; ++j)
      out[i * BS + j] = acc[value_index<BS, PER_BLOCK>(0, k, x, i, j)];
}

template <int BS, bool PER_BLOCK>
__device__ __forceinline__ void store_block(double* __restrict__ acc, int k, int x, const double (&v)[BS * BS])
{
#pragma unroll
  for (int i = 0; i < BS; ++i)
#pragma unroll
    for (int j = 0; j < BS; ++j)
      acc[value_index<BS, PER_BLOCK>(0, k, x, i, j)] = v[i * BS + j];
}

// 0: the pivot change is selected at every step; 1: behind a warp vote; 2 (default): behind bit 12 of the step word, which
// the plan sets when some row of the warp changes its pivot at that step (0.980 -> 0.965 ms at N = 203 against the vote)
#ifndef B2F_FAN_UNIFORM_PIVOT
#define B2F_FAN_UNIFORM_PIVOT 2
#endif
// One step of the walk.  Window = (N, O): N the vertex that entered last, O the other one (tets only).  (eC, accC, sC)
// receive the vertex that enters (eC is filled by the caller: edge vector C - row node for tets, the vertex for triangles); the caller alternates two register sets between the roles N and C, so the "C becomes
// N" renaming costs no instruction.  Tets carry m = eN x eO.
template <class Op, bool PER_BLOCK, bool SKIP_DIAG, bool SRC, class AccT, class ColsT>
__device__ __forceinline__ void fan_step(const Op& op, const CoordView& coords, const double (&X0)[3], AccT acc, ColsT cols,
                                         int k, uint32_t wc, uint32_t first_mask, double cell, double (&eN)[3], double (&accN)[Op::BS * Op::BS], int& sN,
                                         double (&eO)[3], double (&accO)[Op::BS * Op::BS], int& sO, double (&m)[3], double (&eC)[3],
                                         double (&accC)[Op::BS * Op::BS], int& sC, double (&diag)[Op::BS * Op::BS], double& measure)
{
  constexpr int NV = Op::NV, BS = Op::BS, BS2 = BS * BS;
  sC = (int)(wc & 1023u);
  if (wc & FAN_START) {
    // flush the whole window before loading anything: a vertex of the new window may still be held by the old one
    if (sN)
      store_block<BS, PER_BLOCK>(acc, k, sN, accN);
    if constexpr (NV == 4) {
      if (sO)
        store_block<BS, PER_BLOCK>(acc, k, sO, accO);
    }
    // first_mask = 8192 (bit 13) when the plan's first-visit marks apply: START words carry them in bits 20 / 21 (= mask << 7, << 8)
    if (wc & (first_mask << 7)) {
#pragma unroll
      for (int e = 0; e < BS2; ++e)
        accC[e] = 0.0;
    }
    else
      load_block<BS, PER_BLOCK>(acc, k, sC, accC);
    if constexpr (NV == 4) {
      sO = (int)((wc >> 10) & 1023u);
      if (wc & (first_mask << 8)) {
#pragma unroll
        for (int e = 0; e < BS2; ++e)
          accO[e] = 0.0;
      }
      else
        load_block<BS, PER_BLOCK>(acc, k, sO, accO);
      double XB[3];
      load_xyz(coords, cols[sO], XB);
#pragma unroll
      for (int d = 0; d < 3; ++d)
        eO[d] = XB[d] - X0[d];
      m[0] = eC[1] * eO[2] - eC[2] * eO[1];
      m[1] = eC[2] * eO[0] - eC[0] * eO[2];
      m[2] = eC[0] * eO[1] - eC[1] * eO[0];
    }
    return;
  }
  const bool odd = (wc >> 11) & 1u;
  if (wc & first_mask) { // first visit of the slot in this row: the accumulator starts from zero, the slab was not cleared
#pragma unroll
    for (int e = 0; e < BS2; ++e)
      accC[e] = 0.0;
  }
  else
    load_block<BS, PER_BLOCK>(acc, k, sC, accC);
  double out[BS2];
  if constexpr (NV == 4) {
    const bool dropN = (wc >> 10) & 1u;
    // local vertices (0, 1, 2, 3) = (row node, N, O, C): g1 = eO x eC, g2 = eC x eN, g3 = eN x eO = m (carried).
    // h = -g1 = eC x eO is what the next step carries when O stays, so it is the product that is formed; the negations
    // below fold into operand modifiers of the consumers.
    double h[3];
    h[0] = eC[1] * eO[2] - eC[2] * eO[1];
    h[1] = eC[2] * eO[0] - eC[0] * eO[2];
    h[2] = eC[0] * eO[1] - eC[1] * eO[0];
    TetGeom T;
    T.g[2][0] = eC[1] * eN[2] - eC[2] * eN[1];
    T.g[2][1] = eC[2] * eN[0] - eC[0] * eN[2];
    T.g[2][2] = eC[0] * eN[1] - eC[1] * eN[0];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      T.g[1][d] = -h[d];
      T.g[NV - 1][d] = m[d];
      T.g[0][d] = h[d] - T.g[2][d] - m[d];
    }
    const double det = eC[0] * m[0] + eC[1] * m[1] + eC[2] * m[2];
    T.vol6 = fabs(det);
    T.s = fast_rcp(6.0 * T.vol6);
    if constexpr (SRC)
      measure += T.vol6; // 6 * volume of the incident cell (computeVolumeTetra4, ArcaneFemFunctionsGpu.h:110-122)
    const auto G = tet_finish(op, T, cell, odd);
    if constexpr (Op::LINEAR_POST) {
      // plain outer products (s*g0) (x) g_b; the operator's linear map is applied once per stored block at the end
      double u[BS];
      op.raw_begin(G, u);
      op.raw_add(G, u, 1, accN);
      op.raw_add(G, u, 2, accO);
      op.raw_add(G, u, NV - 1, accC);
    }
    else {
    op.block(G, 0, 1, out);
#pragma unroll
    for (int e = 0; e < BS2; ++e)
      accN[e] += out[e];
    op.block(G, 0, 2, out);
#pragma unroll
    for (int e = 0; e < BS2; ++e)
      accO[e] += out[e];
    op.block(G, 0, NV - 1, out);
#pragma unroll
    for (int e = 0; e < BS2; ++e)
      accC[e] += out[e];
    }
    if constexpr (!SKIP_DIAG) {
      op.block(G, 0, 0, out);
#pragma unroll
      for (int e = 0; e < BS2; ++e)
        diag[e] += out[e];
    }
    else if constexpr (row_sum_mass<Op>::value)
      diag[0] += op.mass(G); // the diagonal register carries the sum of the mass terms until the end of the row
    // one vertex leaves: its accumulator is flushed (plain store: it was loaded on entry).  C becomes N (renaming by
    // the caller); the vertex that stays becomes / remains O.  New m: (eC x eN) = g2 if N stays, (eC x eO) = -g1 if O stays.
    double leave[BS2];
#pragma unroll
    for (int e = 0; e < BS2; ++e)
      leave[e] = dropN ? accN[e] : accO[e];
    store_block<BS, PER_BLOCK>(acc, k, dropN ? sN : sO, leave);
#if B2F_FAN_UNIFORM_PIVOT == 2
    // the plan says whether any row of this warp changes its pivot at this step (bit 12, sparsity.cu k_fan_pivot_hint):
    // same branch as below, taken on plain data - no vote, no convergence check
    if (wc & 4096u) {
#elif B2F_FAN_UNIFORM_PIVOT
    // Rows of a warp mostly change their pivot O at the same steps (identical link structure on structured meshes), so
    // the pivot change sits behind a warp-uniform branch: when no lane changes, only the renaming m = h remains.
    if (__any_sync(__activemask(), !dropN)) {
#endif
      if (!dropN) {
        sO = sN;
#pragma unroll
        for (int e = 0; e < BS2; ++e)
          accO[e] = accN[e];
      }
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        eO[d] = dropN ? eO[d] : eN[d];
        m[d] = dropN ? h[d] : T.g[2][d]; // new window (C, O'): eC x eO = h if O stays, eC x eN = g2 if N stays
      }
#if B2F_FAN_UNIFORM_PIVOT
    }
    else {
#pragma unroll
      for (int d = 0; d < 3; ++d)
        m[d] = h[d];
    }
#endif
  }
  else {
    const auto G = tri_prepare(op, X0, eN, eC, cell, odd);
    if constexpr (SRC) {
      // 2 * area with the 3D cross product of computeAreaTria3 (ArcaneFemFunctionsGpu.h:75-83; z takes part)
      const double ax = eN[0] - X0[0], ay = eN[1] - X0[1], az = eN[2] - X0[2];
      const double bx = eC[0] - X0[0], by = eC[1] - X0[1], bz = eC[2] - X0[2];
      const double cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
      measure += sqrt(cx * cx + cy * cy + cz * cz);
    }
    if constexpr (Op::LINEAR_POST) {
      double u[BS];
      op.raw_begin(G, u);
      op.raw_add(G, u, 1, accN);
      op.raw_add(G, u, 2, accC);
    }
    else {
      op.block(G, 0, 1, out);
#pragma unroll
      for (int e = 0; e < BS2; ++e)
        accN[e] += out[e];
      op.block(G, 0, 2, out);
#pragma unroll
      for (int e = 0; e < BS2; ++e)
        accC[e] += out[e];
    }
    if constexpr (!SKIP_DIAG) {
      op.block(G, 0, 0, out);
#pragma unroll
      for (int e = 0; e < BS2; ++e)
        diag[e] += out[e];
    }
    else if constexpr (row_sum_mass<Op>::value)
      diag[0] += op.mass(G);
    store_block<BS, PER_BLOCK>(acc, k, sN, accN);
  }
}

#ifndef B2F_FAN_ROWS
#define B2F_FAN_ROWS 128
#endif
#ifndef B2F_FAN_ROWS_BS3
#define B2F_FAN_ROWS_BS3 32
#endif
template <class Op>
struct FanCfg : SlotsCfg<Op>
{
  // BS = 3: 9 doubles per block make shared memory the occupancy limit (1.2 KB per row); 32-row CTAs pack 6 CTAs = 6 warps
  // per SM where 64-row CTAs fit 2 CTAs = 4 warps (elasticity3d 2.51 -> 2.07 ms)
  static constexpr int ROWS = (Op::BS == 1) ? B2F_FAN_ROWS : (Op::BS == 3 ? B2F_FAN_ROWS_BS3 : 64); // block rows (= threads) per CTA
  static_assert(ROWS == 128 || ROWS == 64 || ROWS == 32, "tile extents exist for 128, 64 and 32 rows");
};
#ifndef B2F_FAN_TILE_AHEAD
#define B2F_FAN_TILE_AHEAD 0
#endif
// 0: one tile per CTA.  n > 0: persistent CTAs, n x (SMs x resident CTAs per SM) of them, each walking every gridDim.x-th tile
#ifndef B2F_FAN_PERSIST
#define B2F_FAN_PERSIST 0
#endif
#ifndef B2F_FAN_WORD_AHEAD
#define B2F_FAN_WORD_AHEAD 0
#endif
#ifndef B2F_FAN_CARVEOUT
#define B2F_FAN_CARVEOUT -1
#endif
// DIAG_BY_ROW_SUM (operators whose element matrices have zero row sums, BS = 1, overwrite mode): the diagonal entry is
// not accumulated along the walk but formed at the end as minus the sum of the row's off-diagonal entries.
// SRC: the constant-source right-hand side (applyConstantSourceToRhsBase, ArcaneFemFunctionsGpu.h:402-429) rides along:
// the walk sums the measures of the incident cells and the row adds qdot * sum / (NV * NV!) to rhs[dof(row, 0)] with a
// plain store, no atomics and no second pass over the cells.
// LOCKSTEP (tets): the walk runs on three fixed register sets and the step word names the variant of the step body to
// run (which set receives the entering vertex, which of the other two leaves): no selects and no register moves, at the
// price of one multi-way branch per step.  Rows of a warp that sit in different variants are serialised.  Opt-in
// (B2F_FAN_LOCKSTEP, plan converted by k_fan_lockstep in sparsity.cu): on B200 it measured slower than the select walk.
// PERM: the reorder mode (b2f_options.reorder) as its own instantiation, so that the default kernel carries none of its state.
template <class Op, bool PER_BLOCK, bool DIAG_BY_ROW_SUM, bool SRC, bool LOCKSTEP, bool PERM = false>
__global__ void __launch_bounds__(FanCfg<Op>::ROWS, Op::FAN_WARPS_PER_SM * 32 / FanCfg<Op>::ROWS) k_node_fan(AssembleArgs A, Op op, int acc_blocks, int step_cap, long long nb_tiles)
{
  using C = FanCfg<Op>;
  constexpr int NV = Op::NV, BS = Op::BS, BS2 = BS * BS;
  // POST (operators that are a linear map of plain outer-product sums, element_ops.cuh LINEAR_POST): the walk accumulates
  // the outer products from zero, the diagonal block is minus the sum of the row's other blocks (gradients of a cell sum
  // to zero), the map runs once per stored block after the walk, and accumulate mode adds the old values at the store.
  constexpr bool POST = Op::LINEAR_POST;
  const uint32_t fmask = LOCKSTEP ? 0u : A.fan_first_mask; // 8192 or 0, see AssembleArgs
#if B2F_CHECKED
  const CoordView CV{ A.coords, A.nb_node, A.counters };
#else
  const CoordView CV{ A.coords };
#endif
  extern __shared__ __align__(16) unsigned char b2f_smem[];
  unsigned long long* bar = reinterpret_cast<unsigned long long*>(b2f_smem);
#if B2F_FAN_BULK_STORE
  double* const acc_base = reinterpret_cast<double*>(b2f_smem + 16);
  double* acc_s = acc_base; // moved back by one element for tiles whose first value sits at an odd index (below)
  int32_t* cols_s = reinterpret_cast<int32_t*>(acc_base + (size_t)acc_blocks * BS2);
#else
  double* acc_s = reinterpret_cast<double*>(b2f_smem + 16);
  int32_t* cols_s = reinterpret_cast<int32_t*>(acc_s + (size_t)acc_blocks * BS2);
#endif
  uint32_t* steps_s = reinterpret_cast<uint32_t*>(cols_s + acc_blocks + C::PAD);
#if B2F_FAN_STAGE_CELLS
  // operators with per-cell data: the cell id of every step is staged beside the step words (third bulk copy) instead of being
  // read from global memory inside the walk
  uint32_t* cells_s = steps_s + step_cap + C::PAD;
#endif

  const int t = threadIdx.x;
  // Persistent form (B2F_FAN_PERSIST, launch_assemble_layout): the grid holds one CTA per resident slot of the device and
  // every CTA takes the tiles blockIdx.x, blockIdx.x + gridDim.x, ...; the mbarrier is initialised once and its phase
  // parity alternates.  Non-persistent form: one tile per CTA (nb_tiles == gridDim.x), the loop runs once.
  if (t == 0)
    mbar_init(bar, 1);
  __syncthreads();
#if B2F_FAN_PERSIST
  uint32_t phase = 0;
  for (long long tile = blockIdx.x; tile < nb_tiles; tile += gridDim.x) {
#else
  constexpr uint32_t phase = 0;
  const long long tile = blockIdx.x;
  (void)nb_tiles;
  do { // one tile per CTA; `continue` leaves
#endif
  const long long row0 = (tile + A.first_tile) * C::ROWS;
  const long long left = A.nb_node - row0;
  const int nrows = left < C::ROWS ? (int)left : C::ROWS;
#if B2F_FAN_EARLY_ROW
  // every per-row load that depends on the row id alone is issued here, together with the tile extents: one DRAM round trip in
  // front of the walk instead of three (extents -> walk length -> row extents and coordinates)
  const long long row = row0 + t; // internal row in the reorder mode
  int32_t e_rb = 0, e_rbn = 0, e_qb = 0, e_qbn = 0;
  double X0[3] = { 0.0, 0.0, 0.0 };
  uint8_t e_own = 1;
  if (t < nrows) {
    e_rb = __ldg(A.row_index + row);
    e_rbn = __ldg(A.row_index + row + 1);
    e_qb = __ldg(A.fan_ptr + row);
    e_qbn = __ldg(A.fan_ptr + row + 1);
    load_xyz(CV, (int32_t)row, X0);
    if (!PERM && A.node_is_own != nullptr)
      e_own = A.node_is_own[row];
  }
#endif
  const int32_t vbeg = __ldg(A.row_index + row0), vend = __ldg(A.row_index + row0 + nrows);
  const int32_t abeg = __ldg(A.fan_ptr + row0), aend = __ldg(A.fan_ptr + row0 + nrows);
  const int32_t vbeg_al = vbeg & ~3, abeg_al = abeg & ~3; // 16-byte aligned sources
  const int ncols_cp = ((vend - vbeg_al) + 3) & ~3, nstep_cp = ((aend - abeg_al) + 3) & ~3;
  const bool fits = (vend - vbeg <= acc_blocks) && (ncols_cp <= acc_blocks + C::PAD) && (nstep_cp <= step_cap + C::PAD);
  if (!fits) { // oversized tile (very long rows): search-based path on global memory
    if (t == 0)
      atomicAdd(A.counters + 1, 1ull);
    if constexpr (!PERM) {
      gather_rows<Op, PER_BLOCK, false>(A, op, row0, nrows, vbeg, vend, __ldg(A.nc_ptr + row0), __ldg(A.nc_ptr + row0 + nrows), nullptr, nullptr, nullptr);
      if constexpr (SRC) {
        if (t < nrows && (A.node_is_own == nullptr || A.node_is_own[row0 + t] != 0))
          A.rhs[(row0 + t) * BS] += A.rhs_factor * row_measure<NV>(A, row0 + t);
      }
    }
    else { // reorder mode: the same rows in the caller's numbering, on the caller's arrays
      AssembleArgs C = A;
      C.row_index = A.out_row_index;
      C.columns = A.caller_columns;
      C.coords = A.caller_coords;
      gather_rows<Op, PER_BLOCK, false>(C, op, row0, nrows, 0, 0, 0, 0, nullptr, nullptr, nullptr, A.row_perm);
      if constexpr (SRC) {
        if (t < nrows) {
          const long long r = __ldg(A.row_perm + row0 + t);
          if (A.node_is_own == nullptr || A.node_is_own[r] != 0)
            A.rhs[r * BS] += A.rhs_factor * row_measure<NV>(C, r);
        }
      }
    }
    continue;
  }

  if (t == 0) {
    // the previous tile's reads of the staging areas (generic proxy) are ordered before the bulk copies that overwrite
    // them (async proxy); the __syncthreads at the end of the previous iteration has collected those reads
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    mbar_expect_tx(bar, (uint32_t)(ncols_cp + nstep_cp + ((B2F_FAN_STAGE_CELLS && Op::NEEDS_CELL) ? nstep_cp : 0)) * 4u);
    if (ncols_cp > 0)
      tma_load_1d_stream(cols_s, A.columns + vbeg_al, (uint32_t)ncols_cp * 4u, bar);
    if (nstep_cp > 0)
      tma_load_1d_stream(steps_s, A.fan_steps + abeg_al, (uint32_t)nstep_cp * 4u, bar);
#if B2F_FAN_STAGE_CELLS
    if (Op::NEEDS_CELL && nstep_cp > 0)
      tma_load_1d_stream(cells_s, A.fan_cell + abeg_al, (uint32_t)nstep_cp * 4u, bar);
#endif
  }
  const int nval = (vend - vbeg) * BS2;
#if B2F_FAN_BULK_STORE
  // the slab takes the 16-byte phase of its destination (values is 16-byte aligned at element 0; bytes 8..15 of the dynamic
  // shared memory are spare): element i of the slab and value vbeg * BS2 + i then sit at the same offset modulo 16
  acc_s = acc_base - (int)(((long long)vbeg * BS2) & 1);
#endif
  // per-row setup overlaps the bulk copies
#if !B2F_FAN_EARLY_ROW
  const long long row = row0 + t; // internal row in the reorder mode
#endif
  constexpr bool permuted = PERM;
  long long out_row = row; // the caller's row: ownership, right-hand side, destination of the values
  int32_t rb = 0, qb = 0;
  int k = 0, nstep = 0;
  long long out_base = 0; // first value of the row in the caller's value array (reorder mode)
#if !B2F_FAN_EARLY_ROW
  double X0[3] = { 0.0, 0.0, 0.0 };
#endif
  if (t < nrows) {
#if B2F_FAN_EARLY_ROW
    rb = e_rb;
    k = e_rbn - rb;
#else
    rb = __ldg(A.row_index + row);
    k = __ldg(A.row_index + row + 1) - rb;
#endif
    if (permuted) {
      out_row = __ldg(A.row_perm + row);
      out_base = (long long)__ldg(A.out_row_index + out_row) * BS2;
    }
  }
  if (A.accumulate && !POST) {
    if (!permuted) {
      for (int q = t; q < nval; q += C::ROWS)
        acc_s[q] = A.values[(long long)vbeg * BS2 + q];
    }
    else if (t < nrows) { // block x of the internal row is block slot_map[x] of the caller's row
      for (int x = 0; x < k; ++x) {
        const int xs = (int)__ldg(A.slot_map + rb + x);
#pragma unroll
        for (int i = 0; i < BS; ++i)
#pragma unroll
          for (int j = 0; j < BS; ++j)
            acc_s[(size_t)(rb - vbeg) * BS2 + value_index<BS, PER_BLOCK>(0, k, x, i, j)] = A.values[out_base + value_index<BS, PER_BLOCK>(0, k, xs, i, j)];
      }
    }
  }
  else if (fmask == 0u) { // with the plan's first-visit marks every slot of a walked row is stored before it is read: no clearing pass
    for (int q = t; q < nval; q += C::ROWS)
      acc_s[q] = 0.0;
  }
  if (t < nrows) {
#if B2F_FAN_EARLY_ROW
    const bool own = PERM ? ((A.node_is_own == nullptr) || (A.node_is_own[out_row] != 0)) : (e_own != 0);
    qb = e_qb;
    nstep = own ? (e_qbn - qb) : 0;
#else
    const bool own = (A.node_is_own == nullptr) || (A.node_is_own[out_row] != 0);
    qb = __ldg(A.fan_ptr + row);
    nstep = own ? (__ldg(A.fan_ptr + row + 1) - qb) : 0;
    load_xyz(CV, (int32_t)row, X0);
#endif
  }
#if B2F_FAN_TILE_AHEAD > 0
  // Streaming data of the tile that will start when this one retires (about one wave of CTAs ahead) is pulled into L2
  // now, so that tile's dependent prologue loads (extents -> bulk copies -> first gathers) pay L2 instead of DRAM latency.
  if (t == 32) {
    const long long fr0 = row0 + (long long)B2F_FAN_TILE_AHEAD * C::ROWS;
    if (fr0 < A.nb_node) {
      const long long fr1 = (fr0 + C::ROWS < A.nb_node) ? fr0 + C::ROWS : A.nb_node;
      const int32_t fv0 = __ldg(A.row_index + fr0) & ~3, fv1 = __ldg(A.row_index + fr1);
      const int32_t fa0 = __ldg(A.fan_ptr + fr0) & ~3, fa1 = __ldg(A.fan_ptr + fr1);
      bulk_prefetch_l2(A.columns + fv0, (uint32_t)(((fv1 - fv0) + 3) & ~3) * 4u);
      bulk_prefetch_l2(A.fan_steps + fa0, (uint32_t)(((fa1 - fa0) + 3) & ~3) * 4u);
      bulk_prefetch_l2(A.coords + 3 * (fr0 & ~1ll), (uint32_t)((((fr1 - (fr0 & ~1ll)) * 24) + 15) & ~15ll));
    }
  }
#endif
  if (fmask != 0u && t < nrows && nstep == 0) { // rows that are not walked (not own, no cell) stay zero
    for (int q = 0; q < k * BS2; ++q)
      acc_s[(size_t)(rb - vbeg) * BS2 + q] = 0.0;
  }
  mbar_wait(bar, phase);
#if B2F_FAN_PERSIST
  phase ^= 1u;
#endif
  __syncthreads();

  if (nstep > 0) {
    // The plan's slots are "diagonal 0, neighbours 1 .. k-1 ascending".  Diagonal-last rows (b2f_pattern_layout) store neighbour x
    // at position x - 1 and the diagonal at k - 1: the row's base addresses are moved back by one slot, so every neighbour access
    // of the walk is unchanged, and the diagonal is addressed as slot k.
    const int dl = A.diag_last;
    const int acc_shift = dl * (PER_BLOCK ? BS2 : BS);
#define B2F_DIAG_SLOT (A.diag_last ? k : 0) /* re-read from the parameter bank where it is used: no register lives across the walk */
#if B2F_FAN_ASM_SMEM
#if B2F_CHECKED
    SmemF64 acc{ (uint32_t)__cvta_generic_to_shared(acc_s + (size_t)(rb - vbeg) * BS2) - 8u * (uint32_t)acc_shift, acc_shift, acc_shift + k * BS2, A.counters };
    SmemI32 cols{ (uint32_t)__cvta_generic_to_shared(cols_s + (rb - vbeg_al)) - 4u * (uint32_t)dl, dl, dl + k, A.counters };
    SmemU32 st{ (uint32_t)__cvta_generic_to_shared(steps_s + (qb - abeg_al)), 0, nstep, A.counters };
#else
    SmemF64 acc{ (uint32_t)__cvta_generic_to_shared(acc_s + (size_t)(rb - vbeg) * BS2) - 8u * (uint32_t)acc_shift };
    SmemI32 cols{ (uint32_t)__cvta_generic_to_shared(cols_s + (rb - vbeg_al)) - 4u * (uint32_t)dl };
    SmemU32 st{ (uint32_t)__cvta_generic_to_shared(steps_s + (qb - abeg_al)) };
#endif
    // the addresses become opaque here, after the barrier: the asm loads below depend on them and cannot be hoisted above
    // the wait, and the compiler keeps them in registers instead of re-deriving them
    asm volatile("" : "+r"(acc.a), "+r"(cols.a), "+r"(st.a)::"memory");
#else
    double* acc = acc_s + (size_t)(rb - vbeg) * BS2 - acc_shift;
    const int32_t* cols = cols_s + (rb - vbeg_al) - dl;
    const uint32_t* st = steps_s + (qb - abeg_al);
#endif
#if B2F_FAN_STAGE_CELLS
    const uint32_t* cl = cells_s + (qb - abeg_al);
#define B2F_CELL_AT(ptr) (*(ptr))
#else
    const uint32_t* cl = A.fan_cell + qb;
#define B2F_CELL_AT(ptr) __ldg(ptr)
#endif
    // bring the coordinates of every neighbour of the row towards the SM now, in one round trip, instead of meeting
    // the misses one by one along the walk
    for (int x = 1; x < k; ++x)
      prefetch_l2(A.coords + 3ll * cols[x]);
    double diag[BS2];
    if (fmask != 0u) { // the slab was not cleared (first-visit marks): the diagonal starts from zero like every other first load
#pragma unroll
      for (int e = 0; e < BS2; ++e)
        diag[e] = 0.0;
    }
    else
      load_block<BS, PER_BLOCK>(acc, k, B2F_DIAG_SLOT, diag);
    double measure = 0.0;
    if constexpr (LOCKSTEP) {
      static_assert(!LOCKSTEP || NV == 4, "the lockstep walk is for tetrahedra");
      // sets 0, 1, 2: edge vector from the row node, accumulator block, column slot.  m = e[f+1] x e[f+2] for the free set f.
      double e[3][3], av[3][BS2], m[3] = { 0.0, 0.0, 0.0 };
      int sl[3] = { 0, 0, 0 };
      double Xn[3];
      uint32_t w = st[0], wc = 0;
      load_xyz(CV, cols[w & 1023u], Xn);
      double cv = 0.0, cvn = 0.0;
      uint32_t cidn = 0;
      if constexpr (Op::NEEDS_CELL) {
        cv = op.cell_value((long long)B2F_CELL_AT(cl));
        if (nstep > 1)
          cvn = op.cell_value((long long)B2F_CELL_AT(cl + 1));
        if (nstep > 2)
          cidn = B2F_CELL_AT(cl + 2);
      }
      bool more = false;
      // F: set that receives the entering vertex C; X = F+1, Y = F+2 hold the window; LY: Y leaves (else X).
      // Local vertices (0, 1, 2, 3) = (row, X, Y, C): g1 = eY x eC, g2 = eC x eX, g3 = eX x eY = m.
      auto step = [&](auto Fc, auto LYc) {
        constexpr int F = decltype(Fc)::value, X = (F + 1) % 3, Y = (F + 2) % 3;
        constexpr bool LY = decltype(LYc)::value;
#pragma unroll
        for (int d = 0; d < 3; ++d)
          e[F][d] = Xn[d] - X0[d];
        if (more)
          load_xyz(CV, cols[w & 1023u], Xn);
        sl[F] = (int)(wc & 1023u);
        load_block<BS, PER_BLOCK>(acc, k, sl[F], av[F]);
        const bool odd = (wc >> 13) & 1u;
        TetGeom T;
        T.g[1][0] = e[Y][1] * e[F][2] - e[Y][2] * e[F][1];
        T.g[1][1] = e[Y][2] * e[F][0] - e[Y][0] * e[F][2];
        T.g[1][2] = e[Y][0] * e[F][1] - e[Y][1] * e[F][0];
        T.g[2][0] = e[F][1] * e[X][2] - e[F][2] * e[X][1];
        T.g[2][1] = e[F][2] * e[X][0] - e[F][0] * e[X][2];
        T.g[2][2] = e[F][0] * e[X][1] - e[F][1] * e[X][0];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
          T.g[3][d] = m[d];
          T.g[0][d] = -T.g[1][d] - T.g[2][d] - m[d];
        }
        const double det = e[F][0] * m[0] + e[F][1] * m[1] + e[F][2] * m[2];
        T.vol6 = fabs(det);
        T.s = fast_rcp(6.0 * T.vol6);
        if constexpr (SRC)
          measure += T.vol6;
        const auto G = tet_finish(op, T, cv, odd);
        if constexpr (POST) {
          double u[BS];
          op.raw_begin(G, u);
          op.raw_add(G, u, 1, av[X]);
          op.raw_add(G, u, 2, av[Y]);
          op.raw_add(G, u, 3, av[F]);
        }
        else {
          double out[BS2];
          op.block(G, 0, 1, out);
#pragma unroll
          for (int q = 0; q < BS2; ++q)
            av[X][q] += out[q];
          op.block(G, 0, 2, out);
#pragma unroll
          for (int q = 0; q < BS2; ++q)
            av[Y][q] += out[q];
          op.block(G, 0, 3, out);
#pragma unroll
          for (int q = 0; q < BS2; ++q)
            av[F][q] += out[q];
        }
        if constexpr (!DIAG_BY_ROW_SUM && !POST) {
          double out[BS2];
          op.block(G, 0, 0, out);
#pragma unroll
          for (int q = 0; q < BS2; ++q)
            diag[q] += out[q];
        }
        else if constexpr (DIAG_BY_ROW_SUM && row_sum_mass<Op>::value)
          diag[0] += op.mass(G);
        if constexpr (LY) {
          store_block<BS, PER_BLOCK>(acc, k, sl[Y], av[Y]);
#pragma unroll
          for (int d = 0; d < 3; ++d)
            m[d] = T.g[2][d]; // next free set Y: e[Y+1] x e[Y+2] = eC x eX
        }
        else {
          store_block<BS, PER_BLOCK>(acc, k, sl[X], av[X]);
#pragma unroll
          for (int d = 0; d < 3; ++d)
            m[d] = T.g[1][d]; // next free set X: e[X+1] x e[X+2] = eY x eC
        }
      };
      using I0 = std::integral_constant<int, 0>;
      using I1 = std::integral_constant<int, 1>;
      using I2 = std::integral_constant<int, 2>;
      for (int p = 0; p < nstep; ++p) {
        wc = w;
        more = p + 1 < nstep;
        if (more)
          w = st[p + 1];
        if (wc & FAN_START) {
          // bits 20-21: free set of the window that ends here (3: there is none); its two sets are flushed, then the new
          // window goes to the sets 1 (the vertex that came through the prefetch) and 2, free set 0
          const int fp = (int)((wc >> 20) & 3u);
          if (fp != 3) {
#pragma unroll
            for (int q = 0; q < 3; ++q)
              if (q != fp)
                store_block<BS, PER_BLOCK>(acc, k, sl[q], av[q]);
          }
#pragma unroll
          for (int d = 0; d < 3; ++d)
            e[1][d] = Xn[d] - X0[d];
          sl[1] = (int)(wc & 1023u);
          sl[2] = (int)((wc >> 10) & 1023u);
          double XB[3];
          load_xyz(CV, cols[sl[2]], XB);
          if (more)
            load_xyz(CV, cols[w & 1023u], Xn);
          load_block<BS, PER_BLOCK>(acc, k, sl[1], av[1]);
          load_block<BS, PER_BLOCK>(acc, k, sl[2], av[2]);
#pragma unroll
          for (int d = 0; d < 3; ++d)
            e[2][d] = XB[d] - X0[d];
          m[0] = e[1][1] * e[2][2] - e[1][2] * e[2][1];
          m[1] = e[1][2] * e[2][0] - e[1][0] * e[2][2];
          m[2] = e[1][0] * e[2][1] - e[1][1] * e[2][0];
        }
        else {
          switch ((wc >> 10) & 7u) {
          case 0: step(I0{}, std::false_type{}); break;
          case 1: step(I0{}, std::true_type{}); break;
          case 2: step(I1{}, std::false_type{}); break;
          case 3: step(I1{}, std::true_type{}); break;
          case 4: step(I2{}, std::false_type{}); break;
          default: step(I2{}, std::true_type{}); break;
          }
        }
        if constexpr (Op::NEEDS_CELL) {
          cv = cvn;
          if (p + 2 < nstep)
            cvn = op.cell_value((long long)cidn);
          if (p + 3 < nstep)
            cidn = B2F_CELL_AT(cl + p + 3);
        }
      }
      // final window: every set but the free one of the last step
      int fl;
      if (wc & FAN_START)
        fl = 0;
      else {
        const int code = (int)((wc >> 10) & 7u);
        fl = ((code >> 1) + 1 + (code & 1)) % 3;
      }
#pragma unroll
      for (int q = 0; q < 3; ++q)
        if (q != fl)
          store_block<BS, PER_BLOCK>(acc, k, sl[q], av[q]);
    }
    else {
    // two register sets P, Q alternate between "newest window vertex" and "entering vertex"; slot 0 (the diagonal) = empty
    double eP[3], eQ[3], eO[3], m[3], accP[BS2], accQ[BS2], accO[BS2];
    int sP = 0, sQ = 0, sO = 0;
    double Xn[3]; // coordinates of the entering vertex, prefetched one step ahead
    uint32_t w = st[0];
    load_xyz(CV, cols[w & 1023u], Xn);
    // AA: the address chain of the gather (step word -> column id -> coordinates) is spread over three steps: the word is read
    // three steps ahead, the column id two, so every load of a step starts from registers that were filled a step ago.  Two
    // registers; the instantiations that have none to spare (reorder mode, fused source term) keep the short chain.
    constexpr bool AA = B2F_FAN_ADDR_AHEAD && !PERM && !SRC && (!Op::NEEDS_CELL || B2F_FAN_AA_CELL);
    // D2 (HeatTet at 16 warps per SM, ElasticityTet at 5: registers to spare): the coordinates are gathered TWO steps ahead into two register sets that
    // alternate with the halves of the loop (Xn: even steps, Xo: odd steps); the words then run four steps ahead, the column id three.
    constexpr bool D2 = AA && B2F_FAN_DEPTH2 && (Op::NEEDS_CELL || (B2F_FAN_DEPTH2_POST && POST)) && NV == 4;
    uint32_t w1 = 0u, w2 = 0u, w3 = 0u;
    int32_t cn = 0;
    double Xo[3] = { 0.0, 0.0, 0.0 };
    if constexpr (D2) {
      w1 = nstep > 1 ? st[1] : 0u;
      w2 = nstep > 2 ? st[2] : 0u;
      w3 = nstep > 3 ? st[3] : 0u;
      if (nstep > 1)
        load_xyz(CV, cols[w1 & 1023u], Xo);
      cn = nstep > 2 ? cols[w2 & 1023u] : 0;
    }
    else if constexpr (AA) {
      w1 = nstep > 1 ? st[1] : 0u;
      w2 = nstep > 2 ? st[2] : 0u;
      cn = nstep > 1 ? cols[w1 & 1023u] : 0;
    }
    int p = 0;
    bool newest_is_Q = false;
    // The prefetched coordinates are consumed (edge vector) BEFORE the next prefetch is issued into the same registers:
    // the consumer then never waits on a scoreboard that the new loads have just re-armed.
    // per-cell datum of the operator (heat conductivity): cell id fetched two steps ahead, its value one step ahead
    double cv = 0.0, cvn = 0.0;
    uint32_t cidn = 0;
#if B2F_FAN_CELL_DEPTH2
    // HeatTet (registers to spare): the per-cell value runs two steps ahead (cv <- cvn <- cvn2), the cell id three
    constexpr bool CD2 = Op::NEEDS_CELL && NV == 4;
    double cvn2 = 0.0;
#else
    constexpr bool CD2 = false;
#endif
    if constexpr (Op::NEEDS_CELL) {
      cv = op.cell_value((long long)B2F_CELL_AT(cl));
      if (nstep > 1)
        cvn = op.cell_value((long long)B2F_CELL_AT(cl + 1));
      if constexpr (CD2) {
#if B2F_FAN_CELL_DEPTH2
        if (nstep > 2)
          cvn2 = op.cell_value((long long)B2F_CELL_AT(cl + 2));
        if (nstep > 3)
          cidn = B2F_CELL_AT(cl + 3);
#endif
      }
      else if (nstep > 2)
        cidn = B2F_CELL_AT(cl + 2);
    }
    auto next_cell = [&](int pp) { // called after step pp
      if constexpr (CD2) {
#if B2F_FAN_CELL_DEPTH2
        cv = cvn;
        cvn = cvn2;
        if (pp + 3 < nstep)
          cvn2 = op.cell_value((long long)cidn);
        if (pp + 4 < nstep)
          cidn = B2F_CELL_AT(cl + pp + 4);
#endif
      }
      else if constexpr (Op::NEEDS_CELL) {
        cv = cvn;
        if (pp + 2 < nstep)
          cvn = op.cell_value((long long)cidn);
        if (pp + 3 < nstep)
          cidn = B2F_CELL_AT(cl + pp + 3);
      }
    };
    auto enter = [&](double(&e)[3]) {
#pragma unroll
      for (int d = 0; d < 3; ++d)
        e[d] = (NV == 4) ? Xn[d] - X0[d] : Xn[d];
    };
#if B2F_FAN_WORD_AHEAD
    // the step word is fetched two steps ahead, so the address chain of the coordinate prefetch (step word -> column id ->
    // coordinates) starts from a register: one shared-memory round trip less in front of every gather
    uint32_t w2 = nstep > 1 ? st[1] : 0u;
#define B2F_NEXT_WORD(pp)      \
  w = w2;                      \
  if ((pp) + 2 < nstep)        \
    w2 = st[(pp) + 2];
#else
#define B2F_NEXT_WORD(pp) w = st[(pp) + 1];
#endif
    while (true) {
      // step p: P enters (Q is the newest)
      uint32_t wc = w;
      enter(eP);
      if constexpr (D2) {
        if (p + 2 < nstep)
          load_xyz(CV, cn, Xn);
        if (p + 3 < nstep)
          cn = cols[w3 & 1023u];
        w = w1;
        w1 = w2;
        w2 = w3;
        if (p + 4 < nstep)
          w3 = st[p + 4];
      }
      else if constexpr (AA) {
        if (p + 1 < nstep)
          load_xyz(CV, cn, Xn);
        if (p + 2 < nstep)
          cn = cols[w2 & 1023u];
        w = w1;
        w1 = w2;
        if (p + 3 < nstep)
          w2 = st[p + 3];
      }
      else if (p + 1 < nstep) {
        B2F_NEXT_WORD(p)
        load_xyz(CV, cols[w & 1023u], Xn);
      }
#if B2F_FAN_CELL_EARLY
      if (p > 0)
        next_cell(p - 1); // the per-cell value of this step moves in, the next one is requested together with the coordinates above
#endif
      fan_step<Op, PER_BLOCK, DIAG_BY_ROW_SUM || POST, SRC>(op, CV, X0, acc, cols, k, wc, fmask, cv, eQ, accQ, sQ, eO, accO, sO, m, eP, accP, sP, diag, measure);
#if !B2F_FAN_CELL_EARLY
      next_cell(p);
#endif
      newest_is_Q = false;
      if (++p >= nstep)
        break;
      // step p: Q enters (P is the newest)
      wc = w;
      if constexpr (D2) {
#pragma unroll
        for (int d = 0; d < 3; ++d)
          eQ[d] = Xo[d] - X0[d];
        if (p + 2 < nstep)
          load_xyz(CV, cn, Xo);
        if (p + 3 < nstep)
          cn = cols[w3 & 1023u];
        w = w1;
        w1 = w2;
        w2 = w3;
        if (p + 4 < nstep)
          w3 = st[p + 4];
      }
      else
        enter(eQ);
      if constexpr (D2) {
      }
      else if constexpr (AA) {
        if (p + 1 < nstep)
          load_xyz(CV, cn, Xn);
        if (p + 2 < nstep)
          cn = cols[w2 & 1023u];
        w = w1;
        w1 = w2;
        if (p + 3 < nstep)
          w2 = st[p + 3];
      }
      else if (p + 1 < nstep) {
        B2F_NEXT_WORD(p)
        load_xyz(CV, cols[w & 1023u], Xn);
      }
#if B2F_FAN_CELL_EARLY
      next_cell(p - 1);
#endif
      fan_step<Op, PER_BLOCK, DIAG_BY_ROW_SUM || POST, SRC>(op, CV, X0, acc, cols, k, wc, fmask, cv, eP, accP, sP, eO, accO, sO, m, eQ, accQ, sQ, diag, measure);
#if !B2F_FAN_CELL_EARLY
      next_cell(p);
#endif
      newest_is_Q = true;
      if (++p >= nstep)
        break;
    }
#undef B2F_NEXT_WORD
    if (newest_is_Q) {
      if (sQ)
        store_block<BS, PER_BLOCK>(acc, k, sQ, accQ);
    }
    else {
      if (sP)
        store_block<BS, PER_BLOCK>(acc, k, sP, accP);
    }
    if constexpr (NV == 4) {
      if (sO)
        store_block<BS, PER_BLOCK>(acc, k, sO, accO);
    }
    } // select walk
    if constexpr (POST) {
#pragma unroll
      for (int q = 0; q < BS2; ++q)
        diag[q] = 0.0;
      for (int x = 1; x < k; ++x) {
        double a[BS2];
        load_block<BS, PER_BLOCK>(acc, k, x, a);
#pragma unroll
        for (int q = 0; q < BS2; ++q)
          diag[q] -= a[q];
        op.post(a);
        store_block<BS, PER_BLOCK>(acc, k, x, a);
      }
      op.post(diag);
    }
    else if constexpr (DIAG_BY_ROW_SUM) {
      double sum = 0.0;
      for (int x = 1; x < k; ++x)
        sum += acc[x];
      if constexpr (row_sum_mass<Op>::value)
        diag[0] = fma((double)(NV + 1), diag[0], -sum);
      else
        diag[0] = -sum;
    }
    store_block<BS, PER_BLOCK>(acc, k, B2F_DIAG_SLOT, diag);
#undef B2F_DIAG_SLOT
    if constexpr (SRC)
      A.rhs[out_row * BS] += A.rhs_factor * measure;
  }

  __syncthreads();
  if constexpr (PERM) {
    // reorder mode: one thread per row writes the row's blocks where the caller's row_index and column order put them
    // (block x of the internal row is block slot_map[x] of the caller's row).  A warp's stores are strided by the row length -
    // uncoalesced, but 3 instructions per block; the cooperative half-warp copy that coalesces them costs ~8x the instructions
    // in a kernel that is bound by instruction issue (measured: +30 % against +10 %).
    const bool add_old_p = POST && A.accumulate;
    // the tile's slot map goes through shared memory (one coalesced pass; the column ids are dead after the walk, their
    // staging area is reused) instead of a dependent global load per block
    uint16_t* map_s = reinterpret_cast<uint16_t*>(cols_s);
    for (int q = t; q < vend - vbeg; q += C::ROWS)
      map_s[q] = __ldg(A.slot_map + vbeg + q);
    __syncthreads();
    if (t < nrows) {
      const double* src = acc_s + (size_t)(rb - vbeg) * BS2;
      for (int x = 0; x < k; ++x) {
        const int xs = (int)map_s[rb - vbeg + x];
#pragma unroll
        for (int i = 0; i < BS; ++i)
#pragma unroll
          for (int jj = 0; jj < BS; ++jj) {
            const long long d = out_base + value_index<BS, PER_BLOCK>(0, k, xs, i, jj);
            A.values[d] = src[value_index<BS, PER_BLOCK>(0, k, x, i, jj)] + (add_old_p ? A.values[d] : 0.0);
          }
      }
    }
    __syncthreads();
    continue;
  }
  const long long base = (long long)vbeg * BS2;
  const int head = (int)(base & 1); // values is 16-byte aligned at element 0
  const bool add_old = POST && A.accumulate;
  if (t < head && t < nval)
    A.values[base + t] = acc_s[t] + (add_old ? A.values[base + t] : 0.0);
  const int nvec = (nval - head) >> 1;
  double2* dst = reinterpret_cast<double2*>(A.values + base + head);
  if (add_old) {
    for (int q = t; q < nvec; q += C::ROWS) {
      const double2 o = dst[q];
      dst[q] = make_double2(o.x + acc_s[head + 2 * q], o.y + acc_s[head + 2 * q + 1]);
    }
  }
#if B2F_FAN_BULK_STORE
  else {
    // generic-proxy writes of the walk (collected by the __syncthreads above) -> async-proxy read of the bulk copy
    if (t == 0 && nvec > 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      tma_store_1d(dst, acc_s + head, (uint32_t)nvec * 16u);
    }
  }
  if (false) {
#else
  else {
#endif
    for (int q = t; q < nvec; q += C::ROWS) {
#if B2F_FAN_STREAM >= 1
      __stcs(dst + q, make_double2(acc_s[head + 2 * q], acc_s[head + 2 * q + 1]));
#else
      dst[q] = make_double2(acc_s[head + 2 * q], acc_s[head + 2 * q + 1]);
#endif
    }
  }
  const int tail = head + 2 * nvec;
  if (t == 0 && tail < nval)
    A.values[base + tail] = acc_s[tail] + (add_old ? A.values[base + tail] : 0.0);
#if B2F_FAN_BULK_STORE
  if (t == 0 && !add_old && nvec > 0)
    tma_store_wait_read(); // the slab must stay in place until the copy has read it (the CTA's shared memory is released at exit)
#endif
#if B2F_FAN_PERSIST
  __syncthreads(); // the next tile of a persistent CTA clears the accumulators and overwrites the staging areas
  } // tiles of this CTA
#else
  } while (0);
#endif
}

// ---------------------------------------------------------------------------------------------------
// Launch
// ---------------------------------------------------------------------------------------------------
// Operators that only offer prepare()/block() (no finish()/prepare_oriented() for the fan walk) declare
// `static constexpr bool NO_FAN = true`: the node gather then runs on the slot-word / search kernels.
template <class Op, class = void>
struct uses_fan : std::true_type
{};
template <class Op>
struct uses_fan<Op, std::enable_if_t<Op::NO_FAN>> : std::false_type
{};

// Operators whose callback is addressed by the cell id alone and answers in the cell's OWN vertex order (the reference's
// `FixedMatrix f(CellLocalId)`, which gathers through its own connectivity view) declare `static constexpr bool OWN_ORDER =
// true`: the node gather hands cells over in a permuted order, so such operators always run on the cell-wise kernel
// (the reference's own algorithm, femutils/BSRFormat.h:786-898), whatever variant the handle was created with.
template <class Op, class = void>
struct own_order : std::false_type
{};
template <class Op>
struct own_order<Op, std::enable_if_t<Op::OWN_ORDER>> : std::true_type
{};

template <class Op, bool PER_BLOCK>
inline cudaError_t launch_assemble_layout(const AssembleArgs& A, const Op& op, int variant, long long nb_values, cudaStream_t stream)
{
  if constexpr (own_order<Op>::value) {
    if (A.first_tile != 0 || A.nb_tile >= 0)
      return cudaErrorInvalidValue;
    variant = VARIANT_CELL_ATOMIC;
  }
  if constexpr (uses_fan<Op>::value)
  if (variant == VARIANT_NODE_GATHER && A.fan_ok) {
    using C = FanCfg<Op>;
    constexpr bool CAN_ROW_SUM = (Op::ROW_SUM_ZERO || row_sum_mass<Op>::value) && Op::BS == 1;
    const bool row_sum = CAN_ROW_SUM && !A.accumulate;
    constexpr bool LS = Op::NV == 4; // the lockstep walk exists for tets
    const bool ls = LS && A.fan_lockstep;
    auto kern = A.rhs ? (row_sum ? (ls ? k_node_fan<Op, PER_BLOCK, CAN_ROW_SUM, true, LS> : k_node_fan<Op, PER_BLOCK, CAN_ROW_SUM, true, false>)
                                 : (ls ? k_node_fan<Op, PER_BLOCK, false, true, LS> : k_node_fan<Op, PER_BLOCK, false, true, false>))
                      : (row_sum ? (ls ? k_node_fan<Op, PER_BLOCK, CAN_ROW_SUM, false, LS> : k_node_fan<Op, PER_BLOCK, CAN_ROW_SUM, false, false>)
                                 : (ls ? k_node_fan<Op, PER_BLOCK, false, false, LS> : k_node_fan<Op, PER_BLOCK, false, false, false>));
    if (A.r_perm != nullptr) // reorder mode: the select walk only (the plan is never converted to the lockstep form there)
      kern = A.rhs ? (row_sum ? k_node_fan<Op, PER_BLOCK, CAN_ROW_SUM, true, false, true> : k_node_fan<Op, PER_BLOCK, false, true, false, true>)
                   : (row_sum ? k_node_fan<Op, PER_BLOCK, CAN_ROW_SUM, false, false, true> : k_node_fan<Op, PER_BLOCK, false, false, false, true>);
    if (B2F_FAN_CARVEOUT >= 0) {
      cudaError_t ec = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, B2F_FAN_CARVEOUT);
      if (ec != cudaSuccess)
        return ec;
    }
    constexpr int lvl = (C::ROWS == 128) ? 0 : (C::ROWS == 64 ? 1 : 2);
    const int tb = A.tile_extents[2 * lvl], ta = A.tile_extents[2 * lvl + 1];
    int acc_blocks = tb < C::ROWS * C::MAX_BLOCKS_PER_ROW ? tb : C::ROWS * C::MAX_BLOCKS_PER_ROW;
    int step_cap = ta < C::ROWS * C::MAX_ADJ_PER_ROW ? ta : C::ROWS * C::MAX_ADJ_PER_ROW;
    if (const char* cap = getenv("B2F_FAN_TILE_CAP_BLOCKS")) { // test hook: a small tile capacity sends ordinary tiles down the global fallback path
      const int c = atoi(cap);
      if (c > 0 && c < acc_blocks)
        acc_blocks = c;
    }
    acc_blocks = (acc_blocks + 3) & ~3;
    step_cap = (step_cap + 3) & ~3;
    const size_t smem = C::smem_bytes(acc_blocks, step_cap) + ((B2F_FAN_STAGE_CELLS && Op::NEEDS_CELL) ? (size_t)(step_cap + C::PAD) * sizeof(uint32_t) : 0);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess)
      return e;
    const long long all = (A.nb_node + C::ROWS - 1) / C::ROWS;
    long long grid = all - A.first_tile;
    if (A.nb_tile >= 0 && A.nb_tile < grid)
      grid = A.nb_tile;
    if (grid > 0) {
      long long ctas = grid;
      AssembleArgs F = A; // the fan kernel's view: in the reorder mode the internal copies take the place of the caller's arrays
      if (A.r_perm != nullptr) {
        F.row_perm = A.r_perm;
        F.out_row_index = A.row_index;
        F.caller_columns = A.columns;
        F.caller_coords = A.coords;
        F.row_index = A.r_row_index;
        F.columns = A.r_columns;
        F.coords = A.r_coords;
        F.slot_map = A.r_slot_map;
      }
#if B2F_FAN_PERSIST
      // one CTA per resident slot (B2F_FAN_PERSIST = waves of such CTAs; 1 = exactly the resident set)
      static int sm_count = 0;
      if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
      }
      int per_sm = 0;
      cudaError_t eo = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, C::ROWS, smem);
      if (eo != cudaSuccess)
        return eo;
      const long long resident = (long long)sm_count * (per_sm > 0 ? per_sm : 1) * B2F_FAN_PERSIST;
      if (resident < ctas)
        ctas = resident;
#endif
      kern<<<(unsigned)ctas, C::ROWS, smem, stream>>>(F, op, acc_blocks, step_cap, grid);
    }
    return cudaGetLastError();
  }
  if (A.first_tile != 0 || A.nb_tile >= 0)
    return cudaErrorInvalidValue; // only the fan kernel takes tile ranges
  if (variant == VARIANT_NODE_GATHER_SLOTS && A.slots_ok) {
    using C = SlotsCfg<Op>;
    auto kern = k_node_gather_slots<Op, PER_BLOCK>;
    // shared-memory tile sized to what this mesh needs (largest tile of the plan), capped: oversized tiles take the
    // global path, so a few very long rows cannot lower the occupancy of every other tile
    constexpr int lvl = (C::ROWS == 128) ? 0 : (C::ROWS == 64 ? 1 : 2);
    const int tb = A.tile_extents[2 * lvl], ta = A.tile_extents[2 * lvl + 1];
    int acc_blocks = tb < C::ROWS * C::MAX_BLOCKS_PER_ROW ? tb : C::ROWS * C::MAX_BLOCKS_PER_ROW;
    int adj_cap = ta < C::ROWS * C::MAX_ADJ_PER_ROW ? ta : C::ROWS * C::MAX_ADJ_PER_ROW;
    acc_blocks = (acc_blocks + 3) & ~3;
    adj_cap = (adj_cap + 3) & ~3;
    const size_t smem = C::smem_bytes(acc_blocks, adj_cap);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess)
      return e;
    const long long grid = (A.nb_node + C::ROWS - 1) / C::ROWS;
    if (grid > 0)
      kern<<<(unsigned)grid, C::ROWS, smem, stream>>>(A, op, acc_blocks, adj_cap);
    return cudaGetLastError();
  }
  if (variant == VARIANT_NODE_GATHER || variant == VARIANT_NODE_GATHER_SLOTS) {
    using C = GatherCfg<Op>;
    auto kern = k_node_gather<Op, PER_BLOCK>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES);
    if (e != cudaSuccess)
      return e;
    const long long grid = (A.nb_node + C::ROWS - 1) / C::ROWS;
    if (grid > 0)
      kern<<<(unsigned)grid, C::ROWS, C::SMEM_BYTES, stream>>>(A, op);
    return cudaGetLastError();
  }
  if (!A.accumulate) {
    cudaError_t e = cudaMemsetAsync(A.values, 0, sizeof(double) * (size_t)nb_values, stream);
    if (e != cudaSuccess)
      return e;
  }
  const int threads = 256;
  const long long grid = (A.nb_cell + threads - 1) / threads;
  if (grid > 0) {
    if (variant == VARIANT_CELL_ATOMIC_WARPAGG)
      k_cell_atomic<Op, PER_BLOCK, true><<<(unsigned)grid, threads, 0, stream>>>(A, op);
    else
      k_cell_atomic<Op, PER_BLOCK, false><<<(unsigned)grid, threads, 0, stream>>>(A, op);
  }
  return cudaGetLastError();
}

template <class Op>
inline cudaError_t launch_assemble(const AssembleArgs& A, const Op& op, int variant, bool per_block, long long nb_values, cudaStream_t stream)
{
  if (per_block)
    return launch_assemble_layout<Op, true>(A, op, variant, nb_values, stream);
  return launch_assemble_layout<Op, false>(A, op, variant, nb_values, stream);
}

} // namespace b2f
