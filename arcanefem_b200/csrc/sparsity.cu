// sparsity.cu — BSR sparsity build on device (sm_100a) + node->cell lists.
//
// Replaces BSRFormat::computeSparsityAtomic and its helpers (femutils/BSRFormat.h:602-770):
//   K1 computeSortedEdges :602-651   cell -> packed edge keys, device radix sort
//   K3 computeNeighbors   :656-672   degree count on the last key of every run
//   K4 computeRowIndex    :677-688   exclusive scan (here with the closing entry, nb_node+1 values)
//   K5 computeColumns     :703-744   diagonal first, then neighbours
//   K6 computeNzPerRowArray :435-474
// and the host-side edge count of BSRFormat::computeNbEdge (:370-385), which falls out of the scan.
//
// Differences by design (results identical to the reference executed in key order):
//   * keys are packed as (min << nbits) | max with nbits = bit width of nb_node-1, so the radix sort runs over
//     2*nbits significant bits instead of 64 (same order as the reference's (min<<32)|max keys);
//   * the reference fills neighbour slots in atomic-arrival order, which is non-deterministic on a GPU; here every
//     row is then sorted (warp rank sort), giving the canonical order "diagonal, then ascending ids";
//   * all sizes are 64-bit internally; Int32 overflow of a per-rank count is reported, not wrapped.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <utility>
#include <cub/cub.cuh>
#include "internal.h"

namespace b2f {

namespace {

  __device__ __forceinline__ unsigned long long pack_key(int32_t n0, int32_t n1, int shift)
  {
    const unsigned mn = (unsigned)(n0 > n1 ? n1 : n0);
    const unsigned mx = (unsigned)(n0 > n1 ? n0 : n1);
    return ((unsigned long long)mn << shift) | (unsigned long long)mx;
  }

  // K1: one thread per cell (femutils/BSRFormat.h:616-643; edge order (01)(02)(12) / (01)(02)(03)(12)(13)(23)).
  template <int NV>
  __global__ void __launch_bounds__(256) k_expand_edges(const int32_t* __restrict__ cell_node, long long nb_cell, long long nb_node, int shift,
                                                        unsigned long long* __restrict__ keys, unsigned long long* __restrict__ counters)
  {
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= nb_cell)
      return;
    int32_t n[NV];
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
    bool bad = false;
#pragma unroll
    for (int k = 0; k < NV; ++k)
      bad |= (n[k] < 0 || n[k] >= nb_node);
    if (bad) {
      atomicAdd(counters + 2, 1ull);
#pragma unroll
      for (int k = 0; k < NV; ++k)
        n[k] = 0;
    }
    constexpr int EPE = (NV == 3) ? 3 : 6;
    unsigned long long* e = keys + cell * EPE;
    if (NV == 3) {
      e[0] = pack_key(n[0], n[1], shift);
      e[1] = pack_key(n[0], n[2], shift);
      e[2] = pack_key(n[1], n[2], shift);
    }
    else {
      e[0] = pack_key(n[0], n[1], shift);
      e[1] = pack_key(n[0], n[2], shift);
      e[2] = pack_key(n[0], n[NV - 1], shift);
      e[3] = pack_key(n[1], n[2], shift);
      e[4] = pack_key(n[1], n[NV - 1], shift);
      e[EPE - 1] = pack_key(n[2], n[NV - 1], shift);
    }
  }

  __global__ void k_fill_i32(int32_t* p, long long n, int32_t v)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
      p[i] = v;
  }

  // K3: degree count, one thread per sorted key, only the last key of a run counts (BSRFormat.h:661-671).
  __global__ void __launch_bounds__(256) k_count_degrees(const unsigned long long* __restrict__ keys, long long n, int shift, int32_t* __restrict__ deg)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
      return;
    const unsigned long long cur = keys[i];
    if (i == n - 1 || cur != keys[i + 1]) {
      const unsigned long long mask = (1ull << shift) - 1ull;
      atomicAdd(deg + (int32_t)(cur >> shift), 1);
      atomicAdd(deg + (int32_t)(cur & mask), 1);
    }
  }

  // K5: diagonal (BSRFormat.h:712-717) ...
  __global__ void k_fill_diagonal(const int32_t* __restrict__ row_index, long long nb_node, int32_t* __restrict__ columns, int diag_last)
  {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n < nb_node)
      columns[diag_last ? row_index[n + 1] - 1 : row_index[n]] = (int32_t)n;
  }

  // ... then both directions of every unique edge (registerEdgeInColumns, BSRFormat.h:693-698,732-742).
  __global__ void __launch_bounds__(256) k_fill_neighbours(const unsigned long long* __restrict__ keys, long long n, int shift, const int32_t* __restrict__ row_index,
                                                           int32_t* __restrict__ cursor, int32_t* __restrict__ columns)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
      return;
    const unsigned long long cur = keys[i];
    if (i == n - 1 || cur != keys[i + 1]) {
      const unsigned long long mask = (1ull << shift) - 1ull;
      const int32_t n0 = (int32_t)(cur >> shift), n1 = (int32_t)(cur & mask);
      columns[row_index[n0] + atomicAdd(cursor + n0, 1)] = n1;
      columns[row_index[n1] + atomicAdd(cursor + n1, 1)] = n0;
    }
  }

  // Sort seg[0..len) ascending, one warp per segment.  Rank sort: every element counts the elements smaller than
  // itself (ties broken by position) and is written at that rank.
  template <class T>
  __device__ __forceinline__ void warp_sort_segment(T* seg, int len, unsigned lane)
  {
    if (len <= 1)
      return;
    if (len <= 32) {
      const T v = (int)lane < len ? seg[lane] : T(0);
      int rank = 0;
      for (int j = 0; j < len; ++j) {
        const T o = __shfl_sync(0xffffffffu, v, j);
        rank += (o < v) || (o == v && j < (int)lane);
      }
      __syncwarp();
      if ((int)lane < len)
        seg[rank] = v;
      return;
    }
    constexpr int PER = 8;
    if (len <= 32 * PER) {
      T v[PER];
      int rank[PER];
#pragma unroll
      for (int p = 0; p < PER; ++p) {
        const int idx = (int)lane + 32 * p;
        v[p] = idx < len ? seg[idx] : T(0);
        rank[p] = 0;
      }
      for (int j = 0; j < len; ++j) {
        const T o = seg[j];
#pragma unroll
        for (int p = 0; p < PER; ++p) {
          const int idx = (int)lane + 32 * p;
          rank[p] += (o < v[p]) || (o == v[p] && j < idx);
        }
      }
      __syncwarp();
#pragma unroll
      for (int p = 0; p < PER; ++p)
        if ((int)lane + 32 * p < len)
          seg[rank[p]] = v[p];
      return;
    }
    if (lane == 0) { // very long segment: serial insertion sort
      for (int i = 1; i < len; ++i) {
        const T x = seg[i];
        int j = i - 1;
        while (j >= 0 && seg[j] > x) {
          seg[j + 1] = seg[j];
          --j;
        }
        seg[j + 1] = x;
      }
    }
  }

  // Sort the neighbour part of every row (entries 1..k-1; entry 0 is the diagonal) / every node->cell list.
  template <class T>
  __global__ void __launch_bounds__(256) k_sort_segments(const int32_t* __restrict__ ptr, long long nb_seg, int skip, T* __restrict__ data, int skip_back = 0)
  {
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31u;
    if (w >= nb_seg)
      return;
    const int32_t b = ptr[w] + skip, e = ptr[w + 1] - skip_back;
    warp_sort_segment<T>(data + b, e - b, lane);
  }

  // K6 (BSRFormat.h:450-470)
  __global__ void k_nz_per_row(const int32_t* __restrict__ row_index, long long nb_node, int32_t* __restrict__ out)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nb_node)
      out[i] = row_index[i + 1] - row_index[i];
  }

  // node -> cell lists hold the cells that are assembled: ghost cells (cell_is_own == 0) only contribute to the pattern
  __global__ void __launch_bounds__(256) k_count_node_cells(const int32_t* __restrict__ cell_node, const uint8_t* __restrict__ cell_is_own, int nv, long long n_entries,
                                                            int32_t* __restrict__ cnt)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_entries && (cell_is_own == nullptr || cell_is_own[i / nv] != 0))
      atomicAdd(cnt + cell_node[i], 1);
  }

  template <int NV>
  __global__ void __launch_bounds__(256) k_fill_node_cells(const int32_t* __restrict__ cell_node, const uint8_t* __restrict__ cell_is_own, long long nb_cell,
                                                           const int32_t* __restrict__ nc_ptr, int32_t* __restrict__ cursor, uint32_t* __restrict__ nc_list)
  {
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= nb_cell || (cell_is_own != nullptr && cell_is_own[cell] == 0))
      return;
#pragma unroll
    for (int a = 0; a < NV; ++a) {
      const int32_t node = cell_node[cell * NV + a];
      nc_list[nc_ptr[node] + atomicAdd(cursor + node, 1)] = (uint32_t)(cell * 4 + a);
    }
  }

  // Plan of the slots-driven gather kernel: for every node->cell entry, the column slots of the cell's other vertices
  // in the row of that node, in the even-permuted vertex order that puts the node first (tets: n[a^k]; triangles:
  // n[(a+k)%3]).  10 bits per slot; rows longer than 1023 blocks clear the `ok` flag (search-based kernel is used).
  template <int NV>
  __global__ void __launch_bounds__(128) k_build_slots(const int32_t* __restrict__ cell_node, const int32_t* __restrict__ row_index, const int32_t* __restrict__ columns,
                                                       const int32_t* __restrict__ nc_ptr, const uint32_t* __restrict__ nc_list, long long nb_node,
                                                       uint32_t* __restrict__ nc_slots, unsigned long long* __restrict__ counters)
  {
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nb_node)
      return;
    const int32_t rb = row_index[row];
    const int k = row_index[row + 1] - rb;
    if (k > 1023) {
      atomicAdd(counters + 3, 1ull);
      return;
    }
    for (int32_t q = nc_ptr[row]; q < nc_ptr[row + 1]; ++q) {
      const uint32_t e = nc_list[q];
      const long long cell = e >> 2;
      const int a = (int)(e & 3u);
      uint32_t word = 0;
#pragma unroll
      for (int b = 1; b < NV; ++b) {
        const int src = (NV == 4) ? (a ^ b) : ((a + b) % 3);
        const int32_t c = cell_node[cell * NV + src];
        int lo = 1, hi = k;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (columns[rb + mid] < c)
            lo = mid + 1;
          else
            hi = mid;
        }
        word |= (uint32_t)(lo & 1023) << (10 * (b - 1));
      }
      nc_slots[q] = word;
    }
  }

  // Plan of the fan-walk gather kernel (k_node_fan, include/b2f/assemble.cuh).  The cells incident to a node are visited
  // along a path in which consecutive cells share a face through the node (a walk on the dual of the node's link), so
  // every step brings in ONE new vertex: one coordinate gather, one accumulator load and one accumulator flush per
  // cell instead of NV-1 of each, and one cross product carried over.  One 32-bit word per step:
  // The window is (N, O): N the vertex that entered last, O the other one (tets only).
  //   START  bit31 = 1 : (re)load the window; bits 0-9 slot of N, bits 10-19 slot of O (tets only)
  //   EMIT   bit31 = 0 : bits 0-9 slot of the entering vertex C, which becomes N; bit 10 = which window vertex leaves
  //                      afterwards (1: N, so O stays; 0: O, so the old N becomes O);
  //                      bit 11 = the vertex order (row node, N, O, C) is an odd permutation of the cell's
  // The walk is built greedily (next = unvisited face neighbour with the fewest unvisited neighbours, Warnsdorff's
  // rule), restarting with a START word when it gets stuck; nearly every row is a single chain.  Rows with more than
  // FAN_MAXC incident cells or more than 1023 blocks use one START + EMIT per cell / clear the `ok` flag.
#ifndef B2F_FAN_KEEP_PIVOT
#define B2F_FAN_KEEP_PIVOT 1
#endif
  constexpr int FAN_MAXC = 64;
  constexpr unsigned FAN_START = 0x80000000u;

  //
  // Built in one walk per row.  k_build_fan walks and leaves ONE staged word per cell at stage[nc_ptr[row] + visit rank]
  // (an exact bound, so no counting pass): bits 0-7 slot of C, bit 8 N leaves, bit 9 parity, bits 10-17 / 18-25 slots of
  // the window (N, O) a START word has to load before this cell (0 = none; slots of other vertices are >= 1), bits 26-31
  // position of the cell in the node's list.  Walked rows have at most FAN_MAXC cells, hence at most 3 FAN_MAXC + 1 = 193
  // blocks, so 8-bit slots do.  The row's word count (cells + STARTs) goes to fan_count; after the scan k_pack_fan expands
  // the staged words into the step words and cell ids at fan_ptr[row].  Rows that are not walked stage nothing (count =
  // 2 per cell) and get their words from k_pack_fan directly.
  template <int NV>
  __global__ void __launch_bounds__(128) k_build_fan(const int32_t* __restrict__ cell_node, const int32_t* __restrict__ row_index, const int32_t* __restrict__ columns,
                                                     const int32_t* __restrict__ nc_ptr, const uint32_t* __restrict__ nc_list, long long nb_node,
                                                     int32_t* __restrict__ fan_count, uint32_t* __restrict__ stage, unsigned long long* __restrict__ counters, int diag_last)
  {
    constexpr int NO = NV - 1; // other vertices of a cell
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nb_node)
      return;
    const int32_t rb = row_index[row];
    const int k = row_index[row + 1] - rb;
    const int32_t qb = nc_ptr[row];
    const int m = nc_ptr[row + 1] - qb;
    if (k > 1023) {
      atomicAdd(counters + 3, 1ull);
      fan_count[row] = 0;
      return;
    }
    if (m > FAN_MAXC) { // very high valence: no walk, every cell is its own chain (k_pack_fan writes the words)
      fan_count[row] = 2 * m;
      return;
    }
    uint32_t* out = stage + qb;
    int nout = 0, nstart = 0;
    // plan slots: diagonal 0, neighbours 1 .. k-1 ascending, whatever the storage layout (diagonal-last rows hold neighbour x at x - 1)
    auto slot_of = [&](int32_t c) {
      int lo = 1, hi = k;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (columns[rb + mid - diag_last] < c)
          lo = mid + 1;
        else
          hi = mid;
      }
      return lo;
    };
    // parity of the vertex order (row node, w0, [w1,] c) given as positions in the cell
    auto is_odd = [&](const int* pos) {
      int inv = 0;
      for (int i = 0; i < NV; ++i)
        for (int j = i + 1; j < NV; ++j)
          inv += pos[i] > pos[j];
      return inv & 1;
    };
    // The walk is made canonical in the row's own terms: the other vertices of a cell are held in ascending slot order and
    // every choice (start cell, ties between equally good next cells) goes to the smallest slot triple, never to a cell
    // id or a position in the list.  Rows with the same local topology then walk alike whatever the cell numbering is, so
    // the lanes of a warp (consecutive rows) meet "the same" neighbour at the same step and their gathers coalesce
    // (Kuhn cube with a random cell order: 1.98 ms with list-order choices).
    uint16_t v[FAN_MAXC][NO]; // slots of the other vertices, ascending
    uint8_t cpos[FAN_MAXC][NO]; // position in the cell of the vertex held in v[i][f]
    int8_t nb[FAN_MAXC][NO]; // neighbour across the face that omits vertex f (shares all the others), -1 if none
    uint8_t apos[FAN_MAXC];
    for (int i = 0; i < m; ++i) {
      const uint32_t e = nc_list[qb + i];
      const long long cell = e >> 2;
      const int a = (int)(e & 3u);
      apos[i] = (uint8_t)a;
      uint16_t sv[NO];
      uint8_t sp[NO];
      for (int f = 0; f < NO; ++f) {
        sp[f] = (uint8_t)(f < a ? f : f + 1);
        sv[f] = (uint16_t)slot_of(cell_node[cell * NV + sp[f]]);
      }
#pragma unroll
      for (int x = 0; x < NO - 1; ++x)
#pragma unroll
        for (int y = 0; y < NO - 1 - x; ++y)
          if (sv[y] > sv[y + 1]) {
            const uint16_t tv = sv[y];
            sv[y] = sv[y + 1];
            sv[y + 1] = tv;
            const uint8_t tp = sp[y];
            sp[y] = sp[y + 1];
            sp[y + 1] = tp;
          }
      for (int f = 0; f < NO; ++f) {
        v[i][f] = sv[f];
        cpos[i][f] = sp[f];
        nb[i][f] = -1;
      }
    }
    auto key = [&](int i) {
      uint32_t kk = 0;
      for (int f = 0; f < NO; ++f)
        kk = (kk << 10) | v[i][f];
      return kk;
    };
    if (k <= 64) {
      // the cell across the face of i that omits vertex f holds all the other vertices of i: intersect, per other vertex,
      // the set of cells (bit mask) that contain it.  O(m) instead of the pairwise comparison below.
      unsigned long long with[64];
      for (int x = 0; x < k; ++x)
        with[x] = 0ull;
      for (int i = 0; i < m; ++i)
        for (int f = 0; f < NO; ++f)
          with[v[i][f]] |= 1ull << i;
      for (int i = 0; i < m; ++i)
        for (int f = 0; f < NO; ++f) {
          unsigned long long mk = ~(1ull << i);
          for (int g = 0; g < NO; ++g)
            if (g != f)
              mk &= with[v[i][g]];
          nb[i][f] = mk ? (int8_t)(__ffsll((long long)mk) - 1) : (int8_t)-1;
        }
    }
    else
    for (int i = 0; i < m; ++i)
      for (int j = i + 1; j < m; ++j) {
        int common = 0, fi = -1, fj = -1; // fi: the vertex of i not shared, fj: same for j
        bool in_j[NO];
        for (int g = 0; g < NO; ++g)
          in_j[g] = false;
        for (int f = 0; f < NO; ++f) {
          bool hit = false;
          for (int g = 0; g < NO; ++g)
            if (v[i][f] == v[j][g]) {
              hit = true;
              in_j[g] = true;
            }
          if (hit)
            ++common;
          else
            fi = f;
        }
        if (common == NO - 1) {
          for (int g = 0; g < NO; ++g)
            if (!in_j[g])
              fj = g;
          nb[i][fi] = (int8_t)j;
          nb[j][fj] = (int8_t)i;
        }
      }
    unsigned long long visited = 0ull;
    auto unvisited_degree = [&](int i) {
      int d = 0;
      for (int f = 0; f < NO; ++f)
        d += (nb[i][f] >= 0 && !((visited >> nb[i][f]) & 1ull));
      return d;
    };
    // best unvisited neighbour of i: returns the omitted-vertex index f (neighbour = nb[i][f]) or -1.  keep_slot (tets, a
    // continuing step): slot of the newest window vertex N - the neighbour that omits N keeps the pivot O, and a step that
    // keeps its pivot skips the ~15 selects of the pivot change in k_node_fan (46 % of the steps changed pivot on the Kuhn
    // cube with slot-key tie-breaks, tools/fan_plan_stats.py).  Warnsdorff's rule (fewest onward moves) still decides
    // first - preferring the pivot outright left a few rows with a second chain - the pivot breaks its ties, then the slot key.
    auto best_next = [&](int i, uint32_t keep_slot) {
      int best = -1, bd = 99, bp = 9;
      for (int f = 0; f < NO; ++f) {
        const int j = nb[i][f];
        if (j < 0 || ((visited >> j) & 1ull))
          continue;
        const int d = unvisited_degree(j);
#if B2F_FAN_KEEP_PIVOT
        const int pc = (NV == 4 && keep_slot != 0u && v[i][f] == keep_slot) ? 0 : 1;
#else
        const int pc = 1;
#endif
        if (d < bd || (d == bd && (pc < bp || (pc == bp && key(j) < key(nb[i][best]))))) {
          bd = d;
          bp = pc;
          best = f;
        }
      }
      return best;
    };
    int cur = -1, nvis = 0;
    uint32_t win[2] = { 0u, 0u }; // slots currently held as N, O
    while (nvis < m) {
      bool start = false;
      if (cur < 0) {
        int bd = 99;
        for (int i = 0; i < m; ++i)
          if (!((visited >> i) & 1ull)) {
            const int d = unvisited_degree(i);
            if (d < bd || (d == bd && key(i) < key(cur))) {
              bd = d;
              cur = i;
            }
          }
        start = true;
      }
      visited |= 1ull << cur;
      ++nvis;
      int fc = -1; // vertex of cur that enters as C
      if (!start) {
        // continuation: the window must hold every vertex of cur but one (checked, so a non-manifold link cannot
        // desynchronise the walk: it just restarts)
        int inwin = 0;
        for (int f = 0; f < NO; ++f) {
          bool hit = false;
          for (int w = 0; w < NV - 2; ++w)
            hit |= (v[cur][f] == win[w]);
          if (hit)
            ++inwin;
          else
            fc = f;
        }
        if (inwin != NO - 1)
          start = true;
      }
      int fnext = best_next(cur, start ? 0u : win[0]); // vertex of cur that the next cell does not contain (dropped after this cell)
      if (!start && fnext == fc)
        fnext = -1;
      if (start) {
        // order the cell so that the dropped vertex sits in the window and C is shared with the next cell
        fc = (fnext == NO - 1) ? 0 : NO - 1;
        int w = 0;
        if (fnext >= 0)
          win[w++] = v[cur][fnext];
        for (int f = 0; f < NO; ++f)
          if (f != fc && f != fnext)
            win[w++] = v[cur][f];
        ++nstart;
      }
      const uint32_t staged_window = start ? ((win[0] << 10) | (NV == 4 ? (win[1] << 18) : 0u)) : 0u;
      const uint32_t sc = v[cur][fc];
      // positions in the cell of (row node, A, [B,] C)
      int pos[NV];
      pos[0] = apos[cur];
      for (int w = 0; w < NV - 2; ++w)
        for (int f = 0; f < NO; ++f)
          if (v[cur][f] == win[w])
            pos[1 + w] = cpos[cur][f];
      pos[NV - 1] = cpos[cur][fc];
      uint32_t drop_n = 0;
      if (fnext >= 0 && NV == 4)
        drop_n = (v[cur][fnext] == win[0]) ? 1u : 0u;
      out[nout] = sc | (drop_n << 8) | ((uint32_t)is_odd(pos) << 9) | staged_window | ((uint32_t)cur << 26);
      ++nout;
      if (NV == 4 && !drop_n)
        win[1] = win[0];
      win[0] = sc;
      cur = fnext >= 0 ? nb[cur][fnext] : -1;
    }
    fan_count[row] = nout + nstart;
  }

  // Expands the staged words of k_build_fan into the plan: one warp per 32 rows, the lanes take the cells of a row together.
  template <int NV>
  __global__ void __launch_bounds__(256) k_pack_fan(const int32_t* __restrict__ cell_node, const int32_t* __restrict__ row_index, const int32_t* __restrict__ columns,
                                                    const int32_t* __restrict__ nc_ptr, const uint32_t* __restrict__ nc_list, const int32_t* __restrict__ fan_ptr,
                                                    long long nb_node, const uint32_t* __restrict__ stage, uint32_t* __restrict__ fan_steps, uint32_t* __restrict__ fan_cell, int diag_last)
  {
    constexpr int NO = NV - 1;
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const long long r0 = w * 32;
    if (r0 >= nb_node)
      return;
    const long long mine = r0 + lane < nb_node ? r0 + lane : nb_node - 1;
    const int32_t my_dst = fan_ptr[mine];
    const int32_t my_cnt = fan_ptr[mine + 1] - my_dst;
    const int32_t my_qb = nc_ptr[mine];
    const int32_t my_m = nc_ptr[mine + 1] - my_qb;
    const int rows = nb_node - r0 < 32 ? (int)(nb_node - r0) : 32;
    for (int r = 0; r < rows; ++r) {
      const int32_t dst = __shfl_sync(0xffffffffu, my_dst, r);
      const int32_t cnt = __shfl_sync(0xffffffffu, my_cnt, r);
      const int32_t qb = __shfl_sync(0xffffffffu, my_qb, r);
      const int32_t m = __shfl_sync(0xffffffffu, my_m, r);
      if (cnt == 0)
        continue;
      if (m > FAN_MAXC) { // not walked: START + EMIT per cell, vertices in cell order
        const int32_t rb = row_index[r0 + r];
        const int k = row_index[r0 + r + 1] - rb;
        for (int q = lane; q < m; q += 32) {
          const uint32_t e = nc_list[qb + q];
          const long long cell = e >> 2;
          const int a = (int)(e & 3u);
          int pos[NV];
          uint32_t sl[NO];
          pos[0] = a;
          for (int f = 0; f < NO; ++f) {
            pos[f + 1] = f < a ? f : f + 1;
            const int32_t c = cell_node[cell * NV + pos[f + 1]];
            int lo = 1, hi = k;
            while (lo < hi) {
              const int mid = (lo + hi) >> 1;
              if (columns[rb + mid - diag_last] < c)
                lo = mid + 1;
              else
                hi = mid;
            }
            sl[f] = (uint32_t)lo;
          }
          int inv = 0;
          for (int i = 0; i < NV; ++i)
            for (int j = i + 1; j < NV; ++j)
              inv += pos[i] > pos[j];
          fan_steps[dst + 2 * q] = FAN_START | sl[0] | (NV == 4 ? (sl[1] << 10) : 0u);
          fan_cell[dst + 2 * q] = (uint32_t)cell;
          fan_steps[dst + 2 * q + 1] = sl[NO - 1] | ((uint32_t)(inv & 1) << 11);
          fan_cell[dst + 2 * q + 1] = (uint32_t)cell;
        }
        continue;
      }
      int before = 0; // START words of the row so far
      for (int base = 0; base < m; base += 32) {
        const int j = base + lane;
        const uint32_t sw = j < m ? stage[qb + j] : 0u;
        const uint32_t wn = (sw >> 10) & 255u, wo = (sw >> 18) & 255u;
        const unsigned has = __ballot_sync(0xffffffffu, wn != 0u);
        if (j < m) {
          int p = dst + j + before + __popc(has & ((1u << lane) - 1u));
          const uint32_t cell = nc_list[qb + (sw >> 26)] >> 2;
          if (wn != 0u) {
            fan_steps[p] = FAN_START | wn | (wo << 10);
            fan_cell[p] = cell;
            ++p;
          }
          fan_steps[p] = (sw & 255u) | (((sw >> 8) & 1u) << 10) | (((sw >> 9) & 1u) << 11);
          fan_cell[p] = cell;
        }
        before += __popc(has);
      }
    }
  }

  // Largest gather tile of the plan for tiles of 128, 64 and 32 consecutive rows: out = {blocks128, adj128, blocks64, adj64, blocks32, adj32}.
  __global__ void k_tile_extents(const int32_t* __restrict__ row_index, const int32_t* __restrict__ nc_ptr, long long nb_node, int* __restrict__ out)
  {
    const long long tile = (long long)blockIdx.x * blockDim.x + threadIdx.x; // tiles of 32 rows
    const long long r0 = tile * 32;
    if (r0 >= nb_node)
      return;
#pragma unroll
    for (int lvl = 0; lvl < 3; ++lvl) { // 32, 64, 128 rows
      const int rows = 32 << lvl;
      if ((tile & ((1 << lvl) - 1)) != 0)
        continue;
      const long long r1 = r0 + rows < nb_node ? r0 + rows : nb_node;
      atomicMax(out + 2 * (2 - lvl), row_index[r1] - row_index[r0]);
      atomicMax(out + 2 * (2 - lvl) + 1, nc_ptr[r1] - nc_ptr[r0]);
    }
  }

  __global__ void k_narrow_offsets(const long long* __restrict__ in, long long n, int32_t* __restrict__ out)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
      out[i] = (int32_t)in[i];
  }

  __global__ void k_widen_counts(const int32_t* __restrict__ in, long long n, long long* __restrict__ out)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
      out[i] = in[i];
    else if (i == n)
      out[i] = 0; // closing entry: the scan's last output is the total
  }

  inline unsigned grid_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

  // counts[0..n) (int32) -> offsets[0..n] (int32), total returned on the host (synchronises the stream).
  int scan_counts(b2f_context_s* h, const int32_t* counts, long long n, int32_t* offsets, long long* total_out, cudaStream_t stream)
  {
    B2F_CUDA(h->ws_i64.alloc(sizeof(long long) * (size_t)(n + 1)));
    long long* tmp = h->ws_i64.as<long long>();
    k_widen_counts<<<grid_for(n + 1, 256), 256, 0, stream>>>(counts, n, tmp);
    size_t temp_bytes = 0;
    B2F_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, tmp, tmp, n + 1, stream));
    B2F_CUDA(h->ws_tmp.alloc(temp_bytes));
    B2F_CUDA(cub::DeviceScan::ExclusiveSum(h->ws_tmp.p, temp_bytes, tmp, tmp, n + 1, stream));
    long long total = 0;
    B2F_CUDA(cudaMemcpyAsync(&total, tmp + n, sizeof(long long), cudaMemcpyDeviceToHost, stream));
    B2F_CUDA(cudaStreamSynchronize(stream));
    if (total > 2147483647ll)
      return fail(B2F_ERR_OVERFLOW, "per-rank count %lld does not fit Int32 indices (femutils/BSRFormat.h:401-402); split the mesh over more ranks", total);
    k_narrow_offsets<<<grid_for(n + 1, 256), 256, 0, stream>>>(tmp, n + 1, offsets);
    B2F_CUDA(cudaGetLastError());
    *total_out = total;
    return B2F_OK;
  }

} // namespace

// ---------------------------------------------------------------------------------------------------
// Lockstep form of the fan plan (tets).  k_build_fan writes step words for the select walk: window (N, O), bit 10 = N
// leaves, bit 11 = parity of (row, N, O, C).  The lockstep walk of k_node_fan keeps the window in two of three fixed
// register sets; its step word names the variant of the step body: bits 10-12 = 2 * f + ly, f = free set (receives C),
// ly = set f+2 leaves (else f+1), bit 13 = parity of (row, set f+1, set f+2, C); START words gain bits 20-21 = free
// set of the window they end (3 = none) and always load N into set 1, O into set 2 (free set 0).
// One thread per row replays its words.  COUNT: nothing is written; per warp and step the number of distinct variants
// among the rows still walking is added up (= step bodies the warp would execute), next to the number of warp steps.
// ---------------------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(128) k_fan_lockstep(const int32_t* __restrict__ fan_ptr, long long nb_node, uint32_t* __restrict__ steps, unsigned long long* __restrict__ cost)
{
  const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int32_t qb = 0, n = 0;
  if (row < nb_node) {
    qb = fan_ptr[row];
    n = fan_ptr[row + 1] - qb;
  }
  int nmax = n;
  if (COUNT) {
    for (int d = 16; d > 0; d >>= 1)
      nmax = max(nmax, __shfl_xor_sync(0xffffffffu, nmax, d));
  }
  int f = 3, setN = 1, setO = 2; // f = 3: no window yet
  unsigned long long bodies = 0, wsteps = 0;
  for (int p = 0; p < nmax; ++p) {
    int code = 7; // not walking
    if (p < n) {
      const uint32_t w = steps[qb + p];
      uint32_t out;
      if (w & 0x80000000u) {
        out = (w & 0x800fffffu) | ((uint32_t)f << 20);
        f = 0;
        setN = 1;
        setO = 2;
        code = 6;
      }
      else {
        const int X = (f + 1) % 3, Y = (f + 2) % 3;
        const bool dropN = (w >> 10) & 1u;
        bool odd = (w >> 11) & 1u;
        if (setN != X)
          odd = !odd; // (row, X, Y, C) lists N and O the other way round
        const int leaving = dropN ? setN : setO;
        const int ly = (leaving == Y) ? 1 : 0;
        code = 2 * f + ly;
        out = (w & 1023u) | ((uint32_t)code << 10) | ((odd ? 1u : 0u) << 13);
        const int stay = dropN ? setO : setN;
        setN = f; // C
        setO = stay;
        f = leaving;
      }
      if (!COUNT)
        steps[qb + p] = out;
    }
    if (COUNT) {
      int distinct = 0;
#pragma unroll
      for (int c = 0; c < 7; ++c)
        distinct += __any_sync(0xffffffffu, code == c) ? 1 : 0;
      bodies += distinct;
      wsteps += 1;
    }
  }
  if (COUNT && (threadIdx.x & 31) == 0 && nmax > 0) {
    atomicAdd(cost, bodies);
    atomicAdd(cost + 1, wsteps);
  }
}

// Pivot hint for the select walk on tets: bit 12 of a (non-START) step word is set when, at that step index, some row of
// the aligned group of 32 rows that a warp of k_node_fan walks together changes its pivot (O leaves, bit 10 clear).  The
// walk can then branch around the pivot change on plain data instead of voting at every step (B2F_FAN_UNIFORM_PIVOT == 2);
// kernels that vote ignore the bit.  One thread per row, the lanes of a warp replay their rows' words side by side.
__global__ void __launch_bounds__(128) k_fan_pivot_hint(const int32_t* __restrict__ fan_ptr, long long nb_node, uint32_t* __restrict__ steps)
{
  const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int32_t qb = 0, n = 0;
  if (row < nb_node) {
    qb = fan_ptr[row];
    n = fan_ptr[row + 1] - qb;
  }
  int nmax = n;
  for (int d = 16; d > 0; d >>= 1)
    nmax = max(nmax, __shfl_xor_sync(0xffffffffu, nmax, d));
  for (int p = 0; p < nmax; ++p) {
    uint32_t w = 0x80000000u;
    if (p < n)
      w = steps[qb + p];
    const bool emits = !(w & 0x80000000u);
    const unsigned changes = __ballot_sync(0xffffffffu, emits && !((w >> 10) & 1u));
    if (emits && changes != 0u)
      steps[qb + p] = w | 4096u;
  }
}

// First visits.  A step that brings a vertex into the window loads its accumulator from the tile's shared-memory slab; the first
// time a slot is loaded in its row the value is the initial zero.  The plan marks those loads (bit 13 of an emit word, bits 20 / 21
// of a START word for its two slots), so the kernel starts the register from zero instead, and - every slot of a walked row
// being stored at least once - does not have to clear the slab first: two passes less through the L1 data pipe, the unit the
// fan kernel loads most.  One thread per row, 1024-slot bitmap in local memory.
__global__ void __launch_bounds__(128) k_fan_first_flags(const int32_t* __restrict__ fan_ptr, const int32_t* __restrict__ row_index, long long nb_node,
                                                         uint32_t* __restrict__ steps)
{
  __shared__ uint32_t seen_s[8][128]; // 256 slots per thread, one column per thread (no bank conflicts)
  const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= nb_node)
    return;
  const int32_t qb = fan_ptr[row];
  const int n = fan_ptr[row + 1] - qb;
  if (n == 0)
    return;
  const int k = row_index[row + 1] - row_index[row];
  auto mark = [&](uint32_t w, bool f0, bool f1) {
    if (w & 0x80000000u)
      return w | (f0 ? 1u << 20 : 0u) | (f1 ? 1u << 21 : 0u);
    return w | (f0 ? 1u << 13 : 0u);
  };
  if (k <= 256) {
    const int t = threadIdx.x;
#pragma unroll
    for (int i = 0; i < 8; ++i)
      seen_s[i][t] = 0u;
    auto first = [&](uint32_t slot) {
      const uint32_t bit = 1u << (slot & 31u);
      const uint32_t old = seen_s[slot >> 5][t];
      seen_s[slot >> 5][t] = old | bit;
      return (old & bit) == 0u;
    };
    for (int p = 0; p < n; ++p) {
      const uint32_t w = steps[qb + p];
      const bool f0 = first(w & 1023u);
      const uint32_t so = (w & 0x80000000u) ? (w >> 10) & 1023u : 0u; // 0 for triangles (slot 0 is the diagonal, never a window vertex)
      const bool f1 = so != 0u && first(so);
      if (f0 || f1)
        steps[qb + p] = mark(w, f0, f1);
    }
    return;
  }
  // rows with more than 256 blocks (very high valence; not walked, every cell is its own chain): 1024-slot bitmap in local memory
  uint32_t seen[32];
  for (int i = 0; i < 32; ++i)
    seen[i] = 0u;
  auto first = [&](uint32_t slot) {
    const uint32_t bit = 1u << (slot & 31u);
    const bool f = (seen[slot >> 5] & bit) == 0u;
    seen[slot >> 5] |= bit;
    return f;
  };
  for (int p = 0; p < n; ++p) {
    const uint32_t w = steps[qb + p];
    const bool f0 = first(w & 1023u);
    const uint32_t so = (w & 0x80000000u) ? (w >> 10) & 1023u : 0u;
    const bool f1 = so != 0u && first(so);
    if (f0 || f1)
      steps[qb + p] = mark(w, f0, f1);
  }
}

namespace {
  // Cells with a node id outside [0, nb_node): counted before anything indexes with them.
  __global__ void __launch_bounds__(256) k_validate_cells(const int32_t* __restrict__ cell_node, long long n_entries, long long nb_node, unsigned long long* __restrict__ counters)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    bool bad = false;
    if (i < n_entries) {
      const int32_t v = cell_node[i];
      bad = v < 0 || v >= nb_node;
    }
    if (__any_sync(0xffffffffu, bad)) {
      const unsigned m = __ballot_sync(0xffffffffu, bad);
      if ((threadIdx.x & 31) == 0)
        atomicAdd(counters + 2, (unsigned long long)__popc(m));
    }
  }

  // Row-wise pattern build: one thread per row walks the node's cell list and keeps the sorted set of the other vertices
  // it meets (shared memory, one column of RU_CAP entries per thread; the cells of a list come in ascending cell order,
  // so most insertions land at or near the end).  Pass 1 (WRITE = false) counts, pass 2 repeats the walk and writes the
  // row: diagonal first, then ascending - the canonical order of the sort-based build, with no global sort.
  // Same result as K1-K5 (femutils/BSRFormat.h:602-744): two nodes are neighbours iff they share a cell (P1 cells).
  constexpr int RU_ROWS = 128, RU_CAP = 64;
  template <int NV, bool WRITE>
  __global__ void __launch_bounds__(RU_ROWS) k_row_unique(const int32_t* __restrict__ cell_node, const int32_t* __restrict__ nc_ptr, const uint32_t* __restrict__ nc_list,
                                                          long long nb_node, const int32_t* __restrict__ row_index, int32_t* __restrict__ deg,
                                                          int32_t* __restrict__ columns, unsigned long long* __restrict__ counters, int diag_last)
  {
    __shared__ int32_t set[RU_CAP][RU_ROWS];
    const int t = threadIdx.x;
    const long long row = (long long)blockIdx.x * RU_ROWS + t;
    if (row >= nb_node)
      return;
    int cnt = 0;
    bool overflow = false;
    const int32_t qe = nc_ptr[row + 1];
    for (int32_t q = nc_ptr[row]; q < qe; ++q) {
      const long long cell = __ldg(nc_list + q) >> 2;
      int32_t n[NV];
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
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int32_t c = n[k];
        if (c == (int32_t)row)
          continue;
        int i = cnt;
        while (i > 0 && set[i - 1][t] > c)
          --i;
        if (i > 0 && set[i - 1][t] == c)
          continue;
        if (cnt == RU_CAP) {
          overflow = true;
          continue;
        }
        for (int j = cnt; j > i; --j)
          set[j][t] = set[j - 1][t];
        set[i][t] = c;
        ++cnt;
      }
    }
    if (!WRITE) {
      deg[row] = cnt + 1;
      if (overflow)
        atomicAdd(counters + 3, 1ull);
    }
    else {
      const int32_t rb = row_index[row];
      columns[rb + (diag_last ? cnt : 0)] = (int32_t)row;
      for (int i = 0; i < cnt; ++i)
        columns[rb + 1 - diag_last + i] = set[i][t];
    }
  }
} // namespace

// K1-K6 from the node->cell lists (built first): no edge keys, no global sort.  *fallback is set when a row has more than
// RU_CAP neighbours: the caller then runs the sort-based build.
static int pattern_from_lists(b2f_context_s* h, cudaStream_t stream, bool* fallback)
{
  const long long nn = h->mesh.nb_node;
  const int nv = h->nv;
  unsigned long long* counters = h->counters.as<unsigned long long>();
  *fallback = false;
  B2F_CUDA(cudaMemsetAsync(counters + 3, 0, sizeof(unsigned long long), stream));
  B2F_CUDA(h->ws_i32a.alloc(sizeof(int32_t) * (size_t)(nn + 1)));
  int32_t* deg = h->ws_i32a.as<int32_t>();
  const int32_t* ncp = h->nc_ptr.as<int32_t>();
  const uint32_t* ncl = h->nc_list.as<uint32_t>();
  if (nn > 0) {
    if (nv == 3)
      k_row_unique<3, false><<<grid_for(nn, RU_ROWS), RU_ROWS, 0, stream>>>(h->mesh.cell_node, ncp, ncl, nn, nullptr, deg, nullptr, counters, h->diag_last);
    else
      k_row_unique<4, false><<<grid_for(nn, RU_ROWS), RU_ROWS, 0, stream>>>(h->mesh.cell_node, ncp, ncl, nn, nullptr, deg, nullptr, counters, h->diag_last);
    B2F_CUDA(cudaGetLastError());
  }
  B2F_CUDA(h->row_index.alloc(sizeof(int32_t) * (size_t)(nn + 1)));
  long long nb_col = 0;
  int rc = scan_counts(h, deg, nn, h->row_index.as<int32_t>(), &nb_col, stream);
  if (rc != B2F_OK)
    return rc;
  unsigned long long long_rows = 0;
  B2F_CUDA(cudaMemcpyAsync(&long_rows, counters + 3, sizeof(long_rows), cudaMemcpyDeviceToHost, stream));
  B2F_CUDA(cudaStreamSynchronize(stream));
  if (long_rows != 0) {
    *fallback = true;
    return B2F_OK;
  }
  const long long nb_nz = nb_col * h->bs * h->bs;
  if (nb_nz > 2147483647ll)
    return fail(B2F_ERR_OVERFLOW, "nb_non_zero_value=%lld does not fit the reference's Int32 (femutils/BSRFormat.h:402)", nb_nz);
  h->nb_col = nb_col;
  h->nb_nz = nb_nz;
  h->nb_edge = (nb_col - nn) / 2;
  B2F_CUDA(h->columns.alloc(sizeof(int32_t) * (size_t)(nb_col + 8))); // +8: bulk copies round the tile end up to 16 bytes
  if (!h->inner)
    B2F_CUDA(h->values.alloc(sizeof(double) * (size_t)nb_nz));
  if (nn > 0) {
    if (nv == 3)
      k_row_unique<3, true><<<grid_for(nn, RU_ROWS), RU_ROWS, 0, stream>>>(h->mesh.cell_node, ncp, ncl, nn, h->row_index.as<int32_t>(), nullptr, h->columns.as<int32_t>(), counters,
                                                                           h->diag_last);
    else
      k_row_unique<4, true><<<grid_for(nn, RU_ROWS), RU_ROWS, 0, stream>>>(h->mesh.cell_node, ncp, ncl, nn, h->row_index.as<int32_t>(), nullptr, h->columns.as<int32_t>(), counters,
                                                                           h->diag_last);
    B2F_CUDA(cudaGetLastError());
  }
  B2F_CUDA(h->nb_nz_per_row.alloc(sizeof(int32_t) * (size_t)nn));
  k_nz_per_row<<<grid_for(nn, 256), 256, 0, stream>>>(h->row_index.as<int32_t>(), nn, h->nb_nz_per_row.as<int32_t>());
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

// K1-K6 by a global sort of the packed edge keys (the reference's algorithm, femutils/BSRFormat.h:583-770).  Used when
// ghost cells take part in the pattern only (cell_is_own: the node->cell lists then miss them), on request
// (b2f_options.sparsity_build) and when a row has more neighbours than the row-wise build holds.
static int pattern_by_sort(b2f_context_s* h, cudaStream_t stream)
{
  const long long nc = h->mesh.nb_cell, nn = h->mesh.nb_node;
  const int nv = h->nv;
  const int epe = (nv == 3) ? 3 : 6;
  const long long nkeys = nc * epe;
  int shift = 1;
  while ((1ll << shift) < nn)
    ++shift;
  unsigned long long* counters = h->counters.as<unsigned long long>();
  // --- K1: edge keys + radix sort over the 2*shift significant bits
  B2F_CUDA(h->ws_keys_a.alloc(sizeof(unsigned long long) * (size_t)nkeys));
  B2F_CUDA(h->ws_keys_b.alloc(sizeof(unsigned long long) * (size_t)nkeys));
  unsigned long long* ka = h->ws_keys_a.as<unsigned long long>();
  unsigned long long* kb = h->ws_keys_b.as<unsigned long long>();
  if (nc > 0) {
    if (nv == 3)
      k_expand_edges<3><<<grid_for(nc, 256), 256, 0, stream>>>(h->mesh.cell_node, nc, nn, shift, ka, counters);
    else
      k_expand_edges<4><<<grid_for(nc, 256), 256, 0, stream>>>(h->mesh.cell_node, nc, nn, shift, ka, counters);
    B2F_CUDA(cudaGetLastError());
  }
  cub::DoubleBuffer<unsigned long long> keys(ka, kb);
  if (nkeys > 0) {
    size_t temp_bytes = 0;
    B2F_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, temp_bytes, keys, nkeys, 0, 2 * shift, stream));
    B2F_CUDA(h->ws_tmp.alloc(temp_bytes));
    B2F_CUDA(cub::DeviceRadixSort::SortKeys(h->ws_tmp.p, temp_bytes, keys, nkeys, 0, 2 * shift, stream));
  }
  const unsigned long long* sorted = keys.Current();

  // --- K3/K4: degrees (diagonal pre-counted) and row_index
  B2F_CUDA(h->ws_i32a.alloc(sizeof(int32_t) * (size_t)(nn + 1)));
  int32_t* deg = h->ws_i32a.as<int32_t>();
  k_fill_i32<<<grid_for(nn, 256), 256, 0, stream>>>(deg, nn, 1);
  if (nkeys > 0)
    k_count_degrees<<<grid_for(nkeys, 256), 256, 0, stream>>>(sorted, nkeys, shift, deg);
  B2F_CUDA(cudaGetLastError());
  B2F_CUDA(h->row_index.alloc(sizeof(int32_t) * (size_t)(nn + 1)));
  long long nb_col = 0;
  int rc = scan_counts(h, deg, nn, h->row_index.as<int32_t>(), &nb_col, stream);
  if (rc != B2F_OK)
    return rc;
  {
    unsigned long long host_counters[4];
    B2F_CUDA(cudaMemcpyAsync(host_counters, counters, sizeof(host_counters), cudaMemcpyDeviceToHost, stream));
    B2F_CUDA(cudaStreamSynchronize(stream));
    if (host_counters[2] != 0)
      return fail(B2F_ERR_INVALID_ARGUMENT, "cell_node holds %llu cells with node ids outside [0, nb_node)", host_counters[2]);
  }
  const long long nb_nz = nb_col * h->bs * h->bs;
  if (nb_nz > 2147483647ll)
    return fail(B2F_ERR_OVERFLOW, "nb_non_zero_value=%lld does not fit the reference's Int32 (femutils/BSRFormat.h:402)", nb_nz);
  h->nb_col = nb_col;
  h->nb_nz = nb_nz;
  h->nb_edge = (nb_col - nn) / 2;

  // --- K5: columns, then canonical order inside every row
  B2F_CUDA(h->columns.alloc(sizeof(int32_t) * (size_t)(nb_col + 8))); // +8: bulk copies round the tile end up to 16 bytes
  if (!h->inner)
    B2F_CUDA(h->values.alloc(sizeof(double) * (size_t)nb_nz));
  int32_t* cursor = deg; // reuse: offsets.fill(1), BSRFormat.h:724 (diagonal-last rows: the neighbours start at 0, :554-567)
  k_fill_i32<<<grid_for(nn, 256), 256, 0, stream>>>(cursor, nn, 1 - h->diag_last);
  k_fill_diagonal<<<grid_for(nn, 256), 256, 0, stream>>>(h->row_index.as<int32_t>(), nn, h->columns.as<int32_t>(), h->diag_last);
  if (nkeys > 0)
    k_fill_neighbours<<<grid_for(nkeys, 256), 256, 0, stream>>>(sorted, nkeys, shift, h->row_index.as<int32_t>(), cursor, h->columns.as<int32_t>());
  k_sort_segments<int32_t><<<grid_for(nn * 32, 256), 256, 0, stream>>>(h->row_index.as<int32_t>(), nn, 1 - h->diag_last, h->columns.as<int32_t>(), h->diag_last);
  B2F_CUDA(cudaGetLastError());

  // --- K6
  B2F_CUDA(h->nb_nz_per_row.alloc(sizeof(int32_t) * (size_t)nn));
  k_nz_per_row<<<grid_for(nn, 256), 256, 0, stream>>>(h->row_index.as<int32_t>(), nn, h->nb_nz_per_row.as<int32_t>());
  B2F_CUDA(cudaGetLastError());

  B2F_CUDA(cudaStreamSynchronize(stream));
  h->ws_keys_a.release();
  h->ws_keys_b.release();
  return B2F_OK;
}

// all_cells: ghost cells (cell_is_own == 0) are listed too - the lists the row-wise pattern build reads in the exchange mode;
// the lists the gather plan is built from hold the assembled cells only.
static int node_cell_lists(b2f_context_s* h, cudaStream_t stream, bool all_cells = false)
{
  const long long nc = h->mesh.nb_cell, nn = h->mesh.nb_node;
  const int nv = h->nv;
  const uint8_t* cell_mask = all_cells ? nullptr : h->mesh.cell_is_own;
  // --- node -> cell lists (the reference takes them from Arcane's nodeCell() connectivity, BSRFormat.h:942)
  B2F_CUDA(h->ws_i32a.alloc(sizeof(int32_t) * (size_t)(nn + 1)));
  int32_t* cnt = h->ws_i32a.as<int32_t>();
  int32_t* cursor = cnt;
  int rc = B2F_OK;
  B2F_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int32_t) * (size_t)(nn + 1), stream));
  if (nc > 0)
    k_count_node_cells<<<grid_for(nc * nv, 256), 256, 0, stream>>>(h->mesh.cell_node, cell_mask, nv, nc * nv, cnt);
  B2F_CUDA(cudaGetLastError());
  B2F_CUDA(h->nc_ptr.alloc(sizeof(int32_t) * (size_t)(nn + 1)));
  long long nb_adj = 0;
  rc = scan_counts(h, cnt, nn, h->nc_ptr.as<int32_t>(), &nb_adj, stream);
  if (rc != B2F_OK)
    return rc;
  h->nb_pair = nb_adj;
  B2F_CUDA(h->nc_list.alloc(sizeof(uint32_t) * (size_t)nb_adj));
  B2F_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int32_t) * (size_t)(nn + 1), stream));
  if (nc > 0) {
    if (nv == 3)
      k_fill_node_cells<3><<<grid_for(nc, 256), 256, 0, stream>>>(h->mesh.cell_node, cell_mask, nc, h->nc_ptr.as<int32_t>(), cursor, h->nc_list.as<uint32_t>());
    else
      k_fill_node_cells<4><<<grid_for(nc, 256), 256, 0, stream>>>(h->mesh.cell_node, cell_mask, nc, h->nc_ptr.as<int32_t>(), cursor, h->nc_list.as<uint32_t>());
  }
  k_sort_segments<uint32_t><<<grid_for(nn * 32, 256), 256, 0, stream>>>(h->nc_ptr.as<int32_t>(), nn, 0, h->nc_list.as<uint32_t>());
  B2F_CUDA(cudaGetLastError());

  return B2F_OK;
}

int build_pattern(b2f_context_s* h, cudaStream_t stream)
{
  const long long nc = h->mesh.nb_cell, nn = h->mesh.nb_node;
  const int nv = h->nv;
  if (nc * 4 > 4294967295ll || nc * (long long)nv > 2147483647ll)
    return fail(B2F_ERR_OVERFLOW, "nb_cell=%lld is too large for one rank (node->cell entries are 32-bit)", nc);

  h->has_pattern = false;
  h->has_csr = false;
  h->ghost.ready = false; // an exchange plan indexes the values of the pattern it was built for
  h->chunk_row.clear();
  B2F_CUDA(h->counters.alloc(sizeof(unsigned long long) * 12));
  B2F_CUDA(cudaMemsetAsync(h->counters.p, 0, sizeof(unsigned long long) * 4, stream));
  unsigned long long* counters = h->counters.as<unsigned long long>();
  int rc = B2F_OK;

  // node ids are checked before anything indexes with them
  if (nc > 0) {
    k_validate_cells<<<grid_for(nc * nv, 256), 256, 0, stream>>>(h->mesh.cell_node, nc * nv, nn, counters);
    B2F_CUDA(cudaGetLastError());
    unsigned long long bad = 0;
    B2F_CUDA(cudaMemcpyAsync(&bad, counters + 2, sizeof(bad), cudaMemcpyDeviceToHost, stream));
    B2F_CUDA(cudaStreamSynchronize(stream));
    if (bad != 0)
      return fail(B2F_ERR_INVALID_ARGUMENT, "cell_node holds %llu node ids outside [0, nb_node)", bad);
  }

  // Pattern (K1-K6) and node -> cell lists.  Default: lists first, rows built from them (no global sort).  The sort-based
  // build runs when ghost cells only take part in the pattern (the lists hold the assembled cells), on request, or when a
  // row is too long for the row-wise kernel.
  h->pattern_by_lists = false;
  const bool exchange_mode = h->mesh.cell_is_own != nullptr;
  bool lists_first = h->opt.sparsity_build != B2F_SPARSITY_SORT;
  if (lists_first) {
    // exchange mode: the pattern covers the ghost-cell layer, so its lists hold every cell; the lists of the assembled
    // cells are built afterwards (a second 6 ms pass instead of the 25 ms of the global sort)
    rc = node_cell_lists(h, stream, exchange_mode);
    if (rc != B2F_OK)
      return rc;
    bool fallback = false;
    rc = pattern_from_lists(h, stream, &fallback);
    if (rc != B2F_OK)
      return rc;
    if (fallback) {
      rc = pattern_by_sort(h, stream);
      if (rc != B2F_OK)
        return rc;
    }
    else
      h->pattern_by_lists = true;
    if (exchange_mode) {
      rc = node_cell_lists(h, stream, false);
      if (rc != B2F_OK)
        return rc;
    }
  }
  else {
    rc = pattern_by_sort(h, stream);
    if (rc != B2F_OK)
      return rc;
    rc = node_cell_lists(h, stream);
    if (rc != B2F_OK)
      return rc;
  }
  const long long nb_adj = h->nb_pair;

  // --- plan of the gather kernel: fan walk (default) or the column-slot words of the first-generation kernel
  // (b2f_reorder: the plan is built by reorder_plan on the internally renumbered connectivity instead)
  const bool want_reorder = !h->inner && h->opt.reorder != B2F_REORDER_NONE && h->variant == B2F_ASM_NODE_GATHER && h->mesh.cell_is_own == nullptr && h->mesh.coords != nullptr && nn > 0;
  const int32_t* tile_ptr = h->nc_ptr.as<int32_t>();
  B2F_CUDA(cudaMemsetAsync(counters + 3, 0, sizeof(unsigned long long), stream));
  h->fan_ok = false;
  h->slots_ok = false;
  if (h->variant == B2F_ASM_NODE_GATHER_SLOTS) {
    B2F_CUDA(h->nc_slots.alloc(sizeof(uint32_t) * (size_t)(nb_adj + 8)));
    if (nv == 3)
      k_build_slots<3><<<grid_for(nn, 128), 128, 0, stream>>>(h->mesh.cell_node, h->row_index.as<int32_t>(), h->columns.as<int32_t>(), h->nc_ptr.as<int32_t>(),
                                                              h->nc_list.as<uint32_t>(), nn, h->nc_slots.as<uint32_t>(), counters);
    else
      k_build_slots<4><<<grid_for(nn, 128), 128, 0, stream>>>(h->mesh.cell_node, h->row_index.as<int32_t>(), h->columns.as<int32_t>(), h->nc_ptr.as<int32_t>(),
                                                              h->nc_list.as<uint32_t>(), nn, h->nc_slots.as<uint32_t>(), counters);
    B2F_CUDA(cudaGetLastError());
  }
  else if (h->variant == B2F_ASM_NODE_GATHER && !want_reorder) {
    B2F_CUDA(h->ws_i32a.alloc(sizeof(int32_t) * (size_t)(nn + 1)));
    int32_t* fan_count = h->ws_i32a.as<int32_t>();
    const int32_t* ri = h->row_index.as<int32_t>();
    const int32_t* co = h->columns.as<int32_t>();
    const int32_t* ncp = h->nc_ptr.as<int32_t>();
    const uint32_t* ncl = h->nc_list.as<uint32_t>();
    // staging of the walk (one word per adjacency): the value array is idle until the first assembly and large enough on
    // any ordinary mesh, so it lends its memory (and is zero-filled afterwards, BSRFormat.h:136)
    uint32_t* stage = nullptr;
    const size_t stage_bytes = sizeof(uint32_t) * (size_t)(nb_adj + 8);
    if (h->values.bytes >= stage_bytes)
      stage = h->values.as<uint32_t>();
    else {
      B2F_CUDA(h->ws_keys_a.alloc(stage_bytes));
      stage = h->ws_keys_a.as<uint32_t>();
    }
    if (nv == 3)
      k_build_fan<3><<<grid_for(nn, 128), 128, 0, stream>>>(h->mesh.cell_node, ri, co, ncp, ncl, nn, fan_count, stage, counters, h->diag_last);
    else
      k_build_fan<4><<<grid_for(nn, 128), 128, 0, stream>>>(h->mesh.cell_node, ri, co, ncp, ncl, nn, fan_count, stage, counters, h->diag_last);
    B2F_CUDA(cudaGetLastError());
    B2F_CUDA(h->fan_ptr.alloc(sizeof(int32_t) * (size_t)(nn + 1)));
    long long nb_step = 0;
    rc = scan_counts(h, fan_count, nn, h->fan_ptr.as<int32_t>(), &nb_step, stream);
    if (rc != B2F_OK)
      return rc;
    h->nb_fan_step = nb_step;
    B2F_CUDA(h->fan_steps.alloc(sizeof(uint32_t) * (size_t)(nb_step + 8)));
    B2F_CUDA(h->fan_cell.alloc(sizeof(uint32_t) * (size_t)(nb_step + 8)));
    if (nv == 3)
      k_pack_fan<3><<<grid_for(((nn + 31) / 32) * 32, 256), 256, 0, stream>>>(h->mesh.cell_node, ri, co, ncp, ncl, h->fan_ptr.as<int32_t>(), nn, stage,
                                                                            h->fan_steps.as<uint32_t>(), h->fan_cell.as<uint32_t>(), h->diag_last);
    else
      k_pack_fan<4><<<grid_for(((nn + 31) / 32) * 32, 256), 256, 0, stream>>>(h->mesh.cell_node, ri, co, ncp, ncl, h->fan_ptr.as<int32_t>(), nn, stage,
                                                                            h->fan_steps.as<uint32_t>(), h->fan_cell.as<uint32_t>(), h->diag_last);
    B2F_CUDA(cudaGetLastError());
    if (nv == 4 && nb_step > 0) {
      k_fan_pivot_hint<<<grid_for(nn, 128), 128, 0, stream>>>(h->fan_ptr.as<int32_t>(), nn, h->fan_steps.as<uint32_t>());
      B2F_CUDA(cudaGetLastError());
    }
    h->fan_first = false;
    if (nb_step > 0 && h->opt.fan_mode != B2F_FAN_LOCKSTEP) {
      k_fan_first_flags<<<grid_for(nn, 128), 128, 0, stream>>>(h->fan_ptr.as<int32_t>(), ri, nn, h->fan_steps.as<uint32_t>());
      B2F_CUDA(cudaGetLastError());
      h->fan_first = true;
    }
    tile_ptr = h->fan_ptr.as<int32_t>();
    // lockstep form of the step words, on request only (B2F_FAN_LOCKSTEP): measured slower than the select walk even
    // where the rows of a warp walk in step (N=203 Kuhn cube: 1.07 bodies per warp step, 1.08 ms against 0.98 ms) -
    // the six-way indirect branch and the moves of the carried product cost what the selects did
    h->fan_lockstep = false;
    h->fan_lockstep_bodies = 0.0;
    if (nv == 4 && nb_step > 0 && h->opt.fan_mode == B2F_FAN_LOCKSTEP) {
      unsigned long long* cost = counters + 8;
      B2F_CUDA(cudaMemsetAsync(cost, 0, sizeof(unsigned long long) * 2, stream));
      k_fan_lockstep<true><<<grid_for(nn, 128), 128, 0, stream>>>(h->fan_ptr.as<int32_t>(), nn, h->fan_steps.as<uint32_t>(), cost);
      B2F_CUDA(cudaGetLastError());
      unsigned long long hc[2] = { 0, 0 };
      B2F_CUDA(cudaMemcpyAsync(hc, cost, sizeof(hc), cudaMemcpyDeviceToHost, stream));
      B2F_CUDA(cudaStreamSynchronize(stream));
      h->fan_lockstep_bodies = hc[1] ? (double)hc[0] / (double)hc[1] : 0.0;
      k_fan_lockstep<false><<<grid_for(nn, 128), 128, 0, stream>>>(h->fan_ptr.as<int32_t>(), nn, h->fan_steps.as<uint32_t>(), nullptr);
      B2F_CUDA(cudaGetLastError());
      h->fan_lockstep = true;
    }
  }
  {
    int* ext = reinterpret_cast<int*>(counters + 4);
    B2F_CUDA(cudaMemsetAsync(ext, 0, sizeof(int) * 6, stream));
    k_tile_extents<<<grid_for((nn + 31) / 32, 128), 128, 0, stream>>>(h->row_index.as<int32_t>(), tile_ptr, nn, ext);
    B2F_CUDA(cudaGetLastError());
    unsigned long long long_rows = 0;
    B2F_CUDA(cudaMemcpyAsync(&long_rows, counters + 3, sizeof(long_rows), cudaMemcpyDeviceToHost, stream));
    B2F_CUDA(cudaMemcpyAsync(h->tile_extents, ext, sizeof(int) * 6, cudaMemcpyDeviceToHost, stream));
    B2F_CUDA(cudaStreamSynchronize(stream));
    h->slots_ok = (h->variant == B2F_ASM_NODE_GATHER_SLOTS) && (long_rows == 0);
    h->fan_ok = (h->variant == B2F_ASM_NODE_GATHER) && (long_rows == 0) && !want_reorder;
  }

  // b2f_reorder: store the plan in a space-filling order of the nodes (single-rank node gather only)
  h->reordered = false;
  if (want_reorder) {
    rc = reorder_plan(h, stream); // sets fan_ok / reordered when every row fits the walk plan; otherwise the search-based kernel runs
    if (rc != B2F_OK)
      return rc;
  }
  if (h->inner) { // shadow context of reorder_plan: pattern + plan only
    B2F_CUDA(cudaStreamSynchronize(stream));
    h->ws_i64.release();
    h->ws_tmp.release();
    h->ws_i32a.release();
    h->ws_keys_a.release();
    h->has_pattern = true;
    return B2F_OK;
  }

  // BSRMatrix::initialize fills the values with 0 (femutils/BSRFormat.h:136); done last because the plan build borrows the array
  B2F_CUDA(cudaMemsetAsync(h->values.p, 0, sizeof(double) * (size_t)h->nb_nz, stream));

  // --- linear-system state (HypreDoFLinearSystem.cc:224-242)
  h->nb_dof = nn * h->bs;
  B2F_CUDA(h->rhs.alloc(sizeof(double) * (size_t)h->nb_dof));
  B2F_CUDA(h->forced_value.alloc(sizeof(double) * (size_t)h->nb_dof));
  B2F_CUDA(h->elim_value.alloc(sizeof(double) * (size_t)h->nb_dof));
  B2F_CUDA(h->forced_info.alloc((size_t)h->nb_dof));
  B2F_CUDA(h->elim_info.alloc((size_t)h->nb_dof));
  B2F_CUDA(cudaMemsetAsync(h->rhs.p, 0, sizeof(double) * (size_t)h->nb_dof, stream));
  B2F_CUDA(cudaMemsetAsync(h->forced_value.p, 0, sizeof(double) * (size_t)h->nb_dof, stream));
  B2F_CUDA(cudaMemsetAsync(h->elim_value.p, 0, sizeof(double) * (size_t)h->nb_dof, stream));
  B2F_CUDA(cudaMemsetAsync(h->forced_info.p, 0, (size_t)h->nb_dof, stream));
  B2F_CUDA(cudaMemsetAsync(h->elim_info.p, 0, (size_t)h->nb_dof, stream));
  B2F_CUDA(cudaStreamSynchronize(stream));
  h->ws_i64.release();
  h->ws_tmp.release();
  h->ws_i32a.release();
  h->ws_keys_a.release();
  h->has_pattern = true;
  return B2F_OK;
}

// ---------------------------------------------------------------------------------------------------
// b2f_reorder (b2f.h): node locality supplied by the library.  The pattern and the walk plan are built in the caller's
// numbering as always; then the rows are sorted along a Morton curve of the node coordinates and the plan is stored in that
// order: internal row i is the caller's row perm[i], its column ids are translated to internal ids, its step words and
// cell ids move with it unchanged (slots are row-local, so the walk of a row - hence every value - is bit for bit the one
// of the unpermuted plan).  The caller's row_index / columns stay as they are: they are the views, and the destination map
// of the kernel's write-out.
// ---------------------------------------------------------------------------------------------------
namespace {
  __device__ __forceinline__ unsigned long long order_bits(double x)
  {
    unsigned long long u = (unsigned long long)__double_as_longlong(x);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull); // monotone map of doubles to unsigned integers
  }
  __host__ __device__ inline double unorder_bits(unsigned long long u)
  {
    u = (u >> 63) ? (u & 0x7fffffffffffffffull) : ~u;
    double x;
    memcpy(&x, &u, sizeof(x));
    return x;
  }
  // box[0..2] = min, box[3..5] = max (order_bits encoding; box initialised to ~0 / 0)
  __global__ void __launch_bounds__(256) k_bbox(const double* __restrict__ coords, long long nn, unsigned long long* __restrict__ box)
  {
    unsigned long long lo[3] = { ~0ull, ~0ull, ~0ull }, hi[3] = { 0ull, 0ull, 0ull };
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nn; i += (long long)gridDim.x * blockDim.x)
      for (int d = 0; d < 3; ++d) {
        const unsigned long long u = order_bits(coords[3 * i + d]);
        lo[d] = u < lo[d] ? u : lo[d];
        hi[d] = u > hi[d] ? u : hi[d];
      }
    for (int d = 0; d < 3; ++d) {
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long a = __shfl_xor_sync(0xffffffffu, lo[d], o), b = __shfl_xor_sync(0xffffffffu, hi[d], o);
        lo[d] = a < lo[d] ? a : lo[d];
        hi[d] = b > hi[d] ? b : hi[d];
      }
      if ((threadIdx.x & 31) == 0) {
        atomicMin(box + d, lo[d]);
        atomicMax(box + 3 + d, hi[d]);
      }
    }
  }
  __device__ __forceinline__ unsigned long long spread3(unsigned long long v) // 21 bits -> every third bit
  {
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
  }
  // MORTON: 21 bits per axis interleaved
  __global__ void __launch_bounds__(256) k_morton_keys(const double* __restrict__ coords, long long nn, double x0, double y0, double z0, double sx, double sy, double sz,
                                                       unsigned long long* __restrict__ keys, int32_t* __restrict__ ids)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nn)
      return;
    const double q = 2097151.0; // 2^21 - 1
    const unsigned long long a = (unsigned long long)fmin(q, fmax(0.0, (coords[3 * i] - x0) * sx));
    const unsigned long long b = (unsigned long long)fmin(q, fmax(0.0, (coords[3 * i + 1] - y0) * sy));
    const unsigned long long c = (unsigned long long)fmin(q, fmax(0.0, (coords[3 * i + 2] - z0) * sz));
    keys[i] = spread3(a) | (spread3(b) << 1) | (spread3(c) << 2);
    ids[i] = (int32_t)i;
  }
  // PENCIL, one pass of three: key = (group of the node's current rank, coordinate `axis` resolved to 40 bits).  ids_in = the
  // order of the previous pass (null: identity); group = rank / group_size.  Equal-COUNT groups: slabs of nn^(2/3) nodes by z,
  // lines of nn^(1/3) nodes by y inside a slab, then x inside a line - on a structured lattice (jittered by less than half a
  // spacing) exactly its planes and lines, on an unstructured mesh pencils that follow the node density.
  __global__ void __launch_bounds__(256) k_pencil_keys(const double* __restrict__ coords, long long nn, const int32_t* __restrict__ ids_in, long long group_size, int axis,
                                                       double lo, double scale40, unsigned long long* __restrict__ keys, int32_t* __restrict__ ids_out)
  {
    const long long rank = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (rank >= nn)
      return;
    const int32_t node = ids_in ? ids_in[rank] : (int32_t)rank;
    const unsigned long long group = (unsigned long long)(rank / group_size);
    const double q = 1099511627775.0; // 2^40 - 1
    const unsigned long long c = (unsigned long long)fmin(q, fmax(0.0, (coords[3ll * node + axis] - lo) * scale40));
    keys[rank] = (group << 40) | c;
    ids_out[rank] = node;
  }
  __global__ void __launch_bounds__(256) k_invert_perm(const int32_t* __restrict__ perm, long long nn, int32_t* __restrict__ newid)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nn)
      newid[perm[i]] = (int32_t)i;
  }
  __global__ void __launch_bounds__(256) k_renumber_cells(const int32_t* __restrict__ cell_node, long long n, const int32_t* __restrict__ newid, int32_t* __restrict__ out)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
      out[i] = newid[cell_node[i]];
  }
  // One warp per internal row i (caller row r = perm[i]): block x of the internal row is column perm[r_columns[x]] of the
  // caller's row; its slot there is found by a scan of the caller's row (rows are short; both rows hold the same set).
  __global__ void __launch_bounds__(256) k_slot_map(const int32_t* __restrict__ perm, long long nn, const int32_t* __restrict__ r_row_index, const int32_t* __restrict__ r_columns,
                                                    const int32_t* __restrict__ row_index, const int32_t* __restrict__ columns, uint16_t* __restrict__ slot_map,
                                                    unsigned long long* __restrict__ unmatched)
  {
    const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= nn)
      return;
    const int32_t r = perm[i];
    const int32_t ib = r_row_index[i], k = r_row_index[i + 1] - ib, cb = row_index[r], kc = row_index[r + 1] - cb;
    for (int x = lane; x < k; x += 32) {
      const int32_t c = perm[r_columns[ib + x]];
      int found = -1;
      for (int y = 0; y < kc; ++y)
        if (columns[cb + y] == c) {
          found = y;
          break;
        }
      if (found < 0 || k != kc)
        atomicAdd(unmatched, 1ull);
      slot_map[ib + x] = (uint16_t)(found < 0 ? 0 : found);
    }
  }
  __global__ void __launch_bounds__(256) k_permute_coords(const int32_t* __restrict__ perm, long long nn, const double* __restrict__ coords, double* __restrict__ out)
  {
    const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; // one thread per scalar: coalesced stores
    if (q >= 3 * nn)
      return;
    const long long i = q / 3;
    out[q] = __ldg(coords + 3ll * perm[i] + (q - 3 * i));
  }
} // namespace

int permute_coords(b2f_context_s* h, cudaStream_t stream)
{
  const long long nn = h->mesh.nb_node;
  k_permute_coords<<<grid_for(3 * nn, 256), 256, 0, stream>>>(h->perm_rows.as<int32_t>(), nn, h->mesh.coords, h->r_coords.as<double>());
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int reorder_plan(b2f_context_s* h, cudaStream_t stream)
{
  const long long nn = h->mesh.nb_node;
  unsigned long long* counters = h->counters.as<unsigned long long>();
  // 1. bounding box
  DevBuf box_d;
  B2F_CUDA(box_d.alloc(sizeof(unsigned long long) * 6));
  unsigned long long box[6] = { ~0ull, ~0ull, ~0ull, 0ull, 0ull, 0ull };
  B2F_CUDA(cudaMemcpyAsync(box_d.p, box, sizeof(box), cudaMemcpyHostToDevice, stream));
  k_bbox<<<(unsigned)std::min<long long>(1184, (nn + 255) / 256), 256, 0, stream>>>(h->mesh.coords, nn, box_d.as<unsigned long long>());
  B2F_CUDA(cudaGetLastError());
  B2F_CUDA(cudaMemcpyAsync(box, box_d.p, sizeof(box), cudaMemcpyDeviceToHost, stream));
  B2F_CUDA(cudaStreamSynchronize(stream));
  box_d.release();
  double lo[3], sc[3], ext[3];
  for (int d = 0; d < 3; ++d) {
    lo[d] = unorder_bits(box[d]);
    ext[d] = unorder_bits(box[3 + d]) - lo[d];
    sc[d] = ext[d] > 0.0 ? 2097151.0 / ext[d] : 0.0;
  }
  // 2. Morton keys, stable radix sort -> perm (internal -> caller), newid (caller -> internal)
  DevBuf keys_a, keys_b, ids_a, newid;
  B2F_CUDA(keys_a.alloc(sizeof(unsigned long long) * (size_t)nn));
  B2F_CUDA(keys_b.alloc(sizeof(unsigned long long) * (size_t)nn));
  B2F_CUDA(ids_a.alloc(sizeof(int32_t) * (size_t)nn));
  B2F_CUDA(newid.alloc(sizeof(int32_t) * (size_t)nn));
  B2F_CUDA(h->perm_rows.alloc(sizeof(int32_t) * (size_t)nn));
  auto cleanup = [&]() {
    keys_a.release();
    keys_b.release();
    ids_a.release();
    newid.release();
  };
  auto sort_pass = [&](int bits) -> cudaError_t { // (keys_a, ids_a) -> (keys_b, perm_rows), stable
    size_t tb = 0;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, tb, keys_a.as<unsigned long long>(), keys_b.as<unsigned long long>(), ids_a.as<int32_t>(),
                                                    h->perm_rows.as<int32_t>(), nn, 0, bits, stream);
    if (e == cudaSuccess)
      e = h->ws_tmp.alloc(tb);
    if (e == cudaSuccess)
      e = cub::DeviceRadixSort::SortPairs(h->ws_tmp.p, tb, keys_a.as<unsigned long long>(), keys_b.as<unsigned long long>(), ids_a.as<int32_t>(), h->perm_rows.as<int32_t>(), nn, 0,
                                          bits, stream);
    return e;
  };
  cudaError_t es = cudaSuccess;
  if (h->opt.reorder == B2F_REORDER_PENCIL) {
    // axes that have an extent, slowest first (z, y, x); group sizes nn^((a-1)/a) ... so that every level holds equal counts
    int axes[3], na = 0;
    for (int d = 2; d >= 0; --d)
      if (ext[d] > 0.0)
        axes[na++] = d;
    const long long line = std::max<long long>(1, llround(pow((double)nn, na > 0 ? 1.0 / na : 1.0)));
    const int32_t* prev = nullptr;
    for (int lvl = 0; lvl < na && es == cudaSuccess; ++lvl) {
      // pass 0 sorts everything along the slowest axis; pass 1 sorts inside groups of line^(na-1) consecutive ranks (slabs),
      // pass 2 inside groups of `line` ranks (lines)
      long long group = nn;
      if (lvl > 0) {
        group = 1;
        for (int q = lvl; q < na; ++q)
          group *= line;
      }
      const int d = axes[lvl];
      k_pencil_keys<<<grid_for(nn, 256), 256, 0, stream>>>(h->mesh.coords, nn, prev, group, d, lo[d], 1099511627775.0 / ext[d], keys_a.as<unsigned long long>(), ids_a.as<int32_t>());
      es = sort_pass(64);
      prev = h->perm_rows.as<int32_t>(); // read by the next pass's key kernel before its sort rewrites it
    }
    if (na == 0) { // all nodes coincide: identity
      k_pencil_keys<<<grid_for(nn, 256), 256, 0, stream>>>(h->mesh.coords, nn, nullptr, nn, 0, 0.0, 0.0, keys_a.as<unsigned long long>(), ids_a.as<int32_t>());
      es = sort_pass(64);
    }
  }
  else {
    k_morton_keys<<<grid_for(nn, 256), 256, 0, stream>>>(h->mesh.coords, nn, lo[0], lo[1], lo[2], sc[0], sc[1], sc[2], keys_a.as<unsigned long long>(), ids_a.as<int32_t>());
    es = sort_pass(63);
  }
  if (es != cudaSuccess) {
    cleanup();
    return fail_cuda(es, "reorder_plan: radix sort of the order keys");
  }
  k_invert_perm<<<grid_for(nn, 256), 256, 0, stream>>>(h->perm_rows.as<int32_t>(), nn, newid.as<int32_t>());
  // 3. the connectivity in the internal numbering, and pattern + node->cell lists + walk plan built from it by the ordinary
  //    build in a shadow context (no value array, no system state): the walks are then canonical in the INTERNAL column order,
  //    so the rows a warp walks together (neighbours in space) meet "the same" neighbour at the same step
  const long long nc = h->mesh.nb_cell;
  DevBuf r_cells;
  cudaError_t e = r_cells.alloc(sizeof(int32_t) * (size_t)(nc * h->nv));
  if (e != cudaSuccess) {
    cleanup();
    return fail_cuda(e, "reorder_plan: allocation of the renumbered connectivity");
  }
  k_renumber_cells<<<grid_for(nc * h->nv, 256), 256, 0, stream>>>(h->mesh.cell_node, nc * h->nv, newid.as<int32_t>(), r_cells.as<int32_t>());
  b2f_context_s in;
  in.inner = true;
  in.mesh = h->mesh;
  in.mesh.cell_node = r_cells.as<int32_t>();
  in.mesh.coords = nullptr;
  in.mesh.node_is_own = nullptr;
  in.opt = h->opt;
  in.opt.reorder = B2F_REORDER_NONE;
  in.bs = h->bs;
  in.nv = h->nv;
  in.variant = h->variant;
  in.per_block = h->per_block;
  in.diag_last = h->diag_last;
  in.device = h->device;
  int rc = build_pattern(&in, stream);
  DevBuf* in_all[] = { &in.row_index, &in.columns, &in.nb_nz_per_row, &in.values, &in.nc_ptr, &in.nc_list, &in.nc_slots, &in.fan_ptr, &in.fan_steps, &in.fan_cell,
                       &in.counters, &in.ws_keys_a, &in.ws_keys_b, &in.ws_tmp, &in.ws_i32a, &in.ws_i64 };
  auto drop_inner = [&]() {
    for (DevBuf* b : in_all)
      b->release();
    r_cells.release();
    cleanup();
  };
  if (rc != B2F_OK) {
    drop_inner();
    return rc;
  }
  if (!in.fan_ok || in.nb_col != h->nb_col) { // a row too long for the walk plan: the search-based kernel runs, on the caller's arrays
    drop_inner();
    h->perm_rows.release();
    return B2F_OK;
  }
  // 4. the internal arrays move into the handle; slot map internal block -> slot in the caller's row
  std::swap(h->r_row_index, in.row_index);
  std::swap(h->r_columns, in.columns);
  std::swap(h->fan_ptr, in.fan_ptr);
  std::swap(h->fan_steps, in.fan_steps);
  std::swap(h->fan_cell, in.fan_cell);
  h->nb_fan_step = in.nb_fan_step;
  h->fan_first = in.fan_first;
  for (int q = 0; q < 6; ++q)
    h->tile_extents[q] = in.tile_extents[q];
  e = h->r_slot_map.alloc(sizeof(uint16_t) * (size_t)(h->nb_col + 8));
  if (e == cudaSuccess)
    e = h->r_coords.alloc(sizeof(double) * 3 * (size_t)nn);
  if (e == cudaSuccess) {
    B2F_CUDA(cudaMemsetAsync(counters + 3, 0, sizeof(unsigned long long), stream));
    k_slot_map<<<grid_for(nn * 32, 256), 256, 0, stream>>>(h->perm_rows.as<int32_t>(), nn, h->r_row_index.as<int32_t>(), h->r_columns.as<int32_t>(), h->row_index.as<int32_t>(),
                                                           h->columns.as<int32_t>(), h->r_slot_map.as<uint16_t>(), counters + 3);
    e = cudaGetLastError();
  }
  unsigned long long unmatched = 0;
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(&unmatched, counters + 3, sizeof(unmatched), cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess)
    e = cudaStreamSynchronize(stream);
  drop_inner();
  if (e != cudaSuccess)
    return fail_cuda(e, "reorder_plan: slot map");
  if (unmatched != 0)
    return fail(B2F_ERR_STATE, "reorder_plan: %llu blocks of the internal pattern have no counterpart in the caller's pattern", unmatched);
  h->fan_ok = true;
  h->fan_lockstep = false;
  h->reordered = true;
  return B2F_OK;
}

} // namespace b2f
