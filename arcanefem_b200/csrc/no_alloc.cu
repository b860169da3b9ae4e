// no_alloc.cu — b2f_ij_own_rows_no_alloc: the own-row extraction of convert.cu (to_hypre_ij) for callers that cannot
// allocate device memory.  No counterpart in the reference (include/b2f.h).
//
// The extraction needs two scans over the rows (slot of each own row, offset of its entries).  to_hypre_ij keeps them in
// the handle's workspace (cudaMalloc + cub temporary storage); here they live in ONE file-scope device array of fixed
// size, so the entry point makes no allocation of any kind:
//     words [0, R)            row slot inside its tile        (R = B2F_NO_ALLOC_MAX_ROWS)
//     words [R, 2R)           entry offset inside its tile
//     words [2R, 2R + T)      own rows per tile, then their exclusive sum   (T = R / 1024 tiles)
//     8-byte words after that entries per tile, then their exclusive sum    (T x int64)
// Every GPU that loads the library holds the array (8 MiB + 12 KiB) until the process exits.  All calls on a device
// share it: each call writes every word it later reads (nothing has to be cleared), and calls that overlap on the GPU race.
//
// In host code the array's name is only a stand-in of the symbol; the device address comes from cudaGetSymbolAddress,
// once per device with that device current, below.
#include "internal.h"
#include <cub/block/block_scan.cuh>
#include <mutex>

namespace b2f {

namespace {
  constexpr int NA_TILE = 1024; // rows per CTA of the count pass (256 threads x 4 rows)
  constexpr int NA_MAX_ROWS = B2F_NO_ALLOC_MAX_ROWS;
  constexpr int NA_MAX_TILES = NA_MAX_ROWS / NA_TILE;
  constexpr size_t NA_WORDS = 2 * (size_t)NA_MAX_ROWS + NA_MAX_TILES + 2 * (size_t)NA_MAX_TILES;
  static_assert(NA_MAX_TILES <= 1024, "the tile sums are scanned by one CTA");

  static __device__ __align__(256) int32_t g_no_alloc_scratch[NA_WORDS];

  struct Scratch
  {
    int32_t* slot;
    int32_t* off;
    int32_t* tile_rows;
    long long* tile_vals;
  };

  __host__ __device__ inline Scratch carve(int32_t* base)
  {
    Scratch s;
    s.slot = base;
    s.off = base + NA_MAX_ROWS;
    s.tile_rows = base + 2 * (size_t)NA_MAX_ROWS;
    s.tile_vals = reinterpret_cast<long long*>(base + 2 * (size_t)NA_MAX_ROWS + NA_MAX_TILES);
    return s;
  }

  __device__ __forceinline__ int row_length(const int32_t* __restrict__ row_ptr, const int32_t* __restrict__ row_ncols, long long r)
  {
    return row_ncols ? row_ncols[r] : row_ptr[r + 1] - row_ptr[r];
  }

  // pass 1: own flag and length of 4 consecutive rows per thread, exclusive sums inside the tile, tile totals
  __global__ void __launch_bounds__(256) k_na_count(const uint8_t* __restrict__ node_is_own, const int32_t* __restrict__ row_ptr, const int32_t* __restrict__ row_ncols,
                                                    long long nb_row, int bs, int32_t* __restrict__ base)
  {
    using Scan = cub::BlockScan<int, 256>;
    __shared__ typename Scan::TempStorage tmp_a, tmp_b;
    const Scratch S = carve(base);
    const long long r0 = (long long)blockIdx.x * NA_TILE + threadIdx.x * 4;
    int flag[4], cnt[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const long long r = r0 + q;
      const bool own = r < nb_row && (node_is_own == nullptr || node_is_own[r / bs] != 0);
      flag[q] = own ? 1 : 0;
      cnt[q] = own ? row_length(row_ptr, row_ncols, r) : 0;
    }
    int tot_f, tot_c;
    Scan(tmp_a).ExclusiveSum(flag, flag, tot_f);
    Scan(tmp_b).ExclusiveSum(cnt, cnt, tot_c);
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (r0 + q < nb_row) {
        S.slot[r0 + q] = flag[q];
        S.off[r0 + q] = cnt[q];
      }
    if (threadIdx.x == 0) {
      S.tile_rows[blockIdx.x] = tot_f;
      S.tile_vals[blockIdx.x] = tot_c;
    }
  }

  // pass 2: exclusive sums of the tile totals (one CTA), totals to the caller
  __global__ void __launch_bounds__(1024) k_na_scan_tiles(int nb_tile, int32_t* __restrict__ base, long long* __restrict__ totals_out)
  {
    using ScanI = cub::BlockScan<int, 1024>;
    using ScanL = cub::BlockScan<long long, 1024>;
    __shared__ typename ScanI::TempStorage tmp_a;
    __shared__ typename ScanL::TempStorage tmp_b;
    const Scratch S = carve(base);
    const int t = threadIdx.x;
    int f = t < nb_tile ? S.tile_rows[t] : 0;
    long long c = t < nb_tile ? S.tile_vals[t] : 0;
    int tot_f;
    long long tot_c;
    ScanI(tmp_a).ExclusiveSum(f, f, tot_f);
    ScanL(tmp_b).ExclusiveSum(c, c, tot_c);
    if (t < nb_tile) {
      S.tile_rows[t] = f;
      S.tile_vals[t] = c;
    }
    if (t == 0 && totals_out) {
      totals_out[0] = tot_f;
      totals_out[1] = tot_c;
    }
  }

  // pass 3: copy the own rows in order, 8 lanes per row (same order and values as k_ij_fill of convert.cu)
  __global__ void __launch_bounds__(256) k_na_fill(const uint8_t* __restrict__ node_is_own, const int32_t* __restrict__ row_ptr, const int32_t* __restrict__ row_ncols,
                                                   const int32_t* __restrict__ cols_local, const double* __restrict__ values, const double* __restrict__ rhs,
                                                   const int32_t* __restrict__ numbering, long long nb_row, int bs, const int32_t* __restrict__ base, b2f_ij_buffers out)
  {
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int l = threadIdx.x & 7;
    if (r >= nb_row)
      return;
    if (node_is_own != nullptr && node_is_own[r / bs] == 0)
      return;
    const Scratch S = carve(const_cast<int32_t*>(base));
    const int tile = (int)(r / NA_TILE);
    const long long i = (long long)S.tile_rows[tile] + S.slot[r];
    const long long dst = S.tile_vals[tile] + S.off[r];
    const int k = row_length(row_ptr, row_ncols, r);
    const int32_t src = row_ptr[r];
    if (l == 0 && i < out.capacity_rows) {
      out.ncols[i] = k;
      out.rows[i] = numbering ? numbering[r] : (int32_t)r;
      out.row_local[i] = (int32_t)r;
      if (out.rhs && rhs)
        out.rhs[i] = rhs[r];
    }
    for (int q = l; q < k; q += 8) {
      if (dst + q >= out.capacity_values)
        break;
      const int32_t c = cols_local[src + q];
      out.cols[dst + q] = numbering ? numbering[c] : c;
      out.values[dst + q] = values[src + q];
    }
  }

  // device address of the array, looked up once per device with that device current
  constexpr int NA_MAX_DEVICES = 64;
  std::mutex g_na_mutex;
  int32_t* g_na_base[NA_MAX_DEVICES] = {};

  int scratch_of_current_device(int32_t** base)
  {
    int dev = -1;
    B2F_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NA_MAX_DEVICES)
      return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ij_own_rows_no_alloc: device ordinal %d outside [0, %d)", dev, NA_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (g_na_base[dev] == nullptr) {
      void* p = nullptr;
      B2F_CUDA(cudaGetSymbolAddress(&p, g_no_alloc_scratch));
      if (p == nullptr)
        return fail(B2F_ERR_CUDA, "b2f_ij_own_rows_no_alloc: cudaGetSymbolAddress returned a null address on device %d", dev);
      g_na_base[dev] = static_cast<int32_t*>(p);
    }
    *base = g_na_base[dev];
    return B2F_OK;
  }
} // namespace

int ij_own_rows_no_alloc(const int32_t* row_ptr, const int32_t* row_ncols, const int32_t* cols, const double* values, const double* rhs, const uint8_t* node_is_own,
                         const int32_t* dof_number, long long nb_row, int bs, const b2f_ij_buffers& out, cudaStream_t s)
{
  int32_t* base = nullptr;
  int rc = scratch_of_current_device(&base);
  if (rc != B2F_OK)
    return rc;
  if (nb_row == 0) {
    if (out.totals_out)
      B2F_CUDA(cudaMemsetAsync(out.totals_out, 0, 2 * sizeof(int64_t), s));
    return B2F_OK;
  }
  const int nb_tile = (int)((nb_row + NA_TILE - 1) / NA_TILE);
  k_na_count<<<nb_tile, 256, 0, s>>>(node_is_own, row_ptr, row_ncols, nb_row, bs, base);
  k_na_scan_tiles<<<1, 1024, 0, s>>>(nb_tile, base, reinterpret_cast<long long*>(out.totals_out));
  k_na_fill<<<(unsigned)((nb_row * 8 + 255) / 256), 256, 0, s>>>(node_is_own, row_ptr, row_ncols, cols, values, rhs, dof_number, nb_row, bs, base, out);
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

} // namespace b2f

extern "C" int b2f_ij_own_rows_no_alloc(const int32_t* row_ptr, const int32_t* row_ncols, const int32_t* cols, const double* values, const double* rhs,
                                        const uint8_t* node_is_own, const int32_t* dof_number, int64_t nb_row, int32_t block_size, const b2f_ij_buffers* out,
                                        b2f_stream stream)
{
  using namespace b2f;
  if (!out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ij_own_rows_no_alloc: null output block");
  if (nb_row < 0 || block_size < 1)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ij_own_rows_no_alloc: nb_row %lld, block_size %d", (long long)nb_row, (int)block_size);
  if (nb_row > B2F_NO_ALLOC_MAX_ROWS)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ij_own_rows_no_alloc: %lld rows exceed the fixed scratch array (B2F_NO_ALLOC_MAX_ROWS = %d); use b2f_to_hypre_ij",
                (long long)nb_row, (int)B2F_NO_ALLOC_MAX_ROWS);
  if (out->capacity_rows < 0 || out->capacity_values < 0)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ij_own_rows_no_alloc: negative capacity");
  if (nb_row > 0) {
    if (!row_ptr || !cols || !values)
      return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ij_own_rows_no_alloc: null row_ptr / cols / values");
    if ((out->capacity_rows > 0 && (!out->ncols || !out->rows || !out->row_local)) || (out->capacity_values > 0 && (!out->cols || !out->values)))
      return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ij_own_rows_no_alloc: null output array with a non-zero capacity");
  }
  return ij_own_rows_no_alloc(row_ptr, row_ncols, cols, values, rhs, node_is_own, dof_number, nb_row, block_size, *out, (cudaStream_t)stream);
}
