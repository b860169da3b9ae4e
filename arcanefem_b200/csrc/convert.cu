// convert.cu — BSR -> CSR view on device and the BSR SpMV check (sm_100a).
//   to_csr : BSRMatrix::toCsr, femutils/BSRFormat.h:194-256 — the reference copies four arrays for BS = 1 and expands
//            rows/columns in host loops for BS > 1 (leaving one row pointer unset for BS = 3).  Here BS = 1 is a
//            zero-copy alias and BS > 1 is a device kernel; values always alias (per-row layout == CSR order).
//   spmv   : y = A x on the BSR storage, both layouts (north-star's device check; no reference counterpart).
#include "internal.h"
#include "b2f/assemble.cuh"
#include <cub/device/device_scan.cuh>
#include <algorithm>

namespace b2f {

namespace {
  inline unsigned grid_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

  // 8 lanes per scalar row r = (block row, i)
  template <int BS>
  __global__ void __launch_bounds__(256) k_expand_csr(const int32_t* __restrict__ row_index, const int32_t* __restrict__ columns, long long nb_node, long long nnz,
                                                      int32_t* __restrict__ rows, int32_t* __restrict__ nbcol, int32_t* __restrict__ cols)
  {
    const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int l = threadIdx.x & 7;
    if (g >= nb_node * BS)
      return;
    const long long br = g / BS;
    const int i = (int)(g % BS);
    const int32_t rb = row_index[br];
    const int k = row_index[br + 1] - rb;
    const long long off = (long long)rb * BS * BS + (long long)i * BS * k;
    if (l == 0) {
      rows[g] = (int32_t)off;
      nbcol[g] = k * BS;
      if (g == nb_node * BS - 1)
        rows[g + 1] = (int32_t)nnz;
    }
    for (int q = l; q < k * BS; q += 8)
      cols[off + q] = columns[rb + q / BS] * BS + (q % BS);
  }

  template <int BS, bool PER_BLOCK>
  __global__ void __launch_bounds__(256) k_spmv(const int32_t* __restrict__ row_index, const int32_t* __restrict__ columns, const double* __restrict__ values,
                                                long long nb_node, const double* __restrict__ x, double* __restrict__ y)
  {
    const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int l = threadIdx.x & 7;
    const bool valid = g < nb_node * BS;
    double s = 0.0;
    if (valid) {
      const long long br = g / BS;
      const int i = (int)(g % BS);
      const int32_t rb = row_index[br];
      const int k = row_index[br + 1] - rb;
      for (int q = l; q < k * BS; q += 8) {
        const int xb = q / BS, j = q % BS;
        s += values[value_index<BS, PER_BLOCK>(rb, k, xb, i, j)] * __ldg(x + (long long)columns[rb + xb] * BS + j);
      }
    }
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    if (valid && l == 0)
      y[g] = s;
  }
} // namespace

int to_csr(b2f_context_s* h, cudaStream_t s)
{
  if (h->bs == 1) {
    h->has_csr = true;
    return B2F_OK;
  }
  if (h->per_block)
    return fail(B2F_ERR_STATE, "b2f_to_csr with block_size > 1 needs use_csr = 1 (per-row value layout, femutils/BSRFormat.h:405)");
  const long long nn = h->mesh.nb_node, nrow = nn * h->bs;
  B2F_CUDA(h->csr_rows.alloc(sizeof(int32_t) * (size_t)(nrow + 1)));
  B2F_CUDA(h->csr_nbcol.alloc(sizeof(int32_t) * (size_t)nrow));
  B2F_CUDA(h->csr_cols.alloc(sizeof(int32_t) * (size_t)h->nb_nz));
  if (nrow > 0) {
    if (h->bs == 2)
      k_expand_csr<2><<<grid_for(nrow * 8, 256), 256, 0, s>>>(h->row_index.as<int32_t>(), h->columns.as<int32_t>(), nn, h->nb_nz, h->csr_rows.as<int32_t>(),
                                                              h->csr_nbcol.as<int32_t>(), h->csr_cols.as<int32_t>());
    else
      k_expand_csr<3><<<grid_for(nrow * 8, 256), 256, 0, s>>>(h->row_index.as<int32_t>(), h->columns.as<int32_t>(), nn, h->nb_nz, h->csr_rows.as<int32_t>(),
                                                              h->csr_nbcol.as<int32_t>(), h->csr_cols.as<int32_t>());
  }
  B2F_CUDA(cudaGetLastError());
  h->has_csr = true;
  return B2F_OK;
}

// Local -> global numbering of the CSR view on device.  The reference does this on the host, one entry at a time, right
// before HYPRE_IJMatrixSetValues (femutils/HypreDoFLinearSystem.cc:474-519: rows_index[i] = numbering[dof],
// columns[q] = numbering[local column]).
namespace {
  __global__ void __launch_bounds__(256) k_renumber(const int32_t* __restrict__ in, long long n, const int32_t* __restrict__ numbering, long long nb_dof, int32_t* __restrict__ out,
                                                    unsigned long long* __restrict__ errors)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
      return;
    const int32_t c = in ? in[i] : (int32_t)i;
    if (c < 0 || c >= nb_dof) {
      atomicAdd(errors, 1ull);
      out[i] = -1;
      return;
    }
    out[i] = numbering[c];
  }
} // namespace

int csr_global_numbering(b2f_context_s* h, const int32_t* dof_number, int32_t* rows_global, int32_t* cols_global, cudaStream_t s)
{
  if (!h->has_csr) {
    int rc = to_csr(h, s);
    if (rc != B2F_OK)
      return rc;
  }
  const long long nd = h->nb_dof, nnz = h->nb_nz;
  const int32_t* cols = h->bs == 1 ? h->columns.as<int32_t>() : h->csr_cols.as<int32_t>();
  unsigned long long* err = h->counters.as<unsigned long long>() + 2;
  B2F_CUDA(cudaMemsetAsync(err, 0, sizeof(unsigned long long), s));
  if (rows_global && nd > 0)
    k_renumber<<<grid_for(nd, 256), 256, 0, s>>>(nullptr, nd, dof_number, nd, rows_global, err);
  if (cols_global && nnz > 0)
    k_renumber<<<grid_for(nnz, 256), 256, 0, s>>>(cols, nnz, dof_number, nd, cols_global, err);
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

// Own-row extraction for the solver hand-off.  HypreDoFLinearSystemImpl::solve() walks the own dofs on the host
// (femutils/HypreDoFLinearSystem.cc:474-519): for each own row it records the global row id, the row length and the
// global ids of its columns, then copies the four arrays to the device and calls
// HYPRE_IJMatrixSetValues(ij_A, nb_local_row, ncols, rows, cols, values) (:587-609) and HYPRE_IJVectorSetValues for the
// right-hand side (:617-640).  Here the same arrays are produced on the device: flag own rows, two exclusive scans (row
// slot, value offset), one pass that copies rows in order (8 lanes per row).  Rows of ghost nodes are dropped.
namespace {
  __global__ void __launch_bounds__(256) k_ij_count(const uint8_t* __restrict__ node_is_own, const int32_t* __restrict__ nbcol, long long nb_row, int bs,
                                                    int32_t* __restrict__ flag, int32_t* __restrict__ count)
  {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nb_row)
      return;
    const bool own = node_is_own == nullptr || node_is_own[r / bs] != 0;
    flag[r] = own ? 1 : 0;
    count[r] = own ? nbcol[r] : 0;
  }

  __global__ void __launch_bounds__(256) k_ij_fill(const uint8_t* __restrict__ node_is_own, const int32_t* __restrict__ rows_ptr, const int32_t* __restrict__ nbcol,
                                                   const int32_t* __restrict__ cols_local, const double* __restrict__ values, const double* __restrict__ rhs,
                                                   const int32_t* __restrict__ numbering, long long nb_row, int bs, const int32_t* __restrict__ row_slot,
                                                   const int32_t* __restrict__ value_off, int32_t* __restrict__ out_ncols, int32_t* __restrict__ out_rows,
                                                   int32_t* __restrict__ out_row_local, int32_t* __restrict__ out_cols, double* __restrict__ out_values,
                                                   double* __restrict__ out_rhs)
  {
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int l = threadIdx.x & 7;
    if (r >= nb_row)
      return;
    if (node_is_own != nullptr && node_is_own[r / bs] == 0)
      return;
    const int k = nbcol[r];
    const int32_t src = rows_ptr[r], dst = value_off[r];
    if (l == 0) {
      const int32_t i = row_slot[r];
      out_ncols[i] = k;
      out_rows[i] = numbering ? numbering[r] : (int32_t)r;
      out_row_local[i] = (int32_t)r;
      if (out_rhs)
        out_rhs[i] = rhs[r];
    }
    for (int q = l; q < k; q += 8) {
      const int32_t c = cols_local[src + q];
      out_cols[dst + q] = numbering ? numbering[c] : c;
      out_values[dst + q] = values[src + q];
    }
  }
} // namespace

namespace {
  // in-place exclusive sum of n int32 values (cub; scratch from the handle's workspace)
  int exclusive_scan_i32(b2f_context_s* h, int32_t* data, long long n, cudaStream_t s)
  {
    size_t tb = 0;
    B2F_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, data, data, n, s));
    B2F_CUDA(h->ws_tmp.alloc(tb));
    B2F_CUDA(cub::DeviceScan::ExclusiveSum(h->ws_tmp.p, tb, data, data, n, s));
    return B2F_OK;
  }
} // namespace

int to_hypre_ij(b2f_context_s* h, const int32_t* dof_number, b2f_ij_view* out, cudaStream_t s)
{
  if (!h->has_csr) {
    int rc = to_csr(h, s);
    if (rc != B2F_OK)
      return rc;
  }
  const long long nd = h->nb_dof, nnz = h->nb_nz;
  const int32_t* rows_ptr = h->bs == 1 ? h->row_index.as<int32_t>() : h->csr_rows.as<int32_t>();
  const int32_t* nbcol = h->bs == 1 ? h->nb_nz_per_row.as<int32_t>() : h->csr_nbcol.as<int32_t>();
  const int32_t* cols = h->bs == 1 ? h->columns.as<int32_t>() : h->csr_cols.as<int32_t>();
  *out = b2f_ij_view{};
  if (nd == 0)
    return B2F_OK;
  B2F_CUDA(h->ij_flag.alloc(sizeof(int32_t) * (size_t)(nd + 1)));
  B2F_CUDA(h->ij_off.alloc(sizeof(int32_t) * (size_t)(nd + 1)));
  int32_t* flag = h->ij_flag.as<int32_t>();
  int32_t* off = h->ij_off.as<int32_t>();
  // one extra zero entry so that the exclusive scans also deliver the totals
  B2F_CUDA(cudaMemsetAsync(flag + nd, 0, sizeof(int32_t), s));
  B2F_CUDA(cudaMemsetAsync(off + nd, 0, sizeof(int32_t), s));
  k_ij_count<<<grid_for(nd, 256), 256, 0, s>>>(h->mesh.node_is_own, nbcol, nd, h->bs, flag, off);
  B2F_CUDA(cudaGetLastError());
  int rc = exclusive_scan_i32(h, flag, nd + 1, s);
  if (rc != B2F_OK)
    return rc;
  rc = exclusive_scan_i32(h, off, nd + 1, s);
  if (rc != B2F_OK)
    return rc;
  int32_t totals[2] = { 0, 0 };
  B2F_CUDA(cudaMemcpyAsync(&totals[0], flag + nd, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  B2F_CUDA(cudaMemcpyAsync(&totals[1], off + nd, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  B2F_CUDA(cudaStreamSynchronize(s));
  const long long nrow = totals[0], nval = totals[1];
  B2F_CUDA(h->ij_ncols.alloc(sizeof(int32_t) * (size_t)nrow));
  B2F_CUDA(h->ij_rows.alloc(sizeof(int32_t) * (size_t)nrow));
  B2F_CUDA(h->ij_row_local.alloc(sizeof(int32_t) * (size_t)nrow));
  B2F_CUDA(h->ij_cols.alloc(sizeof(int32_t) * (size_t)nval));
  B2F_CUDA(h->ij_values.alloc(sizeof(double) * (size_t)nval));
  B2F_CUDA(h->ij_rhs.alloc(sizeof(double) * (size_t)nrow));
  k_ij_fill<<<grid_for(nd * 8, 256), 256, 0, s>>>(h->mesh.node_is_own, rows_ptr, nbcol, cols, h->values.as<double>(), h->rhs.as<double>(), dof_number, nd, h->bs, flag, off,
                                                  h->ij_ncols.as<int32_t>(), h->ij_rows.as<int32_t>(), h->ij_row_local.as<int32_t>(), h->ij_cols.as<int32_t>(),
                                                  h->ij_values.as<double>(), h->rhs.p ? h->ij_rhs.as<double>() : nullptr);
  B2F_CUDA(cudaGetLastError());
  (void)nnz;
  out->nb_row = nrow;
  out->nb_value = nval;
  out->ncols = h->ij_ncols.as<int32_t>();
  out->rows = h->ij_rows.as<int32_t>();
  out->row_local = h->ij_row_local.as<int32_t>();
  out->cols = h->ij_cols.as<int32_t>();
  out->values = h->ij_values.as<double>();
  out->rhs = h->rhs.p ? h->ij_rhs.as<double>() : nullptr;
  return B2F_OK;
}

// Row chunks of the host pipeline (b2f_assemble_host): chunk boundaries on multiples of 128 rows, the offset of each
// boundary in the value array and, per chunk, the highest node id its rows read (= highest column; the diagonal makes the
// row itself a column), made monotone so that "coordinates [0, hi]" is a growing prefix.
namespace {
  __global__ void __launch_bounds__(256) k_max_i32(const int32_t* __restrict__ v, long long n, int32_t* __restrict__ out)
  {
    int32_t m = -1;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
      m = max(m, v[i]);
    for (int d = 16; d > 0; d >>= 1)
      m = max(m, __shfl_xor_sync(0xffffffffu, m, d));
    if ((threadIdx.x & 31) == 0 && m >= 0)
      atomicMax(out, m);
  }
} // namespace

int pipeline_plan(b2f_context_s* h, int nchunk, cudaStream_t s)
{
  const long long nn = h->mesh.nb_node;
  const long long tiles = (nn + 127) / 128;
  const long long per = (tiles + nchunk - 1) / nchunk;
  h->chunk_row.clear();
  for (long long c = 0; c * per < tiles; ++c)
    h->chunk_row.push_back(c * per * 128);
  h->chunk_row.push_back(nn);
  const int nc = (int)h->chunk_row.size() - 1;
  std::vector<int32_t> off(nc + 1);
  for (int c = 0; c <= nc; ++c)
    B2F_CUDA(cudaMemcpyAsync(&off[c], h->row_index.as<int32_t>() + h->chunk_row[c], sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  B2F_CUDA(cudaStreamSynchronize(s));
  B2F_CUDA(h->ij_flag.alloc(sizeof(int32_t) * (size_t)(nc + 1)));
  int32_t* hi_d = h->ij_flag.as<int32_t>();
  B2F_CUDA(cudaMemsetAsync(hi_d, 0xff, sizeof(int32_t) * (size_t)nc, s));
  for (int c = 0; c < nc; ++c) {
    const long long n = (long long)off[c + 1] - off[c];
    if (n > 0)
      k_max_i32<<<(unsigned)std::min<long long>(1184, (n + 255) / 256), 256, 0, s>>>(h->columns.as<int32_t>() + off[c], n, hi_d + c);
  }
  B2F_CUDA(cudaGetLastError());
  std::vector<int32_t> hi(nc);
  B2F_CUDA(cudaMemcpyAsync(hi.data(), hi_d, sizeof(int32_t) * (size_t)nc, cudaMemcpyDeviceToHost, s));
  B2F_CUDA(cudaStreamSynchronize(s));
  h->chunk_value_off.assign(nc + 1, 0);
  h->chunk_node_hi.assign(nc, 0);
  int64_t run = -1;
  for (int c = 0; c < nc; ++c) {
    h->chunk_value_off[c] = (int64_t)off[c] * h->bs * h->bs;
    run = std::max<int64_t>(run, hi[c]);
    h->chunk_node_hi[c] = run;
  }
  h->chunk_value_off[nc] = (int64_t)off[nc] * h->bs * h->bs;
  return B2F_OK;
}

int spmv(b2f_context_s* h, const double* x, double* y, cudaStream_t s)
{
  const long long nn = h->mesh.nb_node, nrow = nn * h->bs;
  if (nrow == 0)
    return B2F_OK;
  const int32_t* ri = h->row_index.as<int32_t>();
  const int32_t* co = h->columns.as<int32_t>();
  const double* va = h->values.as<double>();
  const unsigned grid = grid_for(nrow * 8, 256);
  const bool pb = h->per_block;
  switch (h->bs) {
  case 1:
    k_spmv<1, true><<<grid, 256, 0, s>>>(ri, co, va, nn, x, y);
    break;
  case 2:
    if (pb)
      k_spmv<2, true><<<grid, 256, 0, s>>>(ri, co, va, nn, x, y);
    else
      k_spmv<2, false><<<grid, 256, 0, s>>>(ri, co, va, nn, x, y);
    break;
  case 3:
    if (pb)
      k_spmv<3, true><<<grid, 256, 0, s>>>(ri, co, va, nn, x, y);
    else
      k_spmv<3, false><<<grid, 256, 0, s>>>(ri, co, va, nn, x, y);
    break;
  }
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

} // namespace b2f
