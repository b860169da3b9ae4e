// convert.cu — BSR -> CSR view on device and the BSR SpMV check (sm_100a).
//   to_csr : BSRMatrix::toCsr, femutils/BSRFormat.h:194-256 — the reference copies four arrays for BS = 1 and expands
//            rows/columns in host loops for BS > 1 (leaving one row pointer unset for BS = 3).  Here BS = 1 is a
//            zero-copy alias and BS > 1 is a device kernel; values always alias (per-row layout == CSR order).
//   spmv   : y = A x on the BSR storage, both layouts (north-star's device check; no reference counterpart).
#include "internal.h"
#include "b2f/assemble.cuh"

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
