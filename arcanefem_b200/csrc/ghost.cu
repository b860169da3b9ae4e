// ghost.cu — multi-GPU ghost-row exchange over NCCL (one rank per GPU, NVLink 5 / NVSwitch).
//
// Reference scheme (femutils/BSRFormat.h:807-815, femutils/HypreDoFLinearSystem.cc:264-303): every rank loops over all of
// its cells INCLUDING the ghost-cell layer and keeps only the rows of nodes it owns — no assembly-time communication,
// the ghost layer is computed twice.  Here (BASELINE north star) each rank assembles only the cells it owns
// (b2f_mesh_view::cell_is_own) into ALL local rows; the rows of ghost nodes then hold this rank's partial sums, which
// b2f_ghost_exchange sends to the owner rank and adds there.  Owner rows end up identical to the reference scheme's
// (to summation order); ghost rows are zeroed after packing, like the reference leaves them.
//
// Plan (b2f_ghost_plan, once per pattern), everything device side, metadata travels over NCCL too:
//   1. ghost rows grouped by owner rank, ascending global id; per peer counts exchanged
//   2. per sent row: global row id, block count, global column ids  -> owner
//   3. owner: global -> local ids by binary search in its (sorted global id, local id) table, column slot by search in
//      the owner's row; one value index per incoming scalar
// Exchange (every assembly): pack kernel -> grouped ncclSend/ncclRecv with every peer that shares rows -> add kernel.
//
// NCCL is bound at run time (dlopen of libnccl.so.2, the copy the host process already loaded if any), so the library
// itself has no link-time dependency on it and loads on hosts without NCCL.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <type_traits>
#include <dlfcn.h>
#include <nccl.h>
#include <cub/cub.cuh>
#include "internal.h"
#include "b2f/assemble.cuh"

namespace b2f {

namespace {

  struct NcclApi
  {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
  };

  bool nccl_bind(NcclApi& api)
  {
    const char* names[] = { "libnccl.so.2", "libnccl.so" };
    void* lib = nullptr;
    for (const char* n : names) {
      lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (lib)
        break;
    }
    if (!lib)
      return false;
#define B2F_SYM(field, name)                                          \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(lib, name)); \
  if (!api.field)                                                     \
    return false;
    B2F_SYM(GetUniqueId, "ncclGetUniqueId")
    B2F_SYM(CommInitRank, "ncclCommInitRank")
    B2F_SYM(CommDestroy, "ncclCommDestroy")
    B2F_SYM(Send, "ncclSend")
    B2F_SYM(Recv, "ncclRecv")
    B2F_SYM(GroupStart, "ncclGroupStart")
    B2F_SYM(GroupEnd, "ncclGroupEnd")
    B2F_SYM(GetErrorString, "ncclGetErrorString")
#undef B2F_SYM
    api.lib = lib; // published last: a table with a library handle is complete
    return true;
  }

  // Bound once per process (std::call_once: several host threads may create handles concurrently).
  NcclApi* nccl_api()
  {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] { nccl_bind(api); });
    return api.lib ? &api : nullptr;
  }

  int fail_nccl(ncclResult_t r, const char* what)
  {
    NcclApi* api = nccl_api();
    return fail(B2F_ERR_NCCL, "NCCL error %d (%s) in %s", (int)r, api ? api->GetErrorString(r) : "?", what);
  }

#define B2F_NCCL(expr)                          \
  do {                                          \
    ncclResult_t b2f_r__ = (expr);              \
    if (b2f_r__ != ncclSuccess)                 \
      return fail_nccl(b2f_r__, #expr);         \
  } while (0)

  inline unsigned grid_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

  // An open ncclGroupStart() section.  Calls inside it go through call(): the first failure is remembered and the later calls
  // are skipped, and end() ALWAYS closes the group before the error is reported - a return from inside an open group would
  // leave the thread's NCCL state in group mode and make later NCCL calls misbehave or hang.
  struct NcclGroup
  {
    NcclApi* api;
    ncclResult_t first = ncclSuccess;
    const char* what = "";
    bool open = false;
    explicit NcclGroup(NcclApi* a) : api(a)
    {
      first = api->GroupStart();
      what = "ncclGroupStart";
      open = first == ncclSuccess;
    }
    void call(ncclResult_t r, const char* w)
    {
      if (first == ncclSuccess && r != ncclSuccess) {
        first = r;
        what = w;
      }
    }
    bool ok() const { return first == ncclSuccess; }
    int end()
    {
      if (open) {
        open = false;
        const ncclResult_t r = api->GroupEnd();
        if (first == ncclSuccess && r != ncclSuccess) {
          first = r;
          what = "ncclGroupEnd";
        }
      }
      return first == ncclSuccess ? B2F_OK : fail_nccl(first, what);
    }
    ~NcclGroup()
    {
      if (open)
        api->GroupEnd();
    }
  };
#define B2F_IN_GROUP(g, expr) \
  do {                        \
    if ((g).ok())             \
      (g).call((expr), #expr); \
  } while (0)

  // Temporary device buffer of b2f_ghost_plan: released on every return path.
  struct TmpBuf : DevBuf
  {
    TmpBuf() = default;
    TmpBuf(const TmpBuf&) = delete;
    TmpBuf& operator=(const TmpBuf&) = delete;
    ~TmpBuf() { release(); }
  };

  // metadata of the rows this rank sends: global row id, block count, global column ids (block offsets = row_off)
  __global__ void k_ghost_meta(const int32_t* __restrict__ send_rows, const long long* __restrict__ row_off, long long nrows, const int32_t* __restrict__ row_index,
                               const int32_t* __restrict__ columns, const long long* __restrict__ gid, long long* __restrict__ meta_row, long long* __restrict__ meta_k,
                               long long* __restrict__ meta_col)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nrows)
      return;
    const int32_t r = send_rows[i];
    const int32_t rb = row_index[r];
    const int k = row_index[r + 1] - rb;
    meta_row[i] = gid[r];
    meta_k[i] = k;
    long long* out = meta_col + row_off[i];
    for (int x = 0; x < k; ++x)
      out[x] = gid[columns[rb + x]];
  }

  __device__ __forceinline__ long long lower_bound_i64(const long long* __restrict__ a, long long n, long long key)
  {
    long long lo = 0, hi = n;
    while (lo < hi) {
      const long long mid = (lo + hi) >> 1;
      if (a[mid] < key)
        lo = mid + 1;
      else
        hi = mid;
    }
    return lo;
  }

  // value index of every scalar this rank sends, canonical order: row, block x, (i, j)
  template <int BS, bool PER_BLOCK>
  __global__ void k_send_index(const int32_t* __restrict__ send_rows, const long long* __restrict__ row_off, long long nrows, const int32_t* __restrict__ row_index,
                               long long* __restrict__ send_index)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nrows)
      return;
    const int32_t r = send_rows[i];
    const int32_t rb = row_index[r];
    const int k = row_index[r + 1] - rb;
    long long* out = send_index + row_off[i] * (BS * BS);
    for (int x = 0; x < k; ++x)
      for (int a = 0; a < BS; ++a)
        for (int b = 0; b < BS; ++b)
          out[(x * BS + a) * BS + b] = value_index<BS, PER_BLOCK>(rb, k, x, a, b);
  }

  // owner side: one thread per incoming row; translate global ids and find the value slots
  template <int BS, bool PER_BLOCK>
  __global__ void k_recv_index(const long long* __restrict__ meta_row, const long long* __restrict__ meta_k, const long long* __restrict__ row_off,
                               const long long* __restrict__ meta_col, long long nrows, const long long* __restrict__ sorted_gid, const int32_t* __restrict__ sorted_lid,
                               long long nb_node, const int32_t* __restrict__ row_index, const int32_t* __restrict__ columns, const uint8_t* __restrict__ is_ghost,
                               long long* __restrict__ recv_index, unsigned long long* __restrict__ errors, int diag_last)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nrows)
      return;
    const long long g = meta_row[i];
    const int kin = (int)meta_k[i];
    long long* out = recv_index + row_off[i] * (BS * BS);
    const long long pos = lower_bound_i64(sorted_gid, nb_node, g);
    const bool ok_row = pos < nb_node && sorted_gid[pos] == g && !is_ghost[sorted_lid[pos]];
    int32_t rb = 0;
    int k = 0;
    if (ok_row) {
      const int32_t r = sorted_lid[pos];
      rb = row_index[r];
      k = row_index[r + 1] - rb;
    }
    else
      atomicAdd(errors, 1ull);
    const long long* cin = meta_col + row_off[i];
    for (int x = 0; x < kin; ++x) {
      int slot = -1;
      if (ok_row) {
        const long long cp = lower_bound_i64(sorted_gid, nb_node, cin[x]);
        if (cp < nb_node && sorted_gid[cp] == cin[x]) {
          const int32_t c = sorted_lid[cp];
          const int ds = diag_slot(k, diag_last);
          slot = (columns[rb + ds] == c) ? ds : find_slot(columns + rb, k, c, diag_last);
        }
        if (slot < 0)
          atomicAdd(errors, 1ull);
      }
      for (int a = 0; a < BS; ++a)
        for (int b = 0; b < BS; ++b)
          out[(x * BS + a) * BS + b] = slot >= 0 ? value_index<BS, PER_BLOCK>(rb, k, slot, a, b) : -1ll;
    }
  }

  __global__ void k_ghost_pack(const long long* __restrict__ send_index, long long n, double* __restrict__ values, double* __restrict__ sendbuf)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
      return;
    const long long q = send_index[i];
    sendbuf[i] = values[q];
    values[q] = 0.0; // ghost rows end up zero, like the reference leaves them (BSRFormat.h:815)
  }

  // Owner side.  The incoming scalars are visited in the order of a stable sort by destination (built once by the plan):
  // the thread at the head of a run adds the whole run into its entry, peer by peer in ascending rank order (the packed
  // buffer lists the peers in that order) - a fixed summation order, so a node shared by three or more ranks still gives
  // bitwise reproducible owner rows (an atomicAdd per incoming scalar would not).
  __global__ void k_ghost_add(const long long* __restrict__ dst_sorted, const long long* __restrict__ src_sorted, long long n, const double* __restrict__ recvbuf,
                              double* __restrict__ values)
  {
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n)
      return;
    const long long q = dst_sorted[j];
    if (q < 0 || (j > 0 && dst_sorted[j - 1] == q))
      return;
    double v = values[q];
    for (long long t = j; t < n && dst_sorted[t] == q; ++t)
      v += recvbuf[src_sorted[t]];
    values[q] = v;
  }

  __global__ void k_iota_i64(long long* p, long long n)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
      p[i] = i;
  }

  __global__ void k_iota_i32(int32_t* p, long long n)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
      p[i] = (int32_t)i;
  }

} // namespace

namespace {
  template <class F>
  int dispatch_layout(int bs, bool per_block, F&& f)
  {
    switch (bs * 2 + (per_block ? 1 : 0)) {
    case 2: return f(std::integral_constant<int, 1>{}, std::false_type{});
    case 3: return f(std::integral_constant<int, 1>{}, std::true_type{});
    case 4: return f(std::integral_constant<int, 2>{}, std::false_type{});
    case 5: return f(std::integral_constant<int, 2>{}, std::true_type{});
    case 6: return f(std::integral_constant<int, 3>{}, std::false_type{});
    case 7: return f(std::integral_constant<int, 3>{}, std::true_type{});
    }
    return fail(B2F_ERR_INVALID_ARGUMENT, "unsupported block size %d", bs);
  }
} // namespace

int ghost_plan(b2f_context_s* h, void* comm_v, int rank, int world, const int32_t* owner, const int64_t* node_gid, cudaStream_t stream)
{
  NcclApi* api = nccl_api();
  if (!api)
    return fail(B2F_ERR_NCCL, "libnccl.so.2 could not be loaded");
  if (!comm_v || !owner || !node_gid || world < 1 || rank < 0 || rank >= world)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ghost_plan: bad arguments (rank=%d, world=%d)", rank, world);
  ncclComm_t comm = (ncclComm_t)comm_v;
  GhostPlan& P = h->ghost;
  P.ready = false;
  P.rank = rank;
  P.world = world;
  const long long nn = h->mesh.nb_node;
  const int bs2 = h->bs * h->bs;

  // 1. ghost rows by owner rank, ascending global id
  std::vector<std::vector<std::pair<int64_t, int32_t>>> by_peer((size_t)world);
  std::vector<uint8_t> is_ghost((size_t)nn, 0);
  std::vector<int32_t> ncp((size_t)nn + 1); // ghost nodes that no assembled cell touches hold no partial sum: not sent
  B2F_CUDA(cudaMemcpyAsync(ncp.data(), h->nc_ptr.p, sizeof(int32_t) * ncp.size(), cudaMemcpyDeviceToHost, stream));
  B2F_CUDA(cudaStreamSynchronize(stream));
  for (long long i = 0; i < nn; ++i) {
    const int o = owner[i];
    if (o < 0 || o >= world)
      return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ghost_plan: node %lld has owner rank %d outside [0, %d)", i, o, world);
    if (o != rank) {
      is_ghost[(size_t)i] = 1;
      if (ncp[(size_t)i + 1] > ncp[(size_t)i])
        by_peer[(size_t)o].push_back({ node_gid[i], (int32_t)i });
    }
  }
  std::vector<int32_t> send_rows;
  P.send_rows_n.assign((size_t)world, 0);
  for (int p = 0; p < world; ++p) {
    std::sort(by_peer[(size_t)p].begin(), by_peer[(size_t)p].end());
    P.send_rows_n[(size_t)p] = (int64_t)by_peer[(size_t)p].size();
    for (auto& e : by_peer[(size_t)p])
      send_rows.push_back(e.second);
  }
  const long long nsend_rows = (long long)send_rows.size();
  P.send_row_lo = P.send_row_hi = 0;
  if (nsend_rows > 0) {
    const auto mm = std::minmax_element(send_rows.begin(), send_rows.end());
    P.send_row_lo = *mm.first;
    P.send_row_hi = (int64_t)*mm.second + 1;
  }

  // 2. block offsets of the sent rows
  std::vector<int32_t> ri((size_t)nn + 1);
  B2F_CUDA(cudaMemcpyAsync(ri.data(), h->row_index.p, sizeof(int32_t) * ri.size(), cudaMemcpyDeviceToHost, stream));
  B2F_CUDA(cudaStreamSynchronize(stream));
  std::vector<long long> row_off((size_t)nsend_rows + 1, 0);
  for (long long i = 0; i < nsend_rows; ++i)
    row_off[(size_t)i + 1] = row_off[(size_t)i] + (ri[(size_t)send_rows[(size_t)i] + 1] - ri[(size_t)send_rows[(size_t)i]]);
  P.send_blocks.assign((size_t)world, 0);
  P.send_off.assign((size_t)world + 1, 0);
  {
    long long r0 = 0;
    for (int p = 0; p < world; ++p) {
      const long long r1 = r0 + P.send_rows_n[(size_t)p];
      P.send_blocks[(size_t)p] = row_off[(size_t)r1] - row_off[(size_t)r0];
      P.send_off[(size_t)p + 1] = row_off[(size_t)r1];
      r0 = r1;
    }
  }
  const long long nsend_blocks = row_off[(size_t)nsend_rows];

  TmpBuf d_send_rows, d_row_off, d_gid, d_ghost, d_cnt;
  B2F_CUDA(d_send_rows.alloc(sizeof(int32_t) * (size_t)std::max<long long>(nsend_rows, 1)));
  B2F_CUDA(d_row_off.alloc(sizeof(long long) * (size_t)(nsend_rows + 1)));
  B2F_CUDA(d_gid.alloc(sizeof(long long) * (size_t)nn));
  B2F_CUDA(d_ghost.alloc((size_t)nn));
  B2F_CUDA(d_cnt.alloc(sizeof(long long) * 4 * (size_t)world));
  B2F_CUDA(cudaMemcpyAsync(d_send_rows.p, send_rows.data(), sizeof(int32_t) * (size_t)nsend_rows, cudaMemcpyHostToDevice, stream));
  B2F_CUDA(cudaMemcpyAsync(d_row_off.p, row_off.data(), sizeof(long long) * row_off.size(), cudaMemcpyHostToDevice, stream));
  B2F_CUDA(cudaMemcpyAsync(d_gid.p, node_gid, sizeof(long long) * (size_t)nn, cudaMemcpyHostToDevice, stream));
  B2F_CUDA(cudaMemcpyAsync(d_ghost.p, is_ghost.data(), (size_t)nn, cudaMemcpyHostToDevice, stream));

  // 3. counts
  std::vector<long long> cnt((size_t)world * 4, 0); // [2p], [2p+1]: rows, blocks sent to p ; second half: received
  for (int p = 0; p < world; ++p) {
    cnt[(size_t)p * 2] = P.send_rows_n[(size_t)p];
    cnt[(size_t)p * 2 + 1] = P.send_blocks[(size_t)p];
  }
  long long* dc = d_cnt.as<long long>();
  B2F_CUDA(cudaMemcpyAsync(dc, cnt.data(), sizeof(long long) * 2 * (size_t)world, cudaMemcpyHostToDevice, stream));
  B2F_CUDA(cudaMemsetAsync(dc + 2 * world, 0, sizeof(long long) * 2 * (size_t)world, stream));
  {
    NcclGroup g(api);
    for (int p = 0; p < world; ++p) {
      if (p == rank)
        continue;
      B2F_IN_GROUP(g, api->Send(dc + 2 * p, 2, ncclInt64, p, comm, stream));
      B2F_IN_GROUP(g, api->Recv(dc + 2 * world + 2 * p, 2, ncclInt64, p, comm, stream));
    }
    const int rc_g = g.end();
    if (rc_g != B2F_OK)
      return rc_g;
  }
  B2F_CUDA(cudaMemcpyAsync(cnt.data() + 2 * world, dc + 2 * world, sizeof(long long) * 2 * (size_t)world, cudaMemcpyDeviceToHost, stream));
  B2F_CUDA(cudaStreamSynchronize(stream));
  P.recv_rows_n.assign((size_t)world, 0);
  P.recv_blocks.assign((size_t)world, 0);
  P.recv_off.assign((size_t)world + 1, 0);
  long long nrecv_rows = 0;
  for (int p = 0; p < world; ++p) {
    P.recv_rows_n[(size_t)p] = cnt[(size_t)(2 * world + 2 * p)];
    P.recv_blocks[(size_t)p] = cnt[(size_t)(2 * world + 2 * p + 1)];
    P.recv_off[(size_t)p + 1] = P.recv_off[(size_t)p] + P.recv_blocks[(size_t)p];
    nrecv_rows += P.recv_rows_n[(size_t)p];
  }
  const long long nrecv_blocks = P.recv_off[(size_t)world];

  // 4. metadata to the owners
  TmpBuf ms_row, ms_k, ms_col, mr_row, mr_k, mr_col, mr_off;
  B2F_CUDA(ms_row.alloc(sizeof(long long) * (size_t)std::max<long long>(nsend_rows, 1)));
  B2F_CUDA(ms_k.alloc(sizeof(long long) * (size_t)std::max<long long>(nsend_rows, 1)));
  B2F_CUDA(ms_col.alloc(sizeof(long long) * (size_t)std::max<long long>(nsend_blocks, 1)));
  B2F_CUDA(mr_row.alloc(sizeof(long long) * (size_t)std::max<long long>(nrecv_rows, 1)));
  B2F_CUDA(mr_k.alloc(sizeof(long long) * (size_t)(nrecv_rows + 1)));
  B2F_CUDA(mr_off.alloc(sizeof(long long) * (size_t)(nrecv_rows + 1)));
  B2F_CUDA(mr_col.alloc(sizeof(long long) * (size_t)std::max<long long>(nrecv_blocks, 1)));
  if (nsend_rows > 0)
    k_ghost_meta<<<grid_for(nsend_rows, 128), 128, 0, stream>>>(d_send_rows.as<int32_t>(), d_row_off.as<long long>(), nsend_rows, h->row_index.as<int32_t>(),
                                                                h->columns.as<int32_t>(), d_gid.as<long long>(), ms_row.as<long long>(), ms_k.as<long long>(),
                                                                ms_col.as<long long>());
  B2F_CUDA(cudaGetLastError());
  B2F_CUDA(cudaMemsetAsync(mr_k.p, 0, sizeof(long long) * (size_t)(nrecv_rows + 1), stream));
  {
    NcclGroup g(api);
    long long sr = 0, rr = 0;
    for (int p = 0; p < world; ++p) {
      const long long ns = P.send_rows_n[(size_t)p], nr = P.recv_rows_n[(size_t)p];
      if (p != rank && ns > 0) {
        B2F_IN_GROUP(g, api->Send(ms_row.as<long long>() + sr, (size_t)ns, ncclInt64, p, comm, stream));
        B2F_IN_GROUP(g, api->Send(ms_k.as<long long>() + sr, (size_t)ns, ncclInt64, p, comm, stream));
        B2F_IN_GROUP(g, api->Send(ms_col.as<long long>() + P.send_off[(size_t)p], (size_t)P.send_blocks[(size_t)p], ncclInt64, p, comm, stream));
      }
      if (p != rank && nr > 0) {
        B2F_IN_GROUP(g, api->Recv(mr_row.as<long long>() + rr, (size_t)nr, ncclInt64, p, comm, stream));
        B2F_IN_GROUP(g, api->Recv(mr_k.as<long long>() + rr, (size_t)nr, ncclInt64, p, comm, stream));
        B2F_IN_GROUP(g, api->Recv(mr_col.as<long long>() + P.recv_off[(size_t)p], (size_t)P.recv_blocks[(size_t)p], ncclInt64, p, comm, stream));
      }
      sr += ns;
      rr += nr;
    }
    const int rc_g = g.end();
    if (rc_g != B2F_OK)
      return rc_g;
  }

  // 5. owner side: block offsets of the incoming rows, global -> local table, value indices
  {
    size_t tb = 0;
    B2F_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, mr_k.as<long long>(), mr_off.as<long long>(), nrecv_rows + 1, stream));
    B2F_CUDA(h->ws_tmp.alloc(tb));
    B2F_CUDA(cub::DeviceScan::ExclusiveSum(h->ws_tmp.p, tb, mr_k.as<long long>(), mr_off.as<long long>(), nrecv_rows + 1, stream));
  }
  TmpBuf lid_in, gid_sorted, lid_sorted, errs;
  B2F_CUDA(lid_in.alloc(sizeof(int32_t) * (size_t)nn));
  B2F_CUDA(gid_sorted.alloc(sizeof(long long) * (size_t)nn));
  B2F_CUDA(lid_sorted.alloc(sizeof(int32_t) * (size_t)nn));
  B2F_CUDA(errs.alloc(sizeof(unsigned long long)));
  B2F_CUDA(cudaMemsetAsync(errs.p, 0, sizeof(unsigned long long), stream));
  k_iota_i32<<<grid_for(nn, 256), 256, 0, stream>>>(lid_in.as<int32_t>(), nn);
  {
    size_t tb = 0;
    B2F_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, d_gid.as<long long>(), gid_sorted.as<long long>(), lid_in.as<int32_t>(), lid_sorted.as<int32_t>(), nn, 0, 64,
                                             stream));
    B2F_CUDA(h->ws_tmp.alloc(tb));
    B2F_CUDA(cub::DeviceRadixSort::SortPairs(h->ws_tmp.p, tb, d_gid.as<long long>(), gid_sorted.as<long long>(), lid_in.as<int32_t>(), lid_sorted.as<int32_t>(), nn, 0, 64,
                                             stream));
  }
  P.total_send = nsend_blocks * bs2;
  P.total_recv = nrecv_blocks * bs2;
  B2F_CUDA(P.send_index.alloc(sizeof(long long) * (size_t)std::max<long long>(P.total_send, 1)));
  B2F_CUDA(P.recv_index.alloc(sizeof(long long) * (size_t)std::max<long long>(P.total_recv, 1)));
  B2F_CUDA(P.send_vals.alloc(sizeof(double) * (size_t)std::max<long long>(P.total_send, 1)));
  B2F_CUDA(P.recv_vals.alloc(sizeof(double) * (size_t)std::max<long long>(P.total_recv, 1)));
  int rc = dispatch_layout(h->bs, h->per_block, [&](auto bs_c, auto pb_c) -> int {
    constexpr int BS = decltype(bs_c)::value;
    constexpr bool PB = decltype(pb_c)::value;
    if (nsend_rows > 0)
      k_send_index<BS, PB><<<grid_for(nsend_rows, 128), 128, 0, stream>>>(d_send_rows.as<int32_t>(), d_row_off.as<long long>(), nsend_rows, h->row_index.as<int32_t>(),
                                                                          P.send_index.as<long long>());
    if (nrecv_rows > 0)
      k_recv_index<BS, PB><<<grid_for(nrecv_rows, 128), 128, 0, stream>>>(mr_row.as<long long>(), mr_k.as<long long>(), mr_off.as<long long>(), mr_col.as<long long>(),
                                                                          nrecv_rows, gid_sorted.as<long long>(), lid_sorted.as<int32_t>(), nn, h->row_index.as<int32_t>(),
                                                                          h->columns.as<int32_t>(), d_ghost.as<uint8_t>(), P.recv_index.as<long long>(),
                                                                          errs.as<unsigned long long>(), h->diag_last);
    B2F_CUDA(cudaGetLastError());
    return B2F_OK;
  });
  if (rc != B2F_OK)
    return rc;
  // destinations in a stable sorted order (k_ghost_add); entries that found no slot (-1) sort last as unsigned keys
  B2F_CUDA(P.recv_dst.alloc(sizeof(long long) * (size_t)std::max<long long>(P.total_recv, 1)));
  B2F_CUDA(P.recv_src.alloc(sizeof(long long) * (size_t)std::max<long long>(P.total_recv, 1)));
  if (P.total_recv > 0) {
    TmpBuf iota;
    B2F_CUDA(iota.alloc(sizeof(long long) * (size_t)P.total_recv));
    k_iota_i64<<<grid_for(P.total_recv, 256), 256, 0, stream>>>(iota.as<long long>(), P.total_recv);
    B2F_CUDA(cudaGetLastError());
    size_t tb = 0;
    const unsigned long long* kin = reinterpret_cast<const unsigned long long*>(P.recv_index.as<long long>());
    unsigned long long* kout = reinterpret_cast<unsigned long long*>(P.recv_dst.as<long long>());
    B2F_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, kin, kout, iota.as<long long>(), P.recv_src.as<long long>(), P.total_recv, 0, 64, stream));
    B2F_CUDA(h->ws_tmp.alloc(tb));
    B2F_CUDA(cub::DeviceRadixSort::SortPairs(h->ws_tmp.p, tb, kin, kout, iota.as<long long>(), P.recv_src.as<long long>(), P.total_recv, 0, 64, stream));
    long long ends[2] = { 0, -1 };
    B2F_CUDA(cudaMemcpyAsync(&ends[0], P.recv_dst.as<long long>(), sizeof(long long), cudaMemcpyDeviceToHost, stream));
    B2F_CUDA(cudaMemcpyAsync(&ends[1], P.recv_dst.as<long long>() + (P.total_recv - 1), sizeof(long long), cudaMemcpyDeviceToHost, stream));
    B2F_CUDA(cudaStreamSynchronize(stream)); // iota is released at the end of this scope
    P.recv_val_lo = ends[0];
    P.recv_val_hi = ends[1];
  }
  else {
    P.recv_val_lo = 0;
    P.recv_val_hi = -1;
  }
  unsigned long long nerr = 0;
  B2F_CUDA(cudaMemcpyAsync(&nerr, errs.p, sizeof(nerr), cudaMemcpyDeviceToHost, stream));
  B2F_CUDA(cudaStreamSynchronize(stream));
  h->ws_tmp.release();
  if (nerr != 0)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ghost_plan: %llu incoming ghost rows / columns are not owned rows / pattern entries of rank %d (the local mesh must hold "
                "every cell that touches an owned node)", nerr, rank);
  P.ready = true;
  return B2F_OK;
}

// First half of the exchange: pack the ghost rows (zeroing them) and run the grouped send/recv of every neighbour.
int ghost_send_recv(b2f_context_s* h, void* comm_v, cudaStream_t stream)
{
  NcclApi* api = nccl_api();
  if (!api)
    return fail(B2F_ERR_NCCL, "libnccl.so.2 could not be loaded");
  GhostPlan& P = h->ghost;
  if (!P.ready)
    return fail(B2F_ERR_STATE, "b2f_ghost_exchange: b2f_ghost_plan has not been called");
  if (!comm_v)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_ghost_exchange: null communicator");
  ncclComm_t comm = (ncclComm_t)comm_v;
  const int bs2 = h->bs * h->bs;
  if (P.total_send > 0)
    k_ghost_pack<<<grid_for(P.total_send, 256), 256, 0, stream>>>(P.send_index.as<long long>(), P.total_send, h->values.as<double>(), P.send_vals.as<double>());
  B2F_CUDA(cudaGetLastError());
  NcclGroup g(api);
  for (int p = 0; p < P.world; ++p) {
    if (p == P.rank)
      continue;
    if (P.send_blocks[(size_t)p] > 0)
      B2F_IN_GROUP(g, api->Send(P.send_vals.as<double>() + P.send_off[(size_t)p] * bs2, (size_t)(P.send_blocks[(size_t)p] * bs2), ncclDouble, p, comm, stream));
    if (P.recv_blocks[(size_t)p] > 0)
      B2F_IN_GROUP(g, api->Recv(P.recv_vals.as<double>() + P.recv_off[(size_t)p] * bs2, (size_t)(P.recv_blocks[(size_t)p] * bs2), ncclDouble, p, comm, stream));
  }
  return g.end();
}

// Second half: the received partial sums are added into the owner rows.
int ghost_add(b2f_context_s* h, cudaStream_t stream)
{
  GhostPlan& P = h->ghost;
  if (P.total_recv > 0)
    k_ghost_add<<<grid_for(P.total_recv, 256), 256, 0, stream>>>(P.recv_dst.as<long long>(), P.recv_src.as<long long>(), P.total_recv, P.recv_vals.as<double>(),
                                                                 h->values.as<double>());
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int ghost_exchange(b2f_context_s* h, void* comm_v, cudaStream_t stream)
{
  const int rc = ghost_send_recv(h, comm_v, stream);
  return rc != B2F_OK ? rc : ghost_add(h, stream);
}

} // namespace b2f

extern "C" {

int b2f_nccl_unique_id(void* id128_out_host)
{
  using namespace b2f;
  NcclApi* api = nccl_api();
  if (!api)
    return fail(B2F_ERR_NCCL, "libnccl.so.2 could not be loaded");
  if (!id128_out_host)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_nccl_unique_id: null output");
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  B2F_NCCL(api->GetUniqueId(reinterpret_cast<ncclUniqueId*>(id128_out_host)));
  return B2F_OK;
}

int b2f_nccl_comm_init(const void* id128_host, int32_t rank, int32_t world, void** comm_out)
{
  using namespace b2f;
  NcclApi* api = nccl_api();
  if (!api)
    return fail(B2F_ERR_NCCL, "libnccl.so.2 could not be loaded");
  if (!id128_host || !comm_out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_nccl_comm_init: null argument");
  ncclUniqueId id;
  memcpy(&id, id128_host, sizeof(id));
  ncclComm_t comm = nullptr;
  B2F_NCCL(api->CommInitRank(&comm, world, id, rank));
  *comm_out = comm;
  return B2F_OK;
}

int b2f_nccl_comm_destroy(void* comm)
{
  using namespace b2f;
  NcclApi* api = nccl_api();
  if (!api)
    return fail(B2F_ERR_NCCL, "libnccl.so.2 could not be loaded");
  if (comm)
    B2F_NCCL(api->CommDestroy((ncclComm_t)comm));
  return B2F_OK;
}
}
