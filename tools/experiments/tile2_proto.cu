// Feasibility experiment, NOT product code (last paragraph): "every tet's geometry computed ONCE per
// CTA and shared" instead of once per vertex, WITHOUT shared-memory atomics.  Structured Kuhn cube only: cells, halo and the
// gather plan follow from index arithmetic, so no plan is read at all - a LOWER bound of what a plan-driven version for general
// meshes would cost (that one adds ~1 GB of plan traffic per launch).
//   tile    = TX x TY x TZ nodes, one CTA; halo hexes around it are recomputed by the neighbouring tiles
//   phase 1 = one thread per tet of the (TX+1)(TY+1)(TZ+1) hexes: coordinates from shared memory, the 6 off-diagonal entries of
//             Ke (edge values; Ke is symmetric, the diagonal follows from the zero row sums) -> shared memory, tet-private slots
//   phase 2 = one thread per row: every off-diagonal entry gathers its 4 or 6 edge values (fixed order: deterministic), the
//             diagonal is minus their sum; 15 values per row staged in shared memory and written with coalesced stores
// Output layout: out[node * 15 + slot] (dense 15 slots per node; slot 0 diagonal).  Checked against a one-thread-per-cell
// atomic kernel on the same layout.  build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tile2_proto tile2_proto.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int KUHN[6][4] = { { 0, 1, 3, 7 }, { 0, 5, 1, 7 }, { 0, 3, 2, 7 }, { 0, 2, 6, 7 }, { 0, 4, 5, 7 }, { 0, 6, 4, 7 } };
__constant__ int c_kuhn[6][4];
struct Contrib { short dh; short q; short p; short pad; }; // hex offset (linear, tile-local), tet type, pair index 0..5
__constant__ Contrib c_contrib[14][6];
__constant__ int c_ncontrib[14];
__constant__ int c_slot_of_pair[6][6][2]; // reference kernel: tet type, pair -> slot in row a / slot in row b

__host__ __device__ inline int pair_index(int a, int b) // a < b
{
  return a == 0 ? b - 1 : (a == 1 ? b + 1 : 5);
}

template <int TX, int TY, int TZ>
__global__ void __launch_bounds__(TX* TY* TZ) k_tile2(const double* __restrict__ coords, double* __restrict__ out, int N)
{
  constexpr int ROWS = TX * TY * TZ, HX = TX + 1, HY = TY + 1, HZ = TZ + 1, NHEX = HX * HY * HZ, NTET = 6 * NHEX;
  constexpr int PX = TX + 2, PY = TY + 2, PZ = TZ + 2, NPT = PX * PY * PZ;
  extern __shared__ double sm[];
  double* xs = sm; // [NPT][3]
  double* ke = sm + 3 * NPT; // [6][NTET]: pair-major, conflict-free stores
  double* stage = ke; // phase 2 output staging reuses nothing of ke that is still read: it lives after it
  stage = ke + 6 * NTET; // [ROWS][15]
  const int n1 = N + 1, ntx = (n1 + TX - 1) / TX, nty = (n1 + TY - 1) / TY;
  const int tx = blockIdx.x % ntx, ty = (blockIdx.x / ntx) % nty, tz = blockIdx.x / (ntx * nty);
  const int ox = tx * TX - 1, oy = ty * TY - 1, oz = tz * TZ - 1; // origin of the point box (one halo layer)
  const int t = threadIdx.x;
  for (int q = t; q < NPT; q += ROWS) {
    const int i = q % PX, j = (q / PX) % PY, k = q / (PX * PY);
    const int gx = ox + i, gy = oy + j, gz = oz + k;
    double x = 0, y = 0, z = 0;
    if (gx >= 0 && gx < n1 && gy >= 0 && gy < n1 && gz >= 0 && gz < n1) {
      const long long g = gx + (long long)n1 * (gy + (long long)n1 * gz);
      x = coords[3 * g];
      y = coords[3 * g + 1];
      z = coords[3 * g + 2];
    }
    xs[3 * q] = x;
    xs[3 * q + 1] = y;
    xs[3 * q + 2] = z;
  }
  __syncthreads();
  // phase 1
  for (int item = t; item < NTET; item += ROWS) {
    const int q = item / NHEX, h = item - q * NHEX; // tet-type major: a warp holds one tet type of consecutive hexes
    const int i = h % HX, j = (h / HX) % HY, k = h / (HX * HY);
    const int hx = ox + i, hy = oy + j, hz = oz + k;
    double v[6] = { 0, 0, 0, 0, 0, 0 };
    if (hx >= 0 && hx < N && hy >= 0 && hy < N && hz >= 0 && hz < N) {
      double X[4][3];
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int c = c_kuhn[q][a];
        const int p = (i + (c & 1)) + PX * ((j + ((c >> 1) & 1)) + PY * (k + ((c >> 2) & 1)));
        X[a][0] = xs[3 * p];
        X[a][1] = xs[3 * p + 1];
        X[a][2] = xs[3 * p + 2];
      }
      double e[3][3], g[4][3];
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int d = 0; d < 3; ++d)
          e[a][d] = X[a + 1][d] - X[0][d];
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const double* u = e[(a + 1) % 3];
        const double* w = e[(a + 2) % 3];
        g[a + 1][0] = u[1] * w[2] - u[2] * w[1];
        g[a + 1][1] = u[2] * w[0] - u[0] * w[2];
        g[a + 1][2] = u[0] * w[1] - u[1] * w[0];
      }
#pragma unroll
      for (int d = 0; d < 3; ++d)
        g[0][d] = -(g[1][d] + g[2][d] + g[3][d]);
      const double det = e[0][0] * g[1][0] + e[0][1] * g[1][1] + e[0][2] * g[1][2];
      const double x6 = 6.0 * fabs(det);
      double y;
      asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x6));
      const double er = fma(-x6, y, 1.0);
      const double s = fma(y, fma(er, er, er), y);
      int p = 0;
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = a + 1; b < 4; ++b)
          v[p++] = s * (g[a][0] * g[b][0] + g[a][1] * g[b][1] + g[a][2] * g[b][2]);
    }
#pragma unroll
    for (int p = 0; p < 6; ++p)
      ke[p * NTET + item] = v[p];
  }
  __syncthreads();
  // phase 2: row t = local node (1 + lx, 1 + ly, 1 + lz) of the point box; its hexes are the 8 around it
  {
    const int lx = t % TX, ly = (t / TX) % TY, lz = t / (TX * TY);
    const int hbase = lx + HX * (ly + HY * lz); // hex whose corner 7 is the node: local hex (lx, ly, lz); offsets dh add to it
    double diag = 0.0;
#pragma unroll
    for (int s = 0; s < 14; ++s) {
      double acc = 0.0;
      const int n = c_ncontrib[s];
      for (int c = 0; c < n; ++c) {
        const Contrib cc = c_contrib[s][c];
        acc += ke[cc.p * NTET + cc.q * NHEX + hbase + cc.dh];
      }
      stage[t * 15 + 1 + s] = acc;
      diag -= acc;
    }
    stage[t * 15] = diag;
  }
  __syncthreads();
  // coalesced write-out: TX consecutive nodes along x are contiguous in the output (TX * 15 doubles)
  for (int q = t; q < ROWS * 15; q += ROWS) {
    const int row = q / 15, x = q - row * 15;
    const int lx = row % TX, ly = (row / TX) % TY, lz = row / (TX * TY);
    const int gx = ox + 1 + lx, gy = oy + 1 + ly, gz = oz + 1 + lz;
    if (gx < n1 && gy < n1 && gz < n1)
      out[(gx + (long long)n1 * (gy + (long long)n1 * gz)) * 15 + x] = stage[q];
  }
}

// reference: one thread per cell, fp64 atomics into the same dense layout
__global__ void k_ref(const double* __restrict__ coords, double* __restrict__ out, int N)
{
  const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nc = 6ll * N * N * N;
  if (cell >= nc)
    return;
  const int q = (int)(cell % 6);
  const long long h = cell / 6;
  const int n1 = N + 1;
  const int i = (int)(h % N), j = (int)((h / N) % N), k = (int)(h / ((long long)N * N));
  long long nd[4];
  double X[4][3];
  for (int a = 0; a < 4; ++a) {
    const int c = c_kuhn[q][a];
    nd[a] = (i + (c & 1)) + (long long)n1 * ((j + ((c >> 1) & 1)) + (long long)n1 * (k + ((c >> 2) & 1)));
    for (int d = 0; d < 3; ++d)
      X[a][d] = coords[3 * nd[a] + d];
  }
  double e[3][3], g[4][3];
  for (int a = 0; a < 3; ++a)
    for (int d = 0; d < 3; ++d)
      e[a][d] = X[a + 1][d] - X[0][d];
  for (int a = 0; a < 3; ++a) {
    const double* u = e[(a + 1) % 3];
    const double* w = e[(a + 2) % 3];
    g[a + 1][0] = u[1] * w[2] - u[2] * w[1];
    g[a + 1][1] = u[2] * w[0] - u[0] * w[2];
    g[a + 1][2] = u[0] * w[1] - u[1] * w[0];
  }
  for (int d = 0; d < 3; ++d)
    g[0][d] = -(g[1][d] + g[2][d] + g[3][d]);
  const double det = e[0][0] * g[1][0] + e[0][1] * g[1][1] + e[0][2] * g[1][2];
  const double s = 1.0 / (6.0 * fabs(det));
  for (int a = 0; a < 4; ++a)
    for (int b = a + 1; b < 4; ++b) {
      const double v = s * (g[a][0] * g[b][0] + g[a][1] * g[b][1] + g[a][2] * g[b][2]);
      const int p = pair_index(a, b);
      atomicAdd(out + nd[a] * 15 + c_slot_of_pair[q][p][0], v);
      atomicAdd(out + nd[b] * 15 + c_slot_of_pair[q][p][1], v);
      atomicAdd(out + nd[a] * 15, -v);
      atomicAdd(out + nd[b] * 15, -v);
    }
}

static int slot_of_offset(int dx, int dy, int dz) // 14 neighbour offsets -> slots 1..14
{
  const int code = abs(dx) + 2 * abs(dy) + 4 * abs(dz); // 1..7
  return (dx < 0 || dy < 0 || dz < 0) ? 7 + code : code;
}

template <int TX, int TY, int TZ>
static float run_tile(const double* dc, double* dout, int N, int reps)
{
  constexpr int ROWS = TX * TY * TZ, HX = TX + 1, HY = TY + 1, NHEX = HX * HY * (TZ + 1), NTET = 6 * NHEX, NPT = (TX + 2) * (TY + 2) * (TZ + 2);
  // contribution table for this tile shape: row node = corner 7 of local hex (lx, ly, lz); a tet of type q whose vertex a is the
  // node sits in hex (node - corner(c[q][a])); in hex-local linear terms the offset from `hbase` is +(1 - cx) + HX ((1 - cy) + HY (1 - cz))
  Contrib tab[14][6];
  int cnt[14] = { 0 };
  for (int q = 0; q < 6; ++q)
    for (int a = 0; a < 4; ++a)
      for (int b = 0; b < 4; ++b) {
        if (a == b)
          continue;
        const int ca = KUHN[q][a], cb = KUHN[q][b];
        const int dx = (cb & 1) - (ca & 1), dy = ((cb >> 1) & 1) - ((ca >> 1) & 1), dz = ((cb >> 2) & 1) - ((ca >> 2) & 1);
        const int s = slot_of_offset(dx, dy, dz) - 1;
        const int hx = 1 - (ca & 1), hy = 1 - ((ca >> 1) & 1), hz = 1 - ((ca >> 2) & 1);
        Contrib c;
        c.dh = (short)(hx + HX * (hy + HY * hz));
        c.q = (short)q;
        c.p = (short)(a < b ? pair_index(a, b) : pair_index(b, a));
        c.pad = 0;
        tab[s][cnt[s]++] = c;
      }
  cudaMemcpyToSymbol(c_contrib, tab, sizeof(tab));
  cudaMemcpyToSymbol(c_ncontrib, cnt, sizeof(cnt));
  const size_t smem = sizeof(double) * (3 * NPT + 6 * NTET + ROWS * 15);
  cudaFuncSetAttribute(k_tile2<TX, TY, TZ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int n1 = N + 1;
  const int grid = ((n1 + TX - 1) / TX) * ((n1 + TY - 1) / TY) * ((n1 + TZ - 1) / TZ);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int w = 0; w < 3; ++w)
    k_tile2<TX, TY, TZ><<<grid, ROWS, smem>>>(dc, dout, N);
  cudaEventRecord(e0);
  for (int r = 0; r < reps; ++r)
    k_tile2<TX, TY, TZ><<<grid, ROWS, smem>>>(dc, dout, N);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_tile2<TX, TY, TZ>, ROWS, smem);
  printf("{\"experiment\": \"tile2_proto\", \"tile\": \"%dx%dx%d\", \"threads\": %d, \"smem_bytes\": %zu, \"ctas_per_sm\": %d, \"tets_per_tile\": %d, \"recompute_factor\": %.3f, "
         "\"n\": %d, \"ms_per_launch\": %.4f, \"cuda\": \"%s\"",
         TX, TY, TZ, ROWS, smem, occ, NTET, (double)NTET / (ROWS * 6.0), N, ms / reps, cudaGetErrorString(err));
  return ms / reps;
}

int main(int argc, char** argv)
{
  const int N = argc > 1 ? atoi(argv[1]) : 203;
  const int n1 = N + 1;
  const long long nn = (long long)n1 * n1 * n1;
  std::vector<double> h((size_t)nn * 3);
  unsigned long long st = 88172645463325252ull;
  auto rnd = [&]() { st ^= st << 13; st ^= st >> 7; st ^= st << 17; return (double)(st >> 11) / 9007199254740992.0; };
  const double hh = 1.0 / N;
  for (int k = 0; k < n1; ++k)
    for (int j = 0; j < n1; ++j)
      for (int i = 0; i < n1; ++i) {
        const size_t g = (size_t)i + (size_t)n1 * (j + (size_t)n1 * k);
        const bool in = i > 0 && i < N && j > 0 && j < N && k > 0 && k < N;
        h[3 * g] = i * hh + (in ? (rnd() - 0.5) * 0.5 * hh : 0.0);
        h[3 * g + 1] = j * hh + (in ? (rnd() - 0.5) * 0.5 * hh : 0.0);
        h[3 * g + 2] = k * hh + (in ? (rnd() - 0.5) * 0.5 * hh : 0.0);
      }
  int sop[6][6][2];
  for (int q = 0; q < 6; ++q)
    for (int a = 0; a < 4; ++a)
      for (int b = a + 1; b < 4; ++b) {
        const int ca = KUHN[q][a], cb = KUHN[q][b];
        const int dx = (cb & 1) - (ca & 1), dy = ((cb >> 1) & 1) - ((ca >> 1) & 1), dz = ((cb >> 2) & 1) - ((ca >> 2) & 1);
        sop[q][pair_index(a, b)][0] = slot_of_offset(dx, dy, dz);
        sop[q][pair_index(a, b)][1] = slot_of_offset(-dx, -dy, -dz);
      }
  cudaMemcpyToSymbol(c_kuhn, KUHN, sizeof(KUHN));
  cudaMemcpyToSymbol(c_slot_of_pair, sop, sizeof(sop));
  double *dc, *dout, *dref;
  cudaMalloc(&dc, sizeof(double) * nn * 3);
  cudaMalloc(&dout, sizeof(double) * nn * 15);
  cudaMalloc(&dref, sizeof(double) * nn * 15);
  cudaMemcpy(dc, h.data(), sizeof(double) * nn * 3, cudaMemcpyHostToDevice);
  cudaMemset(dref, 0, sizeof(double) * nn * 15);
  const long long nc = 6ll * N * N * N;
  k_ref<<<(unsigned)((nc + 255) / 256), 256>>>(dc, dref, N);
  cudaDeviceSynchronize();
  std::vector<double> ref((size_t)nn * 15), got((size_t)nn * 15);
  cudaMemcpy(ref.data(), dref, sizeof(double) * nn * 15, cudaMemcpyDeviceToHost);
  auto check = [&]() {
    cudaMemcpy(got.data(), dout, sizeof(double) * nn * 15, cudaMemcpyDeviceToHost);
    double worst = 0.0;
    for (long long n = 0; n < nn; ++n) {
      double scale = 0.0;
      for (int x = 0; x < 15; ++x)
        scale = fmax(scale, fabs(ref[(size_t)n * 15 + x]));
      for (int x = 0; x < 15; ++x)
        worst = fmax(worst, fabs(got[(size_t)n * 15 + x] - ref[(size_t)n * 15 + x]) / scale);
    }
    printf(", \"max_row_relative_difference_vs_cell_atomic\": %.3e}\n", worst);
  };
  cudaMemset(dout, 0, sizeof(double) * nn * 15);
  run_tile<8, 8, 4>(dc, dout, N, 10);
  check();
  cudaMemset(dout, 0, sizeof(double) * nn * 15);
  run_tile<8, 4, 4>(dc, dout, N, 10);
  check();
  cudaMemset(dout, 0, sizeof(double) * nn * 15);
  run_tile<16, 4, 4>(dc, dout, N, 10);
  check();
  cudaMemset(dout, 0, sizeof(double) * nn * 15);
  run_tile<32, 4, 2>(dc, dout, N, 10);
  check();
  return 0;
}
