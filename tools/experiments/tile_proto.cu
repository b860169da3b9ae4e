// Feasibility experiment, NOT product code: "every tet computed once" assembly in row-owned 3D tiles, on the structured
// Kuhn cube only (node ids, cells and column slots follow from index arithmetic, so no plan is read).  It answers one
// question for DESIGN.md section 8: how fast can a cell-parallel kernel with shared-memory accumulation run on B200,
// compared with the fan walk that visits each tet once per vertex?
//   tile  = 8 x 8 x 8 nodes (512 rows, 15 slots each: diagonal + the 14 Kuhn neighbours), one CTA of 512 threads
//   cells = the tets of the 9 x 9 x 9 hexes that touch the tile (halo cells are recomputed by the neighbouring tiles: x1.42)
//   MODE 0: shared-memory atomicAdd(double) (CAS loop, correct); MODE 1: plain += (racy, wrong values: cost bound of a
//   conflict-free schedule)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tile_proto tile_proto.cu ; run: ./tile_proto N coords.bin
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

__constant__ int c_kuhn[6][4] = { { 0, 1, 3, 7 }, { 0, 5, 1, 7 }, { 0, 3, 2, 7 }, { 0, 2, 6, 7 }, { 0, 4, 5, 7 }, { 0, 6, 4, 7 } };

__constant__ int c_slot[6][4][4];

template <int MODE>
__global__ void __launch_bounds__(512, 2) k_tile(const double* __restrict__ coords, double* __restrict__ out, int N)
{
  extern __shared__ double sm[];
  double* xs = sm; // [1000][3] region coordinates
  double* acc = sm + 3000; // [512][15]
  const int n1 = N + 1, nt = (n1 + 7) / 8;
  const int tx = blockIdx.x % nt, ty = (blockIdx.x / nt) % nt, tz = blockIdx.x / (nt * nt);
  const int ox = tx * 8 - 1, oy = ty * 8 - 1, oz = tz * 8 - 1;
  const int t = threadIdx.x;
  for (int q = t; q < 1000; q += 512) {
    const int i = q % 10, j = (q / 10) % 10, k = q / 100;
    const int gx = ox + i, gy = oy + j, gz = oz + k;
    if (gx >= 0 && gx < n1 && gy >= 0 && gy < n1 && gz >= 0 && gz < n1) {
      const long long g = gx + (long long)n1 * (gy + (long long)n1 * gz);
      xs[3 * q] = coords[3 * g];
      xs[3 * q + 1] = coords[3 * g + 1];
      xs[3 * q + 2] = coords[3 * g + 2];
    }
  }
  for (int q = t; q < 512 * 15; q += 512)
    acc[q] = 0.0;
  __syncthreads();
  for (int item = t; item < 729 * 6; item += 512) {
    const int tet = item / 729, hq = item % 729; // tet-major: the lanes of a warp hold the same tet type of consecutive hexes, so their targets are distinct
    const int i = hq % 9, j = (hq / 9) % 9, k = hq / 81;
    const int hx = ox + i, hy = oy + j, hz = oz + k;
    if (hx < 0 || hx >= N || hy < 0 || hy >= N || hz < 0 || hz >= N)
      continue;
    int lx[4], ly[4], lz[4];
    double X[4][3];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int c = c_kuhn[tet][a];
      lx[a] = i + (c & 1);
      ly[a] = j + ((c >> 1) & 1);
      lz[a] = k + ((c >> 2) & 1);
      const int q = lx[a] + 10 * (ly[a] + 10 * lz[a]);
      X[a][0] = xs[3 * q];
      X[a][1] = xs[3 * q + 1];
      X[a][2] = xs[3 * q + 2];
    }
    double e[3][3], g[4][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int d = 0; d < 3; ++d)
        e[a][d] = X[a + 1][d] - X[0][d];
    g[1][0] = e[1][1] * e[2][2] - e[1][2] * e[2][1];
    g[1][1] = e[1][2] * e[2][0] - e[1][0] * e[2][2];
    g[1][2] = e[1][0] * e[2][1] - e[1][1] * e[2][0];
    g[2][0] = e[2][1] * e[0][2] - e[2][2] * e[0][1];
    g[2][1] = e[2][2] * e[0][0] - e[2][0] * e[0][2];
    g[2][2] = e[2][0] * e[0][1] - e[2][1] * e[0][0];
    g[3][0] = e[0][1] * e[1][2] - e[0][2] * e[1][1];
    g[3][1] = e[0][2] * e[1][0] - e[0][0] * e[1][2];
    g[3][2] = e[0][0] * e[1][1] - e[0][1] * e[1][0];
#pragma unroll
    for (int d = 0; d < 3; ++d)
      g[0][d] = -(g[1][d] + g[2][d] + g[3][d]);
    const double det = e[0][0] * g[1][0] + e[0][1] * g[1][1] + e[0][2] * g[1][2];
    const double s = 1.0 / (6.0 * fabs(det));
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const bool in_tile = lx[a] >= 1 && lx[a] <= 8 && ly[a] >= 1 && ly[a] <= 8 && lz[a] >= 1 && lz[a] <= 8;
      if (!in_tile)
        continue;
      const int row = (lx[a] - 1) + 8 * ((ly[a] - 1) + 8 * (lz[a] - 1));
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        if (b == a)
          continue;
        const int slot = c_slot[tet][a][b];
        const double v = s * (g[a][0] * g[b][0] + g[a][1] * g[b][1] + g[a][2] * g[b][2]);
        if (MODE == 0)
          atomicAdd(&acc[slot * 512 + row], v);
        else
          acc[slot * 512 + row] += v;
      }
    }
  }
  __syncthreads();
  {
    double d = 0.0;
    for (int x = 1; x < 15; ++x)
      d -= acc[x * 512 + t];
    acc[t] = d;
  }
  __syncthreads();
  for (int q = t; q < 512 * 15; q += 512) {
    const int row = q / 15, x = q % 15;
    const int gx = ox + 1 + (row & 7), gy = oy + 1 + ((row >> 3) & 7), gz = oz + 1 + (row >> 6);
    if (gx < n1 && gy < n1 && gz < n1)
      out[(gx + (long long)n1 * (gy + (long long)n1 * gz)) * 15 + x] = acc[x * 512 + row];
  }
}

int main(int argc, char** argv)
{
  const int N = argc > 1 ? atoi(argv[1]) : 203;
  const int n1 = N + 1, nt = (n1 + 7) / 8;
  const long long nn = (long long)n1 * n1 * n1;
  std::vector<double> h((size_t)nn * 3);
  FILE* f = fopen(argc > 2 ? argv[2] : "coords.bin", "rb");
  if (!f || fread(h.data(), sizeof(double), h.size(), f) != h.size()) {
    fprintf(stderr, "cannot read coordinates\n");
    return 1;
  }
  fclose(f);
  {
    const int kuhn[6][4] = { { 0, 1, 3, 7 }, { 0, 5, 1, 7 }, { 0, 3, 2, 7 }, { 0, 2, 6, 7 }, { 0, 4, 5, 7 }, { 0, 6, 4, 7 } };
    int slot[6][4][4];
    for (int q = 0; q < 6; ++q)
      for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 4; ++b) {
          const int ca = kuhn[q][a], cb = kuhn[q][b];
          const int dx = (cb & 1) - (ca & 1), dy = ((cb >> 1) & 1) - ((ca >> 1) & 1), dz = ((cb >> 2) & 1) - ((ca >> 2) & 1);
          const int code = abs(dx) + 2 * abs(dy) + 4 * abs(dz);
          slot[q][a][b] = (dx < 0 || dy < 0 || dz < 0) ? 7 + code : code;
        }
    cudaMemcpyToSymbol(c_slot, slot, sizeof(slot));
  }
  double *dc, *dout;
  cudaMalloc(&dc, sizeof(double) * nn * 3);
  cudaMalloc(&dout, sizeof(double) * nn * 15);
  cudaMemcpy(dc, h.data(), sizeof(double) * nn * 3, cudaMemcpyHostToDevice);
  const size_t smem = sizeof(double) * (3000 + 512 * 15);
  cudaFuncSetAttribute(k_tile<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(k_tile<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode) {
    cudaMemset(dout, 0, sizeof(double) * nn * 15);
    for (int w = 0; w < 3; ++w) {
      if (mode == 0)
        k_tile<0><<<nt * nt * nt, 512, smem>>>(dc, dout, N);
      else
        k_tile<1><<<nt * nt * nt, 512, smem>>>(dc, dout, N);
    }
    cudaEventRecord(e0);
    for (int r = 0; r < 10; ++r) {
      if (mode == 0)
        k_tile<0><<<nt * nt * nt, 512, smem>>>(dc, dout, N);
      else
        k_tile<1><<<nt * nt * nt, 512, smem>>>(dc, dout, N);
    }
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    std::vector<double> o((size_t)nn * 15);
    cudaMemcpy(o.data(), dout, sizeof(double) * nn * 15, cudaMemcpyDeviceToHost);
    long double cs = 0;
    for (double v : o)
      cs += fabsl(v);
    printf("{\"experiment\": \"tile_proto\", \"mode\": \"%s\", \"n\": %d, \"ms_per_launch\": %.4f, \"abs_checksum\": %.10Lf, \"cuda\": \"%s\"}\n",
           mode == 0 ? "smem atomicAdd (CAS)" : "plain += (racy, cost bound only)", N, ms / 10, cs, cudaGetErrorString(err));
  }
  return 0;
}
