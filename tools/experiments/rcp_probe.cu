// rcp_probe.cu — accuracy of the fp64 reciprocal used by the element operators (include/b2f/element_ops.cuh fast_rcp):
// seed rcp.approx.ftz.f64 (MUFU.RCP64H), one cubic correction y (1 + e + e^2) against two Newton steps, over 2^26 samples of
// [2^-40, 2^40).  Prints the largest relative errors (vs 1.0 / x, correctly rounded).  nvcc -arch=sm_100a rcp_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ double seed(double x)
{
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  return y;
}

__global__ void probe(double* out)
{
  __shared__ double s[3][256];
  double m0 = 0, m1 = 0, m2 = 0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < (1ull << 26); i += (uint64_t)gridDim.x * blockDim.x) {
    // exponent in [-40, 40), mantissa from a multiplicative hash
    const uint64_t h = i * 0x9E3779B97F4A7C15ull;
    const int e = (int)(h >> 58) + (int)((h >> 52) & 15) - 40;
    const double x = ldexp(1.0 + (double)(h & ((1ull << 52) - 1)) * 0x1p-52, e);
    const double exact = 1.0 / x;
    const double y0 = seed(x);
    double er = fma(-x, y0, 1.0);
    const double yc = fma(y0, fma(er, er, er), y0);
    double y1 = fma(y0, er, y0);
    er = fma(-x, y1, 1.0);
    const double y2 = fma(y1, er, y1);
    m0 = fmax(m0, fabs(y0 - exact) / exact);
    m1 = fmax(m1, fabs(yc - exact) / exact);
    m2 = fmax(m2, fabs(y2 - exact) / exact);
  }
  s[0][threadIdx.x] = m0, s[1][threadIdx.x] = m1, s[2][threadIdx.x] = m2;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 256; ++k)
      for (int q = 0; q < 3; ++q)
        s[q][0] = fmax(s[q][0], s[q][k]);
    for (int q = 0; q < 3; ++q)
      out[blockIdx.x * 3 + q] = s[q][0];
  }
}

int main()
{
  const int nb = 592;
  double* d = nullptr;
  if (cudaMalloc(&d, sizeof(double) * 3 * nb) != cudaSuccess) {
    std::printf("no device\n");
    return 3;
  }
  probe<<<nb, 256>>>(d);
  static double h[3 * 592];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  double m[3] = { 0, 0, 0 };
  for (int b = 0; b < nb; ++b)
    for (int q = 0; q < 3; ++q)
      m[q] = h[3 * b + q] > m[q] ? h[3 * b + q] : m[q];
  std::printf("rcp_probe: seed %.3e  cubic %.3e  two-newton %.3e (max relative error)\n", m[0], m[1], m[2]);
  return 0;
}
