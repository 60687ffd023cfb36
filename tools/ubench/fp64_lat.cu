// Micro-benchmarks behind DESIGN.md's role-A cycle budget: dependent-issue latency and per-SM throughput of the FP64
// instructions on the panel step's critical path (DFMA, rsqrt, DMMA m8n8k4), one CTA on one SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_lat fp64_lat.cu && ./fp64_lat
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double rsqrt_nr(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double t = x * y, h = 0.5 * y;
  double e = fma(-t, y, 1.0);
  y = fma(h, e, y);
  t = x * y; h = 0.5 * y;
  e = fma(-t, y, 1.0);
  return fma(h, e, y);
}
template <int MODE, int CHAINS>
__global__ void k(double* out, long long* cyc, int iters, double seed) {
  double a[CHAINS], c1[CHAINS];
  for (int i = 0; i < CHAINS; ++i) { a[i] = seed + i + threadIdx.x * 1e-3; c1[i] = 0.0; }
  const double b = seed * 0.5 + 1e-9 * threadIdx.x;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
      if (MODE == 0) a[i] = fma(a[i], b, 1.0);
      if (MODE == 1) a[i] = rsqrt(a[i]) + 2.0;
      if (MODE == 2) a[i] = rsqrt_nr(a[i]) + 2.0;
      if (MODE == 3) dmma(a[i], c1[i], b, b);
      if (MODE == 4) a[i] = sqrt(a[i]) + 2.0;
      if (MODE == 5) a[i] = 1.0 / a[i] + 2.0;
    }
  }
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < CHAINS; ++i) s += a[i] + c1[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MODE, int CHAINS>
void run(const char* name, int threads) {
  double* out; long long* cyc;
  cudaMalloc(&out, 4096 * sizeof(double)); cudaMalloc(&cyc, 8 * sizeof(long long));
  const int iters = 2000;
  k<MODE, CHAINS><<<1, threads>>>(out, cyc, iters, 1.25);
  k<MODE, CHAINS><<<1, threads>>>(out, cyc, iters, 1.25);
  long long h; cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-28s warps=%2d chains=%d : %7.1f cycles per round, %6.1f per instruction-per-warp\n", name, threads / 32, CHAINS,
         (double)h / iters, (double)h / iters / CHAINS);
  cudaFree(out); cudaFree(cyc);
}
int main() {
  run<0, 1>("DFMA dependent", 32);
  run<0, 8>("DFMA 8 chains", 32);
  run<0, 8>("DFMA 8 chains", 256);
  run<1, 1>("rsqrt(double) dependent", 32);
  run<2, 1>("rsqrt.approx+2 Newton dep.", 32);
  run<4, 1>("sqrt(double) dependent", 32);
  run<5, 1>("1/x dependent", 32);
  run<3, 1>("DMMA m8n8k4 dependent", 32);
  run<3, 4>("DMMA 4 chains", 32);
  run<3, 8>("DMMA 8 chains", 32);
  run<3, 8>("DMMA 8 chains", 128);
  run<3, 8>("DMMA 8 chains", 256);
  run<3, 8>("DMMA 8 chains", 512);
  double x = 3.7, y; 
  return 0;
}
