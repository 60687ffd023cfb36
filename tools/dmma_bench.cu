// Throughput of the FP64 tensor-core instruction shapes on this GPU (mma.sync, independent accumulator chains).
#include <cstdio>
#include <cuda_runtime.h>
template <int SHAPE>
__global__ void k(double* out, int iters) {
  double c[8][4];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.0;
  double a[8], b[4];
  for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3 + i;
  for (int i = 0; i < 4; ++i) b[i] = threadIdx.x * 1e-4 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      if (SHAPE == 0)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                     : "+d"(c[q][0]), "+d"(c[q][1]) : "d"(a[0]), "d"(b[0]));
      else if (SHAPE == 1)
        asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                     : "+d"(c[q][0]), "+d"(c[q][1]), "+d"(c[q][2]), "+d"(c[q][3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
      else if (SHAPE == 2)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                     : "+d"(c[q][0]), "+d"(c[q][1]), "+d"(c[q][2]), "+d"(c[q][3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
      else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
                     : "+d"(c[q][0]), "+d"(c[q][1]), "+d"(c[q][2]), "+d"(c[q][3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                       "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
    }
  }
  double s = 0;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// plain DFMA throughput for comparison
__global__ void kf(double* out, int iters) {
  double c[16];
  for (int i = 0; i < 16; ++i) c[i] = i;
  double a = threadIdx.x * 1e-3, b = 1.0000001;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int q = 0; q < 16; ++q) c[q] = fma(c[q], b, a);
  double s = 0;
  for (int i = 0; i < 16; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  double* out; cudaMalloc(&out, 148 * 8 * 1024 * sizeof(double));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  const double flops[4] = {512, 1024, 2048, 4096};
  for (int warps = 4; warps <= 16; warps *= 2)
    for (int s = 0; s < 4; ++s) {
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        if (s == 0) k<0><<<148 * 2, warps * 32>>>(out, iters);
        if (s == 1) k<1><<<148 * 2, warps * 32>>>(out, iters);
        if (s == 2) k<2><<<148 * 2, warps * 32>>>(out, iters);
        if (s == 3) k<3><<<148 * 2, warps * 32>>>(out, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
      }
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      printf("shape %d warps/CTA %2d (2 CTAs/SM): %.2f TFLOP/s  err=%s\n", s, warps, 148.0 * 2 * warps * iters * 8 * flops[s] / ms / 1e9,
             cudaGetErrorString(cudaGetLastError()));
    }
  for (int rep = 0; rep < 2; ++rep) { cudaEventRecord(e0); kf<<<148 * 2, 512>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1); }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("DFMA: %.2f TFLOP/s\n", 148.0 * 2 * 512 * iters * 16 * 2 / ms / 1e9);
  return 0;
}
