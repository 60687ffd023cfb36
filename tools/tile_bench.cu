// Inner loop of tile_update in isolation (operands resident in shared memory, no global traffic): which warp tiling and
// FP64 mma shape keeps the DMMA pipe busy with 8 or 16 warps per SM.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int NB = 32, TW = 64, LDT = NB + 4;
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
// V0: 8 warps 2x4, warp tile 32x16, m8n8k4
// V1: 8 warps 2x4, warp tile 32x16, m16n8k8 (2 m-tiles x 2 n-tiles, 4 k-steps of 8)
// V2: 4 warps 2x2, warp tile 32x32, m8n8k4 (4x4 tiles: 8 loads per 16 mma)
template <int V>
__global__ void __launch_bounds__(256) k(double* out, int chunks) {
  extern __shared__ double sm[];
  double* X = sm; double* Y = sm + TW * LDT;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  for (int i = tid; i < 2 * TW * LDT; i += blockDim.x) sm[i] = 1e-3 * (i % 97);
  __syncthreads();
  double acc[4][4][2];
  for (int a = 0; a < 4; ++a) for (int b = 0; b < 4; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
  for (int ch = 0; ch < chunks; ++ch) {
    if (V == 0) {
      const int wr = warp >> 2, wc = warp & 3;
#pragma unroll
      for (int ks = 0; ks < NB; ks += 4) {
        double af[4], bf[2];
#pragma unroll
        for (int a = 0; a < 4; ++a) af[a] = X[(wr * 32 + a * 8 + g) * LDT + ks + t];
#pragma unroll
        for (int b = 0; b < 2; ++b) bf[b] = Y[(wc * 16 + b * 8 + g) * LDT + ks + t];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 2; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
      }
    } else if (V == 1) {
      const int wr = warp >> 2, wc = warp & 3;
#pragma unroll
      for (int ks = 0; ks < NB; ks += 8) {
        double af[2][4], bf[2][2];
#pragma unroll
        for (int a = 0; a < 2; ++a) {
          af[a][0] = X[(wr * 32 + a * 16 + g) * LDT + ks + t];
          af[a][1] = X[(wr * 32 + a * 16 + g + 8) * LDT + ks + t];
          af[a][2] = X[(wr * 32 + a * 16 + g) * LDT + ks + t + 4];
          af[a][3] = X[(wr * 32 + a * 16 + g + 8) * LDT + ks + t + 4];
        }
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          bf[b][0] = Y[(wc * 16 + b * 8 + g) * LDT + ks + t];
          bf[b][1] = Y[(wc * 16 + b * 8 + g) * LDT + ks + t + 4];
        }
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
          for (int b = 0; b < 2; ++b) dmma1688(&acc[a][b * 2][0], af[a], bf[b]);   // 4 contiguous doubles: acc[a][2b][0..1], acc[a][2b+1][0..1]
      }
    } else {
      if (warp < 4) {
        const int wr = warp >> 1, wc = warp & 1;
#pragma unroll
        for (int ks = 0; ks < NB; ks += 4) {
          double af[4], bf[4];
#pragma unroll
          for (int a = 0; a < 4; ++a) af[a] = X[(wr * 32 + a * 8 + g) * LDT + ks + t];
#pragma unroll
          for (int b = 0; b < 4; ++b) bf[b] = Y[(wc * 32 + b * 8 + g) * LDT + ks + t];
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
        }
      }
    }
    if ((ch & 7) == 7) __syncthreads();
  }
  double s = 0;
  for (int a = 0; a < 4; ++a) for (int b = 0; b < 4; ++b) s += acc[a][b][0] + acc[a][b][1];
  out[blockIdx.x * blockDim.x + tid] = s;
}
int main() {
  double* out; cudaMalloc(&out, 148 * 4 * 256 * sizeof(double));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int chunks = 4000, smem = 2 * TW * LDT * 8;
  for (int v = 0; v < 3; ++v)
    for (int per_sm = 1; per_sm <= 4; per_sm *= 2) {
      float ms = 0;
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        if (v == 0) k<0><<<148 * per_sm, 256, smem>>>(out, chunks);
        if (v == 1) k<1><<<148 * per_sm, 256, smem>>>(out, chunks);
        if (v == 2) k<2><<<148 * per_sm, 256, smem>>>(out, chunks);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
      }
      printf("variant %d, %d CTA/SM: %.2f TFLOP/s (%s)\n", v, per_sm, 148.0 * per_sm * chunks * 64.0 * 64 * 32 * 2 / ms / 1e9,
             cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
