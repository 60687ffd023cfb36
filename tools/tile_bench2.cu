// tile_update of eliminate.cuh in isolation: one 64 x 64 Schur tile per CTA, K = 512, operands streamed from a
// column-major front in global memory (L2-resident), cp.async ring of kStages K-chunks.  Which part of the kernel's
// 3.6k cycles per chunk (vs 2.05k for the MMA loop alone) is staging?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define GTB_TILE_BENCH 1
#include "../gtsam_petercdev_b200/csrc/eliminate.cuh"
using namespace gtb;
__global__ void __launch_bounds__(256, 2) k(double* A, int ld, int dim, int F, int ntile_side) {
  extern __shared__ __align__(16) double sm[];
  const int t = blockIdx.x % (ntile_side * ntile_side);
  const int ty = t / ntile_side, tx = t % ntile_side;
  tile_update(A, ld, dim, F + ty * TW, F + tx * TW, dim, 0, F, sm);
}
int main() {
  const int F = 512, S = 1055, dim = F + S + 1, ld = (dim + 3) & ~3;
  double* A;
  cudaMalloc(&A, sizeof(double) * (size_t)ld * dim);
  cudaMemset(A, 0, sizeof(double) * (size_t)ld * dim);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmemTile);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int grid : {74, 148, 153, 296, 592, 1184}) {
    float ms = 0;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      k<<<grid, 256, kStepSmemTile>>>(A, ld, dim, F, 12);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      cudaEventElapsedTime(&ms, e0, e1);
    }
    printf("grid %4d: %.1f us, %.2f TFLOP/s, %.0f cycles per chunk per CTA-wave (%s)\n", grid, ms * 1e3,
           grid * 64.0 * 64 * F * 2 / ms / 1e9, ms * 1e-3 * 1.965e9 / (F / 32) / ((grid + 295) / 296 > 0 ? 1 : 1),
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
