// Latency probes for the role-A critical path of front_step_kernel: dependent DFMA / DMMA / shuffle / rsqrt chains and
// variants of the 32 x 32 in-register Cholesky, one warp, cycles from clock64().
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#include <cmath>
constexpr int NB = 32, LDT = 36;

__device__ __forceinline__ double fast_rsqrt(double a) {
  // float seed + two Newton steps in double
  double x = (double)rsqrtf((float)a);
  const double h = 0.5 * a;
  x = x * (1.5 - h * x * x);
  x = x * (1.5 - h * x * x);
  return x;
}

template <int VAR>
__global__ void potrf(const double* __restrict__ Ain, double* out, long long* cyc) {
  __shared__ double Ds[NB * LDT];
  const int lane = threadIdx.x;
  for (int c = 0; c < NB; ++c) Ds[c * LDT + lane] = Ain[c * NB + lane];
  __syncwarp();
  double d[NB];
#pragma unroll
  for (int i = 0; i < NB; ++i) d[i] = Ds[lane * LDT + i];
  const long long t0 = clock64();
  if (VAR == 0 || VAR == 1) {
    double pk = __shfl_sync(0xffffffffu, d[0], 0);
    double rinv = VAR == 0 ? rsqrt(pk) : fast_rsqrt(pk);
#pragma unroll
    for (int k = 0; k < NB; ++k) {
      if (lane > k) d[k] *= rinv;
      else if (lane == k) d[k] = pk * rinv;
      const double ukl = d[k];
      double pk_next = 1.0, rinv_next = 1.0;
      if (k + 1 < NB) {
        const double uki = __shfl_sync(0xffffffffu, d[k], k + 1);
        if (k + 1 <= lane) d[k + 1] -= uki * ukl;
        pk_next = __shfl_sync(0xffffffffu, d[k + 1], k + 1);
        rinv_next = VAR == 0 ? rsqrt(pk_next) : fast_rsqrt(pk_next);
      }
#pragma unroll
      for (int i = k + 2; i < NB; ++i) {
        const double uki = __shfl_sync(0xffffffffu, d[k], i);
        if (i <= lane) d[i] -= uki * ukl;
      }
      pk = pk_next;
      rinv = rinv_next;
    }
  } else if (VAR == 2) {
    // left-looking by 8-row blocks through shared memory: lane = column, rows of a block solved with values of the block
    // rows above read from shared memory (broadcast), no shuffles in the long loop
    // (layout: Us[row * LDT + col])
    __shared__ double Us[NB * LDT];
#pragma unroll
    for (int k = 0; k < NB; ++k) {
      // d[k] currently holds A(k, lane) minus contributions of rows < k (applied eagerly below)
      const double pk = __shfl_sync(0xffffffffu, d[k], k);
      const double rinv = fast_rsqrt(pk);
      d[k] = (lane >= k) ? d[k] * rinv : 0.0;
      Us[k * LDT + lane] = d[k];
      __syncwarp();
      const double ukl = d[k];
#pragma unroll
      for (int i = k + 1; i < NB; ++i) {
        const double uki = Us[k * LDT + i];   // broadcast read
        if (i <= lane) d[i] -= uki * ukl;
      }
    }
  }
  const long long t1 = clock64();
#pragma unroll
  for (int i = 0; i < NB; ++i) out[i * NB + lane] = (i <= lane) ? d[i] : 0.0;
  if (lane == 0) cyc[VAR] = t1 - t0;
}

// dependent chains: 0 = DFMA, 1 = shfl double, 2 = rsqrt(double), 3 = fast_rsqrt, 4 = DMMA m8n8k4, 5 = smem store+load
__global__ void chains(double* out, long long* cyc, int n) {
  __shared__ double buf[64];
  const int lane = threadIdx.x;
  double x = 1.0 + lane * 1e-3, y = 1.0000001;
  long long t0, t1;
  t0 = clock64();
  for (int i = 0; i < n; ++i) x = fma(x, y, 1e-9);
  t1 = clock64(); if (lane == 0) cyc[0] = t1 - t0;
  t0 = clock64();
  for (int i = 0; i < n; ++i) x = __shfl_sync(0xffffffffu, x, (lane + 1) & 31);
  t1 = clock64(); if (lane == 0) cyc[1] = t1 - t0;
  x = fabs(x) + 1.0;
  t0 = clock64();
  for (int i = 0; i < n; ++i) x = rsqrt(x) + 1.0;
  t1 = clock64(); if (lane == 0) cyc[2] = t1 - t0;
  t0 = clock64();
  for (int i = 0; i < n; ++i) x = fast_rsqrt(x) + 1.0;
  t1 = clock64(); if (lane == 0) cyc[3] = t1 - t0;
  double c0 = x, c1 = y;
  t0 = clock64();
  for (int i = 0; i < n; ++i)
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(y), "d"(y));
  t1 = clock64(); if (lane == 0) cyc[4] = t1 - t0;
  x = c0 + c1;
  t0 = clock64();
  for (int i = 0; i < n; ++i) { buf[lane] = x; __syncwarp(); x = buf[(lane + 1) & 31] + 1.0; __syncwarp(); }
  t1 = clock64(); if (lane == 0) cyc[5] = t1 - t0;
  t0 = clock64();
  for (int i = 0; i < n; ++i) x = 1.0 / x + 1.0;
  t1 = clock64(); if (lane == 0) cyc[6] = t1 - t0;
  t0 = clock64();
  for (int i = 0; i < n; ++i) x = sqrt(x) + 1.0;
  t1 = clock64(); if (lane == 0) cyc[7] = t1 - t0;
  out[lane] = x;
}

int main() {
  std::vector<double> A(NB * NB), R(NB * NB);
  // SPD: A = M^T M + 32 I, stored column-major (A[c*NB + r]); symmetric so layout does not matter
  std::vector<double> M(NB * NB);
  for (int i = 0; i < NB * NB; ++i) M[i] = sin(0.37 * i) ;
  for (int i = 0; i < NB; ++i) for (int j = 0; j < NB; ++j) {
    double s = (i == j) ? 32.0 : 0.0;
    for (int k = 0; k < NB; ++k) s += M[k * NB + i] * M[k * NB + j];
    A[j * NB + i] = s;
  }
  double *dA, *dO; long long* dC;
  cudaMalloc(&dA, sizeof(double) * NB * NB); cudaMalloc(&dO, sizeof(double) * NB * NB * 4); cudaMalloc(&dC, 64 * 8);
  cudaMemcpy(dA, A.data(), sizeof(double) * NB * NB, cudaMemcpyHostToDevice);
  long long cyc[16];
  std::vector<double> ref;
  for (int rep = 0; rep < 2; ++rep) {
    potrf<0><<<1, 32>>>(dA, dO, dC);
    potrf<1><<<1, 32>>>(dA, dO + NB * NB, dC);
    potrf<2><<<1, 32>>>(dA, dO + 2 * NB * NB, dC);
    cudaDeviceSynchronize();
  }
  cudaMemcpy(cyc, dC, 8 * 3, cudaMemcpyDeviceToHost);
  std::vector<double> O(3 * NB * NB);
  cudaMemcpy(O.data(), dO, sizeof(double) * 3 * NB * NB, cudaMemcpyDeviceToHost);
  // host Cholesky (upper): R^T R = A
  std::vector<double> U(A);
  for (int k = 0; k < NB; ++k) {
    U[k * NB + k] = sqrt(U[k * NB + k]);
    for (int j = k + 1; j < NB; ++j) U[j * NB + k] /= U[k * NB + k];
    for (int j = k + 1; j < NB; ++j) for (int i = k + 1; i <= j; ++i) U[j * NB + i] -= U[i * NB + k] * U[j * NB + k];
  }
  for (int v = 0; v < 3; ++v) {
    double err = 0;
    for (int j = 0; j < NB; ++j) for (int i = 0; i <= j; ++i) err = fmax(err, fabs(O[v * NB * NB + i * NB + j] - U[j * NB + i]));
    printf("potrf variant %d: %lld cycles, max abs err vs host %.3e\n", v, cyc[v], err);
  }
  const int n = 256;
  for (int rep = 0; rep < 2; ++rep) { chains<<<1, 32>>>(dO, dC, n); cudaDeviceSynchronize(); }
  cudaMemcpy(cyc, dC, 8 * 8, cudaMemcpyDeviceToHost);
  const char* names[8] = {"DFMA", "SHFL f64", "rsqrt(f64)+add", "fast_rsqrt+add", "DMMA m8n8k4", "smem st/ld + 2 syncwarp + add", "1/x + add", "sqrt + add"};
  for (int i = 0; i < 8; ++i) printf("dependent %-32s %.1f cycles\n", names[i], (double)cyc[i] / n);
  return 0;
}
