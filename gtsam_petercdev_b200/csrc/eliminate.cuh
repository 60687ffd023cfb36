// eliminate.cuh -- numeric multifrontal elimination and back-substitution on the symbolic plan.
//
// Replaces, per clique, EliminateCholesky (linear/HessianFactor.cpp:515-535): Scatter + HessianFactor assembly
// (:240-253, JacobianFactor::updateHessian linear/JacobianFactor.cpp:586-624, HessianFactor::updateHessian
// :349-373), choleskyPartial (base/cholesky.cpp:108-159) and split (base/SymmetricBlockMatrix.cpp:92-107); and
// OptimizeClique (linear/linearAlgorithms-inst.h:50-117) for the solve.  Cliques are processed level by level
// (all children before their parent) instead of by the reference's recursive post-order walk
// (inference/ClusterTree-inst.h:286-318).
//
// Front storage: one column-major (N+1) x (N+1) matrix per clique, upper triangle only, columns
// [frontals | separator (sorted by key) | rhs], exactly the SymmetricBlockMatrix layout of the reference.  After
// elimination rows 0..F-1 hold the conditional [R S d] and the trailing (S+1) x (S+1) block holds the Schur
// complement that the parent reads in place; nothing is copied.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "plan.h"

namespace gtb {

struct CliqueDev {
  int64_t front_off;
  int32_t F, S, ld, parent;
  int32_t fac_begin, fac_end;
  int32_t child_begin, child_end;
  int32_t pmap_off, gidx_off;
  int32_t first_key_var;  // variable id of the first frontal key (failure reporting)
  int32_t pad;
};

struct SlotDev { int32_t col0, col1, bucket, index; };

struct ElimDev {
  const CliqueDev* cliques;
  const SlotDev* slots;
  const int32_t* children;
  const int32_t* pmap;
  const int32_t* gidx;
  double* arena;
  const double* Ab[7];
  const double* damp;   // per global scalar, or nullptr
  int32_t* fail;        // min failing clique id, initialised to INT_MAX
};

__constant__ int c_M[7] = {3, 6, 2, 3, 6, 9, 3};
__constant__ int c_W[7] = {7, 13, 13, 4, 7, 10, 4};
__constant__ int c_N0[7] = {3, 6, 9, 3, 6, 9, 3};

__device__ __forceinline__ int frexp_exp(double x) {
  int e;
  (void)frexp(x, &e);
  return e;
}

// ----------------------------------------------------------------------------------------------------
// small fronts: one CTA per clique, the whole front lives in shared memory
// ----------------------------------------------------------------------------------------------------
// Adds this clique's own factors, damping and children's Schur complements into front A (column-major, leading
// dimension lds, upper triangle).  Works on shared or global memory; all threads of the CTA participate.
__device__ __forceinline__ void assemble_front(const ElimDev& E, const CliqueDev& ci, double* A, int lds) {
  const int N = ci.F + ci.S;
  const int T = blockDim.x, tid = threadIdx.x;
  // own factors, gather order (node factors, then merged children's factors)
  for (int s = ci.fac_begin; s < ci.fac_end; ++s) {
    const SlotDev sl = E.slots[s];
    const int M = c_M[sl.bucket], W = c_W[sl.bucket], n0 = c_N0[sl.bucket];
    const double* ab = E.Ab[sl.bucket] + (int64_t)sl.index * M * W;
    for (int idx = tid; idx < W * W; idx += T) {
      const int a = idx / W, b = idx - a * W;
      if (a > b) continue;
      double acc = 0.0;
      for (int r = 0; r < M; ++r) acc += ab[r * W + a] * ab[r * W + b];
      int i = (a < n0) ? sl.col0 + a : (a < W - 1 ? sl.col1 + (a - n0) : N);
      int j = (b < n0) ? sl.col0 + b : (b < W - 1 ? sl.col1 + (b - n0) : N);
      if (i > j) { const int t = i; i = j; j = t; }
      A[i + (int64_t)j * lds] += acc;
    }
    __syncthreads();
  }
  // LM damping on the frontal diagonal (replaces the appended prior factors of buildDampedSystem)
  if (E.damp) {
    const int32_t* g = E.gidx + ci.gidx_off;
    for (int f = tid; f < ci.F; f += T) A[f + (int64_t)f * lds] += E.damp[g[f]];
    __syncthreads();
  }
  // children's Schur complements, child order (inference/ClusterTree-inst.h:223-226)
  for (int q = ci.child_begin; q < ci.child_end; ++q) {
    const CliqueDev cc = E.cliques[E.children[q]];
    const int S1 = cc.S + 1;
    const double* src = E.arena + cc.front_off;
    const int32_t* pm = E.pmap + cc.pmap_off;
    for (int idx = tid; idx < S1 * S1; idx += T) {
      const int j = idx / S1, i = idx - j * S1;
      if (i > j) continue;
      const double v = src[(cc.F + i) + (int64_t)(cc.F + j) * cc.ld];
      int pi = pm[i], pj = pm[j];
      if (pi > pj) { const int t = pi; pi = pj; pj = t; }
      A[pi + (int64_t)pj * lds] += v;
    }
    __syncthreads();
  }
}

__global__ void small_front_kernel(ElimDev E, const int32_t* __restrict__ list) {
  extern __shared__ double A[];
  const int c = list[blockIdx.x];
  const CliqueDev ci = E.cliques[c];
  const int F = ci.F, N = F + ci.S, dim = N + 1;
  const int lds = dim | 1;
  const int T = blockDim.x, tid = threadIdx.x;
  for (int i = tid; i < lds * dim; i += T) A[i] = 0.0;
  __syncthreads();
  assemble_front(E, ci, A, lds);

  // right-looking partial Cholesky over the F frontal pivots: A = R^T R on the frontal block, S = R^-T B,
  // C -= S^T S  (base/cholesky.cpp:125-142)
  bool ok = true;
  for (int k = 0; k < F; ++k) {
    const double piv = A[k + k * lds];
    if (!(piv > 0.0)) { ok = false; break; }  // Eigen LLT: NumericalIssue when the pivot is <= 0
    const double r = sqrt(piv);
    __syncthreads();
    for (int j = k + 1 + tid; j < dim; j += T) A[k + j * lds] /= r;
    if (tid == 0) A[k + k * lds] = r;
    __syncthreads();
    const int cnt = dim - k - 1;
    for (int idx = tid; idx < cnt * cnt; idx += T) {
      const int jj = idx / cnt, ii = idx - jj * cnt;
      if (ii > jj) continue;
      const int i = k + 1 + ii, j = k + 1 + jj;
      A[i + j * lds] -= A[k + i * lds] * A[k + j * lds];
    }
    __syncthreads();
  }
  if (ok) {
    // ill-conditioning test on the last two pivots' binary exponents (base/cholesky.cpp:146-158)
    if (F >= 2) ok = frexp_exp(A[(F - 2) + (F - 2) * lds]) - frexp_exp(A[(F - 1) + (F - 1) * lds]) < 12;
    else if (F == 1) ok = frexp_exp(A[0]) > -12;
  }
  if (!ok && tid == 0) atomicMin(E.fail, c);
  double* dst = E.arena + ci.front_off;
  for (int idx = tid; idx < dim * dim; idx += T) {
    const int j = idx / dim, i = idx - j * dim;
    if (i <= j) dst[i + (int64_t)j * ci.ld] = A[i + j * lds];
  }
}

// ----------------------------------------------------------------------------------------------------
// large fronts: front in global memory (L2-resident), several CTAs per front
// ----------------------------------------------------------------------------------------------------
// zeroing of very large fronts is spread over many CTAs
__global__ void __launch_bounds__(256) large_zero_kernel(ElimDev E, const int32_t* __restrict__ list) {
  const int c = list[blockIdx.y];
  const CliqueDev ci = E.cliques[c];
  const int dim = ci.F + ci.S + 1;
  double2* A = reinterpret_cast<double2*>(E.arena + ci.front_off);
  const int64_t total = ((int64_t)ci.ld * dim) >> 1;  // ld is a multiple of 4
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
    A[i] = make_double2(0.0, 0.0);
}
__global__ void __launch_bounds__(256) large_factors_kernel(ElimDev E, const int32_t* __restrict__ list) {
  const int c = list[blockIdx.x];
  CliqueDev own = E.cliques[c];
  own.child_begin = own.child_end = 0;
  assemble_front(E, own, E.arena + own.front_off, own.ld);
}

// extend-add of children into large parents: blockIdx.x = (child, column chunk) pair
struct ChildWork { int32_t child; int32_t col_begin; int32_t col_end; int32_t pad; };
__global__ void __launch_bounds__(256) large_children_kernel(ElimDev E, const ChildWork* __restrict__ work) {
  const ChildWork w = work[blockIdx.x];
  const CliqueDev cc = E.cliques[w.child];
  const CliqueDev cp = E.cliques[cc.parent];
  const double* src = E.arena + cc.front_off;
  double* dst = E.arena + cp.front_off;
  const int32_t* pm = E.pmap + cc.pmap_off;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int j = w.col_begin + warp; j < w.col_end; j += nw) {
    const int pj = pm[j];
    const double* col = src + cc.F + (int64_t)(cc.F + j) * cc.ld;
    for (int i = lane; i <= j; i += 32) {
      int a = pm[i], b = pj;
      if (a > b) { const int t = a; a = b; b = t; }
      atomicAdd(dst + a + (int64_t)b * cp.ld, col[i]);
    }
  }
}

// ---- blocked partial Cholesky ---------------------------------------------------------------------
constexpr int NB = 32;    // panel height
constexpr int TW = 64;    // tile width
constexpr int LDT = NB + 4;  // shared leading dimension of a K x 64 operand tile (K contiguous), 36: conflict-free

// Panel step p, part 1: factor the NB x NB diagonal block in place (one CTA per clique).
__global__ void __launch_bounds__(256) potrf_kernel(ElimDev E, const int32_t* __restrict__ list, int p) {
  const int c = list[blockIdx.x];
  const CliqueDev ci = E.cliques[c];
  const int k0 = p * NB;
  if (k0 >= ci.F) return;
  const int kb = min(NB, ci.F - k0);
  double* A = E.arena + ci.front_off;
  const int ld = ci.ld;
  __shared__ double D[NB][NB + 1];
  const int tid = threadIdx.x;
  for (int idx = tid; idx < kb * kb; idx += blockDim.x) {
    const int j = idx / kb, i = idx - j * kb;
    D[i][j] = (i <= j) ? A[(k0 + i) + (int64_t)(k0 + j) * ld] : 0.0;
  }
  __syncthreads();
  bool ok = true;
  for (int k = 0; k < kb; ++k) {
    const double piv = D[k][k];
    if (!(piv > 0.0)) { ok = false; break; }  // Eigen LLT NumericalIssue
    const double r = sqrt(piv);
    __syncthreads();
    if (tid > k && tid < kb) D[k][tid] /= r;
    if (tid == k) D[k][k] = r;
    __syncthreads();
    const int cnt = kb - k - 1;
    for (int idx = tid; idx < cnt * cnt; idx += blockDim.x) {
      const int jj = idx / cnt, ii = idx - jj * cnt;
      if (ii <= jj) D[k + 1 + ii][k + 1 + jj] -= D[k][k + 1 + ii] * D[k][k + 1 + jj];
    }
    __syncthreads();
  }
  if (!ok) {
    // poison the block so dependants cannot look valid, and report
    for (int idx = tid; idx < kb * kb; idx += blockDim.x) {
      const int j = idx / kb, i = idx - j * kb;
      if (i <= j) A[(k0 + i) + (int64_t)(k0 + j) * ld] = __longlong_as_double(0x7ff8000000000000LL);
    }
    if (tid == 0) atomicMin(E.fail, c);
    return;
  }
  for (int idx = tid; idx < kb * kb; idx += blockDim.x) {
    const int j = idx / kb, i = idx - j * kb;
    if (i <= j) A[(k0 + i) + (int64_t)(k0 + j) * ld] = D[i][j];
  }
  if (k0 + kb == ci.F && tid == 0) {  // last panel: exponent test (base/cholesky.cpp:146-158)
    bool good;
    if (ci.F >= 2) {
      const double r2 = (kb >= 2) ? D[kb - 2][kb - 2] : A[(ci.F - 2) + (int64_t)(ci.F - 2) * ld];
      good = frexp_exp(r2) - frexp_exp(D[kb - 1][kb - 1]) < 12;
    } else {
      good = frexp_exp(D[0][0]) > -12;
    }
    if (!good) atomicMin(E.fail, c);
  }
}

// Panel step p, part 2: row panel R12 = R11^-T A12; blockIdx.y = clique, blockIdx.x = 128-column slice, one
// thread per column, R11 staged in shared memory.
__global__ void __launch_bounds__(128) trsm_kernel(ElimDev E, const int32_t* __restrict__ list, int p) {
  const int c = list[blockIdx.y];
  const CliqueDev ci = E.cliques[c];
  const int k0 = p * NB;
  if (k0 >= ci.F) return;
  const int kb = min(NB, ci.F - k0);
  const int dim = ci.F + ci.S + 1;
  const int j0 = k0 + kb + blockIdx.x * 128;
  if (j0 >= dim) return;
  double* A = E.arena + ci.front_off;
  const int ld = ci.ld;
  __shared__ double D[NB][NB + 1];
  const int tid = threadIdx.x;
  for (int idx = tid; idx < kb * kb; idx += blockDim.x) {
    const int j = idx / kb, i = idx - j * kb;
    D[i][j] = (i <= j) ? A[(k0 + i) + (int64_t)(k0 + j) * ld] : 0.0;
  }
  __syncthreads();
  const int j = j0 + tid;
  if (j >= dim) return;
  double x[NB];
  double* col = A + k0 + (int64_t)j * ld;
#pragma unroll
  for (int i = 0; i < NB; ++i) x[i] = (i < kb) ? col[i] : 0.0;
#pragma unroll
  for (int i = 0; i < NB; ++i) {
    if (i < kb) {
      double acc = x[i];
#pragma unroll
      for (int l = 0; l < NB; ++l)
        if (l < i) acc -= D[l][i] * x[l];
      x[i] = acc / D[i][i];
    }
  }
#pragma unroll
  for (int i = 0; i < NB; ++i)
    if (i < kb) col[i] = x[i];
}

// FP64 tensor-core tile: m8n8k4 DMMA (mma.sync), D = A * B + C
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// Symmetric update of the upper triangle:  C[i, j] -= sum_{k in [ka, kb)} R[k, i] * R[k, j]
// for i in [row_begin, row_end), j in [max(i, row_begin), dim), tiles of 64 x 64, operands staged through shared
// memory in K-chunks of NB, products on the FP64 tensor pipe.
//   mode 0 (frontal trailing update of panel p): ka = p*NB, kb = min(ka+NB, F), rows [kb, F)
//   mode 1 (Schur complement):                   ka = 0,    kb = F,            rows [F, dim)
// grid: x = column tile, y = row tile, z = clique in list.  256 threads = 8 warps as 2 (rows) x 4 (cols); each
// warp owns a 32 x 16 sub-tile = 4 x 2 DMMA tiles.
__global__ void __launch_bounds__(256) syrk_kernel(ElimDev E, const int32_t* __restrict__ list, int mode, int p) {
  const int c = list[blockIdx.z];
  const CliqueDev ci = E.cliques[c];
  const int F = ci.F, dim = F + ci.S + 1, ld = ci.ld;
  int ka, kb, rb, re;
  if (mode == 0) {
    ka = p * NB;
    if (ka >= F) return;
    kb = min(ka + NB, F);
    rb = kb;
    re = F;
  } else {
    ka = 0; kb = F; rb = F; re = dim;
  }
  const int r0 = rb + blockIdx.y * TW;
  const int c0 = rb + blockIdx.x * TW;
  if (blockIdx.x < blockIdx.y || r0 >= re || c0 >= dim) return;
  double* A = E.arena + ci.front_off;

  __shared__ __align__(16) double Xs[TW * LDT];
  __shared__ __align__(16) double Ys[TW * LDT];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int wr = warp >> 2, wc = warp & 3;  // warp row (0..1) and column (0..3)
  const int g = lane >> 2, t = lane & 3;
  double acc[4][2][2];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

  const int lk = tid & 31;          // k within the chunk: one warp reads 32 consecutive doubles of one column
  const int lc0 = tid >> 5;        // first of 8 columns (stride 8) this thread loads
  for (int kc = ka; kc < kb; kc += NB) {
    const int kn = min(NB, kb - kc);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int lcol = lc0 + 8 * q;
      const int xi = r0 + lcol, yj = c0 + lcol;
      Xs[lcol * LDT + lk] = (xi < dim && lk < kn) ? A[kc + lk + (int64_t)xi * ld] : 0.0;
      Ys[lcol * LDT + lk] = (yj < dim && lk < kn) ? A[kc + lk + (int64_t)yj * ld] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int ks = 0; ks < NB; ks += 4) {
      double af[4], bf[2];
#pragma unroll
      for (int a = 0; a < 4; ++a) af[a] = Xs[(wr * 32 + a * 8 + g) * LDT + ks + t];
#pragma unroll
      for (int b = 0; b < 2; ++b) bf[b] = Ys[(wc * 16 + b * 8 + g) * LDT + ks + t];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
    __syncthreads();
  }
  // epilogue: C -= acc on the upper triangle
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int i = r0 + wr * 32 + a * 8 + g;
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = c0 + wc * 16 + b * 8 + 2 * t + e;
        if (i < re && j < dim && i <= j) A[i + (int64_t)j * ld] -= acc[a][b][e];
      }
    }
}

// ----------------------------------------------------------------------------------------------------
// back-substitution: x_F = R^-1 (d - S x_S), one CTA per clique, levels from the roots down
// (linear/linearAlgorithms-inst.h:54-116)
// ----------------------------------------------------------------------------------------------------
__global__ void backsolve_kernel(ElimDev E, const int32_t* __restrict__ list, double* __restrict__ delta) {
  extern __shared__ double sh[];
  const int c = list[blockIdx.x];
  const CliqueDev ci = E.cliques[c];
  const int F = ci.F, S = ci.S, N = F + S, ld = ci.ld;
  const double* A = E.arena + ci.front_off;
  const int32_t* g = E.gidx + ci.gidx_off;
  double* rhs = sh;       // F
  double* xs = sh + F;    // S
  const int T = blockDim.x, tid = threadIdx.x;
  for (int j = tid; j < S; j += T) xs[j] = delta[g[F + j]];
  __syncthreads();
  for (int i = tid; i < F; i += T) {
    double acc = A[i + (int64_t)N * ld];
    for (int j = 0; j < S; ++j) acc -= A[i + (int64_t)(F + j) * ld] * xs[j];
    rhs[i] = acc;
  }
  __syncthreads();
  for (int k = F - 1; k >= 0; --k) {
    const double xk = rhs[k] / A[k + (int64_t)k * ld];
    __syncthreads();
    if (tid == 0) rhs[k] = xk;
    for (int i = tid; i < k; i += T) rhs[i] -= A[i + (int64_t)k * ld] * xk;
    __syncthreads();
  }
  for (int i = tid; i < F; i += T) delta[g[i]] = rhs[i];
}

}  // namespace gtb
