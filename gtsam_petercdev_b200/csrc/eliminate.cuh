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
  int32_t dfac_off;       // large fronts: offset of the factored diagonal blocks' side buffer (doubles)
  int32_t flags;          // kCliqueStar: a point clique whose factors are SFM factors of distinct cameras
  int32_t flag_off;       // large fronts whose Schur step is split over K: first hand-shake flag of its tiles (ElimDev::tile_flags)
};
constexpr int32_t kCliqueStar = 1;
constexpr int32_t kCliqueFuseUp = 2;   // large front whose Schur-complement step adds C - S^T S straight into its (large) parent

struct SlotDev { int32_t col0, col1, bucket, index; };

struct ElimDev {
  const CliqueDev* cliques;
  const SlotDev* slots;
  const int32_t* children;
  const int32_t* pmap;
  const int32_t* gidx;
  double* arena;
  double* dfac;         // side buffer for the factored diagonal blocks of large fronts
  const double* Ab[7];
  const double* damp;   // per global scalar, or nullptr
  int32_t* fail;        // min failure code (fail_code), initialised to INT_MAX
  const int32_t* orig_off;   // per clique: its members (reference cliques merged into it for the numeric factorisation) ...
  const int32_t* orig_end;   // ... as the frontal scalar where each member's pivots end
  long long* dbg;       // optional cycle probe of role A (GTB_DEBUG & 8)
  int32_t dbg_clique, dbg_child;
  int32_t debug;        // experiment switches (GTB_DEBUG env): 1 skip tile tasks, 2 skip role-A column work, 4 skip role A
  // Ordered hand-over (default): every entry of a large front receives at most ONE contribution per launch, or its
  // contributions inside a launch are serialised in a fixed order, so that a factorisation repeats bit for bit.
  int32_t ordered;      // 0 (GTB_HANDOVER_ATOMIC=1): the round-1 behaviour, contributions meet through unordered FP64 atomics
  int32_t epoch32;      // 32 x the factorisation counter: base value of the K-split hand-shake flags of this factorisation
  int32_t* tile_flags;  // one flag per tile of every K-split Schur step (CliqueDev::flag_off)
  double* pair_scratch; // partial 9 x 9 blocks of camera-pair groups that are cut into several segments (kPairSlot doubles each)
};

__constant__ int c_M[7] = {3, 6, 2, 3, 6, 9, 3};
__constant__ int c_W[7] = {7, 13, 13, 4, 7, 10, 4};
__constant__ int c_N0[7] = {3, 6, 9, 3, 6, 9, 3};

// Failure code: clique id and the index of the failing member inside it (a clique that was not amalgamated has one member)
__device__ __forceinline__ int fail_code(int c, int member = 0) { return (c << 8) | member; }

__device__ __forceinline__ int frexp_exp(double x) {
  int e;
  (void)frexp(x, &e);
  return e;
}

// Programmatic dependent launch: the kernels of the level loop form one long dependent chain on one stream.  Each of
// them lets its successor be scheduled as soon as all of its own CTAs have started (trigger), reads its static
// schedule (task list, clique descriptors), and only then waits for the predecessor grid to finish and flush.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Fire-and-forget FP64 add to global memory (RED.E.ADD.F64): the thread does not wait for the result.  Written as PTX
// because the compiler turns an unused-result atomicAdd into the waiting form (ATOM) as soon as a fence follows it in
// the same function -- seen in the SASS of schur_step_kernel, where it cost 1.1 ms per step.  A later __threadfence()
// still orders these adds before whatever the thread writes after it.
// 1 / sqrt(x) without the special-case branch of the library routine: the branch ends the basic block, and with it the
// overlap of the Newton steps with the independent updates around them (the 8 x 8 register Cholesky of the panel step is one
// dependent chain of eight of these).  Hardware seed (2^-22) and two Newton steps; x <= 0, NaN and denormals give NaN or
// infinity, which the pivot test that precedes every use reports as a failed factorisation.
__device__ __forceinline__ double rsqrt_chain(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double t = x * y, h = 0.5 * y;
  double e = fma(-t, y, 1.0);
  y = fma(h, e, y);
  t = x * y;
  h = 0.5 * y;
  e = fma(-t, y, 1.0);
  return fma(h, e, y);
}
__device__ __forceinline__ void red_add(double* p, double v) {
  asm volatile("red.global.add.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

// one 32-byte column of a point front in a single request (LDG.256, read-only path)
struct Col4 { double x, y, z, w; };
__device__ __forceinline__ Col4 ld_col(const double* p) {
  Col4 c;
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(c.x), "=d"(c.y), "=d"(c.z), "=d"(c.w) : "l"(p));
  return c;
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

  // partial Cholesky over the F frontal pivots: A = R^T R on the frontal block, S = R^-T B, C -= S^T S
  // (base/cholesky.cpp:125-142).  The pivot loop only touches the F frontal rows (2-D thread grid, no integer division,
  // two barriers per pivot); the Schur complement is one barrier-free pass in which every entry subtracts its F products
  // in pivot order, which is the rounding sequence of a right-looking update of the whole trailing matrix.
  bool ok = true;
  const int tx = tid & 31, ty = tid >> 5, ny = T >> 5;
  for (int k = 0; k < F; ++k) {
    __syncthreads();                           // the update by pivot k-1 (or the assembly) is complete
    const double piv = A[k + k * lds];
    if (!(piv > 0.0)) { ok = false; break; }  // Eigen LLT: NumericalIssue when the pivot is <= 0 (uniform: all threads leave)
    const double r = sqrt(piv);
    for (int j = k + 1 + tid; j < dim; j += T) A[k + j * lds] /= r;
    __syncthreads();                           // row k is scaled and everybody has read the pivot
    if (tid == 0) A[k + k * lds] = r;
    for (int i = k + 1 + ty; i < F; i += ny) {
      const double aki = A[k + i * lds];
      for (int j = i + tx; j < dim; j += 32) A[i + j * lds] -= aki * A[k + j * lds];
    }
  }
  __syncthreads();
  if (ok) {
    for (int j = F + ty; j < dim; j += ny) {
      for (int i = F + tx; i <= j; i += 32) {
        double acc = A[i + j * lds];
        for (int k = 0; k < F; ++k) acc -= A[k + i * lds] * A[k + j * lds];
        A[i + j * lds] = acc;
      }
    }
  }
  __syncthreads();
  if (ok) {
    // ill-conditioning test on the last two pivots' binary exponents (base/cholesky.cpp:146-158)
    if (F >= 2) ok = frexp_exp(A[(F - 2) + (F - 2) * lds]) - frexp_exp(A[(F - 1) + (F - 1) * lds]) < 12;
    else if (F == 1) ok = frexp_exp(A[0]) > -12;
  }
  if (!ok && tid == 0) atomicMin(E.fail, fail_code(c));
  double* dst = E.arena + ci.front_off;
  for (int idx = tid; idx < dim * dim; idx += T) {
    const int j = idx / dim, i = idx - j * dim;
    if (i <= j) dst[i + (int64_t)j * ci.ld] = A[i + j * lds];
  }
}

// ----------------------------------------------------------------------------------------------------
// thin fronts (F <= 12: a point, a pose, a camera): packed upper triangle in shared memory, column loops (no integer
// division), reciprocal pivots.  WARP = true: one warp per clique, kThinWarps cliques per CTA; else one CTA per clique.
// ----------------------------------------------------------------------------------------------------
constexpr int kThinWarps = 8;
__constant__ unsigned char c_pairA[91] = {
    0, 0, 1, 0, 1, 2, 0, 1, 2, 3, 0, 1, 2, 3, 4, 0, 1, 2, 3, 4, 5, 0, 1, 2, 3, 4, 5, 6, 0, 1, 2, 3, 4, 5, 6, 7,
    0, 1, 2, 3, 4, 5, 6, 7, 8, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10,
    0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12};
__constant__ unsigned char c_pairB[91] = {
    0, 1, 1, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 4, 5, 5, 5, 5, 5, 5, 6, 6, 6, 6, 6, 6, 6, 7, 7, 7, 7, 7, 7, 7, 7,
    8, 8, 8, 8, 8, 8, 8, 8, 8, 9, 9, 9, 9, 9, 9, 9, 9, 9, 9, 10, 10, 10, 10, 10, 10, 10, 10, 10, 10, 10,
    11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 12, 12, 12, 12, 12, 12, 12, 12, 12, 12, 12, 12, 12};

// SFM factor (W = 13: camera 0..8, point 9..11, rhs 12) column pairs a <= b: the first 81 involve the camera block (unique to
// the factor inside a point clique), the last 10 only the point and the rhs (shared by all factors of the clique)
__constant__ unsigned char c_starA[91] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 4, 4, 4, 4, 4, 4, 4, 4, 4, 5, 5, 5, 5, 5, 5, 5, 5, 6, 6, 6, 6, 6, 6, 6, 7, 7, 7, 7, 7, 7, 8, 8, 8, 8, 8, 9, 9, 9, 9, 10, 10, 10, 11, 11, 12};
__constant__ unsigned char c_starB[91] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 4, 5, 6, 7, 8, 9, 10, 11, 12, 5, 6, 7, 8, 9, 10, 11, 12, 6, 7, 8, 9, 10, 11, 12, 7, 8, 9, 10, 11, 12, 8, 9, 10, 11, 12, 9, 10, 11, 12, 10, 11, 12, 11, 12, 12};

__device__ __forceinline__ int tri(int i, int j) { return ((j * (j + 1)) >> 1) + i; }   // i <= j

constexpr int kThinStage = 416;   // doubles of factor rows staged per group and pass (16 SFM factors)

// packed index -> column j (largest j with j(j+1)/2 <= idx)
__device__ __forceinline__ int tri_col(int idx) {
  int j = (int)((sqrtf(8.0f * (float)idx + 1.0f) - 1.0f) * 0.5f);
  while (((j + 1) * (j + 2)) >> 1 <= idx) ++j;
  while ((j * (j + 1)) >> 1 > idx) --j;
  return j;
}

template <bool WARP>
__global__ void __launch_bounds__(WARP ? kThinWarps * 32 : 512)
thin_front_kernel(ElimDev E, const int32_t* __restrict__ list, int count, int group_doubles) {
  extern __shared__ double sh[];
  __shared__ unsigned char s_star[2 * 91];   // divergent indexing of a __constant__ table serialises: keep a copy here
  for (int q = threadIdx.x; q < 91; q += blockDim.x) { s_star[q] = c_starA[q]; s_star[91 + q] = c_starB[q]; }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = WARP ? blockIdx.x * kThinWarps + warp : blockIdx.x;
  if (slot >= count) return;
  const int gsize = WARP ? 32 : (int)blockDim.x;
  const int gtid = WARP ? lane : (int)threadIdx.x;
  auto gsync = [&]() { if (WARP) __syncwarp(); else __syncthreads(); };
  double* P = sh + (WARP ? warp * (group_doubles + kThinStage) : 0);
  double* stage = P + group_doubles;
  const int c = list[slot];
  const CliqueDev ci = E.cliques[c];
  const int F = ci.F, N = F + ci.S, dim = N + 1;
  const int plen = (dim * (dim + 1)) >> 1;
  for (int i = gtid; i < plen; i += gsize) P[i] = 0.0;
  const int nfac = ci.fac_end - ci.fac_begin;
  const bool star = (ci.flags & kCliqueStar) != 0;
  if (star) {
    // point clique: every factor is an SFM factor of a different camera, so the 81 column pairs that involve the camera
    // block of a factor touch entries no other factor touches: all factors accumulate at once, no barrier in between.
    // The 10 point / rhs pairs are shared: one thread each, factors in gather order (deterministic sums).
    int* scol = reinterpret_cast<int*>(stage + nfac * 26);   // camera block column of each factor
    gsync();
    for (int t = gtid; t < nfac * 26; t += gsize) {
      const int s = t / 26, r = t - s * 26;
      const SlotDev sl = E.slots[ci.fac_begin + s];
      stage[t] = E.Ab[2][(int64_t)sl.index * 26 + r];
      if (r == 0) scol[s] = sl.col0;
    }
    gsync();
    for (int t = gtid; t < nfac * 81; t += gsize) {
      const int s = t / 81, q = t - s * 81;
      const int a = s_star[q], b = s_star[91 + q];
      const double* ab = stage + s * 26;
      const double acc = ab[a] * ab[b] + ab[13 + a] * ab[13 + b];
      const int i = scol[s] + a;
      const int j = (b < 9) ? scol[s] + b : (b < 12 ? b - 9 : N);
      P[(i <= j) ? tri(i, j) : tri(j, i)] += acc;
    }
    if (gtid < 10) {
      const int a = s_star[81 + gtid], b = s_star[91 + 81 + gtid];
      double acc = 0.0;
      for (int s = 0; s < nfac; ++s) {
        const double* ab = stage + s * 26;
        acc += ab[a] * ab[b] + ab[13 + a] * ab[13 + b];
      }
      const int i = (a < 12) ? a - 9 : N, j = (b < 12) ? b - 9 : N;
      P[tri(i, j)] += acc;
    }
  }
  // own factors, gather order.  The whitened rows of a batch of factors are staged into shared memory with one
  // round of global loads, then accumulated factor by factor (fixed order: deterministic sums).
  for (int s0 = star ? ci.fac_end : ci.fac_begin; s0 < ci.fac_end;) {
    int s1 = s0, used = 0;
    while (s1 < ci.fac_end) {
      const int b = E.slots[s1].bucket;
      const int sz = c_M[b] * c_W[b];
      if (used + sz > kThinStage) break;
      used += sz;
      ++s1;
    }
    gsync();   // previous batch fully consumed (and the zero fill done)
    {
      int off = 0;
      for (int s = s0; s < s1; ++s) {
        const SlotDev sl = E.slots[s];
        const int sz = c_M[sl.bucket] * c_W[sl.bucket];
        const double* ab = E.Ab[sl.bucket] + (int64_t)sl.index * sz;
        for (int i = gtid; i < sz; i += gsize) stage[off + i] = ab[i];
        off += sz;
      }
    }
    gsync();
    int off = 0;
    for (int s = s0; s < s1; ++s) {
      const SlotDev sl = E.slots[s];
      const int M = c_M[sl.bucket], W = c_W[sl.bucket], n0 = c_N0[sl.bucket];
      const double* ab = stage + off;
      const int npairs = (W * (W + 1)) >> 1;
      for (int pidx = gtid; pidx < npairs; pidx += gsize) {
        const int a = c_pairA[pidx], b = c_pairB[pidx];
        double acc = 0.0;
        for (int r = 0; r < M; ++r) acc += ab[r * W + a] * ab[r * W + b];
        int i = (a < n0) ? sl.col0 + a : (a < W - 1 ? sl.col1 + (a - n0) : N);
        int j = (b < n0) ? sl.col0 + b : (b < W - 1 ? sl.col1 + (b - n0) : N);
        if (i > j) { const int t = i; i = j; j = t; }
        P[tri(i, j)] += acc;
      }
      off += M * W;
      gsync();
    }
    s0 = s1;
  }
  gsync();
  if (E.damp) {
    const int32_t* g = E.gidx + ci.gidx_off;
    for (int f = gtid; f < F; f += gsize) P[tri(f, f)] += E.damp[g[f]];
    gsync();
  }
  // children's Schur complements, child order; one child column per pass, rows spread over the lanes
  for (int q = ci.child_begin; q < ci.child_end; ++q) {
    const CliqueDev cc = E.cliques[E.children[q]];
    const int S1 = cc.S + 1;
    const double* src = E.arena + cc.front_off + cc.F + (int64_t)cc.F * cc.ld;
    const int32_t* pm = E.pmap + cc.pmap_off;
    // flattened over the child's packed triangle: every load is independent of the others (a column-by-column walk
    // costs one global round trip per column, which is what a deep pose-graph tree spends its time on)
    const int total = (S1 * (S1 + 1)) >> 1;
    for (int idx = gtid; idx < total; idx += gsize) {
      const int j = tri_col(idx), i = idx - ((j * (j + 1)) >> 1);
      int a = pm[i], b = pm[j];
      if (a > b) { const int t = a; a = b; b = t; }
      P[tri(a, b)] += src[i + (int64_t)j * cc.ld];
    }
    gsync();
  }
  // (1) A_FF = R^T R on the small frontal block (base/cholesky.cpp:125)
  bool ok = true;
  for (int k = 0; k < F; ++k) {
    const double piv = P[tri(k, k)];
    if (!(piv > 0.0)) { ok = false; break; }   // Eigen LLT NumericalIssue
    const double rinv = rsqrt(piv);
    gsync();
    for (int j = k + 1 + gtid; j < F; j += gsize) P[tri(k, j)] *= rinv;
    if (gtid == 0) P[tri(k, k)] = piv * rinv;
    gsync();
    const int cnt = F - k - 1;
    for (int idx = gtid; idx < cnt * cnt; idx += gsize) {
      const int jj = idx / cnt, ii = idx - jj * cnt;
      if (ii <= jj) P[tri(k + 1 + ii, k + 1 + jj)] -= P[tri(k, k + 1 + ii)] * P[tri(k, k + 1 + jj)];
    }
    gsync();
  }
  if (ok) {
    if (F >= 2) ok = frexp_exp(P[tri(F - 2, F - 2)]) - frexp_exp(P[tri(F - 1, F - 1)]) < 12;
    else if (F == 1) ok = frexp_exp(P[0]) > -12;
  }
  if (!ok) {
    if (gtid == 0) atomicMin(E.fail, fail_code(c));
    return;
  }
  // (2) S = R^-T [A_FS | b]: one separator column per thread, forward substitution in place (:136)
  double* rdiag = stage;   // the staging area is free by now
  if (gtid < F) rdiag[gtid] = 1.0 / P[tri(gtid, gtid)];
  gsync();
  for (int j = F + gtid; j < dim; j += gsize) {
    const int base = (j * (j + 1)) >> 1;
    for (int r = 0; r < F; ++r) {
      double x = P[base + r];
      for (int l = 0; l < r; ++l) x -= P[tri(l, r)] * P[base + l];
      P[base + r] = x * rdiag[r];
    }
  }
  gsync();
  gsync();
  double* dst = E.arena + ci.front_off;
  // write-out.  Conditional rows [R S d] always go to the clique's own front (back-substitution reads them).  The Schur
  // complement goes to the own front too when the parent assembles in shared memory; when the parent is a large front
  // (already cleared and holding its own factors) it is added there directly, through the child -> parent column map.
  bool fused = false;
  double* pdst = nullptr;
  int pld = 0;
  int32_t* pm_s = reinterpret_cast<int32_t*>(stage);   // S + 1 <= 208 entries; the staging area is free
  if (ci.parent >= 0) {
    const CliqueDev cp = E.cliques[ci.parent];
    // ordered hand-over: several thin children of one parent may sit in the same launch, so the child keeps its Schur
    // complement in its own front and extend_add_kernel adds it at the parent's level, child by child
    fused = !E.ordered && is_fused_child(F, dim, cp.F, cp.F + cp.S + 1);
    if (fused) {
      pdst = E.arena + cp.front_off;
      pld = cp.ld;
      gsync();   // rdiag (same storage) no longer needed
      const int32_t* pm = E.pmap + ci.pmap_off;
      for (int q = gtid; q <= ci.S; q += gsize) pm_s[q] = pm[q];
      gsync();
    }
  }
  {
    int idx = gtid;
    int j = tri_col(idx), i = idx - ((j * (j + 1)) >> 1);
    for (; idx < plen; idx += gsize) {
      // (3) C -= S^T S on the packed upper triangle, every element independent (base/cholesky.cpp:142), fused with the
      // write-out: rows < F of P are final and nothing below them is read again
      double v = P[idx];
      if (i >= F) {
        const int bi = (i * (i + 1)) >> 1, bj = (j * (j + 1)) >> 1;
        double acc = 0.0;
        if (F == 3) {   // points: the bulk of a bundle-adjustment problem
          acc = fma(P[bi + 2], P[bj + 2], fma(P[bi + 1], P[bj + 1], P[bi] * P[bj]));   // the loop's rounding sequence
        } else {
          for (int r = 0; r < F; ++r) acc += P[bi + r] * P[bj + r];
        }
        v -= acc;
      }
      if (i < F || !fused) {
        dst[i + (int64_t)j * ci.ld] = v;
      } else {
        int a = pm_s[i - F], b = pm_s[j - F];
        if (a > b) { const int t = a; a = b; b = t; }
        atomicAdd(pdst + a + (int64_t)b * pld, v);
      }
      i += gsize;
      while (i > j) { i -= j + 1; ++j; }
    }
  }
}

// ----------------------------------------------------------------------------------------------------
// point fronts: the bulk of a bundle-adjustment problem.  A point eliminated alone whose factors are projections
// (GeneralSFMFactor, W = 13) from k <= 15 distinct cameras and whose parent is a large front.  Nothing of the
// (9k+4)^2 / 2 packed front is materialised: with J_c (2 x 9), J_p (2 x 3), b (2) of every factor staged in shared
// memory, the frontal block V = sum J_p^T J_p (+ damping) is factored in registers, the conditional rows
// y_j = R^-T (J_p^T [J_c | b])_j are computed per separator column, written to the clique's own front (what
// back-substitution reads) and kept in shared memory, and every entry of the Schur complement
//   C_ij = [same camera] (J_c^T J_c)_ij - y_i . y_j          (camera - camera blocks of different cameras are pure -y_i.y_j)
// goes straight into the parent front through the child -> parent column map (FP64 atomics).  One warp per point,
// 2.4 KB of shared memory per warp at k = 5 instead of 13 KB: four times the warps in flight of the generic thin kernel.
// ----------------------------------------------------------------------------------------------------
constexpr int kPointMaxCams = 64;   // most cameras of a point clique on the point path (pair-gathered Schur complement)
__host__ __device__ __forceinline__ int point_front_doubles(int kmax, bool schur) {
  const int nmax = 9 * kmax + 1;
  if (!schur) return 26 * kmax + ((2 * kmax + 2) >> 1) + 2;                       // ab | scol, sfac (ints) | pad
  return 26 * kmax + 3 * nmax + ((nmax + 2) >> 1) + ((2 * kmax + 2) >> 1) + 2;   // ab | Y | pm (ints) | scol, sfac (ints) | pad
}
__global__ void __launch_bounds__(kThinWarps * 32) point_front_kernel(ElimDev E, const int32_t* __restrict__ list, int count,
                                                                       int kmax, int schur) {
  extern __shared__ double sh[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = blockIdx.x * kThinWarps + warp;
  if (slot >= count) return;
  const int nmax = 9 * kmax + 1;
  double* ab = sh + warp * point_front_doubles(kmax, schur != 0);   // ab[26 s + 13 r + col]: rows of factor s
  double* Y = ab + 26 * kmax;                           // Y[a * nmax + j] = y_j[a]   (Y and pm_s only when schur)
  int* pm_s = reinterpret_cast<int*>(Y + 3 * nmax);     // parent column of separator scalar q (and of the rhs)
  int* scol = schur ? pm_s + ((nmax + 2) & ~1) : reinterpret_cast<int*>(Y);   // separator-relative first column of factor s's camera
  int* sfac = scol + kmax;                              // camera block index -> factor
  const int c = list[slot];
  const CliqueDev ci = E.cliques[c];
  const int k = ci.fac_end - ci.fac_begin, n = ci.S + 1, nc = ci.S;   // nc = 9 k camera scalars, column nc is the rhs
  for (int t = lane; t < 26 * k; t += 32) {
    const int s = t / 26, r = t - s * 26;
    const SlotDev sl = E.slots[ci.fac_begin + s];
    ab[t] = E.Ab[2][(int64_t)sl.index * 26 + r];
    if (r == 0) { scol[s] = sl.col0 - 3; sfac[(sl.col0 - 3) / 9] = s; }
  }
  if (schur)
    for (int q = lane; q < n; q += 32) pm_s[q] = E.pmap[ci.pmap_off + q];
  __syncwarp();
  // frontal block and its right-hand side, factors in gather order (deterministic sums); every lane keeps a copy
  double v00 = 0, v01 = 0, v02 = 0, v11 = 0, v12 = 0, v22 = 0, g0 = 0, g1 = 0, g2 = 0, bb = 0;
  for (int s = 0; s < k; ++s) {
    const double* f = ab + 26 * s;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const double p0 = f[13 * r + 9], p1 = f[13 * r + 10], p2 = f[13 * r + 11], br = f[13 * r + 12];
      v00 += p0 * p0; v01 += p0 * p1; v02 += p0 * p2; v11 += p1 * p1; v12 += p1 * p2; v22 += p2 * p2;
      g0 += p0 * br; g1 += p1 * br; g2 += p2 * br; bb += br * br;
    }
  }
  if (E.damp) {
    const int32_t* g = E.gidx + ci.gidx_off;
    v00 += E.damp[g[0]]; v11 += E.damp[g[1]]; v22 += E.damp[g[2]];
  }
  // 3 x 3 Cholesky, upper, same pivot rule and reciprocal form as thin_front_kernel (base/cholesky.cpp:125-158)
  bool ok = v00 > 0.0;
  const double i0 = rsqrt(v00);
  const double r00 = v00 * i0, r01 = v01 * i0, r02 = v02 * i0;
  v11 -= r01 * r01; v12 -= r01 * r02; v22 -= r02 * r02;
  ok = ok && v11 > 0.0;
  const double i1 = rsqrt(v11);
  const double r11 = v11 * i1, r12 = v12 * i1;
  v22 -= r12 * r12;
  ok = ok && v22 > 0.0;
  const double i2 = rsqrt(v22);
  const double r22 = v22 * i2;
  if (ok) ok = frexp_exp(r11) - frexp_exp(r22) < 12;
  if (!ok) {
    if (lane == 0) atomicMin(E.fail, fail_code(c));
    return;
  }
  const double d0 = 1.0 / r00, d1 = 1.0 / r11, d2 = 1.0 / r22;
  double* dst = E.arena + ci.front_off;
  if (lane == 0) {
    dst[0] = r00;
    dst[ci.ld] = r01; dst[1 + ci.ld] = r11;
    dst[2 * (int64_t)ci.ld] = r02; dst[1 + 2 * (int64_t)ci.ld] = r12; dst[2 + 2 * (int64_t)ci.ld] = r22;
  }
  // conditional rows: y_j = R^-T w_j, w_j = J_p^T (column j of [J_c | b]) of the camera's factor, or g for the rhs
  for (int j = lane; j < n; j += 32) {
    double w0, w1, w2;
    if (j < nc) {
      const int blk = j / 9, col = j - blk * 9;
      const double* f = ab + 26 * sfac[blk];
      const double c0 = f[col], c1 = f[13 + col];
      w0 = f[9] * c0 + f[22] * c1;
      w1 = f[10] * c0 + f[23] * c1;
      w2 = f[11] * c0 + f[24] * c1;
    } else {
      w0 = g0; w1 = g1; w2 = g2;
    }
    const double y0 = w0 * d0;
    const double y1 = (w1 - r01 * y0) * d1;
    const double y2 = (w2 - r02 * y0 - r12 * y1) * d2;
    if (schur) { Y[j] = y0; Y[nmax + j] = y1; Y[2 * nmax + j] = y2; }
    double* col3 = dst + (int64_t)(3 + j) * ci.ld;
    col3[0] = y0; col3[1] = y1; col3[2] = y2;
  }
  if (!schur) return;   // the Schur complement is gathered per camera pair by point_pair_kernel / point_diag_kernel
  __syncwarp();
  // Schur complement entries (i <= j over the separator scalars and the rhs) straight into the parent
  const CliqueDev cp = E.cliques[ci.parent];
  double* pdst = E.arena + cp.front_off;
  const int pld = cp.ld;
  const int total = (n * (n + 1)) >> 1;
  int idx = lane;
  int j = tri_col(idx), i = idx - ((j * (j + 1)) >> 1);
  for (; idx < total; idx += 32) {
    double h = 0.0;
    const int bi = i / 9;
    if (j < nc) {
      if (bi == j / 9) {
        const double* f = ab + 26 * sfac[bi];
        const int a = i - bi * 9, b = j - bi * 9;
        h = f[a] * f[b] + f[13 + a] * f[13 + b];
      }
    } else if (i < nc) {
      const double* f = ab + 26 * sfac[bi];
      const int a = i - bi * 9;
      h = f[a] * f[12] + f[13 + a] * f[25];
    } else {
      h = bb;
    }
    const double v = h - (Y[i] * Y[j] + Y[nmax + i] * Y[nmax + j] + Y[2 * nmax + i] * Y[2 * nmax + j]);
    int a = pm_s[i], b = pm_s[j];
    if (a > b) { const int t = a; a = b; b = t; }
    atomicAdd(pdst + a + (int64_t)b * pld, v);
    i += 32;
    while (i > j) { i -= j + 1; ++j; }
  }
}

// Conditional rows only (the Schur complement is gathered per camera pair, below): LPP lanes per point.  LPP = 8 packs four
// points into a warp (up to 8 cameras each: the frontal 3 x 3 block, its Cholesky and the descriptor loads are per point,
// not per lane, and 9 k + 1 <= 73 columns keep eight lanes busy), LPP = 32 takes the points with more cameras.  Same
// arithmetic and summation order as point_front_kernel.
__host__ __device__ __forceinline__ int point_rows_doubles(int kmax) { return 26 * kmax + ((kmax + 1) >> 1) + 1; }
template <int LPP>
__global__ void __launch_bounds__(kThinWarps * 32) point_rows_kernel(ElimDev E, const int32_t* __restrict__ list, int count, int kmax) {
  extern __shared__ double sh[];
  constexpr int PPW = 32 / LPP;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sl = lane % LPP, sw = lane / LPP;
  const int slot = (blockIdx.x * kThinWarps + warp) * PPW + sw;
  if (slot >= count) return;
  double* ab = sh + (warp * PPW + sw) * point_rows_doubles(kmax);   // ab[26 s + 13 r + col]: rows of factor s
  int* sfac = reinterpret_cast<int*>(ab + 26 * kmax);               // camera block index -> factor
  const int c = list[slot];
  const CliqueDev ci = E.cliques[c];
  const int k = ci.fac_end - ci.fac_begin, n = ci.S + 1, nc = ci.S;
  // A lane's time is its chain of dependent round trips: the slot descriptors of the point's factors are loaded by one
  // lane each (and the positions of its damping terms), then every row load of the point is issued in one batch (and
  // the damping terms with it).  (A `for t: slot -> row` loop compiles to two round trips per 4 entries and two per
  // entry of the remainder: 8 instead of 2 for a point seen by 2..4 cameras.)
  const unsigned gmask = LPP == 32 ? 0xffffffffu : (((1u << LPP) - 1u) << (sw * LPP));
  const int gbase = sw * LPP;
  int idx0 = 0, idx1 = 0, gq = 0;
  if (E.damp) gq = E.gidx[ci.gidx_off + (sl < 3 ? sl : 0)];
  if (sl < k) {
    const SlotDev sd = E.slots[ci.fac_begin + sl];
    idx0 = sd.index;
    sfac[(sd.col0 - 3) / 9] = sl;
  }
  if (LPP == 32 && sl + 32 < k) {
    const SlotDev sd = E.slots[ci.fac_begin + sl + 32];
    idx1 = sd.index;
    sfac[(sd.col0 - 3) / 9] = sl + 32;
  }
  const double dq = E.damp ? E.damp[gq] : 0.0;
  constexpr int kBatch = 13;   // one batch covers a point seen by LPP / 2 cameras
  for (int tb = 0; tb < 26 * k; tb += kBatch * LPP) {   // same trip count in every lane of the group (shuffles inside)
    const int t0 = tb + sl;
    double v[kBatch];
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int t = t0 + u * LPP;
      const int s = min(t / 26, k - 1), r = t - (t / 26) * 26;
      int idx = __shfl_sync(gmask, idx0, gbase + (s & (LPP - 1)));
      if (LPP == 32) {
        const int hi = __shfl_sync(gmask, idx1, s & 31);
        if (s >= 32) idx = hi;
      }
      v[u] = t < 26 * k ? E.Ab[2][(int64_t)idx * 26 + r] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int t = t0 + u * LPP;
      if (t < 26 * k) ab[t] = v[u];
    }
  }
  __syncwarp();
  double v00 = 0, v01 = 0, v02 = 0, v11 = 0, v12 = 0, v22 = 0, g0 = 0, g1 = 0, g2 = 0;
  for (int s = 0; s < k; ++s) {
    const double* f = ab + 26 * s;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const double p0 = f[13 * r + 9], p1 = f[13 * r + 10], p2 = f[13 * r + 11], br = f[13 * r + 12];
      v00 += p0 * p0; v01 += p0 * p1; v02 += p0 * p2; v11 += p1 * p1; v12 += p1 * p2; v22 += p2 * p2;
      g0 += p0 * br; g1 += p1 * br; g2 += p2 * br;
    }
  }
  {
    const double q0 = __shfl_sync(gmask, dq, gbase), q1 = __shfl_sync(gmask, dq, gbase + 1), q2 = __shfl_sync(gmask, dq, gbase + 2);
    if (E.damp) { v00 += q0; v11 += q1; v22 += q2; }
  }
  bool ok = v00 > 0.0;
  const double i0 = rsqrt(v00);
  const double r00 = v00 * i0, r01 = v01 * i0, r02 = v02 * i0;
  v11 -= r01 * r01; v12 -= r01 * r02; v22 -= r02 * r02;
  ok = ok && v11 > 0.0;
  const double i1 = rsqrt(v11);
  const double r11 = v11 * i1, r12 = v12 * i1;
  v22 -= r12 * r12;
  ok = ok && v22 > 0.0;
  const double i2 = rsqrt(v22);
  const double r22 = v22 * i2;
  if (ok) ok = frexp_exp(r11) - frexp_exp(r22) < 12;
  if (!ok) {
    if (sl == 0) atomicMin(E.fail, fail_code(c));
    return;
  }
  const double d0 = 1.0 / r00, d1 = 1.0 / r11, d2 = 1.0 / r22;
  double2* dst = reinterpret_cast<double2*>(E.arena + ci.front_off);   // ld = 4: one 32-byte column = two 16-byte stores
  if (sl == 0) {
    dst[0] = make_double2(r00, 0.0);
    dst[2] = make_double2(r01, r11);
    dst[4] = make_double2(r02, r12);
    dst[5] = make_double2(r22, 0.0);
  }
  for (int j = sl; j < n; j += LPP) {
    double w0, w1, w2;
    if (j < nc) {
      const int blk = j / 9, col = j - blk * 9;
      const double* f = ab + 26 * sfac[blk];
      const double c0 = f[col], c1 = f[13 + col];
      w0 = f[9] * c0 + f[22] * c1;
      w1 = f[10] * c0 + f[23] * c1;
      w2 = f[11] * c0 + f[24] * c1;
    } else {
      w0 = g0; w1 = g1; w2 = g2;
    }
    const double y0 = w0 * d0;
    const double y1 = (w1 - r01 * y0) * d1;
    const double y2 = (w2 - r02 * y0 - r12 * y1) * d2;
    dst[2 * (3 + j)] = make_double2(y0, y1);
    dst[2 * (3 + j) + 1] = make_double2(y2, 0.0);
  }
}

// ----------------------------------------------------------------------------------------------------
// Schur complement of the point cliques, gathered per camera pair.  point_front_kernel scatters (9k+1)(9k+2)/2 FP64
// atomics per point into the parent front and is bound by the L2 atomic rate (C4: 1.07 G atomics per step).  Here the
// sum is turned around: the host lists, once per ordering, for every 9 x 9 block (camera A, camera B) of a parent
// front the points that see both cameras, and one thread per (block, 3-column slab) accumulates
//     C_AB = - sum_p Y_p,A^T Y_p,B                                   (different cameras)
//     C_AA = sum_p J_c^T J_c - Y_p,A^T Y_p,A,   C_A,rhs = sum_p J_c^T b - Y_p,A^T y_p,rhs   (same camera: own factor too)
// in registers from the conditional rows the point kernel has just written (ld = 4: the 9 columns of a camera are 288
// contiguous bytes), in a fixed order, and adds the block to the parent once.  Lists longer than kPairSeg items are cut
// into segments whose partial sums meet in the parent through atomics (81 per segment instead of 81 per item).
// ----------------------------------------------------------------------------------------------------
constexpr int kPairSeg = 32;
// ld > 0: the segment is the whole list of its block and adds it to the parent front at arena + dst (one contribution per
// entry and launch: the order of the sum is the list order).  ld == 0 (ordered hand-over, list cut into several
// segments): the partial block goes to pair_scratch + dst (compact: entry (i, c) at i + 9 c, rhs column at 81 + i,
// constant term at 90) and pair_reduce_kernel adds the segments' partials in list order.
constexpr int kPairSlot = 96;
struct PairSeg { int64_t dst; int32_t ld; int32_t begin; int32_t count; int32_t ncol; };   // ncol: columns from the block to the parent's rhs column
struct PairReduce { int64_t dst; int32_t ld, ncol, slot0, count, diag, pad; };
__global__ void __launch_bounds__(128) point_pair_kernel(ElimDev E, const PairSeg* __restrict__ segs, int nseg,
                                                         const uint2* __restrict__ items) {
  const int gt = blockIdx.x * 128 + threadIdx.x;
  const int s = gt / 3, slab = gt - 3 * s;
  if (s >= nseg) return;
  const PairSeg sg = segs[s];
  double acc[9][3];
#pragma unroll
  for (int i = 0; i < 9; ++i) acc[i][0] = acc[i][1] = acc[i][2] = 0.0;
  const double* arena = E.arena;
  for (int q = sg.begin; q < sg.begin + sg.count; ++q) {
    const uint2 it = __ldg(items + q);
    const double* ya = arena + (int64_t)it.x * 4;
    const double* yb = arena + (int64_t)it.y * 4 + 12 * slab;
    Col4 b[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) b[c] = ld_col(yb + 4 * c);
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      const Col4 u = ld_col(ya + 4 * i);
#pragma unroll
      for (int c = 0; c < 3; ++c) acc[i][c] += u.x * b[c].x + u.y * b[c].y + u.z * b[c].z;
    }
  }
  if (sg.ld == 0) {
    double* out = E.pair_scratch + sg.dst + 27 * slab;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int i = 0; i < 9; ++i) out[i + 9 * c] = -acc[i][c];
    return;
  }
  double* dst = E.arena + sg.dst + (int64_t)(3 * slab) * sg.ld;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int i = 0; i < 9; ++i) atomicAdd(dst + i + (int64_t)c * sg.ld, -acc[i][c]);
}
// items: 3 words each: first column of the camera's block in the point's front (32-byte units), factor index in the
// projection bucket, the point's rhs column (32-byte units; bit 31: this is the point's first camera, which carries the
// - y_rhs . y_rhs of the constant term).  Threads 0..2 of a segment own 3 columns of the 9 x 9 block, thread 3 the rhs column.
__global__ void __launch_bounds__(128) point_diag_kernel(ElimDev E, const PairSeg* __restrict__ segs, int nseg,
                                                         const uint32_t* __restrict__ items) {
  const int gt = blockIdx.x * 128 + threadIdx.x;
  const int s = gt >> 2, slab = gt & 3;
  if (s >= nseg) return;
  const PairSeg sg = segs[s];
  double acc[9][3];
#pragma unroll
  for (int i = 0; i < 9; ++i) acc[i][0] = acc[i][1] = acc[i][2] = 0.0;
  double cst = 0.0;
  const double* arena = E.arena;
  const double* Ab = E.Ab[2];
  for (int q = sg.begin; q < sg.begin + sg.count; ++q) {
    const uint32_t iy = __ldg(items + 3 * (int64_t)q), ifac = __ldg(items + 3 * (int64_t)q + 1), ir = __ldg(items + 3 * (int64_t)q + 2);
    const double* ya = arena + (int64_t)iy * 4;
    // the factor's two rows [J_c (9) | J_p (3) | b]: 26 doubles, 16-byte aligned; J_c and b as 13 16-byte loads
    const double2* f2 = reinterpret_cast<const double2*>(Ab + (int64_t)ifac * 26);
    double r0[14], r1[14];   // r0[i] = f[i], r1[i] = f[13 + i - 1 + 1]: rows 0 and 1 (row 1 starts at the odd offset 13)
#pragma unroll
    for (int h = 0; h < 5; ++h) { const double2 v = __ldg(f2 + h); r0[2 * h] = v.x; r0[2 * h + 1] = v.y; }
    { const double2 v = __ldg(f2 + 6); r0[12] = v.x; r1[0] = v.y; }   // f[12] = b_0, f[13] = J_c(1, 0)
#pragma unroll
    for (int h = 0; h < 4; ++h) { const double2 v = __ldg(f2 + 7 + h); r1[1 + 2 * h] = v.x; r1[2 + 2 * h] = v.y; }   // f[14..21]
    const double br1 = __ldg(Ab + (int64_t)ifac * 26 + 25);
    if (slab < 3) {
      Col4 b[3];
      double j0[3], j1[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        b[c] = ld_col(ya + 12 * slab + 4 * c);
        j0[c] = slab == 0 ? r0[c] : (slab == 1 ? r0[3 + c] : r0[6 + c]);
        j1[c] = slab == 0 ? r1[c] : (slab == 1 ? r1[3 + c] : r1[6 + c]);
      }
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        const Col4 u = ld_col(ya + 4 * i);
#pragma unroll
        for (int c = 0; c < 3; ++c) acc[i][c] += (r0[i] * j0[c] + r1[i] * j1[c]) - (u.x * b[c].x + u.y * b[c].y + u.z * b[c].z);
      }
    } else {
      const Col4 r = ld_col(arena + (int64_t)(ir & 0x7fffffffu) * 4);
      const double br0 = r0[12];
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        const Col4 u = ld_col(ya + 4 * i);
        acc[i][0] += (r0[i] * br0 + r1[i] * br1) - (u.x * r.x + u.y * r.y + u.z * r.z);
      }
      cst += br0 * br0 + br1 * br1;
      if (ir >> 31) cst -= r.x * r.x + r.y * r.y + r.z * r.z;
    }
  }
  if (sg.ld == 0) {
    double* out = E.pair_scratch + sg.dst;
    if (slab < 3) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int i = 0; i < 9; ++i) out[i + 9 * (3 * slab + c)] = acc[i][c];
    } else {
#pragma unroll
      for (int i = 0; i < 9; ++i) out[81 + i] = acc[i][0];
      out[90] = cst;
    }
    return;
  }
  if (slab < 3) {
    double* dst = E.arena + sg.dst + (int64_t)(3 * slab) * sg.ld;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int i = 0; i < 9; ++i)
        if (i <= 3 * slab + c) atomicAdd(dst + i + (int64_t)c * sg.ld, acc[i][c]);
  } else {
    double* dst = E.arena + sg.dst + (int64_t)sg.ncol * sg.ld;
#pragma unroll
    for (int i = 0; i < 9; ++i) atomicAdd(dst + i, acc[i][0]);
    atomicAdd(dst + sg.ncol, cst);
  }
}

// Ordered sum of the partial blocks of a camera-pair (or same-camera) group that was cut into several segments: one
// thread per block entry adds the segments' partials in list order and hands the sum to the parent once.
__global__ void __launch_bounds__(2 * kPairSlot) pair_reduce_kernel(ElimDev E, const PairReduce* __restrict__ items, int n) {
  const int it = blockIdx.x * 2 + threadIdx.x / kPairSlot, e = threadIdx.x % kPairSlot;
  if (it >= n) return;
  const PairReduce r = items[it];
  const int i = e % 9, c = e / 9;
  int64_t off = 0;
  bool use = false;
  if (e < 81) { use = !r.diag || i <= c; off = i + (int64_t)c * r.ld; }
  else if (e < 90) { use = r.diag != 0; off = (e - 81) + (int64_t)r.ncol * r.ld; }
  else if (e == 90) { use = r.diag != 0; off = r.ncol + (int64_t)r.ncol * r.ld; }
  if (!use) return;
  const double* sp = E.pair_scratch + (int64_t)r.slot0 * kPairSlot + e;
  double acc = 0.0;
  for (int k = 0; k < r.count; ++k) acc += sp[(int64_t)k * kPairSlot];
  red_add(E.arena + r.dst + off, acc);   // the only contribution to this entry in this launch
}

// ----------------------------------------------------------------------------------------------------
// large fronts: front in global memory (L2-resident), several CTAs per front
// ----------------------------------------------------------------------------------------------------
// zeroing of very large fronts is spread over many CTAs
__global__ void __launch_bounds__(256) large_zero_kernel(ElimDev E, const int32_t* __restrict__ list) {
  const int c = list[blockIdx.x];   // the front on x: the y extent of a grid stops at 65,535
  const CliqueDev ci = E.cliques[c];
  const int dim = ci.F + ci.S + 1;
  // upper triangle only (nothing reads below the diagonal): column j, rows 0..j, as 16-byte stores (ld is a multiple of 4)
  double* A = E.arena + ci.front_off;
  for (int j = blockIdx.y; j < dim; j += gridDim.y) {
    double2* col = reinterpret_cast<double2*>(A + (int64_t)j * ci.ld);
    const int n2 = (j + 2) >> 1;
    for (int i = threadIdx.x; i < n2; i += blockDim.x) col[i] = make_double2(0.0, 0.0);
  }
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
  pdl_trigger();
  const ChildWork w = work[blockIdx.x];
  const CliqueDev cc = E.cliques[w.child];
  const CliqueDev cp = E.cliques[cc.parent];
  pdl_wait();
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
      atomicAdd(dst + a + (int64_t)b * cp.ld, col[i]);   // (a plain read-modify-write measured 2.7x slower)
    }
  }
}

// Ordered extend-add (default): the Schur complements of the children of a large front that do not hand over by
// themselves (thin and dense small children, subtree roots owned by this rank under the shared top) are added by the
// PARENT, child after child in the reference's child order (inference/ClusterTree-inst.h:223-226).  A CTA owns a range of
// parent columns [col_begin, col_end); for every child it finds the child columns that land there through the child's
// separator SORTED BY PARENT COLUMN (spm[v] = parent column, sg[v] = child separator index, ascending spm): entry (u, v),
// u <= v, of the sorted child goes to parent (spm[u], spm[v]), so a parent entry has one owner thread per child and one
// barrier per child keeps the children in order.  The adds are fire-and-forget reductions.
struct GatherWork { int32_t parent, col_begin, col_end, list_begin, list_end, pad; };
__global__ void __launch_bounds__(256) extend_add_kernel(ElimDev E, const GatherWork* __restrict__ work,
                                                         const int32_t* __restrict__ gchild, const int64_t* __restrict__ gsmap_off,
                                                         const int32_t* __restrict__ smap) {
  pdl_trigger();
  const GatherWork w = work[blockIdx.x];
  const CliqueDev cp = E.cliques[w.parent];
  pdl_wait();
  double* dst = E.arena + cp.front_off;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int q = w.list_begin; q < w.list_end; ++q) {
    const CliqueDev cc = E.cliques[gchild[q]];
    const int S1 = cc.S + 1;
    const int32_t* spm = smap + gsmap_off[q];
    const int32_t* sg = spm + S1;
    int lo = 0, hi = S1;   // first sorted position with spm >= col_begin
    while (lo < hi) { const int m = (lo + hi) >> 1; if (spm[m] < w.col_begin) lo = m + 1; else hi = m; }
    const int v0 = lo;
    hi = S1;
    while (lo < hi) { const int m = (lo + hi) >> 1; if (spm[m] < w.col_end) lo = m + 1; else hi = m; }
    const int v1 = lo;
    const double* src = E.arena + cc.front_off + cc.F + (int64_t)cc.F * cc.ld;
    for (int v = v0 + warp; v < v1; v += nw) {
      const int pb = spm[v], jv = sg[v];
      for (int u = lane; u <= v; u += 32) {
        const int iu = sg[u];
        const int ci = min(iu, jv), cj = max(iu, jv);
        red_add(dst + spm[u] + (int64_t)pb * cp.ld, src[ci + (int64_t)cj * cc.ld]);
      }
    }
    __syncthreads();   // the next child's contribution to an entry may come from another thread of this CTA
  }
}

// ---- blocked partial Cholesky ---------------------------------------------------------------------
// Right-looking with one-panel look-ahead, ONE launch per panel step for all large fronts of a level:
//   step p, role A (per clique, per 256-column slice): bring row panel p up to date with panel p-1, factor its
//                   32 x 32 diagonal block (every slice redundantly, in shared memory), solve the slice
//   step p, role B: trailing update by panel p-1 of the frontal rows below row panel p (64 x 64 DMMA tiles)
//   step np (after the clique's last panel): Schur complement C -= S^T S with K = F, and the factored diagonal
//                   blocks move from their side buffer into the front
// Role A of step p and role B of step p touch disjoint rows, and both only read panel p-1, so no CTA waits on
// another inside a launch.
constexpr int NB = 32;       // K-chunk of the tile update and side of the in-register Cholesky block
constexpr int PW = 2 * NB;   // panel height: 64 pivots per launch (two NB blocks factored back to back inside role A)
constexpr int TW = 64;       // tile width
constexpr int LDT = NB + 4;  // shared leading dimension of a K x 64 operand tile (K contiguous), 36: conflict-free
constexpr int LDP = PW + 4;  // shared leading dimension of the role-A operands (68: conflict-free for the DMMA fragments)
constexpr int CWID = 8;      // columns per role-A column warp (one DMMA n-tile)
constexpr int SLICE = 7 * CWID;   // columns per role-A CTA: warp 0 factors the diagonal block, warps 1..7 own 8 columns each
constexpr int kStepSmemTile = 2 * 3 * TW * LDT * (int)sizeof(double);                   // two operands, 3-deep ring of K-chunks
// role A: Ps, Ds, Ws (8 x 8 inverses), dinv, flag, warp 0's scratch, then Xs + Ts per column warp
constexpr int kRoleAWarp0 = 2 * PW * LDP + PW * 12 + PW + 2;
constexpr int kRoleAColWarp = CWID * LDP + CWID * 12;
constexpr int kStepSmemA = (kRoleAWarp0 + 7 * kRoleAColWarp) * (int)sizeof(double);
constexpr int kStepSmem = kStepSmemA > kStepSmemTile ? kStepSmemA : kStepSmemTile;

// FP64 tensor-core tile: m8n8k4 DMMA (mma.sync), D = A * B + C
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// 16-byte asynchronous global -> shared copy; bytes beyond src_bytes are zero-filled (src_bytes = 0, 8 or 16)
__device__ __forceinline__ void cp_async16(double* smem_dst, const double* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// C[i, j] -= sum_{k in [ka, kend)} R[k, i] * R[k, j] for the 64 x 64 tile at (r0, c0), rows < re, upper triangle.
// 256 threads = 8 warps as 2 (rows) x 4 (cols); each warp owns 32 x 16 = 4 x 2 DMMA tiles.  Operands are staged through
// a kStages-deep ring of shared-memory K-chunks of 32 filled with cp.async (two chunks in flight behind the one being
// multiplied: a lone CTA on an SM otherwise waits on L2/HBM every chunk).  ka and the front's leading dimension are
// even, so every 16-byte copy is aligned; ragged edges are zero-filled by the copy itself.
constexpr int kStages = 3;
// MODE 0: C -= product in place.  MODE 1: the same through atomics (K split over several CTAs).  MODE 2: the updated tile
// (C itself only when add_c: one K slice carries it) is added into the parent front PA through the child -> parent
// column map pm instead: the extend-add of a large child fused into its Schur-complement step (a plain read-modify-write
// for an only child measured 10 % slower than the atomics).
template <int MODE = 0>
__device__ __forceinline__ void tile_update(double* __restrict__ A, int ld, int dim, int r0, int c0, int re, int ka,
                                            int kend, double* sm, const int32_t* __restrict__ pm = nullptr,
                                            double* __restrict__ PA = nullptr, int pld = 0, int F = 0, bool add_c = false,
                                            int* flag = nullptr, int turn = 0, int dbg = 0) {
  double* Xs = sm;                        // [kStages][TW * LDT]
  double* Ys = sm + kStages * TW * LDT;   // [kStages][TW * LDT]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int wr = warp >> 2, wc = warp & 3;
  const int g = lane >> 2, t = lane & 3;
  double acc[4][2][2];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
  // copy assignment: thread -> (column lcol = tid / 16 + 16 q, k pair kp = tid % 16), q = 0..3
  const int kp = (tid & 15) * 2, lc0 = tid >> 4;
  auto fetch = [&](int kc, int buf) {
    const int kn = kend - kc - kp;   // valid k from this pair on
    const int nb = kn >= 2 ? 16 : (kn == 1 ? 8 : 0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int lcol = lc0 + 16 * q;
      const int xi = r0 + lcol, yj = c0 + lcol;
      // out-of-range columns read the front's first column with size 0: pure zero fill, no access
      cp_async16(Xs + buf * TW * LDT + lcol * LDT + kp, A + kc + kp + (int64_t)(xi < dim ? xi : 0) * ld, xi < dim ? nb : 0);
      cp_async16(Ys + buf * TW * LDT + lcol * LDT + kp, A + kc + kp + (int64_t)(yj < dim ? yj : 0) * ld, yj < dim ? nb : 0);
    }
  };
  const int nchunk = (kend - ka + NB - 1) / NB;
#pragma unroll
  for (int s = 0; s < kStages - 1; ++s) {
    if (s < nchunk) fetch(ka + s * NB, s);
    cp_async_commit();
  }
  // This thread's 16 entries of C are loaded here, behind the first operand copies, and the product is accumulated onto
  // -C.  Loading them in the epilogue (C[i, j] -= acc) is 16 dependent round trips to L2 per thread: the compiler may
  // not move an entry's load above the previous entry's store or atomic.
  int prow[4], pcol[2][2];   // MODE 2: parent positions of this thread's rows and columns (same reason: 8 loads here
                             // instead of two in front of every atomic)
  if (MODE == 2) {
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int i = r0 + wr * 32 + a * 8 + g;
      prow[a] = i < re ? pm[i - F] : 0;
    }
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = c0 + wc * 16 + b * 8 + 2 * t + e;
        pcol[b][e] = j < dim ? pm[j - F] : 0;
      }
  }
  if (MODE == 0 || (MODE == 2 && add_c)) {
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int i = r0 + wr * 32 + a * 8 + g;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = c0 + wc * 16 + b * 8 + 2 * t + e;
          if (i < re && j < dim && i <= j) acc[a][b][e] = -A[i + (int64_t)j * ld];
        }
      }
  }
  for (int ch = 0; ch < ((dbg & 64) ? 1 : nchunk); ++ch) {   // GTB_DEBUG=64 (timing experiment): one K chunk only
    cp_async_wait<kStages - 2>();   // chunk ch has landed (for this thread's copies)
    __syncthreads();                // ... and for everybody's; also: everybody is done with the buffer refilled below
    if (ch + kStages - 1 < nchunk) fetch(ka + (ch + kStages - 1) * NB, (ch + kStages - 1) % kStages);
    cp_async_commit();
    const double* X = Xs + (ch % kStages) * TW * LDT;
    const double* Y = Ys + (ch % kStages) * TW * LDT;
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
  }
  cp_async_wait<0>();
  if (dbg & 32) return;   // GTB_DEBUG=32 (timing experiment, wrong results): no epilogue
  // Ordered K split: the slices of a tile hand their partial products over one after the other (slice 0 first): a slice
  // waits until the tile's flag says that its predecessor's adds are visible.  The CTAs of earlier slices come earlier in
  // the task list, so they are resident (or done) whenever a later slice spins.
  if (flag && MODE != 0) {
    if (tid == 0) {
      while (*reinterpret_cast<volatile int*>(flag) != turn) __nanosleep(40);
      __threadfence();
    }
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int i = r0 + wr * 32 + a * 8 + g;
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = c0 + wc * 16 + b * 8 + 2 * t + e;
        if (i < re && j < dim && i <= j) {
          if (MODE == 2) {
            int pa = prow[a], pb = pcol[b][e];
            if (pa > pb) { const int tt = pa; pa = pb; pb = tt; }
            red_add(PA + pa + (int64_t)pb * pld, -acc[a][b][e]);
          } else if (MODE == 1) {
            red_add(A + i + (int64_t)j * ld, -acc[a][b][e]);
          } else {
            A[i + (int64_t)j * ld] = -acc[a][b][e];
          }
        }
      }
    }
  if (flag && MODE != 0) {
    __threadfence();
    __syncthreads();
    if (tid == 0) *reinterpret_cast<volatile int*>(flag) = turn + 1;
  }
}

// number of tile tasks of a trapezoid: nrt tile rows, nct tile columns, only tiles with tx >= ty
__host__ __device__ __forceinline__ int trapezoid_tiles(int nrt, int nct) {
  if (nrt <= 0 || nct <= 0) return 0;
  if (nrt > nct) nrt = nct;
  return nrt * nct - nrt * (nrt - 1) / 2;
}
// Tile t of a trapezoid -> (row ty, column tx), rows ty = 0, 1, ... hold nct - ty tiles (tx >= ty).  Row ty starts at
// off(ty) = ty * nct - ty (ty - 1) / 2: solved in floating point and corrected by at most one step (a loop over the rows
// costs ~20 cycles per row on a dependent integer chain: 600 cycles in front of a 64 x 64 tile of a 25-row trapezoid).
__device__ __forceinline__ void trapezoid_decode(int t, int nct, int& ty, int& tx) {
  const float w = 2.0f * nct + 1.0f;
  float root;   // the approximate instruction: no special-case branch, and the correction below absorbs its last bits
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(root) : "f"(fmaxf(w * w - 8.0f * t, 0.0f)));
  ty = (int)((w - root) * 0.5f);
  ty = max(0, min(ty, nct - 1));
  int off = ty * nct - ((ty * (ty - 1)) >> 1);
  if (off > t) { --ty; off -= nct - ty; }
  else if (t - off >= nct - ty) { off += nct - ty; ++ty; }
  tx = ty + (t - off);
}
// tasks of clique (F, dim) at panel step p (host mirrors this to size the grid)
// spc = column slices per role-A CTA of this launch (the host raises it when a launch would hold more role-A CTAs than the
// machine has SMs: every role-A CTA factors the diagonal block on its own, so a second wave of them repeats that work)
__host__ __device__ __forceinline__ int role_a_ctas(int dim, int rest0, int spc) {
  const int nA = (dim - rest0 + SLICE - 1) / SLICE;
  return (nA + spc - 1) / spc;
}
__host__ __device__ __forceinline__ int step_tasks(int F, int dim, int p, int spc = 1) {
  const int np = (F + PW - 1) / PW;
  if (p < np) {
    const int k0 = p * PW, kb = (F - k0 < PW) ? F - k0 : PW, rest0 = k0 + kb;
    int n = role_a_ctas(dim, rest0, spc);
    if (p >= 1) n += trapezoid_tiles((F - rest0 + TW - 1) / TW, (dim - rest0 + TW - 1) / TW);
    return n;
  }
  if (p == np) {
    const int nt = (dim - F + TW - 1) / TW;
    return nt * (nt + 1) / 2;
  }
  return 0;
}

struct StepTask { int32_t clique; int32_t task; };
// Schur-step tasks pack (tile, K slice, number of K slices): a front with too few tiles to fill the machine splits the
// K = F sum of every tile over several CTAs, whose partial products meet in C through FP64 atomics.
constexpr int kSchurTileBits = 20, kSchurSliceBits = 5;
__host__ __device__ __forceinline__ int schur_task(int tile, int slice, int nsplit) {
  return tile | (slice << kSchurTileBits) | (nsplit << (kSchurTileBits + kSchurSliceBits));
}

// Blocked forward substitution Y = R^-T X on NT*8 columns held as X[c * LDP + i] (i = row), one warp, stages
// [b_begin, b_end) of 8 rows each.  R is the upper factor in Ds (Ds[i * LDP + l] = R(l, i)); Ws holds the inverses of
// its 8 x 8 diagonal blocks (Ws[i * 12 + r] = inv(R_bb)(r, i % 8)).  Rows of earlier stages must already be solved.
template <int NT, int NSTAGE>
__device__ __forceinline__ void trsm_stages(const double* __restrict__ Ds, const double* __restrict__ Ws, double* X,
                                            double* Ts, int b_end, int g, int t) {
#pragma unroll
  for (int bi = 0; bi < NSTAGE; ++bi) {
    if (bi >= b_end) break;
    double acc[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) acc[nt][0] = acc[nt][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < bi * 8; ks += 4) {
      const double a = Ds[(bi * 8 + g) * LDP + ks + t];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) dmma884(acc[nt][0], acc[nt][1], a, X[(nt * 8 + g) * LDP + ks + t]);
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int cc = nt * 8 + 2 * t + e;
        Ts[cc * 12 + g] = X[cc * LDP + bi * 8 + g] - acc[nt][e];
      }
    __syncwarp();
    double y[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) y[nt][0] = y[nt][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < 8; ks += 4) {
      // A operand = inv(R_bb)^T: element [g][ks + t] = inv(R_bb)(ks + t, g)
      const double a = Ws[(bi * 8 + g) * 12 + ks + t];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) dmma884(y[nt][0], y[nt][1], a, Ts[(nt * 8 + g) * 12 + ks + t]);
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) X[(nt * 8 + 2 * t + e) * LDP + bi * 8 + g] = y[nt][e];
    __syncwarp();
  }
}

// Factored diagonal block b: side buffer (column-major PW x PW) -> front.  Runs one launch after the block was factored,
// when no CTA reads the unfactored block from the front any more.  256 threads, 16 independent copies each.
__device__ __forceinline__ void copy_diag_block(double* __restrict__ A, int ld, int F, const double* __restrict__ dfac, int b,
                                                int tid) {
  const double* src = dfac + (int64_t)b * PW * PW;
  double v[16];
#pragma unroll
  for (int q = 0; q < 16; ++q) v[q] = src[tid + 256 * q];
#pragma unroll
  for (int q = 0; q < 16; ++q) {
    const int idx = tid + 256 * q, j = idx / PW, i = idx - j * PW;
    const int gi = b * PW + i, gj = b * PW + j;
    if (i <= j && gj < F) A[gi + (int64_t)gj * ld] = v[q];
  }
}

__global__ void __launch_bounds__(256) front_step_kernel(ElimDev E, const StepTask* __restrict__ tasks, int p, int spc) {
  extern __shared__ __align__(16) double sm[];
  pdl_trigger();
  const StepTask me = tasks[blockIdx.x];   // compact list: role-A slices (the critical path) first, no empty CTAs
  const int c = me.clique;
  const CliqueDev ci = E.cliques[c];
  pdl_wait();
  const int F = ci.F, dim = F + ci.S + 1, ld = ci.ld;
  const int np = (F + PW - 1) / PW;
  double* A = E.arena + ci.front_off;
  double* dfac = E.dfac + ci.dfac_off;   // np blocks of PW x PW, column-major
  int task = me.task;
  const int tid = threadIdx.x;
  if (p >= np) return;   // the Schur complement step of this clique runs in schur_step_kernel
  const int k0 = p * PW, kb = min(PW, F - k0), rest0 = k0 + kb;
  const int nA = (dim - rest0 + SLICE - 1) / SLICE;   // column slices of this row panel
  const int nAc = (nA + spc - 1) / spc;               // role-A CTAs: CTA `task` takes slices task, task + nAc, ...
  if (task >= nAc) {
    if (p == 0) return;
    task -= nAc;
    const int nrt = (F - rest0 + TW - 1) / TW, nct = (dim - rest0 + TW - 1) / TW;
    if (task >= trapezoid_tiles(nrt, nct)) return;
    int ty, tx;
    trapezoid_decode(task, nct, ty, tx);
    if (!(E.debug & 1)) tile_update(A, ld, dim, rest0 + ty * TW, rest0 + tx * TW, F, k0 - PW, k0, sm);
    if (task == 0) copy_diag_block(A, ld, F, dfac, p - 1, tid);   // off the critical path: tile CTAs finish early
    return;
  }
  // ---- role A: slice `task` of row panel p (64 pivots) ----
  // Every role-A CTA factors the 64 x 64 diagonal block redundantly (cooperative 8 x 8-block Cholesky below), so no CTA
  // waits on another inside a launch.  Warps 1..7 own 8 columns each: look-ahead update by panel p-1 in the slack of the
  // factorisation, then the blocked triangular solve.  All GEMM-shaped work runs on the FP64 tensor pipe.
  if (E.debug & 4) return;
  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  double* Ps = sm;                    // Ps[i * LDP + k] = R[k0-64+k, k0+i]           (panel p-1, columns of D)
  double* Ds = Ps + PW * LDP;         // Ds[i * LDP + l] = D(l, i), l <= i            (column i of the block)
  double* Ws = Ds + PW * LDP;         // Ws[i * 12 + r] = inv(D_bb)(r, i % 8), b = i / 8
  double* dinv = Ws + PW * 12;        // 1 / D(i, i)
  int* s_ok = reinterpret_cast<int*>(dinv + PW);
  double* Xs = sm + kRoleAWarp0 + (warp > 0 ? warp - 1 : 0) * kRoleAColWarp;   // Xs[c * LDP + i]
  double* Ts = Xs + CWID * LDP;                                                 // Ts[c * 12 + ii]
  const bool probe = E.dbg && p == 2 && task == 0 && c == E.dbg_clique && lane == 0;   // GTB_DEBUG=8: cycle stamps
  auto stamp = [&](int k) { if (probe) E.dbg[k] = clock64(); };
  if (warp == 0) stamp(0);
  if (task == 0 && p >= 1 && F - rest0 <= 0) copy_diag_block(A, ld, F, dfac, p - 1, tid);   // no tile task to do it
  // Every global load of the CTA is issued before the first use.  The diagonal block and the matching columns of panel p-1
  // go first (one column = 64 contiguous doubles per two warp instructions): the Cholesky chain starts from them, while the
  // column warps' own slice (fragment layout) is not needed before the look-ahead update.
  const int jbase = rest0 + task * SLICE + (warp - 1) * CWID;
  const bool colwarp = warp > 0 && jbase < dim;
  double qf[16], b0[8][2];
  {
    double vp[8][2], vd[8][2];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int col = warp + 8 * q;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int l = lane + 32 * h;
        vp[q][h] = (p >= 1 && col < kb) ? A[(k0 - PW + l) + (int64_t)(k0 + col) * ld] : 0.0;
        vd[q][h] = (l <= col && col < kb) ? A[(k0 + l) + (int64_t)(k0 + col) * ld] : ((l == col) ? 1.0 : 0.0);
      }
    }
    if (colwarp) {
      const int j = jbase + g;
#pragma unroll
      for (int ks = 0; ks < 16; ++ks) qf[ks] = (p >= 1 && j < dim) ? A[(k0 - PW + ks * 4 + t) + (int64_t)j * ld] : 0.0;
#pragma unroll
      for (int mt = 0; mt < 8; ++mt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int i = mt * 8 + g, jj = jbase + 2 * t + e;
          b0[mt][e] = (jj < dim && i < kb) ? A[(k0 + i) + (int64_t)jj * ld] : 0.0;
        }
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int col = warp + 8 * q;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        Ps[col * LDP + lane + 32 * h] = vp[q][h];
        Ds[col * LDP + lane + 32 * h] = vd[q][h];
      }
    }
  }
  __syncthreads();
  if (warp == 0) stamp(1);
  // look-ahead on the diagonal block, shared by the 8 warps: D -= P^T P on the 36 upper 8 x 8 tile pairs
  if (p >= 1) {
    double acc[5][2];
    int pm[5], pn[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
      acc[q][0] = acc[q][1] = 0.0;
      // pair index -> (mt, nt), mt <= nt < 8, row-major over the upper triangle (indices >= 36 are masked below)
      trapezoid_decode(min(warp + 8 * q, 35), 8, pm[q], pn[q]);
    }
    if (warp == 0) stamp(4);
#pragma unroll 4
    for (int ks = 0; ks < PW; ks += 4) {
#pragma unroll
      for (int q = 0; q < 5; ++q)
        if (warp + 8 * q < 36)
          dmma884(acc[q][0], acc[q][1], Ps[(pm[q] * 8 + g) * LDP + ks + t], Ps[(pn[q] * 8 + g) * LDP + ks + t]);
    }
    if (warp == 0) stamp(6);
#pragma unroll
    for (int q = 0; q < 5; ++q)
      if (warp + 8 * q < 36) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int i = pm[q] * 8 + g, jj = pn[q] * 8 + 2 * t + e;
          if (i <= jj && jj < kb) Ds[jj * LDP + i] -= acc[q][e];
        }
      }
    if (warp == 0) stamp(7);
    __syncthreads();
  }
  if (warp == 0) stamp(2);
  // Cooperative 64 x 64 Cholesky in 8 x 8 blocks, right-looking, WITH LOOK-AHEAD: warp 0 alone walks the dependent chain
  //   factor block b (8 pivots, the whole 8 x 8 block held redundantly in the registers of every lane: pivot -> rsqrt ->
  //   scale -> update never leaves the register file) -> solve the 8 columns of the NEXT diagonal tile against it (one lane
  //   per column) -> bring that tile up to date (two DMMAs) -> factor block b + 1 ...
  // while warps 1..7 (the helpers) trail one block behind with everything that is not on the chain: the row solve of the
  // other columns (one thread per column) and the rank-8 DMMA updates of the other trailing tiles.  Hand-shakes are named
  // barriers: 1 = "row block b of the next tile's columns is solved" (warp 0 arrives, helpers wait), 2 = "bulk update of block
  // b is done" (helpers arrive, warp 0 waits one block later), 3 = the helpers among themselves.  The column warps' look-ahead
  // X = B - P^T Q (needs only panel p-1) is cut into one slice per block and runs in the helpers' slack.
  // (Before: every block went factor -> CTA barrier -> row solve -> CTA barrier -> update with warp 0 waiting at both barriers,
  // 2,450 cycles per block; the in-kernel stamps of the root front are in profiles/r02/launch_trace_c3_metis.txt.)
  {
    auto bar_sync = [](int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); };
    auto bar_arrive = [](int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); };
    const int nb8 = (kb + 7) >> 3;
    bool okw = true;
    int bad_pivot = -1;
    if (warp == 0) {
      double a[8][8], rv[8];
#pragma unroll
      for (int cc = 0; cc < 8; ++cc)
#pragma unroll
        for (int r = 0; r <= cc; ++r) a[r][cc] = Ds[cc * LDP + r];
      for (int b = 0; b < nb8; ++b) {
        const int o = 8 * b;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const double piv = a[k][k];
          if (okw && !(piv > 0.0)) bad_pivot = k0 + o + k;
          okw = okw && (piv > 0.0);   // Eigen LLT: NumericalIssue on a non-positive pivot (base/cholesky.cpp:126-128)
          const double rinv = rsqrt_chain(piv);
          rv[k] = rinv;
          a[k][k] = piv * rinv;
#pragma unroll
          for (int cc = k + 1; cc < 8; ++cc) a[k][cc] *= rinv;
#pragma unroll
          for (int r = k + 1; r < 8; ++r)
#pragma unroll
            for (int cc = r; cc < 8; ++cc) a[r][cc] -= a[k][r] * a[k][cc];
        }
#pragma unroll
        for (int cc = 0; cc < 8; ++cc)
#pragma unroll
          for (int r = 0; r <= cc; ++r) Ds[(o + cc) * LDP + o + r] = a[r][cc];   // every lane writes the same values
#pragma unroll
        for (int k = 0; k < 8; ++k) dinv[o + k] = rv[k];
        if (b == 0) stamp(3);
        if (b + 1 < nb8) {
          if (b >= 1) bar_sync(2, 256);   // bulk(b-1) is done: row block b and tile (b+1, b+1) carry every earlier update
          if (lane < 8) {                 // u = R_bb^-T x for the 8 columns of the next diagonal tile, factor from registers
            double* col = Ds + (o + 8 + lane) * LDP + o;
            double x[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) x[r] = col[r];
#pragma unroll
            for (int l = 0; l < 8; ++l) {
              x[l] *= rv[l];
#pragma unroll
              for (int r = l + 1; r < 8; ++r) x[r] -= a[l][r] * x[l];
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) col[r] = x[r];
          }
          __syncwarp();
          bar_arrive(1, 256);
          // tile (b+1, b+1) -= U^T U, straight into the register copy of the next block
          double c0 = 0.0, c1 = 0.0;
#pragma unroll
          for (int ks = 0; ks < 8; ks += 4) {
            const double f = Ds[(o + 8 + g) * LDP + o + ks + t];
            dmma884(c0, c1, f, f);
          }
          // this lane holds C[g][2t], C[g][2t+1] of the product; the tile lives at Ds[col * LDP + row]
          if (g <= 2 * t) Ds[(o + 8 + 2 * t) * LDP + o + 8 + g] -= c0;
          if (g <= 2 * t + 1) Ds[(o + 8 + 2 * t + 1) * LDP + o + 8 + g] -= c1;
          __syncwarp();
#pragma unroll
          for (int cc = 0; cc < 8; ++cc)
#pragma unroll
            for (int r = 0; r <= cc; ++r) a[r][cc] = Ds[(o + 8 + cc) * LDP + o + 8 + r];
        }
      }
      if (nb8 >= 2) bar_sync(2, 256);   // the last bulk update (balances the helpers' arrival)
    } else {
      // helpers; the column warps among them carry the look-ahead accumulators through the loop
      double acc[8][2];
#pragma unroll
      for (int mt = 0; mt < 8; ++mt) acc[mt][0] = acc[mt][1] = 0.0;
      auto lookahead = [&](int ks0, int ks1) {
        if (!colwarp || p < 1) return;
#pragma unroll
        for (int ks = 0; ks < 16; ++ks) {
          if (ks < ks0 || ks >= ks1) continue;
#pragma unroll
          for (int mt = 0; mt < 8; ++mt) dmma884(acc[mt][0], acc[mt][1], Ps[(mt * 8 + g) * LDP + ks * 4 + t], qf[ks]);
        }
      };
      const int ht = tid - 32;   // 0 .. 223
      for (int b = 0; b + 1 < nb8; ++b) {
        const int o = 8 * b;
        bar_sync(1, 256);   // R_bb, its reciprocal diagonal and the solved block (b, b+1) are published
        const int ncol = 8 * nb8 - (o + 16);
        if (ht < ncol) {    // row block b of the columns beyond the next tile, one column per thread
          double* col = Ds + (o + 16 + ht) * LDP + o;
          double x[8];
#pragma unroll
          for (int r = 0; r < 8; ++r) x[r] = col[r];
#pragma unroll
          for (int l = 0; l < 8; ++l) {
            x[l] *= dinv[o + l];
#pragma unroll
            for (int r = l + 1; r < 8; ++r) x[r] -= Ds[(o + r) * LDP + o + l] * x[l];
          }
#pragma unroll
          for (int r = 0; r < 8; ++r) col[r] = x[r];
        }
        bar_sync(3, 224);
        // trailing tiles (mt, nt), b < mt <= nt < nb8, without (b+1, b+1) which warp 0 keeps: C -= U(:, mt)^T U(:, nt)
        const int m = nb8 - 1 - b, npair = (m * (m + 1)) >> 1;
        for (int q = warp; q < npair; q += 7) {
          int mt, nt;
          trapezoid_decode(q, m, mt, nt);
          const int ct = b + 1 + nt, rt = b + 1 + mt;
          double c0 = 0.0, c1 = 0.0;
#pragma unroll
          for (int ks = 0; ks < 8; ks += 4) dmma884(c0, c1, Ds[(8 * rt + g) * LDP + o + ks + t], Ds[(8 * ct + g) * LDP + o + ks + t]);
          Ds[(8 * ct + 2 * t) * LDP + 8 * rt + g] -= c0;
          Ds[(8 * ct + 2 * t + 1) * LDP + 8 * rt + g] -= c1;
        }
        __syncwarp();
        bar_arrive(2, 256);
        lookahead(2 * b, 2 * b + 2);
      }
      lookahead(nb8 >= 2 ? 2 * (nb8 - 1) : 0, 16);   // what the loop did not reach (ragged last panel)
      if (colwarp) {
#pragma unroll
        for (int mt = 0; mt < 8; ++mt)
#pragma unroll
          for (int e = 0; e < 2; ++e) Xs[(2 * t + e) * LDP + mt * 8 + g] = b0[mt][e] - acc[mt][e];
        if (warp == 1) stamp(8);
      }
    }
    __syncthreads();
    if (lane < 8 && warp < nb8) {
      // inverses of the 8 x 8 diagonal blocks for the blocked solve below: one block per warp, one column per lane,
      // back-substitution within the block
      const int b8 = warp * 8, jj = lane, cidx = b8 + jj;
      const double myinv = dinv[cidx];
      double w[8];
#pragma unroll
      for (int i = 7; i >= 0; --i) {
        double sacc = 0.0;
#pragma unroll
        for (int k = i + 1; k < 8; ++k)
          if (k <= jj) sacc += Ds[(b8 + k) * LDP + b8 + i] * w[k];
        w[i] = (i == jj) ? myinv : ((i < jj) ? -sacc * dinv[b8 + i] : 0.0);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) Ws[cidx * 12 + i] = w[i];
    }
    if (warp == 0) {
      stamp(5);
      if (lane == 0) *s_ok = okw ? 1 : 0;
      if (task == 0 && !okw && lane == 0) {
        // the member (reference clique) whose pivot failed
        const int m0 = E.orig_off[c], m1 = E.orig_off[c + 1];
        int m = m0;
        while (m + 1 < m1 && E.orig_end[m] <= bad_pivot) ++m;
        atomicMin(E.fail, fail_code(c, m - m0));
      }
    }
  }
  __syncthreads();
  if (!*s_ok) return;
  if (task == 0) {
    // factored diagonal block -> side buffer (other CTAs of this launch still read the unfactored block from the front)
    double* db = dfac + p * PW * PW;
    for (int idx = tid; idx < PW * PW; idx += blockDim.x) {
      const int j = idx / PW, i = idx - j * PW;   // column-major
      db[idx] = (i <= j) ? Ds[j * LDP + i] : 0.0;
    }
    if (tid == 0) {
      // Exponent test on the last two pivots of a clique (base/cholesky.cpp:146-158), for every member of this front
      // whose pivots end inside this panel: an amalgamated front must fail exactly where the reference's cliques do.
      const int m0 = E.orig_off[c], m1 = E.orig_off[c + 1];
      int begin = 0;
      for (int m = m0; m < m1; ++m) {
        const int end = E.orig_end[m];
        if (end > k0 && end <= k0 + kb) {
          const int l = end - 1 - k0;   // last pivot of the member, inside this panel
          const double dlast = Ds[l * LDP + l];
          bool good;
          if (end - begin >= 2) {
            const double r2 = (l >= 1) ? Ds[(l - 1) * LDP + l - 1] : dfac[(p - 1) * PW * PW + (PW - 1) * PW + (PW - 1)];
            good = frexp_exp(r2) - frexp_exp(dlast) < 12;
          } else {
            good = frexp_exp(dlast) > -12;
          }
          if (!good) atomicMin(E.fail, fail_code(c, m - m0));
        }
        begin = end;
      }
    }
  }
  if (!colwarp) return;
  if (warp == 1) stamp(9);
  int jb = jbase;
  for (int sl = task;;) {
    // blocked forward substitution Y = R^-T X in 8-row stages
    trsm_stages<1, 8>(Ds, Ws, Xs, Ts, (kb + 7) / 8, g, t);
    if (warp == 1 && sl == task) stamp(10);
    for (int cc = 0; cc < CWID; ++cc) {
      const int j = jb + cc;
      if (j < dim) {
        if (lane < kb) A[(k0 + lane) + (int64_t)j * ld] = Xs[cc * LDP + lane];
        if (lane + NB < kb) A[(k0 + lane + NB) + (int64_t)j * ld] = Xs[cc * LDP + lane + NB];
      }
    }
    if (warp == 1 && sl == task) stamp(11);
    // further slices of this CTA (wide launches only): the factored block is reused, each column warp repeats its
    // look-ahead update and solve on the next 8 columns it owns; everything is warp-local
    sl += nAc;
    if (sl >= nA) break;
    jb = rest0 + sl * SLICE + (warp - 1) * CWID;
    if (jb >= dim) break;
    __syncwarp();
    {
      const int j = jb + g;
#pragma unroll
      for (int ks = 0; ks < 16; ++ks) qf[ks] = (p >= 1 && j < dim) ? A[(k0 - PW + ks * 4 + t) + (int64_t)j * ld] : 0.0;
#pragma unroll
      for (int mt = 0; mt < 8; ++mt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int i = mt * 8 + g, jj = jb + 2 * t + e;
          b0[mt][e] = (jj < dim && i < kb) ? A[(k0 + i) + (int64_t)jj * ld] : 0.0;
        }
      double acc[8][2];
#pragma unroll
      for (int mt = 0; mt < 8; ++mt) acc[mt][0] = acc[mt][1] = 0.0;
      if (p >= 1) {
#pragma unroll
        for (int ks = 0; ks < 16; ++ks)
#pragma unroll
          for (int mt = 0; mt < 8; ++mt) dmma884(acc[mt][0], acc[mt][1], Ps[(mt * 8 + g) * LDP + ks * 4 + t], qf[ks]);
      }
#pragma unroll
      for (int mt = 0; mt < 8; ++mt)
#pragma unroll
        for (int e = 0; e < 2; ++e) Xs[(2 * t + e) * LDP + mt * 8 + g] = b0[mt][e] - acc[mt][e];
    }
    __syncwarp();
  }
}

// Step np of a clique (after its last panel): Schur complement C -= S^T S with K = F on 64 x 64 DMMA tiles, and the last
// factored diagonal block moves from the side buffer into the front.  A kernel of its own so that it is not held to one
// CTA per SM by role A's registers: 70 % of the large-front flops of a bundle-adjustment tree are in these launches.
__global__ void __launch_bounds__(256, 2) schur_step_kernel(ElimDev E, const StepTask* __restrict__ tasks) {
  extern __shared__ __align__(16) double sm[];
  pdl_trigger();
  const StepTask me = tasks[blockIdx.x];
  const CliqueDev ci = E.cliques[me.clique];
  pdl_wait();
  const int F = ci.F, dim = F + ci.S + 1, ld = ci.ld;
  double* A = E.arena + ci.front_off;
  const int nt = (dim - F + TW - 1) / TW;
  const int tile = me.task & ((1 << kSchurTileBits) - 1), slice = (me.task >> kSchurTileBits) & ((1 << kSchurSliceBits) - 1);
  const int nsplit = me.task >> (kSchurTileBits + kSchurSliceBits);
  int ty, tx;
  trapezoid_decode(tile, nt, ty, tx);
  const bool probe = E.dbg && me.clique == E.dbg_child && threadIdx.x == 0 && (me.task == 0 || me.task == 77);
  long long t0 = 0, g0 = 0;
  if (probe) { t0 = clock64(); asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g0)); }
  // a root's trailing block is the 1 x 1 constant term: nothing reads it
  if (!(E.debug & 1) && ci.parent >= 0) {
    const int nchunk = (F + NB - 1) / NB;
    const int ka = nsplit <= 1 ? 0 : (int)((int64_t)nchunk * slice / nsplit) * NB;
    const int kend = nsplit <= 1 ? F : min(F, (int)((int64_t)nchunk * (slice + 1) / nsplit) * NB);
    // hand-shake flag of this tile when its K sum is split (ordered hand-over): value epoch32 + s lets slice s in
    int* flag = (E.ordered && nsplit > 1) ? E.tile_flags + ci.flag_off + tile : nullptr;
    if (flag && slice == 0 && threadIdx.x == 0) *reinterpret_cast<volatile int*>(flag) = E.epoch32;   // (no predecessor)
    if (ci.flags & kCliqueFuseUp) {
      const CliqueDev cp = E.cliques[ci.parent];
      tile_update<2>(A, ld, dim, F + ty * TW, F + tx * TW, dim, ka, kend, sm, E.pmap + ci.pmap_off, E.arena + cp.front_off, cp.ld, F,
                     slice == 0, flag, E.epoch32 + slice, E.debug);
    } else if (nsplit <= 1) {
      tile_update<0>(A, ld, dim, F + ty * TW, F + tx * TW, dim, 0, F, sm, nullptr, nullptr, 0, 0, false, nullptr, 0, E.debug);
    } else {
      tile_update<1>(A, ld, dim, F + ty * TW, F + tx * TW, dim, ka, kend, sm, nullptr, nullptr, 0, 0, false, flag, E.epoch32 + slice, E.debug);
    }
  }
  if (probe) {
    long long g1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g1));
    const int o = me.task == 0 ? 12 : 14;
    E.dbg[o] = clock64() - t0;
    E.dbg[o + 1] = g1 - g0;
  }
  if (tile == 0 && slice == 0) copy_diag_block(A, ld, F, E.dfac + ci.dfac_off, (F + PW - 1) / PW - 1, threadIdx.x);
}

// ----------------------------------------------------------------------------------------------------
// back-substitution: x_F = R^-1 (d - S x_S), levels from the roots down (linear/linearAlgorithms-inst.h:54-116)
// WARP = true : one warp per clique (F <= 32, F+S <= 160), kBsWarps cliques per CTA
// WARP = false: one CTA per clique; blocked TRSV, 32 pivots per step, diagonal block solved by warp 0 in registers
// ----------------------------------------------------------------------------------------------------
constexpr int kBsWarps = 8;
constexpr int kBsCtaThreads = 512;
constexpr int kBsWarpDoubles = 32 + 160 + 32;   // rhs, xs, reduction scratch per warp

// S x_S for CTA-class cliques, spread over many CTAs: blockIdx.x = clique slot, y = 128-row block, z = 128-column
// chunk.  Partial sums go to scratch[off + chunk * F + i]; backsolve_kernel<false> adds them in chunk order
// (deterministic, no atomics).
constexpr int kGemvRows = 128, kGemvCols = 32;
__global__ void __launch_bounds__(kGemvRows)
backsolve_gemv_kernel(ElimDev E, const int32_t* __restrict__ list, const int32_t* __restrict__ scratch_off,
                      const double* __restrict__ delta, double* __restrict__ scratch) {
  pdl_trigger();
  const CliqueDev ci = E.cliques[list[blockIdx.x]];
  const int F = ci.F, S = ci.S, ld = ci.ld;
  const int i0 = blockIdx.y * kGemvRows, j0 = blockIdx.z * kGemvCols;
  if (i0 >= F || j0 >= S) return;
  {
    // The CTA that solves this clique next reads the triangle of R alone, one memory latency per 32-row block: ask L2
    // for it from here, a few lines per thread over all the CTAs of the clique (R does not depend on the level above,
    // so this runs before the wait).
    const int lines = (F + 15) / 16;   // 128-byte lines per column of R
    const int nz = (S + kGemvCols - 1) / kGemvCols, ny = (F + kGemvRows - 1) / kGemvRows;
    const double* R = E.arena + ci.front_off;
    for (int q = (blockIdx.y * nz + blockIdx.z) * kGemvRows + threadIdx.x; q < F * lines; q += ny * nz * kGemvRows) {
      const int col = q / lines, r0 = (q - col * lines) * 16;
      if (r0 <= col) asm volatile("prefetch.global.L2 [%0];" ::"l"(R + r0 + (int64_t)col * ld));
    }
  }
  pdl_wait();
  __shared__ double xs[kGemvCols];
  const int nj = min(kGemvCols, S - j0);
  const int32_t* g = E.gidx + ci.gidx_off + F + j0;
  for (int j = threadIdx.x; j < nj; j += kGemvRows) xs[j] = delta[g[j]];
  __syncthreads();
  const int i = i0 + threadIdx.x;
  if (i >= F) return;
  const double* row = E.arena + ci.front_off + i + (int64_t)(F + j0) * ld;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  int j = 0;
  for (; j + 32 <= nj; j += 32) {   // 32 independent loads in flight per thread: the kernel is latency-bound
    double v[32];
#pragma unroll
    for (int u = 0; u < 32; ++u) v[u] = row[(int64_t)(j + u) * ld];
#pragma unroll
    for (int u = 0; u < 32; u += 4) {
      a0 += v[u] * xs[j + u];
      a1 += v[u + 1] * xs[j + u + 1];
      a2 += v[u + 2] * xs[j + u + 2];
      a3 += v[u + 3] * xs[j + u + 3];
    }
  }
  if (j < nj) {
    double v[32];
#pragma unroll
    for (int u = 0; u < 32; ++u) v[u] = (j + u < nj) ? row[(int64_t)(j + u) * ld] : 0.0;
#pragma unroll
    for (int u = 0; u < 32; ++u)
      if (j + u < nj) a0 += v[u] * xs[j + u];
  }
  scratch[scratch_off[blockIdx.x] + (int64_t)blockIdx.z * F + i] = (a0 + a1) + (a2 + a3);
}

// Inverses of the 32 x 32 diagonal blocks of R of the CTA-class cliques, one warp per block, once per factorisation:
// with them a block of the back-substitution is a 32 x 32 matrix-vector product instead of a chain of 32 dependent
// pivots.  task = (clique, block, destination block); out block layout [column k][row l], 1024 doubles, zero below
// the diagonal, identity on the padding of a ragged last block.  Lane j solves R x = e_j from the bottom up with 32
// running sums (right-looking), so the dependent chain is one multiply-add per row.
constexpr int kInvWarps = 4;
__global__ void __launch_bounds__(kInvWarps * 32)
diag_inverse_kernel(ElimDev E, const int32_t* __restrict__ tasks, int count, double* __restrict__ rinv) {
  __shared__ double us[kInvWarps][32 * 33 + 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int t = blockIdx.x * kInvWarps + warp;
  pdl_trigger();
  if (t >= count) return;
  const int32_t c = tasks[3 * t], bi = tasks[3 * t + 1], dstb = tasks[3 * t + 2];
  const CliqueDev ci = E.cliques[c];
  pdl_wait();
  const int ld = ci.ld, b0 = bi * 32, kb = min(32, ci.F - b0);
  const double* A = E.arena + ci.front_off;
  double* u = us[warp];            // u[k * 33 + i] = R(b0 + i, b0 + k), i <= k
  double* rd = u + 32 * 33;        // reciprocal diagonal
  double sum[32], x[32];
#pragma unroll
  for (int k = 0; k < 32; ++k) {   // all 32 loads in flight before the first store
    x[k] = (k == lane) ? 1.0 : 0.0;
    if (k < kb && lane <= k) x[k] = A[(b0 + lane) + (int64_t)(b0 + k) * ld];
  }
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    u[k * 33 + lane] = x[k];
    if (k == lane) rd[lane] = 1.0 / x[k];
  }
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 32; ++i) sum[i] = 0.0;
#pragma unroll
  for (int k = 31; k >= 0; --k) {
    const double xk = ((k == lane ? 1.0 : 0.0) - sum[k]) * rd[k];
    x[k] = (k <= lane) ? xk : 0.0;
#pragma unroll
    for (int i = 0; i < k; ++i) sum[i] += u[k * 33 + i] * x[k];
  }
  // column `lane` of the inverse -> out[lane * 32 + i], through shared memory so that the stores are coalesced
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 32; ++i) u[lane * 33 + i] = x[i];
  __syncwarp();
  double* out = rinv + (int64_t)dstb * 1024;
  for (int k = 0; k < 32; ++k) out[k * 32 + lane] = u[k * 33 + lane];
}

// Point cliques with at most 8 cameras, four per warp (8 lanes each): x_F = R^-1 (d - S x_S) over the three conditional rows
// stored as one 32-byte column per separator scalar.  A warp per point leaves three quarters of its lanes idle on the 9 k <=
// 72 columns of such a point; this is the bulk of the cliques of a bundle-adjustment problem (C4: 993 k of them).
__global__ void __launch_bounds__(kBsWarps * 32)
backsolve_points_kernel(ElimDev E, const int32_t* __restrict__ list, int count, double* __restrict__ delta) {
  constexpr int LPP = 8;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sl = lane % LPP, sw = lane / LPP;
  const int slot = (blockIdx.x * kBsWarps + warp) * (32 / LPP) + sw;
  pdl_trigger();
  const bool active = slot < count;
  const CliqueDev ci = E.cliques[list[active ? slot : count - 1]];   // idle groups shadow the last point: the warp stays converged
  pdl_wait();
  const int S = ci.S;
  const double* A = E.arena + ci.front_off;
  const int32_t* g = E.gidx + ci.gidx_off;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0;
#pragma unroll 3
  for (int j = sl; j < S; j += LPP) {
    const double x = delta[g[3 + j]];
    const Col4 col = ld_col(A + (int64_t)(3 + j) * 4);
    a0 += col.x * x; a1 += col.y * x; a2 += col.z * x;
  }
#pragma unroll
  for (int o = LPP / 2; o > 0; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
    a2 += __shfl_xor_sync(0xffffffffu, a2, o);
  }
  if (active && sl == 0) {
    const double* d = A + (int64_t)(3 + S) * 4;
    const double x2 = (d[2] - a2) * (1.0 / A[2 + 2 * 4]);
    const double x1 = (d[1] - a1 - A[1 + 2 * 4] * x2) * (1.0 / A[1 + 1 * 4]);
    const double x0 = (d[0] - a0 - A[0 + 1 * 4] * x1 - A[0 + 2 * 4] * x2) * (1.0 / A[0]);
    delta[g[0]] = x0; delta[g[1]] = x1; delta[g[2]] = x2;
  }
}

template <bool WARP>
__global__ void __launch_bounds__(WARP ? kBsWarps * 32 : kBsCtaThreads)
backsolve_kernel(ElimDev E, const int32_t* __restrict__ list, int count, double* __restrict__ delta,
                 const int32_t* __restrict__ scratch_off, const double* __restrict__ scratch,
                 const int32_t* __restrict__ rinv_off, const double* __restrict__ rinv) {
  extern __shared__ double sh[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = WARP ? blockIdx.x * kBsWarps + warp : blockIdx.x;
  if (slot >= count) return;   // WARP: whole warp exits together; CTA: never true
  const int gsize = WARP ? 32 : (int)blockDim.x;
  const int gtid = WARP ? lane : (int)threadIdx.x;
  pdl_trigger();
  const CliqueDev ci = E.cliques[list[slot]];
  pdl_wait();
  const int F = ci.F, S = ci.S, N = F + S, ld = ci.ld;
  const double* A = E.arena + ci.front_off;
  const int32_t* g = E.gidx + ci.gidx_off;
  double* rhs = WARP ? sh + warp * kBsWarpDoubles : sh;     // F (WARP: 32)
  double* xs = WARP ? rhs + 32 : sh + F;                    // S (WARP: 160)
  double* red = WARP ? xs + 160 : sh + N;                   // gsize
  double* xblk = WARP ? rhs : red + gsize;                  // 32 (CTA only; WARP has a single block)
  auto gsync = [&]() { if (WARP) __syncwarp(); else __syncthreads(); };

  if (!WARP) {
    // rhs = d - (S x_S): partial products were computed by backsolve_gemv_kernel, summed here in chunk order
    const int nchunk = (S + kGemvCols - 1) / kGemvCols;
    const double* part = scratch + scratch_off[slot];
    for (int i = gtid; i < F; i += gsize) {
      double a = A[i + (int64_t)N * ld];
      for (int q = 0; q < nchunk; ++q) a -= part[(int64_t)q * F + i];
      rhs[i] = a;
    }
  } else if (F == 3 && ld == 4) {
    // point front (point_front_kernel): three conditional rows stored as one 32-byte column per separator scalar
    double a0 = 0.0, a1 = 0.0, a2 = 0.0;
    for (int j = lane; j < S; j += 32) {
      const double x = delta[g[3 + j]];
      const Col4 col = ld_col(A + (int64_t)(3 + j) * 4);   // one 32-byte request per column
      a0 += col.x * x; a1 += col.y * x; a2 += col.z * x;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a0 += __shfl_xor_sync(0xffffffffu, a0, o);
      a1 += __shfl_xor_sync(0xffffffffu, a1, o);
      a2 += __shfl_xor_sync(0xffffffffu, a2, o);
    }
    if (lane == 0) {
      const double* d = A + (int64_t)N * 4;
      const double x2 = (d[2] - a2) * (1.0 / A[2 + 2 * 4]);
      const double x1 = (d[1] - a1 - A[1 + 2 * 4] * x2) * (1.0 / A[1 + 1 * 4]);
      const double x0 = (d[0] - a0 - A[0 + 1 * 4] * x1 - A[0 + 2 * 4] * x2) * (1.0 / A[0]);
      delta[g[0]] = x0; delta[g[1]] = x1; delta[g[2]] = x2;
    }
    return;
  } else {
    for (int j = gtid; j < S; j += gsize) xs[j] = delta[g[F + j]];
    gsync();
    int Fp = 1;
    while (Fp < F) Fp <<= 1;
    const int G = gsize / Fp;
    const int i = gtid & (Fp - 1), jg = gtid / Fp;
    double acc = 0.0;
    if (i < F) {
      const double* row = A + i + (int64_t)F * ld;
      for (int j = jg; j < S; j += G) acc -= row[(int64_t)j * ld] * xs[j];
    }
    red[gtid] = acc;
    gsync();
    if (gtid < F) {
      double a = A[gtid + (int64_t)N * ld];
      for (int q = 0; q < G; ++q) a += red[q * Fp + gtid];
      rhs[gtid] = a;
    }
  }
  gsync();
  // blocked back-substitution with R (upper triangular), bottom block first
  if (WARP) {
    // F <= 32: a single block
    const int kb = F;
    double r[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) r[k] = (k < kb && lane <= k && lane < kb) ? A[lane + (int64_t)k * ld] : 1.0;
    double my = (lane < kb) ? rhs[lane] : 0.0;
    // reciprocal of this lane's diagonal entry: one division per lane instead of a chain of 32
    double dg = 1.0;
#pragma unroll
    for (int k = 0; k < 32; ++k)
      if (k == lane) dg = r[k];
    const double rd = 1.0 / dg;
#pragma unroll
    for (int k = 31; k >= 0; --k) {
      if (k < kb) {
        const double xk = __shfl_sync(0xffffffffu, my, k) * __shfl_sync(0xffffffffu, rd, k);
        if (lane == k) my = xk;
        else if (lane < k) my -= r[k] * xk;
      }
    }
    if (lane < kb) rhs[lane] = my;
  } else {
    // CTA variant, software-pipelined over the 32-row blocks: ONE barrier per block.  Warp 0 walks the dependent chain
    // alone: x_b = inv(R[b, b]) r_b with the inverse block diag_inverse_kernel left (a 32 x 32 matrix-vector product,
    // no pivot chain), publish x_b, and fold x_b's contribution to block b-1 (the 32 x 32 block R[b-1, b], prefetched
    // into registers before the barrier) into its next right-hand side itself.  The other warps run one block behind:
    // after the barrier they apply x_b to every row above block b-1 (32 loads in flight per row) and stage the inverse
    // block warp 0 needs two steps later (its loads are issued before the barrier too), while warp 0 is already on
    // block b-1.  A row's sum still runs over the blocks in descending order, as in the plain right-looking loop.
    double* xbuf = xblk;             // [2][32]: x of the block just solved (double-buffered)
    double* dblk = xblk + 64;        // [2][32 * 33]: dblk[k * 33 + l] = inv(R[b, b])(l, k) of the block to solve
    const int nb = (F + 31) / 32;
    const double* inv = rinv + (int64_t)rinv_off[slot] * 1024;
    constexpr int kStagePer = (1024 + kBsCtaThreads - 32 - 1) / (kBsCtaThreads - 32);   // elements per staging thread
    for (int bq = nb - 1; bq >= max(0, nb - 2); --bq) {
      double* dst = dblk + (bq & 1) * (32 * 33);
      for (int q = gtid; q < 1024; q += gsize) dst[(q >> 5) * 33 + (q & 31)] = inv[(int64_t)bq * 1024 + q];
    }
    __syncthreads();
    double carry = 0.0;   // warp 0: (R[b, b+1] x_{b+1})(lane)
    for (int bi = nb - 1; bi >= 0; --bi) {
      const int b0 = bi * 32, kb = min(32, F - b0);
      double w[32];       // warp 0: R[b-1, b], row b0 - 32 + lane; other warps: their first row of this block's update.
                          // Loaded before the barrier: in flight while warp 0 is on block b.
      double sd[kStagePer];   // other warps: their elements of the inverse block b-2
      const int i1 = warp == 0 ? b0 - 32 + lane : gtid - 32;
      const bool have = warp == 0 ? bi >= 1 : i1 < b0 - 32;
      {
        const double* row = A + (have ? i1 : 0) + (int64_t)b0 * ld;
#pragma unroll
        for (int k = 0; k < 32; ++k) w[k] = row[(int64_t)min(k, kb - 1) * ld];   // plain loads: a select behind each
                                                                                   // would wait for it here (measured:
                                                                                   // 3.0k of 4.5k cycles per block);
                                                                                   // x_b is 0 beyond kb
      }
      if (warp > 0 && bi >= 2) {
#pragma unroll
        for (int n = 0; n < kStagePer; ++n) {
          const int q = gtid - 32 + n * (kBsCtaThreads - 32);
          sd[n] = inv[(int64_t)(bi - 2) * 1024 + min(q, 1023)];
        }
      }
      if (warp == 0) {
        const double* db = dblk + (bi & 1) * (32 * 33);
        const double r = (lane < kb) ? rhs[b0 + lane] - carry : 0.0;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
        for (int k = 0; k < 32; k += 4) {
          s0 += db[k * 33 + lane] * __shfl_sync(0xffffffffu, r, k);
          s1 += db[(k + 1) * 33 + lane] * __shfl_sync(0xffffffffu, r, k + 1);
          s2 += db[(k + 2) * 33 + lane] * __shfl_sync(0xffffffffu, r, k + 2);
          s3 += db[(k + 3) * 33 + lane] * __shfl_sync(0xffffffffu, r, k + 3);
        }
        const double my = (s0 + s1) + (s2 + s3);
        if (lane < kb) rhs[b0 + lane] = my;
        double* xb = xbuf + (bi & 1) * 32;
        xb[lane] = (lane < kb) ? my : 0.0;
        __syncwarp();
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int k = 0; k < 32; k += 2) {
          a0 += w[k] * xb[k];
          a1 += w[k + 1] * xb[k + 1];
        }
        carry = a0 + a1;
      }
      __syncthreads();   // x_b is published; the other warps have finished applying x_{b+1} and staging block b-1
      if (warp > 0) {
        const double* xb = xbuf + (bi & 1) * 32;
        if (have) {
          double a0 = 0.0, a1 = 0.0;
#pragma unroll
          for (int k = 0; k < 32; k += 2) {
            a0 += w[k] * xb[k];
            a1 += w[k + 1] * xb[k + 1];
          }
          rhs[i1] -= a0 + a1;
        }
        if (bi >= 2) {   // into the buffer of block b, which warp 0 has left
          double* dst = dblk + (bi & 1) * (32 * 33);
#pragma unroll
          for (int n = 0; n < kStagePer; ++n) {
            const int q = gtid - 32 + n * (kBsCtaThreads - 32);
            if (q < 1024) dst[(q >> 5) * 33 + (q & 31)] = sd[n];
          }
        }
        for (int i = i1 + gsize - 32; i < b0 - 32; i += gsize - 32) {
          const double* row = A + i + (int64_t)b0 * ld;
          double u[32];
#pragma unroll
          for (int k = 0; k < 32; ++k) u[k] = row[(int64_t)min(k, kb - 1) * ld];
          double a0 = 0.0, a1 = 0.0;
#pragma unroll
          for (int k = 0; k < 32; k += 2) {
            a0 += u[k] * xb[k];
            a1 += u[k + 1] * xb[k + 1];
          }
          rhs[i] -= a0 + a1;
        }
      }
    }
  }
  gsync();
  for (int i = gtid; i < F; i += gsize) delta[g[i]] = rhs[i];
}

}  // namespace gtb
