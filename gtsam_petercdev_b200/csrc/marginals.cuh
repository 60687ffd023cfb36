// marginals.cuh -- marginal covariances from the device Bayes tree.
//
// Replaces Marginals::marginalCovariance (nonlinear/Marginals.cpp:108-150), which for every queried variable walks the
// Bayes tree with BayesTree::marginalFactor (inference/BayesTree-inst.h, shortcuts + a dense elimination per query).
// Here ONE sweep from the roots to the leaves computes the covariance of every clique's variables from its conditional
// [R S d] and its parent's covariance (the recursion behind sparse selective inversion): with T = R^-1 S and
// Sigma_SS gathered from the parent through the child -> parent column map,
//     Sigma_FS = -T Sigma_SS,      Sigma_FF = R^-1 R^-T + T Sigma_SS T^T.
// One CTA per clique, levels in launch order; a clique with children keeps its full (F+S)^2 block for them, a leaf only
// its F x F block.  All FP64, fixed summation order (deterministic).  Not on the LM hot path: plain global-memory loops.
#pragma once
#include "eliminate.cuh"

namespace gtb {

struct MargClique {
  int64_t sigma_off;     // this clique's covariance block (ld = sld), column-major, full symmetric
  int64_t scratch_off;   // Rinv (F x F), T (F x S), U (F x S)
  int32_t sld;           // F + S with children, F for a leaf
  int32_t pad;
};

__global__ void __launch_bounds__(1024) marginal_kernel(ElimDev E, const MargClique* __restrict__ mc, const int32_t* __restrict__ list,
                                                       double* __restrict__ sigma, double* __restrict__ scratch) {
  const int c = list[blockIdx.x];
  const CliqueDev ci = E.cliques[c];
  const MargClique m = mc[c];
  const int F = ci.F, S = ci.S, ld = ci.ld, sld = m.sld;
  const double* A = E.arena + ci.front_off;       // R = A[0:F, 0:F] (upper), S block = A[0:F, F:F+S]
  double* Sg = sigma + m.sigma_off;
  double* Ri = scratch + m.scratch_off;           // Rinv, column-major F x F (upper)
  double* T = Ri + (int64_t)F * F;                // F x S, column-major
  double* U = T + (int64_t)F * S;                 // F x S
  const int tid = threadIdx.x, nt = blockDim.x;
  // 1. Rinv: column j solves R x = e_j by back substitution (x_i = 0 for i > j)
  for (int j = tid; j < F; j += nt) {
    double* x = Ri + (int64_t)j * F;
    for (int i = F - 1; i > j; --i) x[i] = 0.0;
    x[j] = 1.0 / A[j + (int64_t)j * ld];
    for (int i = j - 1; i >= 0; --i) {
      double s = 0.0;
      for (int k = i + 1; k <= j; ++k) s += A[i + (int64_t)k * ld] * x[k];
      x[i] = -s / A[i + (int64_t)i * ld];
    }
  }
  __syncthreads();
  // 2. T = Rinv S
  for (int64_t q = tid; q < (int64_t)F * S; q += nt) {
    const int i = (int)(q % F), j = (int)(q / F);
    double s = 0.0;
    for (int k = i; k < F; ++k) s += Ri[i + (int64_t)k * F] * A[k + (int64_t)(F + j) * ld];
    T[q] = s;
  }
  __syncthreads();
  // 3. U = T Sigma_SS, Sigma_SS gathered from the parent's block
  if (S > 0) {
    const CliqueDev cp = E.cliques[ci.parent];
    const MargClique mp = mc[ci.parent];
    const double* P = sigma + mp.sigma_off;
    const int32_t* pm = E.pmap + ci.pmap_off;
    for (int64_t q = tid; q < (int64_t)F * S; q += nt) {
      const int i = (int)(q % F), j = (int)(q / F);
      const int64_t pj = (int64_t)pm[j] * mp.sld;
      double s = 0.0;
      for (int k = 0; k < S; ++k) s += T[i + (int64_t)k * F] * P[pm[k] + pj];
      U[q] = s;
    }
    (void)cp;
  }
  __syncthreads();
  // 4. Sigma_FF = Rinv Rinv^T + U T^T (symmetric), Sigma_FS = -U, and with children also Sigma_SF, Sigma_SS
  for (int64_t q = tid; q < (int64_t)F * F; q += nt) {
    const int i = (int)(q % F), j = (int)(q / F);
    if (i > j) continue;
    double s = 0.0;
    for (int k = j; k < F; ++k) s += Ri[i + (int64_t)k * F] * Ri[j + (int64_t)k * F];
    for (int k = 0; k < S; ++k) s += U[i + (int64_t)k * F] * T[j + (int64_t)k * F];
    Sg[i + (int64_t)j * sld] = s;
    Sg[j + (int64_t)i * sld] = s;
  }
  if (sld > F) {
    const MargClique mp = mc[ci.parent];
    const double* P = sigma + mp.sigma_off;
    const int32_t* pm = E.pmap + ci.pmap_off;
    for (int64_t q = tid; q < (int64_t)F * S; q += nt) {
      const int i = (int)(q % F), j = (int)(q / F);
      const double v = -U[q];
      Sg[i + (int64_t)(F + j) * sld] = v;
      Sg[(F + j) + (int64_t)i * sld] = v;
    }
    for (int64_t q = tid; q < (int64_t)S * S; q += nt) {
      const int i = (int)(q % S), j = (int)(q / S);
      Sg[(F + i) + (int64_t)(F + j) * sld] = P[pm[i] + (int64_t)pm[j] * mp.sld];
    }
  }
}

// packs the d x d diagonal blocks the caller asked for: blockIdx.x = query
struct MargQuery { int64_t src; int64_t dst; int32_t sld; int32_t d; };
__global__ void marginal_gather_kernel(const MargQuery* __restrict__ q, const double* __restrict__ sigma, double* __restrict__ out) {
  const MargQuery m = q[blockIdx.x];
  for (int t = threadIdx.x; t < m.d * m.d; t += blockDim.x) {
    const int r = t / m.d, c = t - r * m.d;
    out[m.dst + t] = sigma[m.src + r + (int64_t)c * m.sld];
  }
}

}  // namespace gtb
