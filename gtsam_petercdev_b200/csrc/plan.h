// plan.h -- symbolic elimination plan (host), flattened to index arrays for the device.
//
// Built once per ordering by build_plan() (symbolic.cpp).  The reference rebuilds the same structure on every
// solve (inference/EliminateableFactorGraph-inst.h:131-137); it is invariant across LM iterations and lambda.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace gtb {

struct SymbolicInput {
  int32_t n_vars = 0;                 // variables, ids = rank in ascending key order
  std::vector<int32_t> var_dim;       // tangent dimension per variable
  std::vector<int64_t> var_off;       // scalar offset of each variable in the key-sorted delta vector
  int64_t n_factors = 0;              // factors in graph order
  std::vector<int32_t> fac_nkeys;     // 1 or 2
  std::vector<int32_t> fac_var;       // 2 per factor (second = -1 for unary)
  std::vector<int32_t> order;         // elimination position -> variable id
};

// One factor as seen by the clique that assembles it.
struct FactorSlot {
  int32_t factor;   // graph index
  int32_t col0;     // first front column of key 0's block
  int32_t col1;     // first front column of key 1's block, -1 for unary factors
  int32_t pad;
};

struct Plan {
  int32_t n_cliques = 0;
  // cliques are numbered in DFS post-order (children in stored order, roots in order)
  std::vector<int32_t> parent;          // clique id or -1
  std::vector<int32_t> level;           // 0 = leaves; parent level > every child's level
  std::vector<int32_t> n_frontal_vars;  // number of frontal variables
  std::vector<int32_t> n_vars;          // frontal + separator variables
  std::vector<int64_t> vars_off;        // [n_cliques+1] into vars
  std::vector<int32_t> vars;            // frontal variables in elimination order, then separator ascending
  std::vector<int32_t> F, S;            // scalar dims
  std::vector<int32_t> ld;              // leading dimension of the front (>= F+S+1)
  std::vector<int64_t> front_off;       // offset (doubles) of the front in the arena
  int64_t arena_doubles = 0;
  std::vector<int64_t> child_off;       // [n_cliques+1] into children
  std::vector<int32_t> children;        // child clique ids in stored order
  std::vector<int64_t> fac_off;         // [n_cliques+1] into slots
  std::vector<FactorSlot> slots;        // own factors in gather order
  std::vector<int64_t> pmap_off;        // [n_cliques] into pmap; S+1 entries per clique
  std::vector<int32_t> pmap;            // front column in the PARENT of each separator scalar, then of the rhs
  std::vector<int64_t> gidx_off;        // [n_cliques] into gidx; F+S entries per clique
  std::vector<int32_t> gidx;            // global (key-sorted) scalar index of every front column
  std::vector<int64_t> cond_off;        // [n_cliques+1] offsets of row-major [R S d] in the conditional dump
  // level schedule
  int32_t n_levels = 0;
  std::vector<int64_t> level_off;       // [n_levels+1] into level_cliques
  std::vector<int32_t> level_cliques;   // cliques of each level, small fronts first (ascending id within class)
  std::vector<int32_t> level_n_small;   // number of small-front cliques at each level
  // statistics
  double flops = 0, flops_large = 0, assembly_bytes = 0;
  double flops_syrk_frontal = 0;   // large fronts: trailing updates inside the frontal rows (syrk_kernel mode 0)
  double flops_syrk_schur = 0;     // large fronts: C -= S^T S (syrk_kernel mode 1)
  double bytes_small = 0;          // small fronts: factor rows read + child Schur complements read + front written
  double bytes_backsolve = 0;      // [R S d] read once + delta written
  int32_t max_front = 0;
  int64_t n_small = 0, n_large = 0;
  // Numeric plan only (relaxed amalgamation, symbolic.cpp): the reference cliques every numeric clique holds, in elimination
  // order, and where each one's frontal scalars end inside the merged front; and the inverse map.
  std::vector<int64_t> orig_off;        // [n_cliques+1] into orig_clique / orig_end
  std::vector<int32_t> orig_clique;     // reference clique id (numbering of the reported plan)
  std::vector<int32_t> orig_end;        // frontal scalar index (exclusive) where that clique's pivots end
  std::vector<int32_t> num_of;          // per reference clique: numeric clique id
  std::vector<int32_t> num_foff;        // per reference clique: first frontal scalar inside the numeric front
};

// Front classes.  "Small" fronts are assembled and eliminated entirely in shared memory by one warp or CTA:
//   thin  : few frontal scalars (a point or a pose), packed upper triangle in shared memory
//   dense : any F, full square in shared memory
// everything else is a "large" front that lives in global memory and is factored by the blocked panel kernels.
#ifdef __CUDACC__
#define GTB_HD __host__ __device__
#else
#define GTB_HD
#endif
constexpr int32_t kThinMaxF = 12, kThinMaxDim = 208, kDenseMaxDim = 128;
GTB_HD inline bool is_thin_front(int32_t F, int32_t dim) { return F <= kThinMaxF && dim <= kThinMaxDim; }
GTB_HD inline bool is_small_front(int32_t F, int32_t dim) { return dim <= kDenseMaxDim || is_thin_front(F, dim); }
// A thin front whose parent is a large front adds its Schur complement straight into the parent (atomics) when it is
// eliminated and keeps only its conditional rows: nothing re-reads the child later.
GTB_HD inline bool is_fused_child(int32_t F, int32_t dim, int32_t parent_F, int32_t parent_dim) {
  return is_thin_front(F, dim) && !is_small_front(parent_F, parent_dim);
}

// Throws std::runtime_error with a message on inconsistent input.
// small_max: fronts with F+S+1 <= small_max are handled by the shared-memory kernel.
// out: the reference's clique layout, bit-exact.  numeric (may be null): the plan the device factorises, after relaxed
// amalgamation of chains of large fronts (amalg_ratio <= 1 disables it: numeric = out plus the identity member lists).
// Elimination position -> variable id by automatic nested dissection (ordering.cpp); subsets of at most leaf_size variables are
// ordered by degree.
std::vector<int32_t> nested_dissection_order(const SymbolicInput& in, int leaf_size = 2);

void build_plan(const SymbolicInput& in, int32_t small_max, Plan* out, Plan* numeric = nullptr, double amalg_ratio = 0.0,
                int32_t amalg_max_F = 0, int32_t amalg_min_S = 0);

}  // namespace gtb
