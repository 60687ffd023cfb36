// engine.cu -- the engine behind the C ABI (include/gtsam_b200.h): problem tables, device buffers, the level
// schedule that drives the kernels in factors.cuh / eliminate.cuh, and the LM / GN control flow.
//
// Host control flow mirrors the reference exactly where it decides anything:
//   LevenbergMarquardtOptimizer::iterate / tryLambda   nonlinear/LevenbergMarquardtOptimizer.cpp:121-304
//   LevenbergMarquardtState::increase/decreaseLambda   nonlinear/internal/LevenbergMarquardtState.h:70-94
//   NonlinearOptimizer::defaultOptimize, checkConvergence   nonlinear/NonlinearOptimizer.cpp:62-117, 182-231
//   GaussNewtonOptimizer::iterate                      nonlinear/GaussNewtonOptimizer.cpp:44-66
// There is no CPU fallback: every numeric entry point needs a CUDA device.
#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <mutex>
#include <queue>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/gtsam_b200.h"
#include "eliminate.cuh"
#include "marginals.cuh"
#include "factors.cuh"
#include "plan.h"
#include "layout.h"

namespace gtb {

struct CudaError : std::runtime_error { using std::runtime_error::runtime_error; };
#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t _e = (call);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      throw CudaError(std::string(#call) + ": " + cudaGetErrorString(_e));                         \
  } while (0)

struct StatusError : std::runtime_error {
  gtb_status st;
  StatusError(gtb_status s, const std::string& m) : std::runtime_error(m), st(s) {}
};

// Device and pinned-host blocks of destroyed engines are kept for the next engine instead of going back to the driver: the
// ISAM2 layer (isam2.cpp) builds a fresh engine at every update, and cudaMalloc / cudaFree / cudaMallocHost / cudaFreeHost of
// an engine's ~40 buffers cost several milliseconds, ten times the numeric update of a small problem.  Process-wide, per
// device; a block is reused for a request of at least half its size.  GTB_BLOCK_CACHE_MB bounds what is held (default 256,
// 0 = off: every release goes to cudaFree as before); a single block above a quarter of the bound is never kept.
struct BlockCache {
  static constexpr int kSlots = 65;   // devices 0..63, slot 64 = pinned host memory
  std::mutex mu;
  std::multimap<size_t, void*> blocks[kSlots];
  size_t held = 0;
  static size_t bound() {
    static const size_t b = (size_t)(getenv("GTB_BLOCK_CACHE_MB") ? strtoull(getenv("GTB_BLOCK_CACHE_MB"), nullptr, 10) : 256) << 20;
    return b;
  }
  void* take(int slot, size_t bytes, size_t* cap) {
    if (!bound() || slot < 0 || slot >= kSlots) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    auto it = blocks[slot].lower_bound(bytes);
    if (it == blocks[slot].end() || it->first > 2 * bytes + 4096) return nullptr;
    void* p = it->second;
    *cap = it->first;
    held -= it->first;
    blocks[slot].erase(it);
    return p;
  }
  bool give(int slot, void* p, size_t cap) {
    if (slot < 0 || slot >= kSlots || cap > bound() / 4) return false;
    std::lock_guard<std::mutex> lock(mu);
    if (held + cap > bound()) return false;
    blocks[slot].emplace(cap, p);
    held += cap;
    return true;
  }
};
static BlockCache g_blocks;

static void* pinned_alloc(size_t bytes, size_t* cap) {
  bytes = (bytes + 255) & ~(size_t)255;
  if (bytes <= BlockCache::bound() / 4) {
    size_t c = 512;
    while (c < bytes) c <<= 1;
    bytes = c;
  }
  void* p = g_blocks.take(BlockCache::kSlots - 1, bytes, cap);
  if (p) return p;
  CK(cudaMallocHost(&p, bytes));
  *cap = bytes;
  return p;
}
static void pinned_free(void* p, size_t cap) {
  if (p && !g_blocks.give(BlockCache::kSlots - 1, p, cap)) cudaFreeHost(p);
}

template <class T> struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  size_t cap = 0;   // bytes of the block behind p
  int dev = 0;      // device the block lives on
  // zero = true: the buffer starts at zero whether it is fresh from the driver or reused (reused blocks are dirty)
  void alloc(size_t count, bool zero = true) {
    release();
    n = count;
    if (!count) return;
    size_t bytes = (count * sizeof(T) + 255) & ~(size_t)255;
    if (bytes <= BlockCache::bound() / 4) {   // cacheable: power-of-two size classes, so that a problem that grows by one
      size_t c = 512;                         // variable per update (ISAM2) finds its blocks again
      while (c < bytes) c <<= 1;
      bytes = c;
    }
    CK(cudaGetDevice(&dev));
    p = static_cast<T*>(g_blocks.take(dev, bytes, &cap));
    if (p) {
      if (zero) {
        CK(cudaMemsetAsync(p, 0, bytes, 0));
        CK(cudaStreamSynchronize(0));   // the engine's stream does not wait for the legacy stream
      }
      return;
    }
    CK(cudaMalloc(&p, bytes));
    cap = bytes;
  }
  void upload(const std::vector<T>& h, cudaStream_t s) {
    if (h.size() != n || (!p && n)) alloc(h.size(), false);
    if (n) CK(cudaMemcpyAsync(p, h.data(), n * sizeof(T), cudaMemcpyHostToDevice, s));
  }
  void release() {
    if (p) {
      bool kept = false;
      if (BlockCache::bound() && cap <= BlockCache::bound() / 4) {
        int cur = 0;
        cudaGetDevice(&cur);
        if (cur != dev) cudaSetDevice(dev);
        // what cudaFree guarantees implicitly: nothing in flight still touches the block
        kept = cudaDeviceSynchronize() == cudaSuccess && g_blocks.give(dev, p, cap);
        if (cur != dev) cudaSetDevice(cur);
      }
      if (!kept) cudaFree(p);
    }
    p = nullptr;
    n = cap = 0;
  }
  ~DevBuf() { release(); }
};

struct LevelSched {
  // small fronts split into thread classes; each class is a contiguous range of the level list
  struct Cls { int64_t begin; int32_t count; int32_t threads; int32_t smem; int32_t kind; int32_t group_doubles; };
  std::vector<Cls> small;
  int64_t large_begin = 0;
  int32_t n_large = 0;
  int32_t max_F = 0, max_dim = 0, min_F = 0;
  int64_t work_begin = 0;     // child extend-add work items for large parents
  int32_t n_work = 0;
  int32_t n_all = 0;          // all cliques at the level (back-substitution)
  int64_t bs_begin = 0;       // into the back-substitution list: warp-class cliques first, then CTA-class
  int32_t n_bs_warp = 0, n_bs_cta = 0, bs_cta_smem = 0;
  int32_t n_bs_pt = 0;        // leading part of the warp class: point cliques with <= 8 cameras (four per warp)
  int32_t bs_max_F = 0, bs_max_S = 0;   // CTA class: grid of the partial-GEMV kernel
  int64_t zero_blocks = 0;
  std::vector<int64_t> step_begin;  // per panel step 0..max_np: range of the compact task list (panel-step tasks)
  std::vector<int32_t> step_count;
  std::vector<int32_t> step_spc;    // per step: column slices per role-A CTA of that launch
  std::vector<int64_t> schur_begin; // per step: tile tasks of the cliques whose Schur-complement step this is
  std::vector<int32_t> schur_count;
  // ordered hand-over: the Schur-step tasks of a step are cut into one launch per rank among same-level siblings that hand
  // over to the same parent, so that a parent entry receives at most one contribution per launch
  struct Run { int64_t begin; int32_t count; };
  std::vector<std::vector<Run>> schur_runs;
  int64_t gather_begin = 0;   // ordered extend-add work items of this level's large parents (extend_add_kernel)
  int32_t n_gather = 0;
  int64_t red_begin = 0;      // camera-pair groups cut into several segments: ordered sums (pair_reduce_kernel)
  int32_t n_red = 0;
  // Schur complement of this level's point cliques, gathered per camera pair (point_pair_kernel / point_diag_kernel)
  int64_t pair_seg_begin = 0, diag_seg_begin = 0;
  int32_t n_pair_seg = 0, n_diag_seg = 0;
};

}  // namespace gtb

using namespace gtb;

struct gtb_engine {
  int device = -1;
  bool has_device = false;
  cudaStream_t stream = nullptr;
  mutable std::string err;
  int64_t launches = 0;

  // ---- host problem ----
  struct VarsHost { int64_t n = 0; std::vector<uint64_t> keys; std::vector<double> state; } vars[4];
  struct FacHost {
    int64_t n = 0;
    std::vector<uint64_t> keys;
    std::vector<double> z;       // device layout
    std::vector<int8_t> nk;      // resolved kinds
    std::vector<double> noise;   // n x m*m
    bool any_noise = false;
    int32_t loss = 0;
    double loss_k = 0;
  } fac[7];
  std::vector<int64_t> graph_order;
  bool finalized = false;

  // variable table, ascending key
  std::vector<uint64_t> vkeys;
  std::vector<int8_t> vtype;
  std::vector<int32_t> vrow, vdim;
  std::vector<int64_t> voff;
  int64_t n_scalars = 0;
  std::vector<int32_t> type_var[4];   // row -> variable id

  SymbolicInput sym;
  std::vector<int32_t> g_bucket, g_index;   // graph position -> (bucket, index)
  Plan plan;    // the NUMERIC plan: what the device factorises (the reference's tree after relaxed amalgamation of chains of large fronts)
  Plan rplan;   // the REPORTED plan: the reference's clique layout, bit-exact (gtb_plan_get, conditionals, failure keys)
  bool has_plan = false;
  std::vector<LevelSched> sched[2];   // phase 0: this rank's subtrees, phase 1: the shared top of the tree (rank 0)
  std::vector<int32_t> sched_lists;   // clique lists of both phases, level by level
  std::vector<ChildWork> child_work;
  // ordered hand-over (default; GTB_HANDOVER_ATOMIC=1 restores the unordered atomics of round 1)
  bool ordered = true;
  int32_t epoch = 0;                       // factorisation counter (K-split hand-shake flags)
  std::vector<GatherWork> gather_work;     // per level, then the top work of this rank
  std::vector<int32_t> gather_child;       // child lists of the work items
  std::vector<int64_t> gather_smap_off;    // per list entry: offset of the child's sorted separator map in smap
  std::vector<int32_t> smap;               // per gathered child: parent columns ascending (S + 1), then the child separator index of each
  std::vector<int64_t> smap_of;            // per clique: offset in smap or -1
  int64_t top_gather_begin = 0;
  int32_t n_top_gather = 0;
  std::vector<PairReduce> pair_reduce;
  int64_t pair_slots = 0, tile_flags = 0;
  std::vector<int32_t> flag_off;           // per clique: first K-split flag or 0
  // ---- distribution over ranks: one process per GPU ----
  int rank = 0, world = 1;
  gtb_comm_fn comm = nullptr;
  void* comm_user = nullptr;
  std::vector<int32_t> owner;         // per clique: owning rank, or -1 for the top of the tree
  std::vector<int64_t> front_off;     // per clique: offset in this engine's arena (top fronts live in one tail region)
  std::vector<int32_t> ld;            // per clique: leading dimension of its front storage (point fronts keep 3 rows: ld = 4)
  std::vector<int32_t> top_list;
  std::vector<uint8_t> is_point;      // per clique: handled by point_front_kernel (star structure, large parent)
  std::vector<uint8_t> fuse_up;       // per clique: large front whose Schur step adds its Schur complement into the parent itself
  std::vector<int32_t> large_list;    // every large front this rank touches: its own (phase 0) first, then the top
  int32_t n_large_own = 0;
  int64_t top_off = 0, top_doubles = 0, arena_doubles = 0, top_work_begin = 0, top_zero_blocks = 0;
  int32_t n_top_work = 0;
  std::vector<int32_t> bs_list, bs_scratch_off;
  int64_t bs_scratch_doubles = 0;
  // inverse diagonal blocks of the CTA-class cliques (diag_inverse_kernel): first block per list slot, (clique, block,
  // destination) tasks per phase
  std::vector<int32_t> bs_rinv_off, rinv_tasks;
  int64_t rinv_blocks = 0;
  int64_t rinv_task_begin[2] = {0, 0}, rinv_task_count[2] = {0, 0};
  std::vector<StepTask> step_tasks_list;
  bool point_pairs = true;            // false (GTB_POINT_ATOMIC=1): point cliques scatter their Schur complement with atomics
  std::vector<PairSeg> pair_segs;     // segments of both kinds, both phases
  std::vector<uint2> pair_items;
  std::vector<uint32_t> diag_items;   // 3 words per item

  // ---- device ----
  DevBuf<double> d_state[2][4];
  int cur = 0;
  DevBuf<int32_t> d_type_goff[4];
  DevBuf<int32_t> d_row[7][2], d_goff[7][2];
  DevBuf<double> d_z[7], d_noise[7], d_Ab[7];
  DevBuf<int8_t> d_nk[7];
  BucketDev bk[7];
  DevBuf<double> d_partials, d_scalars, d_delta, d_hdiag, d_damp, d_arena, d_dfac;
  int64_t partial_off[7] = {0};
  int64_t n_partials = 0;
  double* h_scalars = nullptr;   // pinned
  size_t h_scalars_cap = 0, h_stage_cap[4] = {0, 0, 0, 0};   // block sizes (BlockCache)
  double* h_stage[4] = {nullptr, nullptr, nullptr, nullptr};   // pinned staging of the per-step state refresh
  cudaEvent_t stage_ev[4] = {nullptr, nullptr, nullptr, nullptr};
  DevBuf<int64_t> d_inc_off;
  DevBuf<int32_t> d_inc_bucket, d_inc_index, d_inc_slot, d_var_goff, d_var_dim;
  DevBuf<CliqueDev> d_cliques;
  DevBuf<SlotDev> d_slots;
  DevBuf<int32_t> d_children, d_pmap, d_gidx, d_level_cliques, d_fail, d_bs_list, d_bs_scratch_off, d_bs_rinv_off, d_rinv_tasks;
  DevBuf<double> d_bs_scratch, d_bs_rinv, d_delta_own;
  DevBuf<int32_t> d_top_list, d_large_list;
  DevBuf<uint8_t> d_own_mask;
  DevBuf<ChildWork> d_child_work;
  DevBuf<GatherWork> d_gather_work;
  DevBuf<int32_t> d_gather_child, d_smap, d_tile_flags;
  DevBuf<int64_t> d_gather_smap_off;
  DevBuf<PairReduce> d_pair_reduce;
  DevBuf<int32_t> d_orig_off, d_orig_end;
  DevBuf<double> d_pair_scratch;
  DevBuf<StepTask> d_step_tasks;
  DevBuf<PairSeg> d_pair_segs;
  DevBuf<uint2> d_pair_items;
  DevBuf<uint32_t> d_diag_items;
  DevBuf<long long> d_dbg;
  // marginal covariances (marginals.cuh): built on first use for the installed plan
  std::vector<MargClique> marg_cliques;
  std::vector<int32_t> marg_lists;            // per level: cliques with F + S <= 256 first, then the others
  std::vector<int64_t> marg_level_off;        // [n_levels + 1]
  std::vector<int32_t> marg_level_small;      // [n_levels]
  std::vector<int32_t> var_clique, var_foff;  // per variable: the clique it is a frontal of, its scalar offset there
  DevBuf<MargClique> d_marg_cliques;
  DevBuf<int32_t> d_marg_lists;
  DevBuf<double> d_sigma, d_marg_scratch;
  bool marg_built = false, marg_valid = false;
  bool linearized = false, factorized = false, solved = false, have_hdiag = false;
  bool retracted = false;          // the scratch copy holds Values::retract(delta) of the current values
  bool lin_attr[7] = {false, false, false, false, false, false, false}, elim_attr = false;   // cudaFuncSetAttribute done on this engine's device

  bool profiling = false;
  std::vector<cudaEvent_t> prof_events;
  std::vector<int> prof_class;
  int64_t prof_used = 0;
  int64_t class_launches[32] = {0};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  double stage_ms[8] = {0};
  bool time_stages = true;

  // LM state across gtb_lm_iterate calls
  bool lm_have_error = false;
  double lm_error = 0;
  int lm_last_accepted = 0;   // the last gtb_lm_iterate ended with an accepted step

  ~gtb_engine() {
    pinned_free(h_scalars, h_scalars_cap);
    for (int t = 0; t < 4; ++t) {
      pinned_free(h_stage[t], h_stage_cap[t]);
      if (stage_ev[t]) cudaEventDestroy(stage_ev[t]);
    }
    for (auto ev_ : prof_events) cudaEventDestroy(ev_);
    if (ev[0]) cudaEventDestroy(ev[0]);
    if (ev[1]) cudaEventDestroy(ev[1]);
    if (stream) cudaStreamDestroy(stream);
  }
};

namespace gtb {

typedef gtb_engine Engine;

static void need_device(Engine* e) {
  if (!e->has_device) throw StatusError(GTB_ERR_NO_DEVICE, "this engine was created without a CUDA device; there is no CPU fallback");
}
static void need(bool cond, const char* msg) {
  if (!cond) throw StatusError(GTB_ERR_INVALID, msg);
}

// ---- noise model resolution: the reference's smart factories (linear/NoiseModel.cpp:84-132,284-315,625-634) ----
static bool is_diagonal(const double* M, int d) {
  for (int i = 0; i < d; ++i)
    for (int j = i + 1; j < d; ++j)
      if (std::abs(M[i * d + j]) > 1e-9) return false;
  return true;
}
static int iso_sigma(double sigma, double* out) {
  if (std::abs(sigma - 1.0) < 1e-9) return NK_UNIT;
  out[0] = 1.0 / sigma;
  return NK_ISO;
}
static int iso_variance(double var, double* out) {
  if (std::abs(var - 1.0) < 1e-9) return NK_UNIT;
  out[0] = 1.0 / std::sqrt(var);
  return NK_ISO;
}
static int diag_variances(const double* v, int d, double* out) {
  bool same = true;
  for (int i = 1; i < d; ++i) same = same && (v[i] == v[0]);
  if (same) return iso_variance(v[0], out);
  for (int i = 0; i < d; ++i) out[i] = 1.0 / std::sqrt(v[i]);
  return NK_DIAG;
}
static int diag_sigmas(const double* s, int d, double* out) {
  for (int i = 0; i < d; ++i)
    if (s[i] < 1e-8)
      throw StatusError(GTB_ERR_NOISE_MODEL, "constrained noise model (sigma < 1e-8) needs the QR path, which is not available on device");
  bool same = true;
  for (int i = 1; i < d; ++i) same = same && (s[i] == s[0]);
  if (same) return iso_sigma(s[0], out);
  for (int i = 0; i < d; ++i) out[i] = 1.0 / s[i];
  return NK_DIAG;
}
static int resolve_noise(int kind, const double* data, int d, double* out) {
  for (int i = 0; i < d * d; ++i) out[i] = 0.0;
  double tmp[16];
  switch (kind) {
    case GTB_NOISE_UNIT: return NK_UNIT;
    case GTB_NOISE_ISO_SIGMA: return iso_sigma(data[0], out);
    case GTB_NOISE_DIAG_SIGMAS: return diag_sigmas(data, d, out);
    case GTB_NOISE_DIAG_VARIANCES: return diag_variances(data, d, out);
    case GTB_NOISE_INFORMATION: {
      if (is_diagonal(data, d)) {
        for (int i = 0; i < d; ++i) tmp[i] = 1.0 / data[i * d + i];
        return diag_variances(tmp, d, out);
      }
      // R = chol_upper(information)  (Eigen::LLT(...).matrixU(), linear/NoiseModel.cpp:108-110)
      std::vector<double> L(d * d, 0.0);
      for (int j = 0; j < d; ++j) {
        double s = data[j * d + j];
        for (int k = 0; k < j; ++k) s -= L[j * d + k] * L[j * d + k];
        if (!(s > 0)) throw StatusError(GTB_ERR_NOISE_MODEL, "information matrix is not positive definite");
        const double ljj = std::sqrt(s);
        L[j * d + j] = ljj;
        for (int i = j + 1; i < d; ++i) {
          double t = data[i * d + j];
          for (int k = 0; k < j; ++k) t -= L[i * d + k] * L[j * d + k];
          L[i * d + j] = t / ljj;
        }
      }
      for (int i = 0; i < d; ++i)
        for (int j = 0; j < d; ++j) out[i * d + j] = L[j * d + i];
      return NK_FULL;
    }
    case GTB_NOISE_SQRT_INFORMATION: {
      if (is_diagonal(data, d)) {
        for (int i = 0; i < d; ++i) tmp[i] = 1.0 / data[i * d + i];
        return diag_sigmas(tmp, d, out);
      }
      for (int i = 0; i < d * d; ++i) out[i] = data[i];
      return NK_FULL;
    }
  }
  throw StatusError(GTB_ERR_NOISE_MODEL, "unknown noise kind");
}

static void to_device_state(int vtype, const double* in, double* out) {
  if (vtype == GTB_POSE2) {  // Pose2(x, y, theta): Rot2::fromAngle (geometry/Pose2.h:74-76)
    out[0] = std::cos(in[2]);
    out[1] = std::sin(in[2]);
    out[2] = in[0];
    out[3] = in[1];
  } else {
    std::memcpy(out, in, sizeof(double) * kStateIn[vtype]);
  }
}

// ---- timing helpers ----
struct StageTimer {
  Engine* e;
  int slot;
  StageTimer(Engine* e_, int s) : e(e_), slot(s) {
    if (e->time_stages) cudaEventRecord(e->ev[0], e->stream);
  }
  void stop() {
    if (!e->time_stages) return;
    cudaEventRecord(e->ev[1], e->stream);
    cudaEventSynchronize(e->ev[1]);
    float ms = 0;
    cudaEventElapsedTime(&ms, e->ev[0], e->ev[1]);
    e->stage_ms[slot] = ms;
  }
};

// kernel classes for launch counting and optional per-launch event timing
enum { K_LINEARIZE = 0, K_ERROR, K_LINERR, K_REDUCE, K_HDIAG, K_DAMP, K_SMALL, K_THIN, K_POINT, K_PAIR, K_ZERO, K_LFACT, K_LCHILD, K_STEP, K_SCHUR,
       K_BACKSOLVE, K_BSGEMV, K_RETRACT, K_OWNDELTA, K_MARGINAL, K_COMM, K_NUM };
static const char* kKernelNames[K_NUM] = {"linearize", "error", "linear_error", "sum_partials", "hessian_diag", "damping",
                                          "small_front", "thin_front", "point_front", "point_schur", "large_zero", "large_factors", "large_children", "front_step", "schur_step",
                                          "backsolve", "backsolve_gemv", "retract", "own_delta", "marginals", "collectives"};

static void prof_begin(Engine* e, int k) {
  if (!e->profiling) return;
  if (e->prof_used + 2 > (int64_t)e->prof_events.size()) {
    const size_t old = e->prof_events.size();
    e->prof_events.resize(old + 8192);
    for (size_t i = old; i < e->prof_events.size(); ++i) cudaEventCreate(&e->prof_events[i]);
  }
  e->prof_class.push_back(k);
  cudaEventRecord(e->prof_events[e->prof_used++], e->stream);
}
static void check_launch(Engine* e, int k) {
  e->launches++;
  e->class_launches[k]++;
  if (e->profiling) cudaEventRecord(e->prof_events[e->prof_used++], e->stream);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) throw CudaError(std::string("kernel launch (") + kKernelNames[k] + "): " + cudaGetErrorString(err));
}
#define LAUNCH(e, k, ...)   \
  do {                      \
    prof_begin(e, k);       \
    __VA_ARGS__;            \
    check_launch(e, k);     \
  } while (0)

// Launch with programmatic stream serialization: the kernel may be scheduled before its predecessor on the stream has
// finished; it calls pdl_wait() before touching anything the predecessor writes (eliminate.cuh).
template <class... KArgs, class... Args>
static void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...));
}

// ------------------------------------------------------------------------------------------------------
// problem set-up
// ------------------------------------------------------------------------------------------------------
static void set_variables(Engine* e, int vt, int64_t n, const uint64_t* keys, const double* state) {
  need(vt >= 0 && vt < 4, "bad variable type");
  need(n >= 0 && (n == 0 || state), "null state");
  auto& v = e->vars[vt];
  if (e->finalized) {
    // state refresh (the per-step host -> device copy)
    need(n == v.n, "variable count changed after gtb_finalize_problem");
    if (e->has_device && n) {
      // A page-locked caller buffer whose rows need no conversion is copied from where it lies (one DMA, no staging pass over
      // 4 MB on the host per C3 step).  The caller keeps it unchanged until the next call that synchronises (header).
      if (kStateIn[vt] == kStateDev[vt] && vt != 0) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, state) == cudaSuccess && at.type == cudaMemoryTypeHost) {
          CK(cudaMemcpyAsync(e->d_state[e->cur][vt].p, state, v.state.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
          e->linearized = e->factorized = e->solved = e->retracted = false;
          e->lm_have_error = false;
          return;
        }
        cudaGetLastError();   // (an unregistered pointer is not an error here)
      }
      // otherwise: convert straight into a pinned staging buffer so that the copy is a real asynchronous DMA
      if (!e->h_stage[vt]) {
        e->h_stage[vt] = static_cast<double*>(pinned_alloc(v.state.size() * sizeof(double), &e->h_stage_cap[vt]));
        CK(cudaEventCreateWithFlags(&e->stage_ev[vt], cudaEventDisableTiming));
      } else {
        CK(cudaEventSynchronize(e->stage_ev[vt]));   // the previous refresh has left the buffer
      }
      double* hs = e->h_stage[vt];
      if (kStateIn[vt] == kStateDev[vt] && vt != 0) std::memcpy(hs, state, v.state.size() * sizeof(double));
      else for (int64_t i = 0; i < n; ++i) to_device_state(vt, state + i * kStateIn[vt], hs + i * kStateDev[vt]);
      CK(cudaMemcpyAsync(e->d_state[e->cur][vt].p, hs, v.state.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
      CK(cudaEventRecord(e->stage_ev[vt], e->stream));
      e->linearized = e->factorized = e->solved = e->retracted = false;
      e->lm_have_error = false;
    }
    return;
  }
  need(n == 0 || keys, "null keys");
  v.n = n;
  v.keys.assign(keys, keys + n);
  v.state.resize(n * kStateDev[vt]);
  for (int64_t i = 0; i < n; ++i) to_device_state(vt, state + i * kStateIn[vt], &v.state[i * kStateDev[vt]]);
}

static void set_factors(Engine* e, int b, int64_t n, const uint64_t* keys, const double* z, const int32_t* nk,
                        const double* noise) {
  need(b >= 0 && b < 7, "bad bucket");
  need(!e->finalized, "gtb_set_factors after gtb_finalize_problem");
  need(n >= 0 && (n == 0 || (keys && z)), "null factor arrays");
  auto& f = e->fac[b];
  const int m = kM[b];
  f.n = n;
  f.keys.assign(keys, keys + n * kNKeys[b]);
  f.z.resize(n * kZDev[b]);
  f.nk.resize(n);
  f.noise.assign(n * m * m, 0.0);
  f.any_noise = false;
  for (int64_t i = 0; i < n; ++i) {
    const double* zi = z + i * kZIn[b];
    double* zo = &f.z[i * kZDev[b]];
    if (b == GTB_BETWEEN2 || b == GTB_PRIOR2) to_device_state(GTB_POSE2, zi, zo);
    else std::memcpy(zo, zi, sizeof(double) * kZIn[b]);
    const int kind = nk ? nk[i] : GTB_NOISE_UNIT;
    need(kind == GTB_NOISE_UNIT || noise, "noise data required for non-unit noise kinds");
    f.nk[i] = (int8_t)resolve_noise(kind, noise ? noise + i * m * m : nullptr, m, &f.noise[i * m * m]);
    if (f.nk[i] != NK_UNIT) f.any_noise = true;
  }
}

static void finalize_problem(Engine* e) {
  need(!e->finalized, "problem already finalized");
  // variable table in ascending key order (the iteration order of gtsam::Values, a std::map)
  int64_t V = 0;
  for (int t = 0; t < 4; ++t) V += e->vars[t].n;
  need(V > 0, "no variables");
  need(V < INT32_MAX, "too many variables");
  std::vector<std::pair<uint64_t, std::pair<int8_t, int32_t>>> all;
  all.reserve(V);
  for (int t = 0; t < 4; ++t)
    for (int64_t i = 0; i < e->vars[t].n; ++i) all.push_back({e->vars[t].keys[i], {(int8_t)t, (int32_t)i}});
  std::sort(all.begin(), all.end());
  e->vkeys.resize(V); e->vtype.resize(V); e->vrow.resize(V); e->vdim.resize(V); e->voff.resize(V);
  for (int t = 0; t < 4; ++t) e->type_var[t].assign(e->vars[t].n, -1);
  int64_t off = 0;
  for (int64_t i = 0; i < V; ++i) {
    if (i && all[i].first == all[i - 1].first) throw StatusError(GTB_ERR_INVALID, "duplicate key in Values (ValuesKeyAlreadyExists)");
    e->vkeys[i] = all[i].first;
    e->vtype[i] = all[i].second.first;
    e->vrow[i] = all[i].second.second;
    e->vdim[i] = kDim[e->vtype[i]];
    e->voff[i] = off;
    off += e->vdim[i];
    e->type_var[e->vtype[i]][e->vrow[i]] = (int32_t)i;
  }
  need(off < INT32_MAX, "too many scalars");
  e->n_scalars = off;

  // graph order
  int64_t Mtot = 0;
  for (int b = 0; b < 7; ++b) Mtot += e->fac[b].n;
  need(Mtot < INT32_MAX, "too many factors");
  if (e->graph_order.empty()) {
    for (int b = 0; b < 7; ++b)
      for (int64_t i = 0; i < e->fac[b].n; ++i) { e->graph_order.push_back(b); e->graph_order.push_back(i); }
  }
  need((int64_t)e->graph_order.size() == 2 * Mtot, "graph order must list every factor exactly once");
  e->g_bucket.resize(Mtot);
  e->g_index.resize(Mtot);
  {
    std::vector<std::vector<uint8_t>> seen(7);
    for (int b = 0; b < 7; ++b) seen[b].assign(e->fac[b].n, 0);
    for (int64_t g = 0; g < Mtot; ++g) {
      const int64_t b = e->graph_order[2 * g], i = e->graph_order[2 * g + 1];
      need(b >= 0 && b < 7 && i >= 0 && i < e->fac[b].n && !seen[b][i], "bad graph order entry");
      seen[b][i] = 1;
      e->g_bucket[g] = (int32_t)b;
      e->g_index[g] = (int32_t)i;
    }
  }

  // key -> variable id (binary search), typed
  auto find_var = [&](uint64_t key, int want_type) -> int32_t {
    auto it = std::lower_bound(e->vkeys.begin(), e->vkeys.end(), key);
    if (it == e->vkeys.end() || *it != key) {
      char buf[96];
      snprintf(buf, sizeof buf, "factor key %llu is not in Values (ValuesKeyDoesNotExist)", (unsigned long long)key);
      throw StatusError(GTB_ERR_KEY_MISSING, buf);
    }
    const int32_t v = (int32_t)(it - e->vkeys.begin());
    if (e->vtype[v] != want_type) throw StatusError(GTB_ERR_INVALID, "factor key has the wrong variable type (ValuesIncorrectType)");
    return v;
  };
  std::vector<int32_t> fvar[7];
  for (int b = 0; b < 7; ++b) {
    const auto& f = e->fac[b];
    fvar[b].resize(f.n * 2, -1);
    for (int64_t i = 0; i < f.n; ++i)
      for (int k = 0; k < kNKeys[b]; ++k) fvar[b][2 * i + k] = find_var(f.keys[i * kNKeys[b] + k], kVT[b][k]);
  }
  SymbolicInput& s = e->sym;
  s.n_vars = (int32_t)V;
  s.var_dim = e->vdim;
  s.var_off = e->voff;
  s.n_factors = Mtot;
  s.fac_nkeys.resize(Mtot);
  s.fac_var.resize(2 * Mtot);
  for (int64_t g = 0; g < Mtot; ++g) {
    const int b = e->g_bucket[g];
    const int64_t i = e->g_index[g];
    s.fac_nkeys[g] = kNKeys[b];
    s.fac_var[2 * g] = fvar[b][2 * i];
    s.fac_var[2 * g + 1] = fvar[b][2 * i + 1];
    if (kNKeys[b] == 2 && s.fac_var[2 * g] == s.fac_var[2 * g + 1])
      throw StatusError(GTB_ERR_INVALID, "binary factor with the same key twice");
  }
  e->finalized = true;
  if (!e->has_device) return;

  // ---- device upload ----
  cudaStream_t st = e->stream;
  for (int t = 0; t < 4; ++t) {
    if (!e->vars[t].n) continue;
    e->d_state[0][t].upload(e->vars[t].state, st);
    e->d_state[1][t].alloc(e->vars[t].state.size());
    std::vector<int32_t> goff(e->vars[t].n);
    for (int64_t i = 0; i < e->vars[t].n; ++i) goff[i] = (int32_t)e->voff[e->type_var[t][i]];
    e->d_type_goff[t].upload(goff, st);
  }
  e->n_partials = 0;
  for (int b = 0; b < 7; ++b) {
    auto& f = e->fac[b];
    BucketDev& bk = e->bk[b];
    std::memset(&bk, 0, sizeof bk);
    e->partial_off[b] = e->n_partials;
    if (!f.n) continue;
    e->n_partials += (f.n + kFactorThreads - 1) / kFactorThreads;
    for (int k = 0; k < kNKeys[b]; ++k) {
      std::vector<int32_t> row(f.n), goff(f.n);
      for (int64_t i = 0; i < f.n; ++i) {
        const int32_t v = fvar[b][2 * i + k];
        row[i] = e->vrow[v];
        goff[i] = (int32_t)e->voff[v];
      }
      e->d_row[b][k].upload(row, st);
      e->d_goff[b][k].upload(goff, st);
    }
    e->d_z[b].upload(f.z, st);
    e->d_nk[b].upload(f.nk, st);
    if (f.any_noise) e->d_noise[b].upload(f.noise, st);
    e->d_Ab[b].alloc((size_t)f.n * kM[b] * kW[b]);
    bk.n = (int32_t)f.n;
    bk.any_noise = f.any_noise ? 1 : 0;
    bk.loss = f.loss; bk.pad = 0; bk.loss_k = f.loss_k;
    bk.row0 = e->d_row[b][0].p; bk.row1 = e->d_row[b][1].p;
    bk.goff0 = e->d_goff[b][0].p; bk.goff1 = e->d_goff[b][1].p;
    bk.z = e->d_z[b].p; bk.nk = e->d_nk[b].p; bk.noise = e->d_noise[b].p; bk.Ab = e->d_Ab[b].p;
  }
  e->d_partials.alloc(2 * std::max<int64_t>(e->n_partials, 1));
  e->d_scalars.alloc(16);
  e->h_scalars = static_cast<double*>(pinned_alloc(16 * sizeof(double), &e->h_scalars_cap));
  e->d_delta.alloc(e->n_scalars);
  e->d_hdiag.alloc(e->n_scalars);
  e->d_damp.alloc(e->n_scalars);
  CK(cudaMemsetAsync(e->d_delta.p, 0, e->n_scalars * sizeof(double), st));

  // variable -> factor incidence in graph order (Hessian diagonal gather)
  {
    std::vector<int64_t> ioff(V + 1, 0);
    for (int64_t g = 0; g < Mtot; ++g)
      for (int k = 0; k < s.fac_nkeys[g]; ++k) ioff[s.fac_var[2 * g + k] + 1]++;
    for (int64_t v = 0; v < V; ++v) ioff[v + 1] += ioff[v];
    std::vector<int32_t> ib(ioff[V]), ii(ioff[V]), is(ioff[V]);
    std::vector<int64_t> fill(ioff.begin(), ioff.end() - 1);
    for (int64_t g = 0; g < Mtot; ++g)
      for (int k = 0; k < s.fac_nkeys[g]; ++k) {
        const int64_t q = fill[s.fac_var[2 * g + k]]++;
        ib[q] = e->g_bucket[g]; ii[q] = e->g_index[g]; is[q] = k;
      }
    e->d_inc_off.upload(ioff, st);
    e->d_inc_bucket.upload(ib, st);
    e->d_inc_index.upload(ii, st);
    e->d_inc_slot.upload(is, st);
    std::vector<int32_t> vg(V);
    for (int64_t v = 0; v < V; ++v) vg[v] = (int32_t)e->voff[v];
    e->d_var_goff.upload(vg, st);
    e->d_var_dim.upload(e->vdim, st);
  }
  CK(cudaStreamSynchronize(st));
}

// ------------------------------------------------------------------------------------------------------
// symbolic plan -> device
// ------------------------------------------------------------------------------------------------------
// Ordered extend-add of the given children (all of one large parent, child order) by the parent: the child's separator
// sorted by parent column, and work items that own column ranges of the parent of roughly equal triangle area.
static void push_gather(Engine* e, int32_t par, const std::vector<int32_t>& kids) {
  if (kids.empty()) return;
  const Plan& pp = e->plan;
  const int64_t lb = (int64_t)e->gather_child.size();
  int64_t entries = 0;
  for (int32_t ch : kids) {
    const int32_t S1 = pp.S[ch] + 1;
    if (e->smap_of[ch] < 0) {
      e->smap_of[ch] = (int64_t)e->smap.size();
      const int32_t* pm = pp.pmap.data() + pp.pmap_off[ch];
      std::vector<std::pair<int32_t, int32_t>> byp(S1);
      for (int32_t i = 0; i < S1; ++i) byp[i] = {pm[i], i};
      std::sort(byp.begin(), byp.end());
      for (int32_t i = 0; i < S1; ++i) e->smap.push_back(byp[i].first);
      for (int32_t i = 0; i < S1; ++i) e->smap.push_back(byp[i].second);
    }
    e->gather_child.push_back(ch);
    e->gather_smap_off.push_back(e->smap_of[ch]);
    entries += (int64_t)S1 * (S1 + 1) / 2;
  }
  const int64_t le = (int64_t)e->gather_child.size();
  const int32_t dimP = pp.F[par] + pp.S[par] + 1;
  const int32_t nchunk = (int32_t)std::max<int64_t>(1, std::min<int64_t>(96, entries / 4096));
  int32_t c0 = 0;
  for (int32_t k2 = 1; k2 <= nchunk; ++k2) {
    int32_t c1 = (k2 == nchunk) ? dimP : (int32_t)std::lround(dimP * std::sqrt((double)k2 / nchunk));
    c1 = std::max(c1, c0);
    if (c1 > c0) e->gather_work.push_back(GatherWork{par, c0, c1, (int32_t)lb, (int32_t)le, 0});
    c0 = c1;
  }
}

// extend-add work items of one child: column chunks of roughly equal triangle area
static void push_child_work(Engine* e, int32_t ch) {
  const Plan& pp = e->plan;
  const int32_t par = pp.parent[ch];
  if (is_fused_child(pp.F[ch], pp.F[ch] + pp.S[ch] + 1, pp.F[par], pp.F[par] + pp.S[par] + 1) || e->is_point[ch]) return;   // added by the child itself
  if (e->fuse_up[ch]) return;                                                                             // ... by its Schur step
  const int32_t S1 = e->plan.S[ch] + 1;
  const int32_t nchunk = std::max(1, std::min(64, (int32_t)(((int64_t)S1 * S1 / 2) / 8192)));
  int32_t c0 = 0;
  for (int32_t k2 = 1; k2 <= nchunk; ++k2) {
    int32_t c1 = (k2 == nchunk) ? S1 : (int32_t)std::lround(S1 * std::sqrt((double)k2 / nchunk));
    c1 = std::max(c1, c0);
    if (c1 > c0) e->child_work.push_back(ChildWork{ch, c0, c1, 0});
    c0 = c1;
  }
}

// Cost model of a front for the partitioner: elimination flops plus a fixed per-front term (launch / latency share).
static double front_cost(const Plan& p, int32_t c) {
  const double F = p.F[c], N1 = p.F[c] + p.S[c] + 1;
  return F * N1 * N1 + 2e4;
}

// Modelled time of one factorisation under an ownership map: every rank walks its levels (a level costs the longest
// chain of dependent panel steps in it or its flops at the sustained tile rate, whichever is larger, plus its small
// fronts), then the top fronts are reduced and rank 0 walks the top.  Constants are the measured C3 figures
// (DESIGN.md section 4); only the ranking of candidate cuts matters.
static double modelled_time(const Plan& p, const std::vector<int32_t>& owner, int world) {
  // r02 calibration (profiles/r02): a panel step 23 us, Schur tiles 12 TFLOP/s sustained, 4 ns per point clique (the whole
  // point path of C3 is 0.63 ms for 156 k points), fan-in reduce ~300 GB/s (119 MB in 0.4 ms at 8 ranks)
  const double tau = 23e-6, rate = 12e12, t_small = 4e-9, bw = 3e11;
  const int L = p.n_levels, G = world + 1;
  std::vector<double> flops((size_t)G * L, 0.0), steps((size_t)G * L, 0.0), smalls((size_t)G * L, 0.0);
  double reduce_bytes = 0;
  for (int32_t c = 0; c < p.n_cliques; ++c) {
    const int g = owner[c] < 0 ? world : owner[c];
    const size_t k = (size_t)g * L + p.level[c];
    const int dim = p.F[c] + p.S[c] + 1;
    if (is_small_front(p.F[c], dim)) {
      smalls[k] += 1;
    } else {
      flops[k] += (double)p.F[c] * dim * dim;
      steps[k] = std::max(steps[k], (double)((p.F[c] + PW - 1) / PW + 1));
    }
    if (owner[c] < 0) reduce_bytes += 8.0 * p.ld[c] * dim;
  }
  double worst = 0, top = 0;
  for (int g = 0; g < G; ++g) {
    double t = 0;
    for (int l = 0; l < L; ++l) {
      const size_t k = (size_t)g * L + l;
      t += std::max(steps[k] * tau, flops[k] / rate) + smalls[k] * t_small;
    }
    if (g < world) worst = std::max(worst, t);
    else top = t;
  }
  return worst + top + (world > 1 ? reduce_bytes / bw : 0.0);
}

// Ownership of cliques for `world` ranks: the top of the tree (owner -1) is assembled by fan-in from all
// ranks and factored on rank 0; every subtree hanging below it belongs to exactly one rank.  Candidate tops grow from
// the roots by opening the heaviest remaining subtree (large fronts only); for each candidate the subtrees are assigned
// heaviest-first to the least loaded rank and the cut with the smallest modelled time is kept (possibly the empty top:
// a chain-like tree stays on one rank).  Deterministic: every rank computes the same map.
static void compute_partition(const Plan& p, int world, std::vector<int32_t>& owner) {
  const int32_t C = p.n_cliques;
  owner.assign(C, 0);
  if (world <= 1) return;
  std::vector<double> sub(C);
  for (int32_t c = 0; c < C; ++c) sub[c] = front_cost(p, c);
  for (int32_t c = 0; c < C; ++c)   // post-order: children before parents
    if (p.parent[c] >= 0) sub[p.parent[c]] += sub[c];
  typedef std::pair<double, int32_t> Item;
  // the opening sequence
  std::vector<int32_t> seq;
  {
    std::priority_queue<Item> open;
    for (int32_t c = 0; c < C; ++c)
      if (p.parent[c] < 0) open.push(Item(sub[c], c));
    while (!open.empty() && seq.size() < 48) {
      const int32_t c = open.top().second;
      open.pop();
      const bool large = !is_small_front(p.F[c], p.F[c] + p.S[c] + 1);
      if (!large || p.child_off[c + 1] == p.child_off[c]) continue;
      seq.push_back(c);
      for (int64_t k = p.child_off[c]; k < p.child_off[c + 1]; ++k) open.push(Item(sub[p.children[k]], p.children[k]));
    }
  }
  auto assign = [&](size_t n_open, std::vector<int32_t>& own) {
    own.assign(C, -2);
    for (size_t k = 0; k < n_open; ++k) own[seq[k]] = -1;
    std::vector<Item> roots;
    for (int32_t c = 0; c < C; ++c)
      if (own[c] != -1 && (p.parent[c] < 0 || own[p.parent[c]] == -1)) roots.push_back(Item(sub[c], c));
    std::sort(roots.begin(), roots.end(), [](const Item& a, const Item& b) {
      return a.first != b.first ? a.first > b.first : a.second < b.second;
    });
    typedef std::pair<double, int> Load;   // (load, rank): least loaded first, ties to the lower rank
    std::priority_queue<Load, std::vector<Load>, std::greater<Load>> ranks;
    for (int r = 0; r < world; ++r) ranks.push(Load(0.0, r));
    for (const Item& it : roots) {
      Load l = ranks.top();
      ranks.pop();
      own[it.second] = l.second;
      l.first += it.first;
      ranks.push(l);
    }
    for (int32_t c = C - 1; c >= 0; --c)   // parents before children
      if (own[c] == -2) own[c] = own[p.parent[c]];
  };
  double best = std::numeric_limits<double>::infinity();
  std::vector<int32_t> cand;
  // GTB_MIN_TOP=k (tests): consider only cuts with at least k top cliques, so that small problems exercise the fan-in
  const char* mt = getenv("GTB_MIN_TOP");
  const size_t min_top = std::min(seq.size(), (size_t)(mt ? std::max(0, atoi(mt)) : 0));
  for (size_t n_open = min_top; n_open <= seq.size(); ++n_open) {
    assign(n_open, cand);
    const double t = modelled_time(p, cand, world);
    if (t < best) { best = t; owner = cand; }
  }
}

// A point eliminated alone, every factor a projection (bucket 2) from a distinct camera, at most kmax of them
// (15: what the thin kernel's staging area holds; 22: the largest thin front, for the point kernel).
static bool is_star_clique(const Engine* e, int32_t c, int kmax, bool any_dim = false) {
  const Plan& p = e->plan;
  const int64_t f0 = p.fac_off[c], f1 = p.fac_off[c + 1];
  bool star = p.F[c] == 3 && f1 > f0 && f1 - f0 <= kmax && (any_dim || is_thin_front(p.F[c], p.F[c] + p.S[c] + 1)) &&
              p.S[c] == 9 * (f1 - f0) && p.child_off[c + 1] == p.child_off[c];
  for (int64_t f = f0; star && f < f1; ++f) {
    const FactorSlot& fs = p.slots[f];
    star = e->g_bucket[fs.factor] == 2 && fs.col1 == 0 && fs.col0 >= 3;
    for (int64_t g2 = f0; star && g2 < f; ++g2) star = p.slots[g2].col0 != fs.col0;
  }
  return star;
}

// Builds the level schedule of one phase: phase 0 = subtree cliques owned by this rank, phase 1 = the shared top of the
// tree (factored on rank 0 after the fan-in reduction; empty when there is one rank).
static void build_phase(Engine* e, int ph, std::vector<int32_t>& lc) {
  const Plan& p = e->plan;
  std::vector<LevelSched>& sched = e->sched[ph];
  sched.assign(p.n_levels, LevelSched());
  e->rinv_task_begin[ph] = (int64_t)e->rinv_tasks.size() / 3;
  auto mine = [&](int32_t c) { return ph == 0 ? e->owner[c] == e->rank : e->owner[c] < 0; };
  for (int32_t l = 0; l < p.n_levels; ++l) {
    LevelSched& L = sched[l];
    // filtered copy of the level list: small fronts first, then large ones (the plan's order)
    const int64_t pb = p.level_off[l], pns = p.level_n_small[l], pen = p.level_off[l + 1];
    const int64_t b = (int64_t)lc.size();
    int64_t ns = 0;
    // (a point clique with more cameras than a thin front holds is large by its dimensions but runs on the point path)
    auto small_q = [&](int64_t q) { return q < pb + pns || e->is_point[p.level_cliques[q]]; };
    for (int64_t q = pb; q < pen; ++q)
      if (mine(p.level_cliques[q]) && small_q(q)) {
        lc.push_back(p.level_cliques[q]);
        ++ns;
      }
    for (int64_t q = pb; q < pen; ++q)
      if (mine(p.level_cliques[q]) && !small_q(q)) lc.push_back(p.level_cliques[q]);
    const int64_t en = (int64_t)lc.size();
    L.n_all = (int32_t)(en - b);
    L.large_begin = b + ns;
    L.n_large = (int32_t)(en - b - ns);
    L.min_F = INT_MAX;
    for (int64_t q = L.large_begin; q < en; ++q) {
      const int32_t c = lc[q];
      L.max_F = std::max(L.max_F, p.F[c]);
      L.min_F = std::min(L.min_F, p.F[c]);
      L.max_dim = std::max(L.max_dim, p.F[c] + p.S[c] + 1);
      L.zero_blocks = std::max<int64_t>(L.zero_blocks, ((int64_t)p.ld[c] * (p.F[c] + p.S[c] + 1) / 2 + 255) / 256);
    }
    L.zero_blocks = std::min<int64_t>(L.zero_blocks, 64);
    if (L.n_large) {
      const int max_np = (L.max_F + PW - 1) / PW;
      for (int sidx = 0; sidx <= max_np; ++sidx) {
        L.step_begin.push_back((int64_t)e->step_tasks_list.size());
        // Column slices per role-A CTA: one, unless the launch would then hold more role-A CTAs than the machine has SMs
        // (one such CTA per SM): every role-A CTA factors its front's diagonal block by itself, so a second wave repeats
        // that work, while a CTA that walks several slices pays for it once.
        static const int sm_target = getenv("GTB_ROLE_A_CTAS") ? atoi(getenv("GTB_ROLE_A_CTAS")) : 148;
        int spc = 1;
        for (; spc < 16; ++spc) {
          int64_t ctas = 0;
          for (int64_t q = L.large_begin; q < en; ++q) {
            const int32_t c = lc[q];
            const int F = p.F[c], dim = F + p.S[c] + 1;
            if (sidx >= (F + PW - 1) / PW) continue;
            ctas += role_a_ctas(dim, sidx * PW + std::min(PW, F - sidx * PW), spc);
          }
          if (ctas <= sm_target) break;
        }
        L.step_spc.push_back(spc);
        // role-A CTAs of every clique first (critical path), then the tile tasks
        for (int pass = 0; pass < 2; ++pass)
          for (int64_t q = L.large_begin; q < en; ++q) {
            const int32_t c = lc[q];
            const int F = p.F[c], dim = F + p.S[c] + 1, np_c = (F + PW - 1) / PW;
            if (sidx >= np_c) continue;
            const int n = step_tasks(F, dim, sidx, spc);
            const int k0 = sidx * PW, kb = std::min(PW, F - k0);
            const int nA = role_a_ctas(dim, k0 + kb, spc);
            const int lo = pass == 0 ? 0 : nA, hi = pass == 0 ? nA : n;
            for (int t = lo; t < hi; ++t) e->step_tasks_list.push_back(StepTask{c, t});
          }
        L.step_count.push_back((int32_t)(e->step_tasks_list.size() - L.step_begin.back()));
        if (getenv("GTB_DUMP_STEPS")) {   // composition of this panel-step launch (planning aid, tools/step_mix.py)
          int nf = 0, na = 0, maxdim = 0, maxF = 0;
          for (int64_t q = L.large_begin; q < en; ++q) {
            const int32_t c = lc[q];
            const int F = p.F[c], dim = F + p.S[c] + 1;
            if (sidx >= (F + PW - 1) / PW) continue;
            const int k0 = sidx * PW, kb = std::min(PW, F - k0);
            ++nf; na += role_a_ctas(dim, k0 + kb, spc); maxdim = std::max(maxdim, dim); maxF = std::max(maxF, F);
          }
          fprintf(stderr, "step level=%d p=%d fronts=%d roleA=%d (x%d slices) tiles=%d maxF=%d maxdim=%d\n", (int)l, sidx, nf, na, spc,
                  (int)L.step_count.back() - na, maxF, maxdim);
        }
        L.schur_begin.push_back((int64_t)e->step_tasks_list.size());
        // a launch with fewer tiles than the machine holds (2 CTAs per SM) splits K = F over several CTAs per tile
        int64_t launch_tiles = 0;
        for (int64_t q = L.large_begin; q < en; ++q) {
          const int32_t c = lc[q];
          if (sidx == (p.F[c] + PW - 1) / PW) launch_tiles += step_tasks(p.F[c], p.F[c] + p.S[c] + 1, sidx);
        }
        static const int split_target = getenv("GTB_SCHUR_SPLIT") ? atoi(getenv("GTB_SCHUR_SPLIT")) : 2 * 148;
        // ordered hand-over: rank of a clique among the same-level siblings that hand over to the same parent by
        // themselves (fuse_up): one launch per rank (the unordered mode keeps one launch)
        auto sib_rank = [&](int32_t c) {
          if (!e->ordered || !e->fuse_up[c]) return 0;
          int r = 0;
          const int32_t par = p.parent[c];
          for (int64_t k = p.child_off[par]; k < p.child_off[par + 1] && p.children[k] != c; ++k) {
            const int32_t sb = p.children[k];
            if (e->fuse_up[sb] && p.level[sb] == p.level[c] && mine(sb)) ++r;
          }
          return r;
        };
        int max_rank = 0;
        for (int64_t q = L.large_begin; q < en; ++q)
          if (sidx == (p.F[lc[q]] + PW - 1) / PW) max_rank = std::max(max_rank, sib_rank(lc[q]));
        L.schur_runs.emplace_back();
        for (int rk = 0; rk <= max_rank; ++rk) {
          const int64_t run_begin = (int64_t)e->step_tasks_list.size();
          int64_t run_tiles = 0;
          for (int64_t q = L.large_begin; q < en; ++q) {
            const int32_t c = lc[q];
            if (sidx == (p.F[c] + PW - 1) / PW && sib_rank(c) == rk) run_tiles += step_tasks(p.F[c], p.F[c] + p.S[c] + 1, sidx);
          }
          if (!e->ordered) run_tiles = launch_tiles;
          for (int64_t q = L.large_begin; q < en; ++q) {
            const int32_t c = lc[q];
            const int F = p.F[c], dim = F + p.S[c] + 1, np_c = (F + PW - 1) / PW;
            if (sidx != np_c || sib_rank(c) != rk) continue;
            const int n = step_tasks(F, dim, sidx);
            need(n < (1 << kSchurTileBits), "Schur-step tile count overflow");
            int nsplit = 1;
            if (run_tiles > 0 && run_tiles < split_target)
              nsplit = (int)std::max<int64_t>(1, std::min<int64_t>({(split_target + run_tiles - 1) / run_tiles, F / 128, 31}));
            if (nsplit > 1 && e->ordered) {   // hand-shake flags: the K slices of a tile add one after the other
              e->flag_off[c] = (int32_t)e->tile_flags;
              e->tile_flags += n;
              need(e->tile_flags < INT32_MAX, "K-split flag overflow");
            }
            for (int sl = 0; sl < nsplit; ++sl)
              for (int t = 0; t < n; ++t) e->step_tasks_list.push_back(StepTask{c, schur_task(t, sl, nsplit)});
          }
          const int32_t cnt = (int32_t)((int64_t)e->step_tasks_list.size() - run_begin);
          if (cnt > 0) L.schur_runs.back().push_back(LevelSched::Run{run_begin, cnt});
        }
        L.schur_count.push_back((int32_t)(e->step_tasks_list.size() - L.schur_begin.back()));
      }
    }
    // extend-add work of this level's large parents.  Phase 1 only adds the children that are top cliques themselves:
    // subtree children were added before the reduction by whichever rank owns them (top_work).
    L.work_begin = (int64_t)e->child_work.size();
    L.gather_begin = (int64_t)e->gather_work.size();
    for (int64_t q = L.large_begin; q < en; ++q) {
      const int32_t c = lc[q];
      std::vector<int32_t> kids;
      for (int64_t k = p.child_off[c]; k < p.child_off[c + 1]; ++k) {
        const int32_t ch = p.children[k];
        if (!(ph == 0 || e->owner[ch] < 0)) continue;
        if (!e->ordered) push_child_work(e, ch);
        else if (!e->is_point[ch] && !e->fuse_up[ch]) kids.push_back(ch);
      }
      push_gather(e, c, kids);
    }
    L.n_work = (int32_t)(e->child_work.size() - L.work_begin);
    L.n_gather = (int32_t)(e->gather_work.size() - L.gather_begin);
    L.bs_begin = (int64_t)e->bs_list.size();
    int32_t maxN = 0;
    for (int pass = 0; pass < 3; ++pass)
      for (int64_t q = b; q < en; ++q) {
        const int32_t c = lc[q];
        const bool warp_class = (p.F[c] <= 32 && p.F[c] + p.S[c] + 1 <= kDenseMaxDim) || e->is_point[c];
        const bool pt8 = e->is_point[c] && p.S[c] <= 72;   // (is_point: handled by the point path, conditional rows with ld = 4)
        if ((pass == 0) != pt8 || (pass < 2) != warp_class) continue;
        e->bs_list.push_back(c);
        if (pt8) L.n_bs_pt++;
        if (warp_class) L.n_bs_warp++;
        else {
          L.n_bs_cta++;
          maxN = std::max(maxN, p.F[c] + p.S[c]);
          L.bs_max_F = std::max(L.bs_max_F, p.F[c]);
          L.bs_max_S = std::max(L.bs_max_S, p.S[c]);
        }
      }
    // scratch offsets of the CTA-class cliques (same order as the list)
    {
      int64_t off = 0;
      e->bs_scratch_off.resize(e->bs_list.size(), 0);
      e->bs_rinv_off.resize(e->bs_list.size(), 0);
      for (int64_t q = L.bs_begin + L.n_bs_warp; q < (int64_t)e->bs_list.size(); ++q) {
        const int32_t c = e->bs_list[q];
        e->bs_scratch_off[q] = (int32_t)off;
        off += (int64_t)p.F[c] * ((p.S[c] + kGemvCols - 1) / kGemvCols);
        e->bs_rinv_off[q] = (int32_t)e->rinv_blocks;
        for (int32_t bi = 0; bi < (p.F[c] + 31) / 32; ++bi) {
          e->rinv_tasks.push_back(c);
          e->rinv_tasks.push_back(bi);
          e->rinv_tasks.push_back((int32_t)e->rinv_blocks++);
        }
      }
      need(e->rinv_blocks < INT32_MAX / 2, "back-substitution inverse blocks overflow");
      need(off < INT32_MAX, "back-substitution scratch overflow");
      e->bs_scratch_doubles = std::max(e->bs_scratch_doubles, off);
    }
    L.bs_cta_smem = (maxN + kBsCtaThreads + 64 + 2 * 32 * 33) * (int32_t)sizeof(double);

    // re-sort the small part of the level by dimension and cut into thread classes
    auto dim_of = [&](int32_t c) { return p.F[c] + p.S[c] + 1; };
    // point fronts first, then the other thin fronts, then dense ones; each group sorted by dimension
    auto point = [&](int32_t c) { return e->is_point[c] != 0; };
    auto thin = [&](int32_t c) { return !point(c) && is_thin_front(p.F[c], dim_of(c)); };
    std::stable_sort(lc.begin() + b, lc.begin() + b + ns, [&](int32_t a, int32_t c2) {
      const int ka = point(a) ? 0 : (thin(a) ? 1 : 2), kc = point(c2) ? 0 : (thin(c2) ? 1 : 2);
      return ka != kc ? ka < kc : dim_of(a) < dim_of(c2);
    });
    int64_t q = b;
    // point classes by number of cameras: shared memory per warp grows with k
    while (q < b + ns && point(lc[q])) {
      const int64_t q0 = q;
      const int k0 = p.S[lc[q]] / 9;
      const int hi = std::max(k0 + 1, k0 * 5 / 4);
      int32_t mk = 0;
      while (q < b + ns && point(lc[q]) && p.S[lc[q]] / 9 <= hi) { mk = std::max(mk, p.S[lc[q]] / 9); ++q; }
      LevelSched::Cls cl;
      cl.begin = q0; cl.count = (int32_t)(q - q0); cl.kind = 3; cl.threads = kThinWarps * 32;
      cl.group_doubles = mk;
      cl.smem = e->point_pairs ? point_rows_doubles(mk) * (int32_t)sizeof(double) * kThinWarps * (mk <= 8 ? 4 : 1)
                               : point_front_doubles(mk, true) * (int32_t)sizeof(double) * kThinWarps;
      L.small.push_back(cl);
    }
    // thin classes are cut adaptively so that the shared memory of a class (sized by its largest member) stays within
    // ~20 % of what its smallest member needs (occupancy is limited by shared memory) -- but only once a class fills the
    // machine: the launches of a level run one after the other, and a pose-graph level cut into 30 classes of 40 fronts is
    // 30 launch latencies for an almost empty GPU.  Warp and CTA kinds stay apart (different kernels).
    constexpr int kMinClass = 2 * 148;   // fronts (CTA kind) or CTAs of 8 fronts (warp kind) before a class may be cut
    while (q < b + ns && thin(lc[q])) {
      const int64_t q0 = q;
      const int d0 = dim_of(lc[q]);
      const bool warp_kind = d0 <= 48;
      const int kind_max = warp_kind ? 48 : kThinMaxDim;
      const int hi = warp_kind ? (d0 <= 32 ? 32 : (d0 <= 40 ? 40 : 48)) : std::max(d0 + 8, d0 * 6 / 5);
      int32_t md = 0;
      while (q < b + ns && thin(lc[q]) && dim_of(lc[q]) <= kind_max &&
             (dim_of(lc[q]) <= hi || (q - q0) < (int64_t)kMinClass * (warp_kind ? kThinWarps : 1))) {
        md = std::max(md, dim_of(lc[q]));
        ++q;
      }
      LevelSched::Cls cl;
      cl.begin = q0; cl.count = (int32_t)(q - q0); cl.kind = warp_kind ? 0 : 1;
      // CTA kind: the front's passes (zero fill, children, write-out with the Schur update) are element-parallel over up to
      // 21 k packed entries and shared memory already limits a big front to one or two CTAs per SM: more threads per CTA
      static const int thin_threads = getenv("GTB_THIN_THREADS") ? atoi(getenv("GTB_THIN_THREADS")) : 0;
      cl.threads = warp_kind ? kThinWarps * 32 : (thin_threads ? thin_threads : (md >= 128 ? 512 : (md >= 80 ? 256 : 128)));
      cl.group_doubles = ((md * (md + 1) / 2) + 1) & ~1;
      cl.smem = (cl.group_doubles + kThinStage) * (int32_t)sizeof(double) * (warp_kind ? kThinWarps : 1);
      L.small.push_back(cl);
    }
    // dense small fronts: thread classes by dimension, merged while a class does not fill the machine
    static const int small_mul = getenv("GTB_SMALL_MUL") ? atoi(getenv("GTB_SMALL_MUL")) : 2;
    const int32_t bounds[4] = {32, 64, 96, kDenseMaxDim};
    const int32_t threads[4] = {32 * small_mul, 64 * small_mul, 128 * small_mul, 256 * small_mul};
    while (q < b + ns) {
      const int64_t q0 = q;
      int32_t md = 0;
      int kcls = 0;
      while (q < b + ns) {
        const int d = dim_of(lc[q]);
        int kq = 0;
        while (kq < 3 && d > bounds[kq]) ++kq;
        if (q > q0 && kq != kcls && (q - q0) >= kMinClass) break;
        kcls = std::max(kcls, kq);
        md = std::max(md, d);
        ++q;
      }
      LevelSched::Cls cl;
      cl.begin = q0; cl.count = (int32_t)(q - q0); cl.threads = threads[kcls]; cl.kind = 2; cl.group_doubles = 0;
      cl.smem = (md | 1) * md * (int32_t)sizeof(double);
      L.small.push_back(cl);
    }
  }
}

// Stable LSD radix sort of records by a 64-bit key (11-bit digits, passes whose digit is constant are skipped).
template <class R> static void radix_sort_by_key(std::vector<R>& v) {
  if (v.size() < 2) return;
  std::vector<R> tmp(v.size());
  uint64_t all_or = 0, all_and = ~0ull;
  for (const R& r : v) { all_or |= r.key; all_and &= r.key; }
  const uint64_t varying = all_or & ~all_and;
  for (int shift = 0; shift < 64; shift += 11) {
    if (!((varying >> shift) & 0x7ff)) continue;
    size_t cnt[2049] = {0};
    for (const R& r : v) cnt[((r.key >> shift) & 0x7ff) + 1]++;
    for (int i = 0; i < 2048; ++i) cnt[i + 1] += cnt[i];
    for (const R& r : v) tmp[cnt[(r.key >> shift) & 0x7ff]++] = r;
    v.swap(tmp);
  }
}

// Camera-pair lists for the Schur complement of the point cliques (eliminate.cuh, point_pair_kernel): for every level
// and phase, the (parent, block row, block column) groups of all its point cliques, items in list order, cut into
// segments of at most kPairSeg items, longest segments first.
static void build_point_pairs(Engine* e) {
  // A thread walks its segment one item at a time (item -> rows: dependent round trips), so the cap sets the length of
  // the longest chain in the launch.  Same-camera lists are long (≈400 points a camera): 16 measured best on C3.
  // GTB_PAIR_SEG / GTB_DIAG_SEG override the caps for experiments.
  const size_t cap_pair = getenv("GTB_PAIR_SEG") ? (size_t)std::max(1, atoi(getenv("GTB_PAIR_SEG"))) : (size_t)kPairSeg;
  const size_t cap_diag = getenv("GTB_DIAG_SEG") ? (size_t)std::max(1, atoi(getenv("GTB_DIAG_SEG"))) : (size_t)kPairSeg / 2;
  const Plan& p = e->plan;
  e->pair_segs.clear();
  e->pair_items.clear();
  e->diag_items.clear();
  e->pair_reduce.clear();
  e->pair_slots = 0;
  if (!e->point_pairs) return;
  struct OffRec { uint64_t key; uint32_t a, b; };
  struct DiagRec { uint64_t key; uint32_t y, fac, rhs; };
  std::vector<OffRec> off;
  std::vector<DiagRec> dg;
  for (int ph = 0; ph < 2; ++ph)
    for (int32_t l = 0; l < p.n_levels; ++l) {
      LevelSched& L = e->sched[ph][l];
      off.clear();
      dg.clear();
      for (const auto& cl : L.small) {
        if (cl.kind != 3) continue;
        for (int64_t q = cl.begin; q < cl.begin + cl.count; ++q) {
          const int32_t c = e->sched_lists[q], P = p.parent[c];
          const int k = p.S[c] / 9;
          const int32_t* pm = p.pmap.data() + p.pmap_off[c];
          need(e->front_off[c] % 4 == 0 && e->front_off[c] / 4 + 4 + 9 * k < (int64_t)0x7fffffff, "point front offset out of range");
          need(P < (1 << 22) && pm[9 * k] < (1 << 21), "point pair key overflow");
          const uint32_t base = (uint32_t)(e->front_off[c] / 4) + 3;   // first separator column, 32-byte units
          const uint32_t rhs = base + 9 * k;
          uint32_t fac_of[kPointMaxCams];
          for (int64_t f = p.fac_off[c]; f < p.fac_off[c + 1]; ++f) fac_of[(p.slots[f].col0 - 3) / 9] = (uint32_t)e->g_index[p.slots[f].factor];
          for (int b1 = 0; b1 < k; ++b1) {
            const uint64_t A = (uint64_t)pm[9 * b1];
            dg.push_back(DiagRec{((uint64_t)P << 42) | (A << 21), base + 9 * b1, fac_of[b1], rhs | (b1 == 0 ? 0x80000000u : 0u)});
            for (int b2 = b1 + 1; b2 < k; ++b2) {
              const uint64_t B = (uint64_t)pm[9 * b2];
              if (A < B) off.push_back(OffRec{((uint64_t)P << 42) | (A << 21) | B, base + 9 * b1, base + 9 * b2});
              else off.push_back(OffRec{((uint64_t)P << 42) | (B << 21) | A, base + 9 * b2, base + 9 * b1});
            }
          }
        }
      }
      if (dg.empty()) continue;
      radix_sort_by_key(off);
      radix_sort_by_key(dg);
      int64_t level_slots = 0;
      L.red_begin = (int64_t)e->pair_reduce.size();
      auto emit = [&](auto& recs, bool diag, int64_t& seg_begin, int32_t& n_seg) {
        std::vector<PairSeg> segs;
        const int64_t item0 = diag ? (int64_t)e->diag_items.size() / 3 : (int64_t)e->pair_items.size();
        for (size_t i = 0; i < recs.size();) {
          size_t j = i;
          while (j < recs.size() && recs[j].key == recs[i].key) ++j;
          const int32_t P = (int32_t)(recs[i].key >> 42), A = (int32_t)((recs[i].key >> 21) & 0x1fffff), B = (int32_t)(recs[i].key & 0x1fffff);
          const int32_t ldp = e->ld[P], NP = p.F[P] + p.S[P];
          const int64_t dst = e->front_off[P] + A + (int64_t)(diag ? A : B) * ldp;
          const size_t cap = diag ? cap_diag : cap_pair;
          if (e->ordered && j - i > cap) {
            // several segments: their partial blocks go to scratch slots and are summed in list order
            const int32_t nsg = (int32_t)((j - i + cap - 1) / cap);
            e->pair_reduce.push_back(PairReduce{dst, ldp, NP - A, (int32_t)level_slots, nsg, diag ? 1 : 0, 0});
            for (size_t s0 = i; s0 < j; s0 += cap)
              segs.push_back(PairSeg{(level_slots++) * kPairSlot, 0, (int32_t)(item0 + s0), (int32_t)std::min<size_t>(cap, j - s0), NP - A});
            need(level_slots < INT32_MAX / kPairSlot, "pair scratch overflow");
          } else {
            for (size_t s0 = i; s0 < j; s0 += cap)
              segs.push_back(PairSeg{dst, ldp, (int32_t)(item0 + s0), (int32_t)std::min<size_t>(cap, j - s0), NP - A});
          }
          i = j;
        }
        need(item0 + (int64_t)recs.size() < INT32_MAX, "point pair list overflow");
        // Longest first, but only inside windows of the key order: neighbouring groups share a camera, so its points'
        // conditional rows stay in L2 across them, while the warps of a window still get segments of similar length.
        static const int64_t window = getenv("GTB_PAIR_WINDOW") ? atoll(getenv("GTB_PAIR_WINDOW")) : 4096;
        for (size_t w0 = 0; w0 < segs.size(); w0 += (window > 0 ? (size_t)window : segs.size())) {
          const size_t w1 = window > 0 ? std::min(segs.size(), w0 + (size_t)window) : segs.size();
          std::stable_sort(segs.begin() + w0, segs.begin() + w1, [](const PairSeg& x, const PairSeg& y) { return x.count > y.count; });
        }
        seg_begin = (int64_t)e->pair_segs.size();
        n_seg = (int32_t)segs.size();
        e->pair_segs.insert(e->pair_segs.end(), segs.begin(), segs.end());
      };
      emit(off, false, L.pair_seg_begin, L.n_pair_seg);
      for (const OffRec& r : off) e->pair_items.push_back(make_uint2(r.a, r.b));
      emit(dg, true, L.diag_seg_begin, L.n_diag_seg);
      for (const DiagRec& r : dg) { e->diag_items.push_back(r.y); e->diag_items.push_back(r.fac); e->diag_items.push_back(r.rhs); }
      L.n_red = (int32_t)((int64_t)e->pair_reduce.size() - L.red_begin);
      e->pair_slots = std::max(e->pair_slots, level_slots);
    }
}

static void install_plan(Engine* e) {
  const Plan& p = e->plan;
  const int32_t C = p.n_cliques;
  compute_partition(p, e->world, e->owner);
  {
    const char* pa = getenv("GTB_POINT_ATOMIC");
    e->point_pairs = !(pa && atoi(pa));
  }
  {
    const char* ha = getenv("GTB_HANDOVER_ATOMIC");
    e->ordered = !(ha && atoi(ha));
  }
  e->child_work.clear();
  e->gather_work.clear();
  e->gather_child.clear();
  e->gather_smap_off.clear();
  e->smap.clear();
  e->smap_of.assign(C, -1);
  e->flag_off.assign(C, 0);
  e->tile_flags = 0;
  e->bs_list.clear();
  e->bs_scratch_off.clear();
  e->bs_scratch_doubles = 0;
  e->bs_rinv_off.clear();
  e->rinv_tasks.clear();
  e->rinv_blocks = 0;
  e->step_tasks_list.clear();
  e->is_point.assign(C, 0);
  for (int32_t c = 0; c < C; ++c) {
    const int32_t par = p.parent[c];
    if (par < 0 || is_small_front(p.F[par], p.F[par] + p.S[par] + 1)) continue;
    // with the pair-gathered Schur complement the point kernel only produces the conditional rows: any camera count
    // up to kPointMaxCams, also points whose front would be large by its dimensions
    if (e->point_pairs ? is_star_clique(e, c, kPointMaxCams, true)
                       : (is_thin_front(p.F[c], p.F[c] + p.S[c] + 1) && is_star_clique(e, c, 22)))
      e->is_point[c] = 1;
  }
  // a large child of a large parent on the same rank and phase adds its Schur complement during its own Schur step
  {
    const char* fu = getenv("GTB_FUSE_UP");
    const bool on = !(fu && !atoi(fu));
    e->fuse_up.assign(C, 0);
    for (int32_t c = 0; c < C && on; ++c) {
      const int32_t par = p.parent[c];
      if (par < 0 || e->is_point[c] || is_small_front(p.F[c], p.F[c] + p.S[c] + 1) || is_small_front(p.F[par], p.F[par] + p.S[par] + 1)) continue;
      if (e->owner[c] == e->owner[par]) e->fuse_up[c] = 1;
    }
  }
  std::vector<int32_t> lc;
  lc.reserve(C);
  build_phase(e, 0, lc);
  build_phase(e, 1, lc);
  e->rinv_task_count[0] = e->rinv_task_begin[1] - e->rinv_task_begin[0];
  e->rinv_task_count[1] = (int64_t)e->rinv_tasks.size() / 3 - e->rinv_task_begin[1];
  e->sched_lists = lc;
  if (getenv("GTB_DUMP_SCHED")) {   // the large-front schedule, one line per level and per Schur-step launch
    for (int ph = 0; ph < 2; ++ph)
      for (int32_t l = 0; l < p.n_levels; ++l) {
        const LevelSched& L = e->sched[ph][l];
        if (!L.n_large) continue;
        fprintf(stderr, "phase %d level %d: %d large fronts:", ph, l, L.n_large);
        for (int64_t q = L.large_begin; q < L.large_begin + L.n_large; ++q) fprintf(stderr, " (F=%d S=%d)", p.F[lc[q]], p.S[lc[q]]);
        fprintf(stderr, "\n");
        for (size_t sx = 0; sx < L.step_count.size(); ++sx)
          if (L.schur_count[sx]) {
            double fl = 0;
            for (int64_t q = L.large_begin; q < L.large_begin + L.n_large; ++q) {
              const int32_t c = lc[q];
              if ((int)sx == (p.F[c] + PW - 1) / PW) fl += (double)p.F[c] * (p.S[c] + 1) * (p.S[c] + 2);
            }
            fprintf(stderr, "  schur launch after step %zu: %d CTAs, %.3f GFLOP (%.1f us at 35 TF/s); panel-step CTAs before it: %d\n", sx, L.schur_count[sx], fl / 1e9,
                    fl / 35e6, sx > 0 ? L.step_count[sx - 1] : 0);
          }
      }
  }
  // the shared top: fronts relocated to one contiguous tail of the arena (identical on every rank) so that the fan-in
  // is a single reduction; assembled before it from own factors (rank 0) and from the subtree children each rank owns
  // Front storage.  The plan's layout gives every clique a full (N+1)^2 front; a point front only ever stores its three
  // conditional rows [R S d] (its Schur complement goes straight into the parent), so it gets ld = 4: one 32-byte
  // column per separator scalar, contiguous.  C3: 4.9 -> 2.0 GB, C4: 29 -> 8 GB.  Same layout on every rank.
  e->front_off.assign(C, 0);
  e->ld = p.ld;
  int64_t off = 0;
  for (int32_t c = 0; c < C; ++c) {
    if (e->owner[c] < 0) continue;
    const int64_t dim = p.F[c] + p.S[c] + 1;
    if (e->is_point[c]) e->ld[c] = 4;
    e->front_off[c] = off;
    off += ((int64_t)e->ld[c] * dim + 15) & ~(int64_t)15;
  }
  e->top_list.clear();
  e->top_off = off;
  int64_t tail = e->top_off;
  e->top_zero_blocks = 0;
  for (int32_t c = 0; c < C; ++c)
    if (e->owner[c] < 0) {
      e->top_list.push_back(c);
      e->front_off[c] = tail;
      const int64_t sz = (int64_t)p.ld[c] * (p.F[c] + p.S[c] + 1);
      tail += (sz + 15) & ~(int64_t)15;
      e->top_zero_blocks = std::max<int64_t>(e->top_zero_blocks, (sz / 2 + 255) / 256);
    }
  e->top_zero_blocks = std::min<int64_t>(e->top_zero_blocks, 64);
  build_point_pairs(e);
  e->large_list.clear();
  for (int32_t c = 0; c < C; ++c)
    if (e->owner[c] == e->rank && !e->is_point[c] && !is_small_front(p.F[c], p.F[c] + p.S[c] + 1)) e->large_list.push_back(c);
  e->n_large_own = (int32_t)e->large_list.size();
  e->large_list.insert(e->large_list.end(), e->top_list.begin(), e->top_list.end());
  e->top_doubles = tail - e->top_off;
  e->arena_doubles = tail;
  e->top_work_begin = (int64_t)e->child_work.size();
  e->top_gather_begin = (int64_t)e->gather_work.size();
  for (int32_t c : e->top_list) {
    std::vector<int32_t> kids;
    for (int64_t k = p.child_off[c]; k < p.child_off[c + 1]; ++k) {
      const int32_t ch = p.children[k];
      if (e->owner[ch] != e->rank) continue;
      if (!e->ordered) push_child_work(e, ch);
      else if (!e->is_point[ch]) kids.push_back(ch);
    }
    push_gather(e, c, kids);
  }
  e->n_top_work = (int32_t)(e->child_work.size() - e->top_work_begin);
  e->n_top_gather = (int32_t)(e->gather_work.size() - e->top_gather_begin);
  // scalars whose solution this rank produces (frontals of its cliques; the top belongs to rank 0)
  std::vector<uint8_t> own(e->n_scalars, 0);
  for (int32_t c = 0; c < C; ++c)
    if (e->owner[c] == e->rank || (e->owner[c] < 0 && e->rank == 0))
      for (int32_t k = 0; k < p.F[c]; ++k) own[p.gidx[p.gidx_off[c] + k]] = 1;
  if (!e->has_device) return;

  cudaStream_t st = e->stream;
  std::vector<CliqueDev> cd(C);
  int64_t dfac_total = 0;
  for (int32_t c = 0; c < C; ++c) {
    CliqueDev& d = cd[c];
    d.front_off = e->front_off[c];
    d.F = p.F[c]; d.S = p.S[c]; d.ld = e->ld[c]; d.parent = p.parent[c];
    d.fac_begin = (int32_t)p.fac_off[c]; d.fac_end = (int32_t)p.fac_off[c + 1];
    d.child_begin = (int32_t)p.child_off[c]; d.child_end = (int32_t)p.child_off[c + 1];
    need(p.pmap_off[c] < INT32_MAX && p.gidx_off[c] < INT32_MAX, "plan index overflow");
    d.pmap_off = (int32_t)p.pmap_off[c]; d.gidx_off = (int32_t)p.gidx_off[c];
    d.first_key_var = p.vars[p.vars_off[c]];
    d.dfac_off = 0;
    d.flag_off = e->flag_off[c];
    d.flags = (is_star_clique(e, c, 15) ? kCliqueStar : 0) | (e->fuse_up[c] ? kCliqueFuseUp : 0);
    if (!e->is_point[c] && !is_small_front(p.F[c], p.F[c] + p.S[c] + 1)) {
      d.dfac_off = (int32_t)dfac_total;
      dfac_total += (int64_t)((p.F[c] + PW - 1) / PW) * PW * PW;
      need(dfac_total < INT32_MAX, "diagonal-block side buffer overflow");
    }
  }
  std::vector<SlotDev> sd(p.slots.size());
  for (size_t i = 0; i < sd.size(); ++i) {
    const FactorSlot& f = p.slots[i];
    sd[i] = SlotDev{f.col0, f.col1, e->g_bucket[f.factor], e->g_index[f.factor]};
  }
  e->d_cliques.upload(cd, st);
  e->d_slots.upload(sd, st);
  e->d_children.upload(p.children, st);
  e->d_pmap.upload(p.pmap, st);
  e->d_gidx.upload(p.gidx, st);
  e->d_level_cliques.upload(e->sched_lists, st);
  e->d_top_list.upload(e->top_list, st);
  e->d_large_list.upload(e->large_list, st);
  e->d_own_mask.upload(own, st);
  e->d_child_work.upload(e->child_work, st);
  e->d_gather_work.upload(e->gather_work, st);
  e->d_gather_child.upload(e->gather_child, st);
  e->d_gather_smap_off.upload(e->gather_smap_off, st);
  e->d_smap.upload(e->smap, st);
  e->d_pair_reduce.upload(e->pair_reduce, st);
  {
    need(p.orig_off.back() < INT32_MAX, "member list overflow");
    std::vector<int32_t> oo(p.orig_off.begin(), p.orig_off.end());
    e->d_orig_off.upload(oo, st);
    e->d_orig_end.upload(p.orig_end, st);
  }
  e->d_pair_scratch.alloc((size_t)e->pair_slots * kPairSlot + 1);
  e->d_tile_flags.alloc((size_t)e->tile_flags + 1);
  CK(cudaMemsetAsync(e->d_tile_flags.p, 0, ((size_t)e->tile_flags + 1) * sizeof(int32_t), st));
  e->epoch = 0;
  e->d_bs_list.upload(e->bs_list, st);
  e->d_bs_scratch_off.upload(e->bs_scratch_off, st);
  e->d_bs_scratch.alloc((size_t)e->bs_scratch_doubles + 1);
  e->d_bs_rinv_off.upload(e->bs_rinv_off, st);
  e->d_rinv_tasks.upload(e->rinv_tasks, st);
  e->d_bs_rinv.alloc((size_t)e->rinv_blocks * 1024 + 1);
  e->d_step_tasks.upload(e->step_tasks_list, st);
  e->d_pair_segs.upload(e->pair_segs, st);
  e->d_pair_items.upload(e->pair_items, st);
  e->d_diag_items.upload(e->diag_items, st);
  e->d_dbg.alloc(16);
  e->d_fail.alloc(1);
  e->d_dfac.alloc((size_t)dfac_total + PW * PW);
  e->d_arena.alloc((size_t)e->arena_doubles + 16);
  if (e->world > 1) e->d_delta_own.alloc((size_t)e->n_scalars + 2);
  CK(cudaStreamSynchronize(st));
  e->factorized = e->solved = false;
  e->marg_built = e->marg_valid = false;
}

static void set_ordering_ids(Engine* e, const std::vector<int32_t>& order) {
  need(e->finalized, "gtb_finalize_problem must be called first");
  e->sym.order = order;
  try {
    // GTB_AMALGAMATE = padded-area ratio up to which a large child is merged into its large parent for the numeric
    // factorisation (0 or 1: off), GTB_AMALG_MAXF = most frontal scalars of a merged front, GTB_AMALG_MINS = fewest
    // separator scalars of a child worth merging
    static const double ratio = getenv("GTB_AMALGAMATE") ? atof(getenv("GTB_AMALGAMATE")) : 1.35;
    static const int max_f = getenv("GTB_AMALG_MAXF") ? atoi(getenv("GTB_AMALG_MAXF")) : 512;
    static const int min_s = getenv("GTB_AMALG_MINS") ? atoi(getenv("GTB_AMALG_MINS")) : 512;
    const auto t0 = std::chrono::steady_clock::now();
    build_plan(e->sym, kDenseMaxDim, &e->rplan, &e->plan, ratio, max_f, min_s);
    const auto t1 = std::chrono::steady_clock::now();
    e->has_plan = true;
    install_plan(e);
    if (getenv("GTB_PLAN_TIMES"))
      fprintf(stderr, "plan times: build_plan %.3f ms, install_plan (schedule + upload) %.3f ms\n",
              std::chrono::duration<double, std::milli>(t1 - t0).count(),
              std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count());
    return;
  } catch (const std::runtime_error& ex) {
    throw StatusError(GTB_ERR_INCONSISTENT, ex.what());
  }
  e->has_plan = true;
  install_plan(e);
}

// ------------------------------------------------------------------------------------------------------
// numeric stages
// ------------------------------------------------------------------------------------------------------
template <int B> static void launch_linearize(Engine* e) {
  typedef BucketTraits<B> T;
  const BucketDev& bk = e->bk[B];
  if (!bk.n) return;
  constexpr int MW = T::M * T::W;
  constexpr int LDS = (MW % 2 == 0) ? MW + 1 : MW;
  const int smem = kFactorThreads * LDS * sizeof(double);
  if (!e->lin_attr[B]) {   // function attributes are per device: one flag per engine, not per process
    CK(cudaFuncSetAttribute(linearize_kernel<B>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    e->lin_attr[B] = true;
  }
  const int grid = (bk.n + kFactorThreads - 1) / kFactorThreads;
  const double* s0 = e->d_state[e->cur][T::T0].p;
  const double* s1 = T::T1 >= 0 ? e->d_state[e->cur][T::T1 >= 0 ? T::T1 : 0].p : nullptr;
  LAUNCH(e, K_LINEARIZE, linearize_kernel<B><<<grid, kFactorThreads, smem, e->stream>>>(bk, s0, s1));
}
static void linearize(Engine* e) {
  need_device(e);
  need(e->finalized, "problem not finalized");
  StageTimer tm(e, 0);
  launch_linearize<0>(e); launch_linearize<1>(e); launch_linearize<2>(e); launch_linearize<3>(e);
  launch_linearize<4>(e); launch_linearize<5>(e); launch_linearize<6>(e);
  tm.stop();
  e->linearized = true;
  e->factorized = e->solved = e->have_hdiag = e->retracted = false;
}

template <int B> static void launch_error(Engine* e, int which) {
  typedef BucketTraits<B> T;
  const BucketDev& bk = e->bk[B];
  if (!bk.n) return;
  const int grid = (bk.n + kFactorThreads - 1) / kFactorThreads;
  const double* s0 = e->d_state[which][T::T0].p;
  const double* s1 = T::T1 >= 0 ? e->d_state[which][T::T1 >= 0 ? T::T1 : 0].p : nullptr;
  LAUNCH(e, K_ERROR, error_kernel<B><<<grid, kFactorThreads, 0, e->stream>>>(bk, s0, s1, e->d_partials.p + e->partial_off[B]));
}
// slot >= 0: the sum goes to d_scalars[slot] and nothing is read back (the caller syncs once for several results)
static double nonlinear_error(Engine* e, int which, int slot = -1) {
  need_device(e);
  need(e->finalized, "problem not finalized");
  StageTimer tm(e, 5);
  const int w = which ? (e->cur ^ 1) : e->cur;
  launch_error<0>(e, w); launch_error<1>(e, w); launch_error<2>(e, w); launch_error<3>(e, w);
  launch_error<4>(e, w); launch_error<5>(e, w); launch_error<6>(e, w);
  LAUNCH(e, K_REDUCE, sum_partials_kernel<<<1, 256, 0, e->stream>>>(e->d_partials.p, (int)e->n_partials, e->d_scalars.p + (slot >= 0 ? slot : 0)));
  if (slot >= 0) { tm.stop(); return 0.0; }
  CK(cudaMemcpyAsync(e->h_scalars, e->d_scalars.p, sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  tm.stop();
  return e->h_scalars[0];
}

template <int B> static void launch_linear_error(Engine* e) {
  const BucketDev& bk = e->bk[B];
  if (!bk.n) return;
  const int grid = (bk.n + kFactorThreads - 1) / kFactorThreads;
  LAUNCH(e, K_LINERR, linear_error_kernel<B><<<grid, kFactorThreads, 0, e->stream>>>(
      bk, e->d_delta.p, e->d_partials.p + e->partial_off[B], e->d_partials.p + e->n_partials + e->partial_off[B]));
}
// e0 == nullptr: the two sums stay in d_scalars[0], [1] and nothing is read back
static void linear_errors(Engine* e, double* e0, double* e1) {
  need_device(e);
  need(e->linearized && e->solved, "gtb_linear_errors needs gtb_linearize and gtb_backsolve first");
  StageTimer tm(e, 3);
  launch_linear_error<0>(e); launch_linear_error<1>(e); launch_linear_error<2>(e); launch_linear_error<3>(e);
  launch_linear_error<4>(e); launch_linear_error<5>(e); launch_linear_error<6>(e);
  LAUNCH(e, K_REDUCE, sum_partials_kernel<<<1, 256, 0, e->stream>>>(e->d_partials.p, (int)e->n_partials, e->d_scalars.p));
  LAUNCH(e, K_REDUCE, sum_partials_kernel<<<1, 256, 0, e->stream>>>(e->d_partials.p + e->n_partials, (int)e->n_partials, e->d_scalars.p + 1));
  if (!e0) { tm.stop(); return; }
  CK(cudaMemcpyAsync(e->h_scalars, e->d_scalars.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  tm.stop();
  *e0 = e->h_scalars[0];
  *e1 = e->h_scalars[1];
}

static void hessian_diagonal(Engine* e) {
  need_device(e);
  need(e->linearized, "gtb_hessian_diagonal needs gtb_linearize first");
  StageTimer tm(e, 7);
  IncidenceDev inc{e->d_inc_off.p, e->d_inc_bucket.p, e->d_inc_index.p, e->d_inc_slot.p,
                   e->d_var_goff.p, e->d_var_dim.p, (int32_t)e->vkeys.size()};
  AbTable tab;
  for (int b = 0; b < 7; ++b) tab.Ab[b] = e->d_Ab[b].p;
  const int grid = ((int)e->vkeys.size() + 3) / 4;
  LAUNCH(e, K_HDIAG, hessian_diag_kernel<<<grid, 128, 0, e->stream>>>(inc, tab, e->d_hdiag.p));
  tm.stop();
  e->have_hdiag = true;
}

static ElimDev elim_dev(Engine* e, bool damp) {
  ElimDev E;
  E.cliques = e->d_cliques.p; E.slots = e->d_slots.p; E.children = e->d_children.p;
  E.pmap = e->d_pmap.p; E.gidx = e->d_gidx.p; E.arena = e->d_arena.p; E.dfac = e->d_dfac.p;
  for (int b = 0; b < 7; ++b) E.Ab[b] = e->d_Ab[b].p;
  E.damp = damp ? e->d_damp.p : nullptr;
  E.fail = e->d_fail.p;
  E.orig_off = e->d_orig_off.p;
  E.orig_end = e->d_orig_end.p;
  E.ordered = e->ordered ? 1 : 0;
  E.epoch32 = (e->epoch & 0x3ffffff) * 32;
  E.tile_flags = e->d_tile_flags.p;
  E.pair_scratch = e->d_pair_scratch.p;
  static const int dbg = getenv("GTB_DEBUG") ? atoi(getenv("GTB_DEBUG")) : 0;
  E.debug = dbg;
  E.dbg = (dbg & 8) ? e->d_dbg.p : nullptr;
  E.dbg_clique = e->plan.n_cliques - 1;   // the root
  E.dbg_child = -1;                       // its widest child (Schur-step probe)
  if (E.dbg) {
    const Plan& p = e->plan;
    const int32_t r = p.n_cliques - 1;
    for (int64_t k = p.child_off[r]; k < p.child_off[r + 1]; ++k)
      if (E.dbg_child < 0 || p.S[p.children[k]] > p.S[E.dbg_child]) E.dbg_child = p.children[k];
  }
  return E;
}

// Failure codes from the device: numeric clique id * 256 + index of the failing member (reference clique) inside it.
static uint64_t failed_key_of(const Engine* e, int32_t code) {
  const int32_t nc = code >> 8, sub = code & 255;
  const Plan& p = e->plan;
  const int64_t m = std::min<int64_t>(p.orig_off[nc] + sub, p.orig_off[nc + 1] - 1);
  const int32_t ce = p.orig_clique[m];
  return e->vkeys[e->rplan.vars[e->rplan.vars_off[ce]]];
}

static void comm(Engine* e, int op, double* buf, int64_t count) {
  if (e->world <= 1) return;
  if (!e->comm) throw StatusError(GTB_ERR_INVALID, "gtb_set_distributed: world > 1 needs a communication callback");
  prof_begin(e, K_COMM);   // a profile class of its own (not a kernel of this library: not counted as a launch)
  const int rc = e->comm(e->comm_user, op, buf, count, e->stream);
  if (e->profiling) cudaEventRecord(e->prof_events[e->prof_used++], e->stream);
  if (rc != 0) throw StatusError(GTB_ERR_CUDA, "communication callback failed");
}

// Elimination of one phase, leaves to roots.  Phase 1 fronts were cleared and partly assembled before the fan-in; only
// the children that are top cliques themselves are added here.
static void factorize_levels(Engine* e, const ElimDev& E, int phase) {
  cudaStream_t st = e->stream;
  const int32_t* lists = e->d_level_cliques.p;
  for (int32_t l = 0; l < e->plan.n_levels; ++l) {
    const LevelSched& L = e->sched[phase][l];
    for (const auto& c : L.small) {
      if (c.kind == 3 && e->point_pairs && c.group_doubles <= 8)
        LAUNCH(e, K_POINT, point_rows_kernel<8><<<(c.count + 4 * kThinWarps - 1) / (4 * kThinWarps), c.threads, c.smem, st>>>(
                               E, lists + c.begin, c.count, c.group_doubles));
      else if (c.kind == 3 && e->point_pairs)
        LAUNCH(e, K_POINT, point_rows_kernel<32><<<(c.count + kThinWarps - 1) / kThinWarps, c.threads, c.smem, st>>>(
                               E, lists + c.begin, c.count, c.group_doubles));
      else if (c.kind == 3)
        LAUNCH(e, K_POINT, point_front_kernel<<<(c.count + kThinWarps - 1) / kThinWarps, c.threads, c.smem, st>>>(
                               E, lists + c.begin, c.count, c.group_doubles, 1));
      else if (c.kind == 0)
        LAUNCH(e, K_THIN, thin_front_kernel<true><<<(c.count + kThinWarps - 1) / kThinWarps, c.threads, c.smem, st>>>(
                              E, lists + c.begin, c.count, c.group_doubles));
      else if (c.kind == 1)
        LAUNCH(e, K_THIN, thin_front_kernel<false><<<c.count, c.threads, c.smem, st>>>(E, lists + c.begin, c.count,
                                                                                       c.group_doubles));
      else
        LAUNCH(e, K_SMALL, small_front_kernel<<<c.count, c.threads, c.smem, st>>>(E, lists + c.begin));
    }
    if (L.n_pair_seg)
      LAUNCH(e, K_PAIR, point_pair_kernel<<<(3 * L.n_pair_seg + 127) / 128, 128, 0, st>>>(E, e->d_pair_segs.p + L.pair_seg_begin, L.n_pair_seg,
                                                                                          e->d_pair_items.p));
    if (L.n_diag_seg)
      LAUNCH(e, K_PAIR, point_diag_kernel<<<(4 * L.n_diag_seg + 127) / 128, 128, 0, st>>>(E, e->d_pair_segs.p + L.diag_seg_begin, L.n_diag_seg,
                                                                                          e->d_diag_items.p));
    if (L.n_red)
      LAUNCH(e, K_PAIR, pair_reduce_kernel<<<(L.n_red + 1) / 2, 2 * kPairSlot, 0, st>>>(E, e->d_pair_reduce.p + L.red_begin, L.n_red));
    if (!L.n_large) continue;
    if (L.n_work) {
      LAUNCH(e, K_LCHILD, launch_pdl(large_children_kernel, L.n_work, 256, 0, st, E, e->d_child_work.p + L.work_begin));
    }
    if (L.n_gather)
      LAUNCH(e, K_LCHILD, launch_pdl(extend_add_kernel, L.n_gather, 256, 0, st, E, e->d_gather_work.p + L.gather_begin,
                                     (const int32_t*)e->d_gather_child.p, (const int64_t*)e->d_gather_smap_off.p, (const int32_t*)e->d_smap.p));
    for (size_t s = 0; s < L.step_count.size(); ++s) {
      if (L.step_count[s] > 0)
        LAUNCH(e, K_STEP, launch_pdl(front_step_kernel, L.step_count[s], 256, kStepSmem, st, E, e->d_step_tasks.p + L.step_begin[s], (int)s,
                                     (int)L.step_spc[s]));
      for (const LevelSched::Run& run : L.schur_runs[s])
        LAUNCH(e, K_SCHUR, launch_pdl(schur_step_kernel, run.count, 256, kStepSmemTile, st, E, e->d_step_tasks.p + run.begin));
    }
  }
}

// the failure flag of the factorisation just enqueued, as a double in d_scalars[8] (agreed over the ranks: every rank must
// take the same branch of the LM control flow); nothing is read back here
// ride = true (a whole trial with several ranks): no collective here; the flag rides on the delta all-reduce of the
// back-substitution as a count of failing ranks (backsolve), and the ranks only agree on the failing clique when that
// count is not zero -- one collective less in every trial.
static void enqueue_fail_flag(Engine* e, bool ride = false) {
  LAUNCH(e, K_REDUCE, fail_flag_kernel<<<1, 1, 0, e->stream>>>(e->d_fail.p, e->d_scalars.p + 8));
  if (e->world > 1 && !ride) comm(e, GTB_COMM_ALLREDUCE_MIN, e->d_scalars.p + 8, 1);
}

// returns failing clique id or -1; defer = true: returns -1 at once, the flag is left in d_scalars[8] (enqueue_fail_flag)
// for a caller that syncs once at the end of a whole trial
static int32_t factorize(Engine* e, double lambda, int diag, double lo, double hi, bool defer = false) {
  need_device(e);
  need(e->has_plan, "gtb_set_ordering must be called first");
  need(e->linearized, "gtb_factorize needs gtb_linearize first");
  cudaStream_t st = e->stream;
  StageTimer tm(e, 1);
  const bool damp = lambda > 0;
  if (damp) {
    if (diag && !e->have_hdiag) {
      const bool keep = e->time_stages;
      e->time_stages = false;
      hessian_diagonal(e);
      e->time_stages = keep;
    }
    const double s = 1.0 / (1.0 / std::sqrt(lambda));   // CachedModel sigma = 1/sqrt(lambda), whitened A = I / sigma
    const int n = (int)e->n_scalars;
    LAUNCH(e, K_DAMP, damping_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, s, diag, lo, hi, e->d_hdiag.p, e->d_damp.p));
  }
  const int32_t init = INT_MAX;
  CK(cudaMemcpyAsync(e->d_fail.p, &init, sizeof(int32_t), cudaMemcpyHostToDevice, st));
  e->epoch++;
  const ElimDev E = elim_dev(e, damp);
  if (!e->elim_attr) {   // per device, hence per engine
    CK(cudaFuncSetAttribute(small_front_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(thin_front_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(point_front_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(point_rows_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(point_rows_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(thin_front_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(front_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmem));
    CK(cudaFuncSetAttribute(schur_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStepSmemTile));
    CK(cudaFuncSetAttribute(backsolve_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    e->elim_attr = true;
  }
  // every large front this rank touches is cleared and given its own factors first: thin children add their Schur
  // complements into their parents as soon as they are eliminated.  The top of the tree is cleared on every rank but
  // gets its factors (and damping) on rank 0 only: the fan-in sums the copies.
  if (!e->large_list.empty()) {
    const int nz = (int)e->large_list.size(), nf = e->rank == 0 ? nz : e->n_large_own;
    LAUNCH(e, K_ZERO, large_zero_kernel<<<dim3(nz, 64), 256, 0, st>>>(E, e->d_large_list.p));
    if (nf) LAUNCH(e, K_LFACT, large_factors_kernel<<<nf, 256, 0, st>>>(E, e->d_large_list.p));
  }
  factorize_levels(e, E, 0);
  if (!e->top_list.empty()) {
    // fan-in: every rank adds what it owns into its copy of the top fronts, the copies are summed onto rank 0
    if (e->n_top_work)
      LAUNCH(e, K_LCHILD, large_children_kernel<<<e->n_top_work, 256, 0, st>>>(E, e->d_child_work.p + e->top_work_begin));
    if (e->n_top_gather)
      LAUNCH(e, K_LCHILD, extend_add_kernel<<<e->n_top_gather, 256, 0, st>>>(E, e->d_gather_work.p + e->top_gather_begin, e->d_gather_child.p,
                                                                             e->d_gather_smap_off.p, e->d_smap.p));
    comm(e, GTB_COMM_REDUCE_SUM_TO_0, e->d_arena.p + e->top_off, e->top_doubles);
    if (e->rank == 0) factorize_levels(e, E, 1);
  }
  enqueue_fail_flag(e, defer);
  if (defer) {
    tm.stop();
    e->factorized = true;   // provisional: the caller checks the flag after its single sync
    e->solved = false;
    e->marg_valid = false;
    return -1;
  }
  CK(cudaMemcpyAsync(e->h_scalars + 5, e->d_scalars.p + 8, sizeof(double), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const int32_t fail = (int32_t)e->h_scalars[5];
  tm.stop();
  e->factorized = (fail == INT_MAX);
  e->solved = false;
  e->marg_valid = false;
  return fail == INT_MAX ? -1 : fail;
}

static void backsolve_levels(Engine* e, const ElimDev& E, int phase) {
  if (e->rinv_task_count[phase] > 0)
    LAUNCH(e, K_BACKSOLVE, launch_pdl(diag_inverse_kernel, (int)((e->rinv_task_count[phase] + kInvWarps - 1) / kInvWarps), kInvWarps * 32, 0,
                                      e->stream, E, (const int32_t*)e->d_rinv_tasks.p + 3 * e->rinv_task_begin[phase],
                                      (int)e->rinv_task_count[phase], e->d_bs_rinv.p));
  for (int32_t l = e->plan.n_levels - 1; l >= 0; --l) {
    const LevelSched& L = e->sched[phase][l];
    const int32_t* bl = e->d_bs_list.p + L.bs_begin;
    const int32_t* so = e->d_bs_scratch_off.p + L.bs_begin;
    const int32_t* ro = e->d_bs_rinv_off.p + L.bs_begin;
    if (L.n_bs_pt)
      LAUNCH(e, K_BACKSOLVE, launch_pdl(backsolve_points_kernel, (L.n_bs_pt + 4 * kBsWarps - 1) / (4 * kBsWarps), kBsWarps * 32, 0, e->stream, E,
                                        bl, L.n_bs_pt, e->d_delta.p));
    if (L.n_bs_warp > L.n_bs_pt)
      LAUNCH(e, K_BACKSOLVE, launch_pdl(backsolve_kernel<true>, (L.n_bs_warp - L.n_bs_pt + kBsWarps - 1) / kBsWarps, kBsWarps * 32,
                                        kBsWarps * kBsWarpDoubles * sizeof(double), e->stream, E, bl + L.n_bs_pt, L.n_bs_warp - L.n_bs_pt, e->d_delta.p,
                                        so + L.n_bs_pt, (const double*)e->d_bs_scratch.p, ro + L.n_bs_pt, (const double*)e->d_bs_rinv.p));
    if (L.n_bs_cta) {
      if (L.bs_max_S > 0)
        LAUNCH(e, K_BSGEMV, launch_pdl(backsolve_gemv_kernel,
                                       dim3(L.n_bs_cta, (L.bs_max_F + kGemvRows - 1) / kGemvRows, (L.bs_max_S + kGemvCols - 1) / kGemvCols),
                                       kGemvRows, 0, e->stream, E, bl + L.n_bs_warp, so + L.n_bs_warp, (const double*)e->d_delta.p,
                                       e->d_bs_scratch.p));
      LAUNCH(e, K_BACKSOLVE, launch_pdl(backsolve_kernel<false>, L.n_bs_cta, kBsCtaThreads, L.bs_cta_smem, e->stream, E, bl + L.n_bs_warp,
                                        L.n_bs_cta, e->d_delta.p, so + L.n_bs_warp, (const double*)e->d_bs_scratch.p, ro + L.n_bs_warp,
                                        (const double*)e->d_bs_rinv.p));
    }
  }
}

static void backsolve(Engine* e) {
  need_device(e);
  need(e->factorized, "gtb_backsolve needs a successful gtb_factorize first");
  StageTimer tm(e, 2);
  const ElimDev E = elim_dev(e, false);
  if (!e->top_list.empty()) {
    if (e->rank == 0) backsolve_levels(e, E, 1);
    comm(e, GTB_COMM_BROADCAST_FROM_0, e->d_delta.p, e->n_scalars);
  }
  backsolve_levels(e, E, 0);
  if (e->world > 1) {
    const int n = (int)e->n_scalars;
    // entry n of the summed vector: how many ranks saw their factorisation fail (d_scalars[8] holds this rank's flag)
    LAUNCH(e, K_OWNDELTA, own_delta_kernel<<<(n + 1 + 255) / 256, 256, 0, e->stream>>>(n, e->d_delta.p, e->d_own_mask.p, e->d_delta_own.p,
                                                                                      e->d_scalars.p + 8));
    comm(e, GTB_COMM_ALLREDUCE_SUM, e->d_delta_own.p, e->n_scalars + 1);
    CK(cudaMemcpyAsync(e->d_delta.p, e->d_delta_own.p, (size_t)n * sizeof(double), cudaMemcpyDeviceToDevice, e->stream));
  }
  tm.stop();
  e->solved = true;
}

// ------------------------------------------------------------------------------------------------------
// marginal covariances (Marginals::marginalCovariance, nonlinear/Marginals.cpp:108-150)
// ------------------------------------------------------------------------------------------------------
static void marginals_build(Engine* e) {
  const Plan& p = e->plan;
  const int32_t C = p.n_cliques;
  e->marg_cliques.assign(C, MargClique());
  int64_t soff = 0;
  for (int32_t c = 0; c < C; ++c) {
    MargClique& m = e->marg_cliques[c];
    const bool leaf = p.child_off[c + 1] == p.child_off[c];
    m.sld = leaf ? p.F[c] : p.F[c] + p.S[c];
    m.sigma_off = soff;
    soff += ((int64_t)m.sld * m.sld + 1) & ~(int64_t)1;
  }
  e->marg_lists.clear();
  e->marg_level_off.assign(p.n_levels + 1, 0);
  e->marg_level_small.assign(p.n_levels, 0);
  int64_t scratch_max = 0;
  for (int32_t l = 0; l < p.n_levels; ++l) {
    e->marg_level_off[l] = (int64_t)e->marg_lists.size();
    int64_t sc = 0;
    for (int pass = 0; pass < 2; ++pass)
      for (int64_t q = p.level_off[l]; q < p.level_off[l + 1]; ++q) {
        const int32_t c = p.level_cliques[q];
        const bool small = p.F[c] + p.S[c] <= 256;
        if (small != (pass == 0)) continue;
        e->marg_lists.push_back(c);
        if (small) e->marg_level_small[l]++;
        e->marg_cliques[c].scratch_off = sc;
        sc += (int64_t)p.F[c] * p.F[c] + 2 * (int64_t)p.F[c] * p.S[c];
      }
    scratch_max = std::max(scratch_max, sc);
  }
  e->marg_level_off[p.n_levels] = (int64_t)e->marg_lists.size();
  e->var_clique.assign(e->vkeys.size(), -1);
  e->var_foff.assign(e->vkeys.size(), 0);
  for (int32_t c = 0; c < C; ++c) {
    int32_t off = 0;
    for (int32_t k = 0; k < p.n_frontal_vars[c]; ++k) {
      const int32_t v = p.vars[p.vars_off[c] + k];
      e->var_clique[v] = c;
      e->var_foff[v] = off;
      off += e->vdim[v];
    }
  }
  e->d_marg_cliques.upload(e->marg_cliques, e->stream);
  e->d_marg_lists.upload(e->marg_lists, e->stream);
  e->d_sigma.alloc((size_t)soff + 2);
  e->d_marg_scratch.alloc((size_t)scratch_max + 2);
  CK(cudaStreamSynchronize(e->stream));
  e->marg_built = true;
}

static void marginals_compute(Engine* e) {
  need_device(e);
  need(e->factorized, "gtb_marginals_compute needs gtb_linearize and a successful gtb_factorize (lambda = 0) first");
  need(e->world == 1, "marginal covariances are computed on one GPU");
  if (!e->marg_built) marginals_build(e);
  const ElimDev E = elim_dev(e, false);
  for (int32_t l = e->plan.n_levels - 1; l >= 0; --l) {
    const int64_t b = e->marg_level_off[l];
    const int32_t ns = e->marg_level_small[l], nb = (int32_t)(e->marg_level_off[l + 1] - b) - ns;
    if (ns) LAUNCH(e, K_MARGINAL, marginal_kernel<<<ns, 128, 0, e->stream>>>(E, e->d_marg_cliques.p, e->d_marg_lists.p + b, e->d_sigma.p, e->d_marg_scratch.p));
    if (nb) LAUNCH(e, K_MARGINAL, marginal_kernel<<<nb, 1024, 0, e->stream>>>(E, e->d_marg_cliques.p, e->d_marg_lists.p + b + ns, e->d_sigma.p, e->d_marg_scratch.p));
  }
  CK(cudaStreamSynchronize(e->stream));
  e->marg_valid = true;
}

static void marginal_covariance(Engine* e, int64_t n, const uint64_t* keys, double* out) {
  need_device(e);
  need(e->marg_valid && e->factorized, "gtb_marginal_covariance needs gtb_marginals_compute after the last gtb_factorize");
  std::vector<MargQuery> q((size_t)n);
  int64_t dst = 0;
  for (int64_t i = 0; i < n; ++i) {
    const auto it = std::lower_bound(e->vkeys.begin(), e->vkeys.end(), keys[i]);
    if (it == e->vkeys.end() || *it != keys[i]) throw StatusError(GTB_ERR_KEY_MISSING, "marginalCovariance: key is not in the problem");
    const int32_t v = (int32_t)(it - e->vkeys.begin()), c = e->var_clique[v];
    need(c >= 0, "variable is not a frontal of any clique");
    const MargClique& m = e->marg_cliques[c];
    q[i] = MargQuery{m.sigma_off + e->var_foff[v] + (int64_t)e->var_foff[v] * m.sld, dst, m.sld, e->vdim[v]};
    dst += (int64_t)e->vdim[v] * e->vdim[v];
  }
  if (!n) return;
  DevBuf<MargQuery> dq;
  DevBuf<double> dout;
  dq.upload(q, e->stream);
  dout.alloc((size_t)dst);
  LAUNCH(e, K_MARGINAL, marginal_gather_kernel<<<(unsigned)n, 32, 0, e->stream>>>(dq.p, e->d_sigma.p, dout.p));
  CK(cudaMemcpyAsync(out, dout.p, (size_t)dst * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
}

template <int VT> static void launch_retract(Engine* e) {
  const int n = (int)e->vars[VT].n;
  if (!n) return;
  LAUNCH(e, K_RETRACT, retract_kernel<VT><<<(n + 127) / 128, 128, 0, e->stream>>>(n, e->d_state[e->cur][VT].p, e->d_type_goff[VT].p,
                                                             e->d_delta.p, e->d_state[e->cur ^ 1][VT].p));
}
template <int VT> static void launch_relinearize(Engine* e, double threshold, uint8_t* d_marks, int* d_count) {
  const int n = (int)e->vars[VT].n;
  if (!n) return;
  relinearize_marked_kernel<VT><<<(n + 127) / 128, 128, 0, e->stream>>>(n, e->d_type_goff[VT].p, e->d_delta.p, threshold,
                                                                        e->d_state[e->cur ^ 1][VT].p, e->d_state[e->cur][VT].p, d_marks,
                                                                        d_count);
  CK(cudaGetLastError());
}
static void retract(Engine* e) {
  need_device(e);
  need(e->solved, "gtb_retract needs gtb_backsolve first");
  StageTimer tm(e, 4);
  launch_retract<0>(e); launch_retract<1>(e); launch_retract<2>(e); launch_retract<3>(e);
  tm.stop();
  e->retracted = true;
}

// One tryLambda worth of device work with ONE host synchronisation (LevenbergMarquardtOptimizer.cpp:121-270): damped
// factorisation, back-substitution, both linear errors, retract into the scratch values and their nonlinear error are
// enqueued back to back; the failure flag and the three scalars come back together.  When the factorisation fails the
// later kernels run on an unusable factor (bounded loops, results ignored): a rare wasted 1 ms against two round trips
// to the host in every trial.
struct TrialResult { int32_t fail; double lin0, lin1, err; };
static TrialResult trial(Engine* e, double lambda, int diag, double lo, double hi) {
  factorize(e, lambda, diag, lo, hi, true);
  backsolve(e);
  linear_errors(e, nullptr, nullptr);
  retract(e);
  nonlinear_error(e, 1, 2);
  CK(cudaMemcpyAsync(e->h_scalars, e->d_scalars.p, 9 * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  if (e->world > 1)
    CK(cudaMemcpyAsync(e->h_scalars + 9, e->d_delta_own.p + e->n_scalars, sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  if (e->world > 1 && e->h_scalars[9] > 0.0) {
    // some rank failed (every rank sees the same count): agree on the first failing clique, as the reference would report
    comm(e, GTB_COMM_ALLREDUCE_MIN, e->d_scalars.p + 8, 1);
    CK(cudaMemcpyAsync(e->h_scalars + 8, e->d_scalars.p + 8, sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
  }
  TrialResult r;
  r.fail = e->h_scalars[8] >= (double)INT_MAX ? -1 : (int32_t)e->h_scalars[8];
  r.lin0 = e->h_scalars[0]; r.lin1 = e->h_scalars[1]; r.err = e->h_scalars[2];
  if (r.fail >= 0) e->factorized = e->solved = e->retracted = false;
  return r;
}

// ------------------------------------------------------------------------------------------------------
// LM / GN drivers
// ------------------------------------------------------------------------------------------------------
struct LMState { double error, lambda, factor; int iterations, inner; };

struct Trace {
  double* rows; int64_t cap; int64_t n = 0;
  void add(const LMState& s) {
    if (rows && n < cap) {
      double* r = rows + 5 * n;
      r[0] = s.inner; r[1] = 0; r[2] = s.error; r[3] = s.lambda; r[4] = s.iterations;
    }
    n++;
  }
};

// one tryLambda; returns true when the lambda search of this outer iteration is over
static bool try_lambda(Engine* e, const gtb_lm_params& P, LMState& S) {
  double modelFidelity = 0.0;
  bool step_ok = false, stop = false;
  double newError = std::numeric_limits<double>::infinity();
  const TrialResult tr = trial(e, S.lambda, P.diagonal_damping, P.min_diagonal, P.max_diagonal);
  if (tr.fail < 0) {
    const double oldLin = tr.lin0, newLin = tr.lin1;
    const double linearizedCostChange = oldLin - newLin;
    if (linearizedCostChange >= 0) {
      newError = tr.err;
      const double costChange = S.error - newError;
      if (linearizedCostChange > std::numeric_limits<double>::epsilon() * oldLin) {
        modelFidelity = costChange / linearizedCostChange;
        step_ok = modelFidelity > P.min_model_fidelity;
      }
      const double minAbsoluteTolerance = P.relative_error_tol * S.error;
      if (std::abs(costChange) < minAbsoluteTolerance) stop = true;
    }
  }
  if (step_ok) {
    // decreaseLambda (LevenbergMarquardtState.h:81-94)
    double newLambda = S.lambda, newFactor = S.factor;
    if (P.use_fixed_lambda_factor) {
      newLambda /= S.factor;
    } else {
      newLambda *= std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * modelFidelity - 1.0, 3));
      newFactor = 2.0 * S.factor;
    }
    newLambda = std::max(P.lambda_lower_bound, newLambda);
    e->cur ^= 1;
    e->linearized = e->factorized = e->solved = e->have_hdiag = e->retracted = false;
    S.error = newError; S.lambda = newLambda; S.factor = newFactor;
    S.iterations += 1; S.inner += 1;
    return true;
  } else if (!stop) {
    S.lambda *= S.factor;
    S.inner += 1;
    if (!P.use_fixed_lambda_factor) S.factor *= 2.0;
    return S.lambda >= P.lambda_upper_bound;
  }
  return true;
}

static void lm_iterate(Engine* e, const gtb_lm_params& P, LMState& S, Trace* tr) {
  linearize(e);
  if (S.inner == 0 && tr) tr->add(S);
  if (P.diagonal_damping) hessian_diagonal(e);
  while (!try_lambda(e, P, S))
    if (tr) tr->add(S);
}

static bool check_convergence(double rel, double abs_, double errThr, double cur, double next) {
  if (next <= errThr) return true;
  const double absoluteDecrease = cur - next;
  const double relativeDecrease = absoluteDecrease / cur;
  return (rel && (relativeDecrease <= rel)) || (absoluteDecrease <= abs_);
}

}  // namespace gtb

// ------------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------------
// The CUDA current device is per host thread: every entry point makes the engine's device current for its own duration
// and restores the caller's (an engine may be driven from another thread than the one that created it, or after the
// caller switched devices).
struct EngineDeviceGuard {
  int prev = -1;
  bool switched = false;
  explicit EngineDeviceGuard(const gtb_engine* e) {
    if (!e || !e->has_device) return;
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != e->device) switched = cudaSetDevice(e->device) == cudaSuccess;
  }
  ~EngineDeviceGuard() { if (switched && prev >= 0) cudaSetDevice(prev); }
};
#define GTB_TRY(e) EngineDeviceGuard _guard(e); try {
#define GTB_CATCH(e)                                                                         \
  }                                                                                          \
  catch (const StatusError& ex) { (e)->err = ex.what(); return ex.st; }                      \
  catch (const CudaError& ex) { (e)->err = ex.what(); return GTB_ERR_CUDA; }                 \
  catch (const std::bad_alloc&) { (e)->err = "out of host memory"; return GTB_ERR_INVALID; } \
  catch (const std::exception& ex) { (e)->err = ex.what(); return GTB_ERR_INVALID; }         \
  return GTB_OK;

extern "C" {

const char* gtb_version(void) { return "gtsam_b200 0.1 (sm_100a)"; }

gtb_status gtb_create(int device, gtb_engine** out) {
  if (!out) return GTB_ERR_INVALID;
  *out = nullptr;
  gtb_engine* e = new gtb_engine();
  e->device = device;
  if (device >= 0) {
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess || device >= count) {
      delete e;
      return GTB_ERR_NO_DEVICE;
    }
    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    const bool ok = cudaSetDevice(device) == cudaSuccess && cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess &&
                    cudaEventCreate(&e->ev[0]) == cudaSuccess && cudaEventCreate(&e->ev[1]) == cudaSuccess;
    if (!ok) {
      delete e;
      if (prev >= 0) cudaSetDevice(prev);
      return GTB_ERR_CUDA;
    }
    e->has_device = true;
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
  }
  *out = e;
  return GTB_OK;
}

void gtb_destroy(gtb_engine* e) {
  EngineDeviceGuard guard(e);
  if (e && e->has_device && e->stream) cudaStreamSynchronize(e->stream);   // its blocks go back to the cache: nothing may be in flight
  delete e;
}
const char* gtb_last_error(const gtb_engine* e) { return e ? e->err.c_str() : "null engine"; }

gtb_status gtb_set_variables(gtb_engine* e, int vtype, int64_t n, const uint64_t* keys, const double* state) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e) set_variables(e, vtype, n, keys, state); GTB_CATCH(e)
}
gtb_status gtb_set_factors(gtb_engine* e, int bucket, int64_t n, const uint64_t* keys, const double* z,
                           const int32_t* noise_kind, const double* noise) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e) set_factors(e, bucket, n, keys, z, noise_kind, noise); GTB_CATCH(e)
}
gtb_status gtb_set_graph_order(gtb_engine* e, int64_t n, const int64_t* pairs) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need(!e->finalized, "gtb_set_graph_order after gtb_finalize_problem");
  e->graph_order.assign(pairs, pairs + 2 * n);
  GTB_CATCH(e)
}
gtb_status gtb_finalize_problem(gtb_engine* e) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e) finalize_problem(e); GTB_CATCH(e)
}

gtb_status gtb_set_ordering(gtb_engine* e, int64_t n, const uint64_t* keys) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need(e->finalized, "gtb_finalize_problem must be called first");
  std::vector<int32_t> order(n);
  for (int64_t j = 0; j < n; ++j) {
    auto it = std::lower_bound(e->vkeys.begin(), e->vkeys.end(), keys[j]);
    if (it == e->vkeys.end() || *it != keys[j]) throw StatusError(GTB_ERR_KEY_MISSING, "ordering key is not a variable");
    order[j] = (int32_t)(it - e->vkeys.begin());
  }
  set_ordering_ids(e, order);
  GTB_CATCH(e)
}
gtb_status gtb_set_ordering_builtin(gtb_engine* e, int which) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need(e->finalized, "gtb_finalize_problem must be called first");
  const int32_t V = (int32_t)e->vkeys.size();
  std::vector<int32_t> order;
  order.reserve(V);
  if (which == 0) {
    for (int32_t v = 0; v < V; ++v) order.push_back(v);
  } else if (which == 1) {
    for (int64_t i = 0; i < e->vars[GTB_POINT3].n; ++i) order.push_back(e->type_var[GTB_POINT3][i]);
    for (int64_t i = 0; i < e->vars[GTB_CAMERA].n; ++i) order.push_back(e->type_var[GTB_CAMERA][i]);
    need((int32_t)order.size() == V, "Schur ordering needs a problem with only cameras and points");
  } else if (which == 2) {
    const auto t0 = std::chrono::steady_clock::now();
    order = nested_dissection_order(e->sym);
    if (getenv("GTB_PLAN_TIMES"))
      fprintf(stderr, "plan times: nested dissection %.3f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    need((int32_t)order.size() == V, "nested dissection lost a variable");
  } else {
    need(false, "unknown builtin ordering");
  }
  set_ordering_ids(e, order);
  GTB_CATCH(e)
}
gtb_status gtb_plan_sizes(const gtb_engine* e, int64_t* n_cliques, int64_t* n_clique_keys, int64_t* n_cond,
                          int64_t* n_scalars) {
  if (!e) return GTB_ERR_INVALID;
  if (!e->has_plan) { e->err = "no plan"; return GTB_ERR_INVALID; }
  if (n_cliques) *n_cliques = e->rplan.n_cliques;
  if (n_clique_keys) *n_clique_keys = (int64_t)e->rplan.vars.size();
  if (n_cond) *n_cond = e->rplan.cond_off[e->rplan.n_cliques];
  if (n_scalars) *n_scalars = e->n_scalars;
  return GTB_OK;
}
gtb_status gtb_plan_get(const gtb_engine* e, int64_t* nfrontal, int64_t* nkeys, uint64_t* keys, int64_t* parent) {
  if (!e) return GTB_ERR_INVALID;
  if (!e->has_plan) { e->err = "no plan"; return GTB_ERR_INVALID; }
  const Plan& p = e->rplan;   // the reference's layout, not the amalgamated numeric plan
  for (int32_t c = 0; c < p.n_cliques; ++c) {
    if (nfrontal) nfrontal[c] = p.n_frontal_vars[c];
    if (nkeys) nkeys[c] = p.n_vars[c];
    if (parent) parent[c] = p.parent[c];
  }
  if (keys)
    for (size_t i = 0; i < p.vars.size(); ++i) keys[i] = e->vkeys[p.vars[i]];
  return GTB_OK;
}
gtb_status gtb_set_robust(gtb_engine* e, int bucket, int loss, double k) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need(bucket >= 0 && bucket < 7, "bad factor bucket");
  need(loss >= GTB_LOSS_NONE && loss <= GTB_LOSS_L2WITHDEADZONE, "unknown m-estimator");
  // the reference's constructors reject a non-positive parameter (linear/LossFunctions.cpp:140-144, 173-177, ...)
  need(loss == GTB_LOSS_NONE || k > 0, "m-estimator parameter must be positive");
  e->fac[bucket].loss = loss;
  e->fac[bucket].loss_k = k;
  e->bk[bucket].loss = loss;
  e->bk[bucket].loss_k = k;
  e->linearized = e->factorized = e->solved = e->have_hdiag = false;
  e->lm_have_error = false;
  GTB_CATCH(e)
}
gtb_status gtb_set_distributed(gtb_engine* e, int32_t rank, int32_t world, gtb_comm_fn fn, void* user) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need(world >= 1 && rank >= 0 && rank < world, "gtb_set_distributed: need 0 <= rank < world");
  need(world == 1 || fn || !e->has_device, "gtb_set_distributed: world > 1 needs a communication callback");
  e->rank = rank; e->world = world; e->comm = fn; e->comm_user = user;
  if (e->has_plan) install_plan(e);
  GTB_CATCH(e)
}
gtb_status gtb_partition_get(const gtb_engine* e, int32_t* owner, double* out) {
  if (!e) return GTB_ERR_INVALID;
  if (!e->has_plan) { e->err = "no plan"; return GTB_ERR_INVALID; }
  const Plan& p = e->plan;
  if (owner)   // per REFERENCE clique (ids as in gtb_plan_get): the owner of the numeric clique that holds it
    for (int32_t c = 0; c < e->rplan.n_cliques; ++c) owner[c] = e->owner[p.num_of[c]];
  if (out) {
    for (int k = 0; k < 16; ++k) out[k] = 0;
    double total = 0;
    for (int32_t c = 0; c < p.n_cliques; ++c) {
      const double w = front_cost(p, c);
      total += w;
      if (e->owner[c] < 0) out[2] += w;
      else if (e->owner[c] < 13) out[3 + e->owner[c]] += w;
    }
    for (int k = 2; k < 16; ++k) out[k] /= total;
    out[0] = (double)e->top_list.size();
    out[1] = (double)e->top_doubles;
  }
  return GTB_OK;
}
// Host-only self-check of the launch schedule install_plan built (works on an engine without a device): every owned
// clique is eliminated exactly once and solved exactly once per phase, every class member fits its class's kernel
// and shared memory, the panel-step / Schur-step task lists hold what step_tasks() says, the inverse-block tasks and
// the camera-pair segments tile their buffers.  out[0..7] = launch classes, panel-step launches, Schur-step launches,
// back-substitution launches, inverse blocks, pair segments, same-camera segments, panel + Schur tasks.
gtb_status gtb_schedule_check(const gtb_engine* e, int64_t* out8) {
  if (!e) return GTB_ERR_INVALID;
  if (!e->has_plan) { e->err = "no plan"; return GTB_ERR_INVALID; }
  const Plan& p = e->plan;
  int64_t cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  auto fail = [&](const std::string& m) { e->err = "schedule check: " + m; return GTB_ERR_INCONSISTENT; };
  auto dim_of = [&](int32_t c) { return p.F[c] + p.S[c] + 1; };
  std::vector<int32_t> seen_elim(p.n_cliques, 0), seen_bs(p.n_cliques, 0);
  std::vector<uint8_t> rinv_used((size_t)e->rinv_blocks, 0);
  int64_t pair_items_seen = 0, diag_items_seen = 0;
  for (int ph = 0; ph < 2; ++ph) {
    if ((int32_t)e->sched[ph].size() != p.n_levels) return fail("level count");
    auto mine = [&](int32_t c) { return ph == 0 ? e->owner[c] == e->rank : e->owner[c] < 0; };
    for (int32_t l = 0; l < p.n_levels; ++l) {
      const LevelSched& L = e->sched[ph][l];
      const int64_t ns = L.n_all - L.n_large, b = L.large_begin - ns;
      int64_t expect = 0;
      for (int64_t q = p.level_off[l]; q < p.level_off[l + 1]; ++q) expect += mine(p.level_cliques[q]) ? 1 : 0;
      if (expect != L.n_all) return fail("level " + std::to_string(l) + ": clique count");
      if (b < 0 || b + L.n_all > (int64_t)e->sched_lists.size()) return fail("level list range");
      // launch classes tile the small part
      int64_t q = b;
      for (const LevelSched::Cls& cl : L.small) {
        ++cnt[0];
        if (cl.begin != q || cl.count <= 0) return fail("class ranges do not tile the level");
        if (cl.smem > 227 * 1024) return fail("class shared memory above 227 KB");
        for (int64_t k = cl.begin; k < cl.begin + cl.count; ++k) {
          const int32_t c = e->sched_lists[k];
          const int32_t d = dim_of(c);
          const int64_t packed = (int64_t)d * (d + 1) / 2;
          bool ok = false;
          switch (cl.kind) {
            case 3: ok = e->is_point[c] && p.S[c] / 9 <= cl.group_doubles && cl.group_doubles <= kPointMaxCams; break;
            case 0: ok = !e->is_point[c] && is_thin_front(p.F[c], d) && d <= 48 && packed <= cl.group_doubles && cl.threads == kThinWarps * 32; break;
            case 1: ok = !e->is_point[c] && is_thin_front(p.F[c], d) && packed <= cl.group_doubles && (cl.threads == 128 || cl.threads == 256 || cl.threads == 512); break;
            case 2: ok = !e->is_point[c] && d <= kDenseMaxDim && (int64_t)(d | 1) * d * 8 <= cl.smem && cl.threads >= 32 && cl.threads <= 1024; break;
            default: break;
          }
          if (!ok) return fail("clique " + std::to_string(c) + " does not fit its class (kind " + std::to_string(cl.kind) + ")");
        }
        q += cl.count;
      }
      if (q != b + ns) return fail("classes do not cover the small part of level " + std::to_string(l));
      for (int64_t k = b; k < b + L.n_all; ++k) {
        const int32_t c = e->sched_lists[k];
        if (c < 0 || c >= p.n_cliques || p.level[c] != l || !mine(c)) return fail("foreign clique in a level list");
        const bool large = k >= L.large_begin;
        if (large && (e->is_point[c] || is_small_front(p.F[c], dim_of(c)))) return fail("small front in the large part");
        if (!large && !e->is_point[c] && !is_small_front(p.F[c], dim_of(c))) return fail("large front in the small part");
        seen_elim[c]++;
      }
      // panel-step and Schur-step task lists
      if (L.n_large) {
        const int max_np = (L.max_F + PW - 1) / PW;
        if ((int)L.step_begin.size() != max_np + 1 || (int)L.schur_begin.size() != max_np + 1) return fail("step table size");
        for (int sidx = 0; sidx <= max_np; ++sidx) {
          int64_t want = 0, want_schur_min = 0;
          for (int64_t k = L.large_begin; k < L.large_begin + L.n_large; ++k) {
            const int32_t c = e->sched_lists[k];
            const int np_c = (p.F[c] + PW - 1) / PW;
            if (sidx < np_c) want += step_tasks(p.F[c], dim_of(c), sidx, L.step_spc[sidx]);
            if (sidx == np_c) want_schur_min += step_tasks(p.F[c], dim_of(c), sidx);
          }
          if (want != L.step_count[sidx]) return fail("panel-step task count at level " + std::to_string(l));
          if (L.schur_count[sidx] < want_schur_min || (want_schur_min == 0) != (L.schur_count[sidx] == 0)) return fail("Schur-step task count");
          cnt[1] += L.step_count[sidx] > 0;
          {   // the per-rank runs tile the step's Schur tasks
            int64_t at = L.schur_begin[sidx];
            if ((int)L.schur_runs.size() != max_np + 1) return fail("Schur run table size");
            for (const LevelSched::Run& run : L.schur_runs[sidx]) {
              if (run.begin != at || run.count <= 0) return fail("Schur runs do not tile the step");
              at += run.count;
            }
            if (at != L.schur_begin[sidx] + L.schur_count[sidx]) return fail("Schur runs do not cover the step");
            if (!e->ordered && L.schur_runs[sidx].size() > 1) return fail("unordered mode with several Schur runs");
          }
          cnt[2] += (int64_t)L.schur_runs[sidx].size();
          cnt[7] += L.step_count[sidx] + L.schur_count[sidx];
          for (int64_t t = L.schur_begin[sidx]; t < L.schur_begin[sidx] + L.schur_count[sidx]; ++t) {
            const StepTask& st = e->step_tasks_list[t];
            const int tile = st.task & ((1 << kSchurTileBits) - 1), slice = (st.task >> kSchurTileBits) & ((1 << kSchurSliceBits) - 1);
            const int nsplit = st.task >> (kSchurTileBits + kSchurSliceBits);
            if (nsplit < 1 || slice >= nsplit || tile >= step_tasks(p.F[st.clique], dim_of(st.clique), sidx)) return fail("Schur-step task encoding");
          }
        }
      }
      // back-substitution list: the same cliques, warp class first
      for (int64_t k = L.bs_begin; k < L.bs_begin + L.n_bs_warp + L.n_bs_cta; ++k) {
        if (k >= (int64_t)e->bs_list.size()) return fail("back-substitution list range");
        const int32_t c = e->bs_list[k];
        if (p.level[c] != l || !mine(c)) return fail("foreign clique in a back-substitution list");
        const bool warp_class = (p.F[c] <= 32 && dim_of(c) <= kDenseMaxDim) || e->is_point[c];
        if (warp_class != (k < L.bs_begin + L.n_bs_warp)) return fail("back-substitution class of clique " + std::to_string(c));
        if ((e->is_point[c] && p.S[c] <= 72) != (k < L.bs_begin + L.n_bs_pt)) return fail("point class of the back-substitution list");
        if (warp_class && !e->is_point[c] && p.S[c] > 160) return fail("warp-class conditional wider than its shared memory");
        if (!warp_class) {
          if ((int64_t)(p.F[c] + p.S[c] + kBsCtaThreads + 64 + 2 * 32 * 33) * 8 > L.bs_cta_smem || L.bs_cta_smem > 200 * 1024) return fail("back-substitution shared memory");
          const int nb = (p.F[c] + 31) / 32;
          for (int bi = 0; bi < nb; ++bi) {
            const int64_t dst = (int64_t)e->bs_rinv_off[k] + bi;
            if (dst < e->rinv_task_begin[ph] || dst >= e->rinv_task_begin[ph] + e->rinv_task_count[ph] || dst >= e->rinv_blocks) return fail("inverse-block range");
            if (e->rinv_tasks[3 * dst] != c || e->rinv_tasks[3 * dst + 1] != bi || e->rinv_tasks[3 * dst + 2] != dst) return fail("inverse-block task");
            if (rinv_used[(size_t)dst]++) return fail("inverse block used twice");
          }
        }
        seen_bs[c]++;
      }
      if (L.n_bs_warp + L.n_bs_cta != L.n_all) return fail("back-substitution count at level " + std::to_string(l));
      cnt[3] += (L.n_bs_pt > 0) + (L.n_bs_warp > L.n_bs_pt) + (L.n_bs_cta > 0) * (1 + (L.bs_max_S > 0));
      // camera-pair segments of the level's point cliques
      for (int pass = 0; pass < 2; ++pass) {
        const int64_t sb = pass ? L.diag_seg_begin : L.pair_seg_begin, sn = pass ? L.n_diag_seg : L.n_pair_seg;
        const int64_t nitems = pass ? (int64_t)e->diag_items.size() / 3 : (int64_t)e->pair_items.size();
        if (sn && sb + sn > (int64_t)e->pair_segs.size()) return fail("segment range");
        for (int64_t k = sb; k < sb + sn; ++k) {
          const PairSeg& sg = e->pair_segs[k];
          if (sg.count < 1 || sg.begin < 0 || (int64_t)sg.begin + sg.count > nitems || sg.dst < 0) return fail("segment contents");
          // ld == 0: a partial block in a scratch slot (ordered hand-over), summed by one of the level's reduce items
          if (sg.ld == 0 ? (!e->ordered || sg.dst % kPairSlot || sg.dst / kPairSlot >= e->pair_slots) : sg.ld < 4) return fail("segment target");
          (pass ? diag_items_seen : pair_items_seen) += sg.count;
        }
        cnt[pass ? 6 : 5] += sn;
      }
      {   // the reduce items of the level cover its scratch slots exactly once
        std::vector<uint8_t> slot_used;
        for (int64_t k = L.red_begin; k < L.red_begin + L.n_red; ++k) {
          const PairReduce& r = e->pair_reduce[k];
          if (r.count < 2 || r.slot0 < 0 || (int64_t)r.slot0 + r.count > e->pair_slots || r.ld < 4) return fail("pair reduce item");
          if ((int64_t)slot_used.size() < r.slot0 + r.count) slot_used.resize(r.slot0 + r.count, 0);
          for (int q = 0; q < r.count; ++q)
            if (slot_used[r.slot0 + q]++) return fail("scratch slot summed twice");
        }
        int64_t partial_segs = 0;
        for (int pass = 0; pass < 2; ++pass) {
          const int64_t sb = pass ? L.diag_seg_begin : L.pair_seg_begin, sn = pass ? L.n_diag_seg : L.n_pair_seg;
          for (int64_t k = sb; k < sb + sn; ++k)
            if (e->pair_segs[k].ld == 0) {
              const int64_t slot = e->pair_segs[k].dst / kPairSlot;
              if (slot >= (int64_t)slot_used.size() || slot_used[slot] != 1) return fail("partial segment without a reduce item");
              ++partial_segs;
            }
        }
        if (partial_segs != (int64_t)slot_used.size()) return fail("scratch slots without a segment");
      }
    }
  }
  for (int32_t c = 0; c < p.n_cliques; ++c) {
    const int want = (e->owner[c] == e->rank || e->owner[c] < 0) ? 1 : 0;
    if (seen_elim[c] != want || seen_bs[c] != want) return fail("clique " + std::to_string(c) + " scheduled " + std::to_string(seen_elim[c]) + " / " + std::to_string(seen_bs[c]) + " times");
  }
  for (size_t k = 0; k < rinv_used.size(); ++k)
    if (!rinv_used[k]) return fail("unused inverse block");
  if (e->ordered) {
    // ordered extend-add: the work items of a parent tile its columns, every listed child hands over to that parent
    // through a sorted map that is a permutation of its separator, and no child is gathered twice
    std::vector<int32_t> gathered(p.n_cliques, 0);
    for (size_t k = 0; k < e->gather_work.size();) {
      const GatherWork& w0 = e->gather_work[k];
      int32_t at = 0;
      size_t k2 = k;
      for (; k2 < e->gather_work.size() && e->gather_work[k2].parent == w0.parent && e->gather_work[k2].list_begin == w0.list_begin; ++k2) {
        const GatherWork& w = e->gather_work[k2];
        if (w.col_begin != at || w.col_end <= w.col_begin || w.list_end != w0.list_end) return fail("gather work items do not tile the parent");
        at = w.col_end;
      }
      if (at != dim_of(w0.parent)) return fail("gather work items do not cover the parent");
      for (int32_t q = w0.list_begin; q < w0.list_end; ++q) {
        const int32_t ch = e->gather_child[q];
        if (p.parent[ch] != w0.parent || e->is_point[ch]) return fail("gathered child of another parent");
        gathered[ch]++;
        const int32_t S1 = p.S[ch] + 1;
        const int32_t* sm = e->smap.data() + e->gather_smap_off[q];
        const int32_t* pm = p.pmap.data() + p.pmap_off[ch];
        for (int32_t v = 0; v < S1; ++v)
          if ((v && sm[v] <= sm[v - 1]) || sm[S1 + v] < 0 || sm[S1 + v] >= S1 || pm[sm[S1 + v]] != sm[v]) return fail("sorted separator map");
      }
      k = k2;
    }
    for (int32_t c = 0; c < p.n_cliques; ++c) {
      const int32_t par = p.parent[c];
      if (par < 0 || is_small_front(p.F[par], dim_of(par)) || e->is_point[par]) { if (gathered[c]) return fail("gathered child of a small parent"); continue; }
      // who adds c into its large parent: the pair lists (point), its own Schur step (fuse_up), or the parent's gather
      const bool here = (e->owner[c] == e->rank && e->owner[par] == e->rank) ||
                        (e->owner[par] < 0 && (e->owner[c] == e->rank || e->owner[c] < 0));
      const int want = (here && !e->is_point[c] && !e->fuse_up[c]) ? 1 : 0;
      if (gathered[c] != want) return fail("clique " + std::to_string(c) + " gathered " + std::to_string(gathered[c]) + " times, expected " + std::to_string(want));
    }
  }
  if (e->point_pairs && (pair_items_seen != (int64_t)e->pair_items.size() || diag_items_seen != (int64_t)e->diag_items.size() / 3)) return fail("segments do not cover the item lists");
  cnt[4] = e->rinv_blocks;
  if (out8) std::memcpy(out8, cnt, sizeof(cnt));
  return GTB_OK;
}
gtb_status gtb_plan_stats(const gtb_engine* e, double* out) {
  if (!e || !out) return GTB_ERR_INVALID;
  if (!e->has_plan) { e->err = "no plan"; return GTB_ERR_INVALID; }
  // Algorithmic work (flops, bytes) is that of the REFERENCE's cliques: the zeros that relaxed amalgamation pads into the
  // numeric fronts are overhead, not work.  Levels, front sizes and the arena are those of the numeric plan (what runs).
  const Plan& p = e->plan;
  const Plan& r = e->rplan;
  out[0] = p.n_levels; out[1] = r.max_front; out[2] = r.flops; out[3] = (double)(e->arena_doubles ? e->arena_doubles : p.arena_doubles);
  out[4] = (double)r.n_small; out[5] = (double)r.n_large; out[6] = r.flops_large; out[7] = r.assembly_bytes;
  double thin_bytes = 0, point_bytes = 0, dense_bytes = 0;
  out[8] = r.flops_syrk_frontal; out[9] = r.flops_syrk_schur; out[10] = r.bytes_backsolve;
  {
    // small fronts also read the whitened rows of their own factors
    static const int kMW[7] = {21, 78, 26, 12, 42, 90, 12};
    double b = p.bytes_small;
    for (int32_t c = 0; c < p.n_cliques; ++c)
      if (is_small_front(p.F[c], p.F[c] + p.S[c] + 1))
        for (int64_t q = p.fac_off[c]; q < p.fac_off[c + 1]; ++q) b += 8.0 * kMW[e->g_bucket[p.slots[q].factor]];
    out[11] = b;
    // share of it in thin fronts (thin_front_kernel): own front written, own factor rows and child Schur blocks read
    double bt = 0;
    for (int32_t c = 0; c < p.n_cliques; ++c) {
      const int dim = p.F[c] + p.S[c] + 1;
      if (!e->is_point.empty() && e->is_point[c]) {
        // point clique: its factors' whitened rows read, the three conditional rows (32-byte columns) written
        for (int64_t q = p.fac_off[c]; q < p.fac_off[c + 1]; ++q) point_bytes += 8.0 * kMW[e->g_bucket[p.slots[q].factor]];
        point_bytes += 32.0 * dim;
        continue;
      }
      if (!is_small_front(p.F[c], dim)) continue;
      double& acc = is_thin_front(p.F[c], dim) ? bt : dense_bytes;
      acc += 8.0 * 0.5 * dim * (dim + 1);
      for (int64_t q = p.fac_off[c]; q < p.fac_off[c + 1]; ++q) acc += 8.0 * kMW[e->g_bucket[p.slots[q].factor]];
      for (int64_t q = p.child_off[c]; q < p.child_off[c + 1]; ++q) {
        const double s1 = p.S[p.children[q]] + 1;
        acc += 8.0 * 0.5 * s1 * (s1 + 1);
      }
    }
    thin_bytes = bt;
  }
  // linearize: variables read once + per-factor indices, measurement, noise + whitened [A|b] written once
  {
    double b = 0;
    for (int t = 0; t < 4; ++t) b += 8.0 * kStateDev[t] * e->vars[t].n;
    for (int k = 0; k < 7; ++k)
      b += (double)e->fac[k].n * (4.0 * kNKeys[k] + 8.0 * kZDev[k] + (e->fac[k].any_noise ? 8.0 * kM[k] * kM[k] + 1 : 0) +
                                  8.0 * kM[k] * kW[k]);
    out[12] = b;
  }
  out[13] = thin_bytes; out[14] = point_bytes; out[15] = dense_bytes;
  return GTB_OK;
}

gtb_status gtb_linearize(gtb_engine* e) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e) linearize(e); GTB_CATCH(e)
}
gtb_status gtb_get_jacobians(gtb_engine* e, int bucket, double* out) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(bucket >= 0 && bucket < 7 && e->linearized, "gtb_get_jacobians needs gtb_linearize first");
  const size_t n = (size_t)e->fac[bucket].n * kM[bucket] * kW[bucket];
  if (n) CK(cudaMemcpyAsync(out, e->d_Ab[bucket].p, n * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  GTB_CATCH(e)
}
gtb_status gtb_hessian_diagonal(gtb_engine* e, double* out) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  hessian_diagonal(e);
  if (out) CK(cudaMemcpyAsync(out, e->d_hdiag.p, e->n_scalars * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  GTB_CATCH(e)
}
gtb_status gtb_factorize(gtb_engine* e, double lambda, int diag, double lo, double hi, uint64_t* failed_key) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  const int32_t fail = factorize(e, lambda, diag, lo, hi);
  if (fail >= 0) {
    const uint64_t key = failed_key_of(e, fail);
    if (failed_key) *failed_key = key;
    char buf[128];
    snprintf(buf, sizeof buf, "indeterminant linear system detected while working near variable %llu", (unsigned long long)key);
    throw StatusError(GTB_ERR_INDETERMINATE, buf);
  }
  GTB_CATCH(e)
}
gtb_status gtb_backsolve(gtb_engine* e) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e) backsolve(e); GTB_CATCH(e)
}
gtb_status gtb_trial(gtb_engine* e, double lambda, int diag, double lo, double hi, double* out3, uint64_t* failed_key) {
  if (!e || !out3) return GTB_ERR_INVALID;
  GTB_TRY(e)
  const TrialResult r = trial(e, lambda, diag, lo, hi);
  out3[0] = r.lin0; out3[1] = r.lin1; out3[2] = r.err;
  if (r.fail >= 0) {
    const uint64_t key = failed_key_of(e, r.fail);
    if (failed_key) *failed_key = key;
    char buf[128];
    snprintf(buf, sizeof buf, "indeterminant linear system detected while working near variable %llu", (unsigned long long)key);
    throw StatusError(GTB_ERR_INDETERMINATE, buf);
  }
  GTB_CATCH(e)
}
gtb_status gtb_get_delta(gtb_engine* e, double* out) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(e->solved && out, "gtb_get_delta needs gtb_backsolve first");
  CK(cudaMemcpyAsync(out, e->d_delta.p, e->n_scalars * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  GTB_CATCH(e)
}
gtb_status gtb_get_conditionals(gtb_engine* e, double* out) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(e->factorized && out, "gtb_get_conditionals needs a successful gtb_factorize first");
  const Plan& p = e->plan;
  const Plan& r = e->rplan;
  // parity dump: not a hot path; stage the whole arena through the host
  // (with several ranks only the cliques this rank factored hold conditionals; the others read back as assembled)
  std::vector<double> arena((size_t)e->arena_doubles);
  CK(cudaMemcpyAsync(arena.data(), e->d_arena.p, arena.size() * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  // A reference clique is a band of rows of the numeric front that holds it: its frontal columns follow its first pivot,
  // its separator scalars are found through their global scalar index (columns the clique does not see are structural
  // zeros of the merged front and are not part of its conditional).
  std::vector<int32_t> col_of((size_t)e->n_scalars, -1);
  for (int32_t nc = 0; nc < p.n_cliques; ++nc) {
    const int Fn = p.F[nc], Nn = Fn + p.S[nc], ld = e->ld[nc];
    const double* A = arena.data() + e->front_off[nc];
    for (int k = 0; k < Nn; ++k) col_of[p.gidx[p.gidx_off[nc] + k]] = k;
    for (int64_t m = p.orig_off[nc]; m < p.orig_off[nc + 1]; ++m) {
      const int32_t c = p.orig_clique[m];
      const int F = r.F[c], N = F + r.S[c], dim = N + 1, r0 = p.num_foff[c];
      double* o = out + r.cond_off[c];
      for (int j = 0; j < dim; ++j) {
        const int jn = j < N ? col_of[r.gidx[r.gidx_off[c] + j]] : Nn;
        for (int i = 0; i < F; ++i) o[(int64_t)i * dim + j] = (j >= i) ? A[(r0 + i) + (int64_t)jn * ld] : 0.0;
      }
    }
  }
  GTB_CATCH(e)
}
gtb_status gtb_marginals_compute(gtb_engine* e) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e) marginals_compute(e); GTB_CATCH(e)
}
gtb_status gtb_marginal_covariance(gtb_engine* e, int64_t n, const uint64_t* keys, double* out) {
  if (!e || n < 0 || (n > 0 && (!keys || !out))) return GTB_ERR_INVALID;
  GTB_TRY(e) marginal_covariance(e, n, keys, out); GTB_CATCH(e)
}
gtb_status gtb_linear_errors(gtb_engine* e, double* e0, double* e1) {
  if (!e || !e0 || !e1) return GTB_ERR_INVALID;
  GTB_TRY(e) linear_errors(e, e0, e1); GTB_CATCH(e)
}
gtb_status gtb_retract(gtb_engine* e) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e) retract(e); CK(cudaStreamSynchronize(e->stream)); GTB_CATCH(e)
}
gtb_status gtb_error(gtb_engine* e, int which, double* error) {
  if (!e || !error) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need(!which || e->retracted, "gtb_error(which = 1) needs gtb_retract first: the scratch copy holds no trial values");
  *error = nonlinear_error(e, which);
  GTB_CATCH(e)
}
gtb_status gtb_accept(gtb_engine* e) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(e->retracted, "gtb_accept needs gtb_retract since the last gtb_linearize / gtb_accept / gtb_set_variables");
  e->cur ^= 1;
  e->linearized = e->factorized = e->solved = e->have_hdiag = e->retracted = false;
  e->lm_have_error = false;
  GTB_CATCH(e)
}
gtb_status gtb_relinearize_marked(gtb_engine* e, double threshold, uint8_t* marks, int64_t* n_marked) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(e->retracted, "gtb_relinearize_marked needs gtb_retract first: the marked variables take their retracted values");
  need(e->world == 1, "gtb_relinearize_marked: single rank only");
  int64_t off[5] = {0, 0, 0, 0, 0};
  for (int t = 0; t < 4; ++t) off[t + 1] = off[t] + e->vars[t].n;
  DevBuf<uint8_t> d_marks;
  DevBuf<int> d_count;
  d_marks.alloc((size_t)off[4] + 1);
  d_count.alloc(1);
  CK(cudaMemsetAsync(d_count.p, 0, sizeof(int), e->stream));
  launch_relinearize<0>(e, threshold, d_marks.p + off[0], d_count.p);
  launch_relinearize<1>(e, threshold, d_marks.p + off[1], d_count.p);
  launch_relinearize<2>(e, threshold, d_marks.p + off[2], d_count.p);
  launch_relinearize<3>(e, threshold, d_marks.p + off[3], d_count.p);
  int count = 0;
  CK(cudaMemcpyAsync(&count, d_count.p, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
  if (marks && off[4]) CK(cudaMemcpyAsync(marks, d_marks.p, (size_t)off[4], cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  if (n_marked) *n_marked = count;
  // the current values moved: everything computed from them is stale
  if (count) e->linearized = e->factorized = e->solved = e->have_hdiag = e->retracted = false;
  e->lm_have_error = false;
  GTB_CATCH(e)
}
gtb_status gtb_get_variables(gtb_engine* e, int vtype, int which, double* state) {
  if (!e) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(vtype >= 0 && vtype < 4 && state, "bad arguments");
  need(!which || e->retracted, "gtb_get_variables(which = 1) needs gtb_retract first: the scratch copy holds no trial values");
  const int w = which ? (e->cur ^ 1) : e->cur;
  const size_t n = (size_t)e->vars[vtype].n * kStateDev[vtype];
  if (n) CK(cudaMemcpyAsync(state, e->d_state[w][vtype].p, n * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  GTB_CATCH(e)
}

void gtb_lm_params_default(gtb_lm_params* p) {
  p->max_iterations = 100; p->relative_error_tol = 1e-5; p->absolute_error_tol = 1e-5; p->error_tol = 0.0;
  p->lambda_initial = 1e-5; p->lambda_factor = 10.0; p->lambda_upper_bound = 1e5; p->lambda_lower_bound = 0.0;
  p->min_model_fidelity = 1e-3; p->diagonal_damping = 0; p->use_fixed_lambda_factor = 1;
  p->min_diagonal = 1e-6; p->max_diagonal = 1e32;
}
void gtb_lm_params_ceres(gtb_lm_params* p) {
  gtb_lm_params_default(p);
  p->max_iterations = 50; p->absolute_error_tol = 0; p->relative_error_tol = 1e-6;
  p->lambda_upper_bound = 1e32; p->lambda_lower_bound = 1e-16; p->lambda_initial = 1e-4; p->lambda_factor = 2.0;
  p->min_model_fidelity = 1e-3; p->diagonal_damping = 1; p->use_fixed_lambda_factor = 0;
}

gtb_status gtb_lm_optimize(gtb_engine* e, const gtb_lm_params* p, gtb_lm_result* result, double* trace,
                           int64_t trace_rows) {
  if (!e || !p || !result) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(e->has_plan, "gtb_set_ordering must be called first");
  const gtb_lm_params& P = *p;
  LMState S;
  S.error = nonlinear_error(e, 0);
  S.lambda = P.lambda_initial; S.factor = P.lambda_factor; S.iterations = 0; S.inner = 0;
  Trace tr{trace, trace_rows};
  result->initial_error = S.error;
  // NonlinearOptimizer::defaultOptimize
  double currentError = S.error;
  if (!(currentError <= P.error_tol) && S.iterations < P.max_iterations) {
    double newError = currentError;
    do {
      currentError = newError;
      lm_iterate(e, P, S, &tr);
      newError = S.error;
    } while (S.iterations < P.max_iterations &&
             !check_convergence(P.relative_error_tol, P.absolute_error_tol, P.error_tol, currentError, newError) &&
             std::isfinite(currentError));
  }
  result->final_error = S.error; result->final_lambda = S.lambda;
  result->iterations = S.iterations; result->inner_iterations = S.inner;
  result->log_rows = (int32_t)std::min<int64_t>(tr.n, trace ? trace_rows : 0);
  e->lm_error = S.error; e->lm_have_error = true;
  GTB_CATCH(e)
}

gtb_status gtb_lm_iterate(gtb_engine* e, const gtb_lm_params* p, double* lambda_io, double* factor_io,
                          double* error_io, int32_t* inner) {
  if (!e || !p || !lambda_io || !factor_io || !error_io) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(e->has_plan, "gtb_set_ordering must be called first");
  LMState S;
  S.error = e->lm_have_error ? e->lm_error : nonlinear_error(e, 0);
  S.lambda = *lambda_io; S.factor = *factor_io; S.iterations = 0; S.inner = 1;  // inner != 0: no initial trace row
  lm_iterate(e, *p, S, nullptr);
  *lambda_io = S.lambda; *factor_io = S.factor; *error_io = S.error;
  if (inner) *inner = S.inner - 1;
  e->lm_last_accepted = S.iterations > 0 ? 1 : 0;
  e->lm_error = S.error; e->lm_have_error = true;
  GTB_CATCH(e)
}

int gtb_lm_last_accepted(const gtb_engine* e) { return e ? e->lm_last_accepted : 0; }

gtb_status gtb_gn_optimize(gtb_engine* e, int32_t max_iterations, double rel, double abs_, gtb_lm_result* result) {
  if (!e || !result) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  need(e->has_plan, "gtb_set_ordering must be called first");
  double err = nonlinear_error(e, 0);
  result->initial_error = err;
  int it = 0;
  double currentError = err;
  if (!(currentError <= 0.0) && it < max_iterations) {
    double newError = currentError;
    do {
      currentError = newError;
      linearize(e);
      const int32_t fail = factorize(e, 0.0, 0, 0, 0);
      if (fail >= 0) {
        char buf[128];
        snprintf(buf, sizeof buf, "indeterminant linear system detected while working near variable %llu",
                 (unsigned long long)failed_key_of(e, fail));
        throw StatusError(GTB_ERR_INDETERMINATE, buf);
      }
      backsolve(e);
      retract(e);
      newError = nonlinear_error(e, 1);
      e->cur ^= 1;
      e->linearized = e->factorized = e->solved = e->have_hdiag = e->retracted = false;
      ++it;
    } while (it < max_iterations && !check_convergence(rel, abs_, 0.0, currentError, newError) && std::isfinite(currentError));
    err = newError;
  }
  result->final_error = err; result->final_lambda = 0; result->iterations = it; result->inner_iterations = 0;
  result->log_rows = 0;
  e->lm_error = err; e->lm_have_error = true;
  GTB_CATCH(e)
}

int64_t gtb_kernel_launches(const gtb_engine* e) { return e ? e->launches : 0; }
gtb_status gtb_stage_ms(gtb_engine* e, double* out) {
  if (!e || !out) return GTB_ERR_INVALID;
  for (int i = 0; i < 8; ++i) out[i] = e->stage_ms[i];
  return GTB_OK;
}
gtb_status gtb_set_stage_timing(gtb_engine* e, int on) {
  if (!e) return GTB_ERR_INVALID;
  e->time_stages = on != 0;
  return GTB_OK;
}
gtb_status gtb_profile(gtb_engine* e, int on) {
  if (!e) return GTB_ERR_INVALID;
  e->profiling = on != 0;
  e->prof_used = 0;
  e->prof_class.clear();
  return GTB_OK;
}
int gtb_profile_classes(void) { return K_NUM; }
const char* gtb_profile_class_name(int k) { return (k >= 0 && k < K_NUM) ? kKernelNames[k] : ""; }
gtb_status gtb_profile_read(gtb_engine* e, double* total_ms, int64_t* launches) {
  if (!e || !total_ms || !launches) return GTB_ERR_INVALID;
  GTB_TRY(e)
  need_device(e);
  CK(cudaStreamSynchronize(e->stream));
  for (int k = 0; k < K_NUM; ++k) { total_ms[k] = 0; launches[k] = 0; }
  for (size_t i = 0; i < e->prof_class.size(); ++i) {
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e->prof_events[2 * i], e->prof_events[2 * i + 1]));
    total_ms[e->prof_class[i]] += ms;
    launches[e->prof_class[i]]++;
  }
  e->prof_used = 0;
  e->prof_class.clear();
  GTB_CATCH(e)
}
int64_t gtb_profile_read_raw(gtb_engine* e, int32_t* klass, float* ms, int64_t cap) {
  if (!e || !e->has_device) return -1;
  EngineDeviceGuard guard(e);
  cudaStreamSynchronize(e->stream);
  const int64_t n = std::min<int64_t>((int64_t)e->prof_class.size(), cap);
  for (int64_t i = 0; i < n; ++i) {
    klass[i] = e->prof_class[i];
    cudaEventElapsedTime(&ms[i], e->prof_events[2 * i], e->prof_events[2 * i + 1]);
  }
  e->prof_used = 0;
  e->prof_class.clear();
  return n;
}
void gtb_debug_read(gtb_engine* e, long long* out8) {
  if (!e || !e->has_device) return;
  EngineDeviceGuard guard(e);
  cudaStreamSynchronize(e->stream);
  cudaMemcpy(out8, e->d_dbg.p, 16 * sizeof(long long), cudaMemcpyDeviceToHost);
}
void* gtb_stream(gtb_engine* e) { return e ? (void*)e->stream : nullptr; }

}  // extern "C"

// ------------------------------------------------------------------------------------------------------
// Allocation-free entry point (include/gtsam_b200.h: gtb_bucket_error_noalloc).  No counterpart in the reference:
// added for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
// loops, harnesses that forbid it).  It is a variant of error_kernel<B> + sum_partials_kernel (nonlinear_error above)
// whose per-CTA partial sums live in a file-scope __device__ array instead of the engine's cudaMalloc'ed d_partials.
//
// The array's NAME is only a host-side stand-in in host code, not a device address: its address is taken once per device,
// with that device current, through cudaGetSymbolAddress, and that pointer is what the kernels get.  Every GPU that
// loads the library holds the array (256 KB) until exit.  All calls share it and it keeps the last call's contents:
// nothing in it has to start at zero (every partial that the final sum reads is written by this call's first kernel), but
// two calls that overlap on the GPU (different streams, no dependency) race on it.
// ------------------------------------------------------------------------------------------------------
namespace gtb {

constexpr int64_t kNoAllocMaxFactors = GTB_NOALLOC_MAX_FACTORS;
constexpr int kNoAllocPartials = (int)((kNoAllocMaxFactors + kFactorThreads - 1) / kFactorThreads);   // 32,768 doubles = 256 KB
static __device__ __align__(256) double g_noalloc_partials[kNoAllocPartials];

constexpr int kNoAllocMaxDevices = 64;
static double* g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};   // device address of the array, per device ordinal
static thread_local std::string g_noalloc_err;

// Restores the caller's current device on scope exit.
struct DeviceRestore {
  int prev = -1;
  bool switched = false;
  cudaError_t enter(int device) {
    cudaError_t st = cudaGetDevice(&prev);
    if (st != cudaSuccess) return st;
    if (prev != device) {
      st = cudaSetDevice(device);
      switched = st == cudaSuccess;
    }
    return st;
  }
  ~DeviceRestore() { if (switched) cudaSetDevice(prev); }
};

// Set-up, once per device, with that device current: the device address of the scratch array.
static gtb_status noalloc_setup(int device, double** out) {
  if (device < 0 || device >= kNoAllocMaxDevices) {
    g_noalloc_err = "device ordinal out of range";
    return GTB_ERR_INVALID;
  }
  if (!g_noalloc_ptr[device]) {
    double* p = nullptr;
    const cudaError_t st = cudaGetSymbolAddress((void**)&p, g_noalloc_partials);
    if (st != cudaSuccess || !p) {
      g_noalloc_err = std::string("cudaGetSymbolAddress(g_noalloc_partials): ") + cudaGetErrorString(st);
      return st == cudaErrorNoDevice || st == cudaErrorInsufficientDriver ? GTB_ERR_NO_DEVICE : GTB_ERR_CUDA;
    }
    g_noalloc_ptr[device] = p;
  }
  if (out) *out = g_noalloc_ptr[device];
  return GTB_OK;
}

template <int B>
static cudaError_t noalloc_launch(const BucketDev& bk, const double* s0, const double* s1, double* partials, double* out,
                                  cudaStream_t st) {
  const int grid = (bk.n + kFactorThreads - 1) / kFactorThreads;
  error_kernel<B><<<grid, kFactorThreads, 0, st>>>(bk, s0, s1, partials);
  sum_partials_kernel<<<1, 256, 0, st>>>(partials, grid, out);
  return cudaGetLastError();
}

}  // namespace gtb

extern "C" {

const char* gtb_noalloc_last_error(void) { return g_noalloc_err.c_str(); }

gtb_status gtb_noalloc_setup(int device) {
  DeviceRestore guard;
  if (guard.enter(device) != cudaSuccess) {
    g_noalloc_err = "cannot make the device current";
    return GTB_ERR_NO_DEVICE;
  }
  return noalloc_setup(device, nullptr);
}

gtb_status gtb_bucket_error_noalloc(int device, int bucket, int64_t n, const double* state0, const double* state1,
                                    const int32_t* row0, const int32_t* row1, const double* z, const int8_t* noise_kind,
                                    const double* noise, double* error_out, void* stream) {
  if (bucket < 0 || bucket >= GTB_NUM_BUCKETS) { g_noalloc_err = "bad factor bucket"; return GTB_ERR_INVALID; }
  if (n < 0) { g_noalloc_err = "negative factor count"; return GTB_ERR_INVALID; }
  if (n > kNoAllocMaxFactors) {
    char buf[160];
    snprintf(buf, sizeof buf, "gtb_bucket_error_noalloc: %lld factors exceed the fixed scratch (GTB_NOALLOC_MAX_FACTORS = %lld)",
             (long long)n, (long long)kNoAllocMaxFactors);
    g_noalloc_err = buf;
    return GTB_ERR_INVALID;
  }
  if (!error_out || (n > 0 && (!state0 || !row0 || !z)) || (n > 0 && kNKeys[bucket] == 2 && (!state1 || !row1)) ||
      ((noise_kind == nullptr) != (noise == nullptr))) {
    g_noalloc_err = "null argument";
    return GTB_ERR_INVALID;
  }
  DeviceRestore guard;
  if (guard.enter(device) != cudaSuccess) {
    g_noalloc_err = "cannot make the device current";
    return GTB_ERR_NO_DEVICE;
  }
  double* partials = nullptr;
  const gtb_status su = noalloc_setup(device, &partials);
  if (su != GTB_OK) return su;
  cudaStream_t st = (cudaStream_t)stream;
  BucketDev bk;
  std::memset(&bk, 0, sizeof bk);
  bk.n = (int32_t)n;
  bk.any_noise = noise_kind ? 1 : 0;
  bk.row0 = row0; bk.row1 = row1; bk.z = z; bk.nk = noise_kind; bk.noise = noise;
  cudaError_t err = cudaSuccess;
  if (n == 0) {
    err = cudaMemsetAsync(error_out, 0, sizeof(double), st);
  } else {
    switch (bucket) {
      case 0: err = noalloc_launch<0>(bk, state0, state1, partials, error_out, st); break;
      case 1: err = noalloc_launch<1>(bk, state0, state1, partials, error_out, st); break;
      case 2: err = noalloc_launch<2>(bk, state0, state1, partials, error_out, st); break;
      case 3: err = noalloc_launch<3>(bk, state0, state1, partials, error_out, st); break;
      case 4: err = noalloc_launch<4>(bk, state0, state1, partials, error_out, st); break;
      case 5: err = noalloc_launch<5>(bk, state0, state1, partials, error_out, st); break;
      default: err = noalloc_launch<6>(bk, state0, state1, partials, error_out, st); break;
    }
  }
  if (err != cudaSuccess) {
    g_noalloc_err = std::string("gtb_bucket_error_noalloc launch: ") + cudaGetErrorString(err);
    return GTB_ERR_CUDA;
  }
  return GTB_OK;
}

}  // extern "C"
