// isam2.cpp -- gtsam::ISAM2::update semantics on top of the batch engine, written against the public C ABI only
// (include/gtsam_b200.h: gtb_isam2_*).  What the reference does per update (nonlinear/ISAM2.cpp:401-480):
//   addVariables(newTheta)                      theta := theta U theta_new, delta_new = 0           :432
//   updateDelta()                               delta of the previous update, complete (wildfire 0) :435-436
//   pushBackFactors                             factors := factors U new                            :439-441
//   gatherRelinearizeKeys + retractMasked       theta_J := theta_J (+) delta_J, |delta_j| >= beta   :456-468
//   linearizeNewFactors, recalculate            re-eliminate what the marked keys touch             :472-478
//   calculateEstimate                           theta (+) delta                                     :725-733
// Here the linear system of an update is rebuilt and solved as a whole by a fresh gtb_engine (linearised at theta); the
// relinearisation test and the masked retract run on the device of the previous update's engine, which still holds theta,
// delta and theta (+) delta.  See the header for what this keeps of the reference and what it does not.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../../include/gtsam_b200.h"
#include "layout.h"

using gtb::kM;
using gtb::kNKeys;
using gtb::kStateDev;
using gtb::kStateIn;
using gtb::kZIn;

namespace {

struct VarTable {
  std::vector<uint64_t> keys;
  std::vector<double> state;   // rows in the caller's layout (kStateIn): this is theta
};
struct FactorTable {
  std::vector<uint64_t> keys;
  std::vector<double> z, noise;
  std::vector<int32_t> kind;
  int64_t n = 0;
};

double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

struct gtb_isam2 {
  int device = 0;
  double threshold = 0.1;    // ISAM2Params::relinearizeThreshold (nonlinear/ISAM2Params.h: default 0.1)
  int32_t skip = 10;         // ISAM2Params::relinearizeSkip (default 10)
  int64_t update_count = 0;
  VarTable vars[GTB_NUM_VTYPES], new_vars[GTB_NUM_VTYPES];
  FactorTable factors[GTB_NUM_BUCKETS], new_factors[GTB_NUM_BUCKETS];
  std::unordered_set<uint64_t> known;    // every key of vars and new_vars
  std::vector<uint64_t> insertion;       // committed keys in the order they were added
  std::vector<uint64_t> ordering;        // caller's elimination order (may name keys that do not exist yet: skipped)
  bool have_ordering = false;
  // Without a caller's order: the library's nested dissection, recomputed every reorder_every updates; in between the last one
  // is kept and the variables added since are eliminated last, in the order they came (the periodic reordering of ISAM,
  // nonlinear/NonlinearISAM.h:30-36; here it only trades ordering time against a slightly taller tree, the estimate is
  // order-independent).
  int32_t reorder_every = 8;
  std::vector<uint64_t> auto_order;
  int64_t auto_order_update = 0;
  gtb_engine* eng = nullptr;
  bool have_delta = false;               // eng holds theta, delta and theta (+) delta of the last update
  std::string err;
};

namespace {

gtb_status fail(gtb_isam2* s, gtb_status st, const std::string& msg) {
  s->err = msg;
  return st;
}
gtb_status pass(gtb_isam2* s, gtb_status st, const char* what) {
  if (st != GTB_OK) s->err = std::string(what) + ": " + (s->eng ? gtb_last_error(s->eng) : "no engine");
  return st;
}

// theta_J := theta_J (+) delta_J for |delta_j|_inf >= threshold, on the device of the last update's engine; the moved rows
// are read back into the host copy of theta.
gtb_status relinearize(gtb_isam2* s, int64_t* n_marked) {
  int64_t total = 0;
  for (int t = 0; t < GTB_NUM_VTYPES; ++t) total += (int64_t)s->vars[t].keys.size();
  std::vector<uint8_t> marks((size_t)total + 1);
  gtb_status st = pass(s, gtb_relinearize_marked(s->eng, s->threshold, marks.data(), n_marked), "gtb_relinearize_marked");
  if (st != GTB_OK || *n_marked == 0) return st;
  int64_t off = 0;
  std::vector<double> rows;
  for (int t = 0; t < GTB_NUM_VTYPES; ++t) {
    const int64_t n = (int64_t)s->vars[t].keys.size();
    if (n && std::any_of(marks.begin() + off, marks.begin() + off + n, [](uint8_t m) { return m != 0; })) {
      rows.resize((size_t)n * kStateDev[t]);
      st = pass(s, gtb_get_variables(s->eng, t, 0, rows.data()), "gtb_get_variables");
      if (st != GTB_OK) return st;
      for (int64_t i = 0; i < n; ++i) {
        if (!marks[off + i]) continue;
        const double* r = rows.data() + i * kStateDev[t];
        double* th = s->vars[t].state.data() + i * kStateIn[t];
        if (t == GTB_POSE2) {   // device rows are (cos, sin, x, y): Rot2::theta() = atan2(s, c), geometry/Rot2.h
          th[0] = r[2]; th[1] = r[3]; th[2] = std::atan2(r[1], r[0]);
        } else {
          std::memcpy(th, r, sizeof(double) * kStateIn[t]);
        }
      }
    }
    off += n;
  }
  return GTB_OK;
}

void commit_staged(gtb_isam2* s) {
  for (int t = 0; t < GTB_NUM_VTYPES; ++t) {
    VarTable& nv = s->new_vars[t];
    s->insertion.insert(s->insertion.end(), nv.keys.begin(), nv.keys.end());
    s->vars[t].keys.insert(s->vars[t].keys.end(), nv.keys.begin(), nv.keys.end());
    s->vars[t].state.insert(s->vars[t].state.end(), nv.state.begin(), nv.state.end());
    nv = VarTable();
  }
  for (int b = 0; b < GTB_NUM_BUCKETS; ++b) {
    FactorTable& nf = s->new_factors[b];
    FactorTable& f = s->factors[b];
    f.keys.insert(f.keys.end(), nf.keys.begin(), nf.keys.end());
    f.z.insert(f.z.end(), nf.z.begin(), nf.z.end());
    f.noise.insert(f.noise.end(), nf.noise.begin(), nf.noise.end());
    f.kind.insert(f.kind.end(), nf.kind.begin(), nf.kind.end());
    f.n += nf.n;
    nf = FactorTable();
  }
}

// GTB_ISAM2_TIMES=1: mean host time of the stages of rebuild() every 500 updates (stderr)
double g_lap[6];
long g_laps;
void lap(int k, double& t) {
  const double n = now_ms();
  g_lap[k] += n - t;
  t = n;
}

// A fresh engine holding the whole problem at theta, planned with the current ordering.
gtb_status rebuild(gtb_isam2* s) {
  static const bool times = getenv("GTB_ISAM2_TIMES") != nullptr;
  double t = now_ms();
  if (times && ++g_laps % 500 == 0) {
    fprintf(stderr, "isam2 rebuild, mean ms over %ld updates: destroy %.3f create %.3f tables %.3f finalize %.3f ordering + plan %.3f\n", g_laps,
            g_lap[0] / g_laps, g_lap[1] / g_laps, g_lap[2] / g_laps, g_lap[3] / g_laps, g_lap[4] / g_laps);
  }
  if (s->eng) gtb_destroy(s->eng);
  s->eng = nullptr;
  s->have_delta = false;
  lap(0, t);
  gtb_status st = gtb_create(s->device, &s->eng);
  lap(1, t);
  if (st != GTB_OK) return fail(s, st, "gtb_create failed (no CUDA device?)");
  for (int vt = 0; vt < GTB_NUM_VTYPES; ++vt) {
    const VarTable& v = s->vars[vt];
    if (v.keys.empty()) continue;
    st = pass(s, gtb_set_variables(s->eng, vt, (int64_t)v.keys.size(), v.keys.data(), v.state.data()), "gtb_set_variables");
    if (st != GTB_OK) return st;
  }
  for (int b = 0; b < GTB_NUM_BUCKETS; ++b) {
    const FactorTable& f = s->factors[b];
    if (!f.n) continue;
    st = pass(s, gtb_set_factors(s->eng, b, f.n, f.keys.data(), f.z.data(), f.kind.data(), f.noise.data()), "gtb_set_factors");
    if (st != GTB_OK) return st;
  }
  lap(2, t);
  st = pass(s, gtb_finalize_problem(s->eng), "gtb_finalize_problem");
  lap(3, t);
  if (st != GTB_OK) return st;
  struct PlanLap { double& t; ~PlanLap() { lap(4, t); } } plan_lap{t};
  if (!s->have_ordering && (s->auto_order.empty() || s->update_count - s->auto_order_update >= s->reorder_every)) {
    st = pass(s, gtb_set_ordering_builtin(s->eng, 2), "gtb_set_ordering_builtin");
    if (st != GTB_OK) return st;
    // keep an equivalent order for the next updates: the frontal keys of the cliques in post-order
    int64_t nc = 0, nk = 0, a = 0, b = 0;
    gtb_plan_sizes(s->eng, &nc, &nk, &a, &b);
    std::vector<int64_t> nfrontal((size_t)nc), nkeys((size_t)nc);
    std::vector<uint64_t> keys((size_t)nk);
    gtb_plan_get(s->eng, nfrontal.data(), nkeys.data(), keys.data(), nullptr);
    s->auto_order.clear();
    size_t off = 0;
    for (int64_t c = 0; c < nc; ++c) {
      s->auto_order.insert(s->auto_order.end(), keys.begin() + off, keys.begin() + off + nfrontal[c]);
      off += (size_t)nkeys[c];
    }
    s->auto_order_update = s->update_count;
    return GTB_OK;
  }
  // the caller's order (or the last computed one) for the keys it names, then the rest in the order they were added (the
  // newest variables last: they are the ones the next update touches, cf. the constrained ordering of ISAM2.cpp:301-340)
  const std::vector<uint64_t>& base = s->have_ordering ? s->ordering : s->auto_order;
  std::vector<uint64_t> order;
  order.reserve(s->insertion.size());
  std::unordered_set<uint64_t> placed;
  for (uint64_t k : base)
    if (s->known.count(k) && placed.insert(k).second) order.push_back(k);
  for (uint64_t k : s->insertion)
    if (placed.insert(k).second) order.push_back(k);
  return pass(s, gtb_set_ordering(s->eng, (int64_t)order.size(), order.data()), "gtb_set_ordering");
}

}  // namespace

extern "C" {

gtb_status gtb_isam2_create(int device, double relinearize_threshold, int32_t relinearize_skip, gtb_isam2** out) {
  if (!out || !(relinearize_threshold >= 0.0) || relinearize_skip < 1) return GTB_ERR_INVALID;
  gtb_isam2* s = new gtb_isam2();
  s->device = device;
  s->threshold = relinearize_threshold;
  s->skip = relinearize_skip;
  if (const char* v = getenv("GTB_ISAM2_REORDER_EVERY")) s->reorder_every = std::max(1, atoi(v));
  *out = s;
  return GTB_OK;
}

void gtb_isam2_destroy(gtb_isam2* s) {
  if (!s) return;
  if (s->eng) gtb_destroy(s->eng);
  delete s;
}

const char* gtb_isam2_last_error(const gtb_isam2* s) { return s ? s->err.c_str() : "null ISAM2 handle"; }

gtb_status gtb_isam2_add_variables(gtb_isam2* s, int vtype, int64_t n, const uint64_t* keys, const double* state) {
  if (!s) return GTB_ERR_INVALID;
  if (vtype < 0 || vtype >= GTB_NUM_VTYPES || n < 0 || (n && (!keys || !state))) return fail(s, GTB_ERR_INVALID, "bad arguments");
  for (int64_t i = 0; i < n; ++i)
    if (s->known.count(keys[i])) return fail(s, GTB_ERR_INVALID, "key already exists (ValuesKeyAlreadyExists): " + std::to_string(keys[i]));
  // (a key repeated inside this call is caught by the engine at the next update)
  VarTable& v = s->new_vars[vtype];
  v.keys.insert(v.keys.end(), keys, keys + n);
  v.state.insert(v.state.end(), state, state + n * kStateIn[vtype]);
  s->known.insert(keys, keys + n);
  return GTB_OK;
}

gtb_status gtb_isam2_add_factors(gtb_isam2* s, int bucket, int64_t n, const uint64_t* keys, const double* z,
                                 const int32_t* noise_kind, const double* noise) {
  if (!s) return GTB_ERR_INVALID;
  if (bucket < 0 || bucket >= GTB_NUM_BUCKETS || n < 0 || (n && (!keys || !z)) || (noise_kind && !noise))
    return fail(s, GTB_ERR_INVALID, "bad arguments");
  FactorTable& f = s->new_factors[bucket];
  const int m = kM[bucket];
  f.keys.insert(f.keys.end(), keys, keys + n * kNKeys[bucket]);
  f.z.insert(f.z.end(), z, z + n * kZIn[bucket]);
  if (noise_kind) {
    f.kind.insert(f.kind.end(), noise_kind, noise_kind + n);
    f.noise.insert(f.noise.end(), noise, noise + n * m * m);
  } else {
    f.kind.insert(f.kind.end(), (size_t)n, (int32_t)GTB_NOISE_UNIT);
    f.noise.insert(f.noise.end(), (size_t)(n * m * m), 0.0);
  }
  f.n += n;
  return GTB_OK;
}

gtb_status gtb_isam2_set_ordering(gtb_isam2* s, int64_t n, const uint64_t* keys) {
  if (!s) return GTB_ERR_INVALID;
  if (n < 0 || (n && !keys)) return fail(s, GTB_ERR_INVALID, "bad arguments");
  s->ordering.assign(keys, keys + n);
  s->have_ordering = n > 0;
  return GTB_OK;
}

gtb_status gtb_isam2_update(gtb_isam2* s, gtb_isam2_result* result, uint64_t* failed_key) {
  if (!s) return GTB_ERR_INVALID;
  s->update_count += 1;
  int64_t relinearized = 0;
  gtb_status st;
  // steps 4 and 6: the variables known before this update, against the delta of the previous one
  if (s->eng && s->have_delta && s->update_count % s->skip == 0) {
    st = relinearize(s, &relinearized);
    if (st != GTB_OK) return st;
  }
  commit_staged(s);
  const double t0 = now_ms();
  st = rebuild(s);
  if (st != GTB_OK) return st;
  const double t1 = now_ms();
  uint64_t fk = 0;
  st = pass(s, gtb_linearize(s->eng), "gtb_linearize");
  if (st == GTB_OK) st = pass(s, gtb_factorize(s->eng, 0.0, 0, 0.0, 0.0, &fk), "gtb_factorize");
  if (st == GTB_ERR_INDETERMINATE && failed_key) *failed_key = fk;
  if (st == GTB_OK) st = pass(s, gtb_backsolve(s->eng), "gtb_backsolve");
  if (st == GTB_OK) st = pass(s, gtb_retract(s->eng), "gtb_retract");
  if (st != GTB_OK) return st;
  const double t2 = now_ms();
  s->have_delta = true;
  if (result) {
    result->update_count = s->update_count;
    result->variables_relinearized = relinearized;
    result->n_variables = (int64_t)s->insertion.size();
    result->n_factors = 0;
    for (int b = 0; b < GTB_NUM_BUCKETS; ++b) result->n_factors += s->factors[b].n;
    int64_t nc = 0, a = 0, b2 = 0, c = 0;
    gtb_plan_sizes(s->eng, &nc, &a, &b2, &c);
    result->n_cliques = nc;
    result->plan_ms = t1 - t0;
    result->solve_ms = t2 - t1;
  }
  return GTB_OK;
}

gtb_status gtb_isam2_estimate(gtb_isam2* s, int vtype, double* state) {
  if (!s) return GTB_ERR_INVALID;
  if (vtype < 0 || vtype >= GTB_NUM_VTYPES || !state) return fail(s, GTB_ERR_INVALID, "bad arguments");
  if (!s->eng || !s->have_delta) return fail(s, GTB_ERR_INVALID, "gtb_isam2_estimate needs a successful gtb_isam2_update first");
  if (s->vars[vtype].keys.empty()) return GTB_OK;
  return pass(s, gtb_get_variables(s->eng, vtype, 1, state), "gtb_get_variables");
}

int64_t gtb_isam2_count(const gtb_isam2* s, int vtype) {
  if (!s || vtype < 0 || vtype >= GTB_NUM_VTYPES) return -1;
  return (int64_t)s->vars[vtype].keys.size();
}

gtb_engine* gtb_isam2_engine(gtb_isam2* s) { return s ? s->eng : nullptr; }

}  // extern "C"
