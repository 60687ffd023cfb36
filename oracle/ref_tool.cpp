// oracle/ref_tool.cpp -- TEST INFRASTRUCTURE (not product code).
//
// Driver around the UNMODIFIED reference (GTSAM 4.3a0 compiled from
// /root/reference by oracle/Makefile into oracle/_ref/libgtsam_ref.so).  It
// reads a flat problem file (GTB1 container, schema in
// gtsam_petercdev_b200/problem.py), builds the reference's own
// NonlinearFactorGraph / Values from it and dumps what the parity tests compare
// against: whitened Jacobians, ordering, clique layout, [R S d] per clique,
// delta, errors, LM traces, and stage timings for the CPU baseline.
//
//   ref_tool <problem.gtb> <out.gtb> key=value ...
//     mode=linearize | solve | lm | gn | time | symbolic | ordering
//     ordering=colamd|metis|schur|natural|file   (file: array "ordering" in problem)
//     lambda=<double> diag=0|1                    (solve: LM damping, 0 => none)
//     tree=0                                      (solve: skip the clique layout, [R S d] and new values: full-size runs)
//     ld=2 ldsweeps=<k>                           (solve: delta_star by long-double iterative refinement; any size)
//     ld=1 ldmax=<n>                              (solve: also delta_star, the same damped system solved in long double)
//     ceres=0|1 maxiter=<n> lambda0=<d>           (lm)
//     iters=<n>                                   (time)
//   ref_tool <city10000 file> <out.gtb> mode=isam2 loops=<n> every=<k>
//     the reference's ISAM2 over a City10000-format file with the update schedule and parameters of
//     examples/ISAM2_City10000.cpp:62-134 (first hypothesis of every line): estimate after every k-th update,
//     seconds per update (the oracle and CPU baseline of the ISAM2 row)
#include <gtsam/geometry/Cal3Bundler.h>
#include <gtsam/geometry/PinholeCamera.h>
#include <gtsam/geometry/Point3.h>
#include <gtsam/geometry/Pose2.h>
#include <gtsam/geometry/Pose3.h>
#include <gtsam/inference/Ordering.h>
#include <gtsam/inference/Symbol.h>
#include <gtsam/linear/GaussianBayesTree.h>
#include <gtsam/linear/GaussianFactorGraph.h>
#include <gtsam/linear/HessianFactor.h>
#include <gtsam/linear/JacobianFactor.h>
#include <gtsam/linear/NoiseModel.h>
#include <gtsam/linear/VectorValues.h>
#include <gtsam/nonlinear/GaussNewtonOptimizer.h>
#include <gtsam/nonlinear/ISAM2.h>
#include <gtsam/nonlinear/LevenbergMarquardtOptimizer.h>
#include <gtsam/nonlinear/Marginals.h>
#include <gtsam/nonlinear/NonlinearFactorGraph.h>
#include <gtsam/nonlinear/PriorFactor.h>
#include <gtsam/nonlinear/Values.h>
#include <gtsam/nonlinear/internal/LevenbergMarquardtState.h>
#include <gtsam/slam/BetweenFactor.h>
#include <gtsam/slam/GeneralSFMFactor.h>
#include <gtsam/sfm/SfmData.h>
#include <gtsam/slam/dataset.h>

#include <chrono>
#include <fstream>
#include <iostream>
#include <sstream>

#include "gtb_io.h"

using namespace gtsam;
typedef PinholeCamera<Cal3Bundler> Camera;
typedef GeneralSFMFactor<Camera, Point3> SfmFactor;

static std::map<std::string, std::string> g_opt;
static std::string opt(const std::string& k, const std::string& d) {
  auto it = g_opt.find(k);
  return it == g_opt.end() ? d : it->second;
}
static double optd(const std::string& k, double d) {
  auto it = g_opt.find(k);
  return it == g_opt.end() ? d : std::stod(it->second);
}

static SharedNoiseModel make_noise(int kind, const double* p, int d) {
  switch (kind) {
    case 0: return noiseModel::Unit::Create(d);
    case 1: return noiseModel::Isotropic::Sigma(d, p[0]);
    case 2: { Vector s(d); for (int i = 0; i < d; ++i) s(i) = p[i]; return noiseModel::Diagonal::Sigmas(s); }
    case 5: { Vector s(d); for (int i = 0; i < d; ++i) s(i) = p[i]; return noiseModel::Diagonal::Variances(s); }
    case 3: case 4: {
      Matrix M(d, d);
      for (int i = 0; i < d; ++i) for (int j = 0; j < d; ++j) M(i, j) = p[i * d + j];
      return kind == 3 ? noiseModel::Gaussian::Information(M) : noiseModel::Gaussian::SqrtInformation(M);
    }
  }
  throw std::runtime_error("bad noise kind");
}
static Pose2 mk_pose2(const double* p) { return Pose2(p[0], p[1], p[2]); }
static Pose3 mk_pose3(const double* p) {
  Matrix3 R;
  R << p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8];
  return Pose3(Rot3(R), Point3(p[9], p[10], p[11]));
}
static Camera mk_cam(const double* p) {
  return Camera(mk_pose3(p), Cal3Bundler(p[12], p[13], p[14], p[15], p[16]));
}

struct Built {
  NonlinearFactorGraph graph;
  Values values;
  // per bucket: graph index of each factor
  std::vector<std::vector<size_t>> where = std::vector<std::vector<size_t>>(7);
};

static const char* kBucket[7] = {"between2", "between3", "sfm", "prior2", "prior3", "priorcam", "priorpt"};
static const int kZ[7] = {3, 12, 2, 3, 12, 17, 3};
static const int kDim[7] = {3, 6, 2, 3, 6, 9, 3};

static NonlinearFactor::shared_ptr make_factor(const gtbio::File& P, int b, size_t i) {
  std::string n = kBucket[b];
  const auto& keys = P.at(n + "_keys");
  const double* z = P.at(n + "_z").f64() + i * kZ[b];
  int d = kDim[b];
  int kind = P.at(n + "_nk").i32()[i];
  const double* np = P.at(n + "_noise").f64() + i * d * d;
  SharedNoiseModel model = make_noise(kind, np, d);
  if (P.has(n + "_robust")) {
    // <bucket>_robust = [m-estimator id, parameter]: ids as GTB_LOSS_* in include/gtsam_b200.h
    const double* rb = P.at(n + "_robust").f64();
    const int id = (int)rb[0];
    const double c = rb[1];
    noiseModel::mEstimator::Base::shared_ptr est;
    switch (id) {
      case 1: est = noiseModel::mEstimator::Fair::Create(c); break;
      case 2: est = noiseModel::mEstimator::Huber::Create(c); break;
      case 3: est = noiseModel::mEstimator::Cauchy::Create(c); break;
      case 4: est = noiseModel::mEstimator::Tukey::Create(c); break;
      case 5: est = noiseModel::mEstimator::Welsch::Create(c); break;
      case 6: est = noiseModel::mEstimator::GemanMcClure::Create(c); break;
      case 7: est = noiseModel::mEstimator::DCS::Create(c); break;
      case 8: est = noiseModel::mEstimator::L2WithDeadZone::Create(c); break;
      case 0: break;
      default: throw std::runtime_error("bad m-estimator id");
    }
    if (est) model = noiseModel::Robust::Create(est, model);
  }
  const uint64_t* k = keys.u64() + i * (b < 3 ? 2 : 1);
  switch (b) {
    case 0: return std::make_shared<BetweenFactor<Pose2>>(k[0], k[1], mk_pose2(z), model);
    case 1: return std::make_shared<BetweenFactor<Pose3>>(k[0], k[1], mk_pose3(z), model);
    case 2: return std::make_shared<SfmFactor>(Point2(z[0], z[1]), model, k[0], k[1]);
    case 3: return std::make_shared<PriorFactor<Pose2>>(k[0], mk_pose2(z), model);
    case 4: return std::make_shared<PriorFactor<Pose3>>(k[0], mk_pose3(z), model);
    case 5: return std::make_shared<PriorFactor<Camera>>(k[0], mk_cam(z), model);
    case 6: return std::make_shared<PriorFactor<Point3>>(k[0], Point3(z[0], z[1], z[2]), model);
  }
  throw std::runtime_error("bad bucket");
}

static Built build(const gtbio::File& P) {
  Built B;
  if (P.has("pose2_keys")) {
    const auto& k = P.at("pose2_keys");
    for (size_t i = 0; i < k.rows(); ++i) B.values.insert(k.u64()[i], mk_pose2(P.at("pose2").f64() + 3 * i));
  }
  if (P.has("pose3_keys")) {
    const auto& k = P.at("pose3_keys");
    for (size_t i = 0; i < k.rows(); ++i) B.values.insert(k.u64()[i], mk_pose3(P.at("pose3").f64() + 12 * i));
  }
  if (P.has("cam_keys")) {
    const auto& k = P.at("cam_keys");
    for (size_t i = 0; i < k.rows(); ++i) B.values.insert(k.u64()[i], mk_cam(P.at("cam").f64() + 17 * i));
  }
  if (P.has("point_keys")) {
    const auto& k = P.at("point_keys");
    for (size_t i = 0; i < k.rows(); ++i) {
      const double* p = P.at("point").f64() + 3 * i;
      B.values.insert(k.u64()[i], Point3(p[0], p[1], p[2]));
    }
  }
  if (P.has("graph_order")) {
    const auto& go = P.at("graph_order");
    for (size_t r = 0; r < go.rows(); ++r) {
      int b = go.i64()[2 * r];
      size_t i = go.i64()[2 * r + 1];
      if (B.where[b].size() <= i) B.where[b].resize(i + 1);
      B.where[b][i] = B.graph.size();
      B.graph.push_back(make_factor(P, b, i));
    }
  } else {
    for (int b = 0; b < 7; ++b) {
      std::string n = kBucket[b];
      if (!P.has(n + "_keys")) continue;
      size_t m = P.at(n + "_keys").rows();
      for (size_t i = 0; i < m; ++i) {
        B.where[b].push_back(B.graph.size());
        B.graph.push_back(make_factor(P, b, i));
      }
    }
  }
  return B;
}

static Ordering make_ordering(const gtbio::File& P, const Built& B) {
  std::string o = opt("ordering", "colamd");
  if (o == "colamd") return Ordering::Colamd(B.graph);
  if (o == "metis") return Ordering::Metis(B.graph);
  if (o == "natural") return Ordering::Natural(B.graph);
  if (o == "schur") {
    Ordering ord;
    if (P.has("point_keys")) for (size_t i = 0; i < P.at("point_keys").rows(); ++i) ord.push_back(P.at("point_keys").u64()[i]);
    if (P.has("cam_keys")) for (size_t i = 0; i < P.at("cam_keys").rows(); ++i) ord.push_back(P.at("cam_keys").u64()[i]);
    return ord;
  }
  if (o == "file") {
    Ordering ord;
    const auto& a = P.at("ordering");
    for (size_t i = 0; i < a.rows(); ++i) ord.push_back(a.u64()[i]);
    return ord;
  }
  throw std::runtime_error("unknown ordering " + o);
}

static double now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// whitened [A|b] per bucket, row-major m x (sum n_j + 1), in bucket order
static void dump_jacobians(const Built& B, const GaussianFactorGraph& gfg, gtbio::File& out) {
  for (int b = 0; b < 7; ++b) {
    if (B.where[b].empty()) continue;
    std::vector<double> buf;
    size_t width = 0;
    for (size_t gi : B.where[b]) {
      auto jf = std::dynamic_pointer_cast<JacobianFactor>(gfg[gi]);
      if (!jf) throw std::runtime_error("not a JacobianFactor");
      Matrix Ab = jf->augmentedJacobian();  // whitened (model is unit/null here)
      width = Ab.rows() * Ab.cols();
      for (int r = 0; r < Ab.rows(); ++r) for (int c = 0; c < Ab.cols(); ++c) buf.push_back(Ab(r, c));
    }
    out.put_f64(std::string(kBucket[b]) + "_Ab", buf, {B.where[b].size(), width});
  }
}

static void dump_ordering(const Ordering& ord, gtbio::File& out) {
  std::vector<uint64_t> k(ord.begin(), ord.end());
  out.put_u64("ordering", k);
}

// Clique layout and (optionally) numeric [R S d] in DFS post-order with children in stored order.
static void dump_bayes_tree(const GaussianBayesTree& bt, gtbio::File& out, bool numeric) {
  typedef GaussianBayesTree::sharedClique C;
  std::vector<C> post;
  std::vector<int64_t> parent;
  {
    struct Item { C c; bool expanded; };
    std::vector<Item> st;
    for (auto it = bt.roots().rbegin(); it != bt.roots().rend(); ++it) st.push_back({*it, false});
    while (!st.empty()) {
      Item& top = st.back();
      if (top.expanded) { post.push_back(top.c); st.pop_back(); continue; }
      top.expanded = true;
      C c = top.c;
      for (auto it = c->children.rbegin(); it != c->children.rend(); ++it) st.push_back({*it, false});
    }
  }
  std::map<const void*, int64_t> index;
  for (size_t i = 0; i < post.size(); ++i) index[post[i].get()] = i;
  std::vector<int64_t> nfront, nkeys;
  std::vector<uint64_t> keys;
  std::vector<double> rsd;
  for (auto& c : post) {
    auto cond = c->conditional();
    nfront.push_back(cond->nrFrontals());
    nkeys.push_back(cond->size());
    for (Key k : cond->keys()) keys.push_back(k);
    parent.push_back(c->parent() ? index.at(c->parent().get()) : -1);
    if (numeric) {
      Matrix Ab = cond->augmentedJacobian();  // [R S d], F x (N+1)
      for (int r = 0; r < Ab.rows(); ++r) for (int cc = 0; cc < Ab.cols(); ++cc) rsd.push_back(Ab(r, cc));
    }
  }
  out.put_i64("clique_nfrontal", nfront);
  out.put_i64("clique_nkeys", nkeys);
  out.put_u64("clique_keys", keys);
  out.put_i64("clique_parent", parent);
  if (numeric) out.put_f64("clique_Rsd", rsd);
}

static void dump_delta(const VectorValues& delta, gtbio::File& out) {
  std::vector<uint64_t> keys;
  std::vector<double> v;
  std::vector<int64_t> dims;
  for (const auto& kv : delta) keys.push_back(kv.first);
  std::sort(keys.begin(), keys.end());
  for (Key k : keys) {
    const Vector& x = delta.at(k);
    dims.push_back(x.size());
    for (int i = 0; i < x.size(); ++i) v.push_back(x(i));
  }
  out.put_u64("delta_keys", keys);
  out.put_i64("delta_dims", dims);
  out.put_f64("delta", v);
}

static void dump_values(const Values& v, gtbio::File& out, const std::string& prefix) {
  std::vector<double> p2, p3, cam, pt;
  std::vector<uint64_t> k2, k3, kc, kp;
  for (const auto& kv : v) {
    Key k = kv.key;
    if (auto g = dynamic_cast<const GenericValue<Pose2>*>(&kv.value)) {
      const Pose2& p = g->value();
      k2.push_back(k); p2.insert(p2.end(), {p.r().c(), p.r().s(), p.x(), p.y()});
    } else if (auto g = dynamic_cast<const GenericValue<Pose3>*>(&kv.value)) {
      const Pose3& p = g->value(); Matrix3 R = p.rotation().matrix();
      k3.push_back(k);
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) p3.push_back(R(r, c));
      p3.insert(p3.end(), {p.x(), p.y(), p.z()});
    } else if (auto g = dynamic_cast<const GenericValue<Camera>*>(&kv.value)) {
      const Camera& cm = g->value(); Matrix3 R = cm.pose().rotation().matrix();
      kc.push_back(k);
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) cam.push_back(R(r, c));
      cam.insert(cam.end(), {cm.pose().x(), cm.pose().y(), cm.pose().z(), cm.calibration().fx(),
                             cm.calibration().k1(), cm.calibration().k2(), cm.calibration().px(),
                             cm.calibration().py()});
    } else if (auto g = dynamic_cast<const GenericValue<Point3>*>(&kv.value)) {
      const Point3& p = g->value();
      kp.push_back(k); pt.insert(pt.end(), {p.x(), p.y(), p.z()});
    }
  }
  if (!k2.empty()) { out.put_u64(prefix + "pose2_keys", k2); out.put_f64(prefix + "pose2_cs", p2, {k2.size(), 4}); }
  if (!k3.empty()) { out.put_u64(prefix + "pose3_keys", k3); out.put_f64(prefix + "pose3", p3, {k3.size(), 12}); }
  if (!kc.empty()) { out.put_u64(prefix + "cam_keys", kc); out.put_f64(prefix + "cam", cam, {kc.size(), 17}); }
  if (!kp.empty()) { out.put_u64(prefix + "point_keys", kp); out.put_f64(prefix + "point", pt, {kp.size(), 3}); }
}

static GaussianFactorGraph damp(const GaussianFactorGraph& gfg, const Values& values, double lambda, bool diag) {
  if (lambda <= 0) return gfg;
  internal::LevenbergMarquardtState st(values, 0.0, lambda, 10.0);
  if (!diag) return st.buildDampedSystem(gfg);
  // mirrors LevenbergMarquardtOptimizer.cpp:293-299 with default min/maxDiagonal
  LevenbergMarquardtParams prm;
  VectorValues sq = gfg.hessianDiagonal();
  for (auto& kv : sq) kv.second = kv.second.cwiseMax(prm.minDiagonal).cwiseMin(prm.maxDiagonal).cwiseSqrt();
  return st.buildDampedSystem(gfg, sq);
}

// Extended-precision arbiter for the linear solve (mode=solve ld=1): the SAME damped system the reference eliminates
// (whitened [A | b] of every factor of `sys`, taken from the reference's own JacobianFactor::augmentedJacobian), solved
// through dense normal equations H = A^T A, g = A^T b with every sum, the Cholesky factorisation and both triangular
// solves in 80-bit long double (unit round-off 5.4e-20, 2048 times finer than FP64).  The result delta* is what both the
// reference's delta and the device's delta are measured against: "the device is no further from the exact solution than
// the reference" is then a statement about two numbers, not about an input-perturbation yardstick.  Dense and O(n^3):
// meant for the golden fixtures (n <= ~7k scalars), skipped above `ldmax` scalars.
static bool solve_long_double(const GaussianFactorGraph& sys, const VectorValues& like, std::vector<double>& out) {
  std::vector<uint64_t> keys;
  for (const auto& kv : like) keys.push_back(kv.first);
  std::sort(keys.begin(), keys.end());
  std::map<Key, size_t> off;
  size_t n = 0;
  for (Key k : keys) { off[k] = n; n += like.at(k).size(); }
  if ((double)n > optd("ldmax", 8000)) return false;
  typedef long double LD;
  std::vector<LD> H(n * n, 0.0L), g(n, 0.0L);   // row-major, lower triangle used
  for (const auto& f : sys) {
    if (!f) continue;
    auto jf = std::dynamic_pointer_cast<JacobianFactor>(f);
    if (!jf) throw std::runtime_error("solve_long_double: not a JacobianFactor");
    const Matrix Ab = jf->augmentedJacobian();   // whitened
    std::vector<size_t> col;
    for (auto it = jf->begin(); it != jf->end(); ++it)
      for (size_t d = 0; d < (size_t)jf->getDim(it); ++d) col.push_back(off.at(*it) + d);
    const size_t w = col.size();
    for (int r = 0; r < Ab.rows(); ++r)
      for (size_t a = 0; a < w; ++a) {
        const LD va = (LD)Ab(r, a);
        g[col[a]] += va * (LD)Ab(r, w);
        for (size_t b2 = 0; b2 < w; ++b2)
          if (col[b2] <= col[a]) H[col[a] * n + col[b2]] += va * (LD)Ab(r, b2);
      }
  }
  // Cholesky-Banachiewicz, H = L L^T, rows contiguous
  for (size_t i = 0; i < n; ++i) {
    LD* Li = &H[i * n];
    for (size_t j = 0; j <= i; ++j) {
      const LD* Lj = &H[j * n];
      LD s = Li[j];
      for (size_t k = 0; k < j; ++k) s -= Li[k] * Lj[k];
      if (j < i) Li[j] = s / Lj[j];
      else {
        if (!(s > 0)) return false;
        Li[i] = sqrtl(s);
      }
    }
  }
  for (size_t i = 0; i < n; ++i) {   // L y = g
    LD s = g[i];
    for (size_t k = 0; k < i; ++k) s -= H[i * n + k] * g[k];
    g[i] = s / H[i * n + i];
  }
  for (size_t ii = n; ii-- > 0;) {   // L^T x = y
    LD s = g[ii];
    for (size_t k = ii + 1; k < n; ++k) s -= H[k * n + ii] * g[k];
    g[ii] = s / H[ii * n + ii];
  }
  out.resize(n);
  for (size_t i = 0; i < n; ++i) out[i] = (double)g[i];
  return true;
}

// The same arbiter at sizes where dense normal equations do not fit (mode=solve ld=2): iterative refinement of the
// reference's own delta.  Each sweep forms the gradient residual rho = A^T (b - A x) of the damped system factor by factor
// in long double (the entries of A and b are the reference's own doubles, exact), and solves H d = rho with the
// reference's own multifrontal Cholesky on the same ordering: the factors keep their A with b = 0, and rho enters as
// one HessianFactor(key, G = 0, g = rho_key, f = 0) per variable.  x += d in long double.  The double-precision solve
// contracts the error by cond * 1e-16 per sweep, so three or four sweeps reach the long-double solution; the sweep
// corrections' norms are returned so that the caller sees the convergence.
static bool refine_long_double(const GaussianFactorGraph& sys, const Ordering& ord, const VectorValues& delta,
                               std::vector<double>& out, std::vector<double>& steps) {
  std::vector<uint64_t> keys;
  for (const auto& kv : delta) keys.push_back(kv.first);
  std::sort(keys.begin(), keys.end());
  std::map<Key, size_t> off;
  std::vector<size_t> dims;
  size_t n = 0;
  for (Key k : keys) { off[k] = n; dims.push_back(delta.at(k).size()); n += delta.at(k).size(); }
  typedef long double LD;
  std::vector<LD> x(n);
  for (size_t i = 0; i < keys.size(); ++i)
    for (size_t d = 0; d < dims[i]; ++d) x[off[keys[i]] + d] = (LD)delta.at(keys[i])(d);
  GaussianFactorGraph zero_b;
  for (const auto& f : sys) {
    auto jf = std::dynamic_pointer_cast<JacobianFactor>(f);
    if (!jf) throw std::runtime_error("refine_long_double: not a JacobianFactor");
    auto copy = std::make_shared<JacobianFactor>(*jf);
    copy->getb().setZero();
    zero_b.push_back(copy);
  }
  const int sweeps = (int)optd("ldsweeps", 4);
  for (int it = 0; it < sweeps; ++it) {
    std::vector<LD> rho(n, 0.0L);
    for (const auto& f : sys) {
      auto jf = std::static_pointer_cast<JacobianFactor>(f);
      const Matrix Ab = jf->augmentedJacobian();
      std::vector<size_t> col;
      for (auto kit = jf->begin(); kit != jf->end(); ++kit)
        for (size_t d = 0; d < (size_t)jf->getDim(kit); ++d) col.push_back(off.at(*kit) + d);
      const size_t w = col.size();
      for (int r = 0; r < Ab.rows(); ++r) {
        LD res = (LD)Ab(r, w);
        for (size_t a = 0; a < w; ++a) res -= (LD)Ab(r, a) * x[col[a]];
        for (size_t a = 0; a < w; ++a) rho[col[a]] += (LD)Ab(r, a) * res;
      }
    }
    GaussianFactorGraph g2 = zero_b;
    for (size_t i = 0; i < keys.size(); ++i) {
      Vector gv(dims[i]);
      for (size_t d = 0; d < dims[i]; ++d) gv(d) = (double)rho[off[keys[i]] + d];
      g2.emplace_shared<HessianFactor>(keys[i], Matrix::Zero(dims[i], dims[i]), gv, 0.0);
    }
    const VectorValues d = g2.eliminateMultifrontal(ord)->optimize();
    LD dn = 0, xn = 0;
    for (size_t i = 0; i < keys.size(); ++i)
      for (size_t q = 0; q < dims[i]; ++q) {
        const LD dv = (LD)d.at(keys[i])(q);
        x[off[keys[i]] + q] += dv;
        dn += dv * dv;
      }
    for (size_t i = 0; i < n; ++i) xn += x[i] * x[i];
    steps.push_back((double)(sqrtl(dn) / sqrtl(xn)));
  }
  out.resize(n);
  for (size_t i = 0; i < n; ++i) out[i] = (double)x[i];
  return true;
}

// mode=isam2: argv[1] is a City10000-format file: "EDGE2 <from> <n> <to> <n> <hypotheses> {x y theta}...".  Schedule and
// parameters follow examples/ISAM2_City10000.cpp:62-134 (no ambiguity: the first hypothesis of every line): prior
// 1e-4 on pose 0 and one update; then one update per line -- an odometry line (to = from + 1) adds the new pose,
// initialised by composing the current estimate of `from` with the measurement, and a BetweenFactor with sigmas
// (1/30, 1/30, 1/100); any other line is a loop closure with sigmas 10, and only those count towards `loops`.
// Gauss-Newton ISAM2 with wildfire threshold 0, relinearizeThreshold 0.01, relinearizeSkip 1.
static int isam2_mode(const std::string& data, const std::string& outpath) {
  const size_t max_loops = (size_t)std::stoll(opt("loops", "1000000000"));
  const size_t every = (size_t)std::max<long long>(1, std::stoll(opt("every", "100")));
  std::ifstream in(data);
  if (!in) { std::cerr << "cannot open " << data << "\n"; return 2; }
  ISAM2Params params;
  params.optimizationParams = ISAM2GaussNewtonParams(0.0);
  params.relinearizeThreshold = 0.01;
  params.relinearizeSkip = 1;
  ISAM2 isam(params);
  const auto prior_model = noiseModel::Diagonal::Sigmas(Vector3(1e-4, 1e-4, 1e-4));
  const auto odo_model = noiseModel::Diagonal::Sigmas(Vector3(1.0 / 30.0, 1.0 / 30.0, 1.0 / 100.0));
  const auto loop_model = noiseModel::Diagonal::Sigmas(Vector3(10.0, 10.0, 10.0));
  NonlinearFactorGraph fresh;
  Values fresh_values;
  fresh_values.insert(Symbol('x', 0), Pose2(0, 0, 0));
  fresh.addPrior<Pose2>(Symbol('x', 0), Pose2(0, 0, 0), prior_model);
  isam.update(fresh, fresh_values);
  Values est = isam.calculateBestEstimate();
  std::vector<double> snap, secs;          // estimates (x, y, theta per pose, poses 0..n-1) at the snapshot updates
  std::vector<int64_t> snap_update, snap_nposes, snap_off, line_kind;
  size_t loops = 0, updates = 0, nposes = 1;
  std::string line;
  while (loops < max_loops && std::getline(in, line)) {
    std::istringstream ls(line);
    std::string tag;
    long long from, to, d0, d1, nh;
    if (!(ls >> tag >> from >> d0 >> to >> d1 >> nh) || nh < 1) continue;
    double x, y, th;
    ls >> x >> y >> th;
    const Pose2 z(x, y, th);
    fresh.resize(0);
    fresh_values.clear();
    const bool odometry = from == to - 1;
    if (odometry) {
      fresh_values.insert(Symbol('x', to), est.at<Pose2>(Symbol('x', from)) * z);
      fresh.emplace_shared<BetweenFactor<Pose2>>(Symbol('x', from), Symbol('x', to), z, odo_model);
      nposes = std::max(nposes, (size_t)to + 1);
    } else {
      fresh.emplace_shared<BetweenFactor<Pose2>>(Symbol('x', from), Symbol('x', to), z, loop_model);
      ++loops;
    }
    const auto t0 = std::chrono::steady_clock::now();
    isam.update(fresh, fresh_values);
    est = isam.calculateBestEstimate();
    secs.push_back(std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    line_kind.push_back(odometry ? 0 : 1);
    ++updates;
    if (updates % every == 0) {
      snap_update.push_back((int64_t)updates);
      snap_nposes.push_back((int64_t)nposes);
      snap_off.push_back((int64_t)snap.size());
      for (size_t i = 0; i < nposes; ++i) {
        const Pose2 q = est.at<Pose2>(Symbol('x', i));
        snap.push_back(q.x()); snap.push_back(q.y()); snap.push_back(q.theta());
      }
    }
  }
  gtbio::File out;
  std::vector<double> fin;
  for (size_t i = 0; i < nposes; ++i) {
    const Pose2 q = est.at<Pose2>(Symbol('x', i));
    fin.push_back(q.x()); fin.push_back(q.y()); fin.push_back(q.theta());
  }
  out.put_f64("final", fin, {(uint64_t)nposes, 3});
  out.put_f64("snapshots", snap);
  out.put_i64("snapshot_update", snap_update);
  out.put_i64("snapshot_nposes", snap_nposes);
  out.put_i64("snapshot_offset", snap_off);
  out.put_f64("update_seconds", secs);
  out.put_i64("update_is_loop", line_kind);
  out.put_i64("counts", {(int64_t)updates, (int64_t)loops, (int64_t)nposes, (int64_t)isam.getFactorsUnsafe().size()});
  out.put_f64("final_error", {isam.getFactorsUnsafe().error(est)});
  out.save(outpath);
  return 0;
}

// mode=readg2o is3d=0|1 kernel=none|huber|tukey / mode=readbal: argv[1] is the DATA file; the reference's own readers
// (readG2o, SfmData::FromBalFile + the graph of tests/testGeneralSFMFactorB.cpp:44-58) build graph and values, which
// are dumped in the problem schema (+ *_info = R^T R of every Gaussian model, pose2_cs = (cos, sin)).
// mode=writeg2o: reads the file with readG2o and writes it back with the reference's writeG2o to <out>.
static int dataset_mode(const std::string& data, const std::string& outpath) {
  const std::string mode = opt("mode", "");
  gtbio::File out;
  NonlinearFactorGraph graph;
  Values values;
  if (mode == "readbal") {
    SfmData db = SfmData::FromBalFile(data);
    for (size_t j = 0; j < db.numberTracks(); j++)
      for (const SfmMeasurement& m : db.tracks[j].measurements)
        graph.emplace_shared<SfmFactor>(m.second, noiseModel::Unit::Create(2), m.first, Symbol('p', j));
    size_t i = 0, j = 0;
    for (const SfmCamera& camera : db.cameras) values.insert(i++, camera);
    for (const SfmTrack& track : db.tracks) values.insert(Symbol('p', j++), track.p);
  } else {
    const bool is3d = opt("is3d", "0") == "1";
    const std::string k = opt("kernel", "none");
    const KernelFunctionType kt = k == "huber" ? KernelFunctionTypeHUBER : (k == "tukey" ? KernelFunctionTypeTUKEY : KernelFunctionTypeNONE);
    auto gv = readG2o(data, is3d, kt);
    graph = *gv.first;
    values = *gv.second;
    if (mode == "writeg2o") {
      writeG2o(graph, values, outpath);
      return 0;
    }
  }
  std::vector<uint64_t> k2, k3, kc, kp;
  std::vector<double> v2, v2cs, v3, vc, vp;
  for (const auto& kv : values.extract<Pose2>()) {
    k2.push_back(kv.first);
    v2.insert(v2.end(), {kv.second.x(), kv.second.y(), kv.second.theta()});
    v2cs.insert(v2cs.end(), {kv.second.r().c(), kv.second.r().s()});
  }
  auto push_pose3 = [](std::vector<double>& v, const Pose3& p) {
    const Matrix3 R = p.rotation().matrix();
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) v.push_back(R(r, c));
    v.insert(v.end(), {p.x(), p.y(), p.z()});
  };
  for (const auto& kv : values.extract<Pose3>()) { k3.push_back(kv.first); push_pose3(v3, kv.second); }
  for (const auto& kv : values.extract<Camera>()) {
    kc.push_back(kv.first);
    push_pose3(vc, kv.second.pose());
    const Cal3Bundler& K = kv.second.calibration();
    vc.insert(vc.end(), {K.fx(), K.k1(), K.k2(), K.px(), K.py()});
  }
  for (const auto& kv : values.extract<Point3>()) { kp.push_back(kv.first); vp.insert(vp.end(), {kv.second.x(), kv.second.y(), kv.second.z()}); }
  std::vector<uint64_t> b2k, b3k, sk;
  std::vector<double> b2z, b2cs, b2i, b3z, b3i, sz;
  auto info_of = [](const SharedNoiseModel& m, std::vector<double>& dst) {
    SharedNoiseModel base = m;
    if (auto rob = std::dynamic_pointer_cast<noiseModel::Robust>(m)) base = rob->noise();
    auto g = std::dynamic_pointer_cast<noiseModel::Gaussian>(base);
    const Matrix I = g->R().transpose() * g->R();
    for (int r = 0; r < I.rows(); ++r) for (int c = 0; c < I.cols(); ++c) dst.push_back(I(r, c));
  };
  for (const auto& f : graph) {
    if (auto b = std::dynamic_pointer_cast<BetweenFactor<Pose2>>(f)) {
      b2k.insert(b2k.end(), {b->key<1>(), b->key<2>()});
      b2z.insert(b2z.end(), {b->measured().x(), b->measured().y(), b->measured().theta()});
      b2cs.insert(b2cs.end(), {b->measured().r().c(), b->measured().r().s()});
      info_of(b->noiseModel(), b2i);
    } else if (auto b3 = std::dynamic_pointer_cast<BetweenFactor<Pose3>>(f)) {
      b3k.insert(b3k.end(), {b3->key<1>(), b3->key<2>()});
      push_pose3(b3z, b3->measured());
      info_of(b3->noiseModel(), b3i);
    } else if (auto sf = std::dynamic_pointer_cast<SfmFactor>(f)) {
      sk.insert(sk.end(), {sf->key<1>(), sf->key<2>()});
      sz.insert(sz.end(), {sf->measured().x(), sf->measured().y()});
    }
  }
  out.put_u64("pose2_keys", k2); out.put_f64("pose2", v2, {k2.size(), 3}); out.put_f64("pose2_cs", v2cs, {k2.size(), 2});
  out.put_u64("pose3_keys", k3); out.put_f64("pose3", v3, {k3.size(), 12});
  out.put_u64("cam_keys", kc); out.put_f64("cam", vc, {kc.size(), 17});
  out.put_u64("point_keys", kp); out.put_f64("point", vp, {kp.size(), 3});
  out.put_u64("between2_keys", b2k, {b2k.size() / 2, 2}); out.put_f64("between2_z", b2z, {b2k.size() / 2, 3});
  out.put_f64("between2_cs", b2cs, {b2k.size() / 2, 2}); out.put_f64("between2_info", b2i, {b2k.size() / 2, 9});
  out.put_u64("between3_keys", b3k, {b3k.size() / 2, 2}); out.put_f64("between3_z", b3z, {b3k.size() / 2, 12});
  out.put_f64("between3_info", b3i, {b3k.size() / 2, 36});
  out.put_u64("sfm_keys", sk, {sk.size() / 2, 2}); out.put_f64("sfm_z", sz, {sk.size() / 2, 2});
  try {
    out.put_f64("error0", {graph.error(values)});
  } catch (const std::exception&) {   // files without vertices (load3D does not create them)
    out.put_f64("error0", {-1.0});
  }
  out.put_i64("n_factors", {(int64_t)graph.size()});
  out.save(outpath);
  return 0;
}

int main(int argc, char** argv) {
  if (argc < 3) { std::cerr << "usage: ref_tool problem.gtb out.gtb key=value...\n"; return 2; }
  for (int i = 3; i < argc; ++i) {
    std::string a = argv[i];
    auto p = a.find('=');
    if (p != std::string::npos) g_opt[a.substr(0, p)] = a.substr(p + 1);
  }
  try {
    if (opt("mode", "solve").rfind("read", 0) == 0 || opt("mode", "solve") == "writeg2o") return dataset_mode(argv[1], argv[2]);
    if (opt("mode", "solve") == "isam2") return isam2_mode(argv[1], argv[2]);
    gtbio::File P, out;
    P.load(argv[1]);
    Built B = build(P);
    std::string mode = opt("mode", "solve");
    out.put_f64("error0", {B.graph.error(B.values)});
    out.put_i64("n_factors", {(int64_t)B.graph.size()});

    if (mode == "ordering") {
      Ordering ord = make_ordering(P, B);
      dump_ordering(ord, out);
    } else if (mode == "marginals") {
      // Marginals(graph, values).marginalCovariance(key) for every variable, ascending key (nonlinear/Marginals.cpp)
      const auto tm0 = std::chrono::steady_clock::now();
      Marginals marginals(B.graph, B.values);
      std::vector<uint64_t> mk;
      std::vector<double> mv;
      std::vector<int64_t> md;
      for (const auto& kv : B.values) {
        const Matrix Cv = marginals.marginalCovariance(kv.key);
        mk.push_back(kv.key);
        md.push_back(Cv.rows());
        for (int r = 0; r < Cv.rows(); ++r) for (int c = 0; c < Cv.cols(); ++c) mv.push_back(Cv(r, c));
      }
      out.put_f64("t_marginals", {std::chrono::duration<double>(std::chrono::steady_clock::now() - tm0).count()});
      out.put_u64("marginal_keys", mk);
      out.put_i64("marginal_dims", md);
      out.put_f64("marginal_cov", mv);
    } else if (mode == "linearize") {
      auto gfg = B.graph.linearize(B.values);
      dump_jacobians(B, *gfg, out);
      VectorValues hd = gfg->hessianDiagonal();
      std::vector<uint64_t> hk; std::vector<double> hv;
      for (auto& kv : hd) { hk.push_back(kv.first); }
      std::sort(hk.begin(), hk.end());
      for (auto k : hk) for (int i = 0; i < hd.at(k).size(); ++i) hv.push_back(hd.at(k)(i));
      out.put_u64("hdiag_keys", hk);
      out.put_f64("hdiag", hv);
    } else if (mode == "symbolic" || mode == "solve") {
      Ordering ord = make_ordering(P, B);
      dump_ordering(ord, out);
      auto gfg = B.graph.linearize(B.values);
      double lambda = optd("lambda", 0.0);
      bool diag = optd("diag", 0) != 0;
      GaussianFactorGraph sys = damp(*gfg, B.values, lambda, diag);
      bool failed = false;
      try {
        auto bt = sys.eliminateMultifrontal(ord);
        if (optd("tree", 1) != 0) dump_bayes_tree(*bt, out, mode == "solve");   // tree=0: full-size runs keep delta and errors only
        if (mode == "solve") {
          VectorValues delta = bt->optimize();
          dump_delta(delta, out);
          out.put_f64("linear_error0", {gfg->error(VectorValues::Zero(delta))});
          out.put_f64("linear_error1", {gfg->error(delta)});
          Values nv = B.values.retract(delta);
          out.put_f64("error1", {B.graph.error(nv)});
          if (optd("tree", 1) != 0) dump_values(nv, out, "new_");
          if (optd("ld", 0) == 1) {
            std::vector<double> star;
            if (solve_long_double(sys, delta, star)) out.put_f64("delta_star", star);
          }
          if (optd("ld", 0) == 2 || optd("ld2", 0) != 0) {
            std::vector<double> star, steps;
            if (refine_long_double(sys, ord, delta, star, steps)) {
              out.put_f64(optd("ld", 0) == 2 ? "delta_star" : "delta_star_refined", star);
              out.put_f64("delta_star_sweeps", steps);
            }
          }
        }
      } catch (const IndeterminantLinearSystemException& e) {
        failed = true;
        out.put_u64("failed_key", {e.nearbyVariable()});
      }
      out.put_i64("failed", {failed ? 1 : 0});
    } else if (mode == "lm" || mode == "gn") {
      Ordering ord = make_ordering(P, B);
      dump_ordering(ord, out);
      std::string log = std::string(argv[2]) + ".lmlog";
      double t0 = now();
      Values result;
      int64_t iters = 0, inner = 0;
      double err = 0, lam = 0;
      if (mode == "lm") {
        LevenbergMarquardtParams prm;
        if (optd("ceres", 0) != 0) LevenbergMarquardtParams::SetCeresDefaults(&prm);
        prm.maxIterations = (int)optd("maxiter", prm.maxIterations);
        prm.lambdaInitial = optd("lambda0", prm.lambdaInitial);
        prm.relativeErrorTol = optd("reltol", prm.relativeErrorTol);
        prm.absoluteErrorTol = optd("abstol", prm.absoluteErrorTol);
        prm.setOrdering(ord);
        prm.logFile = log;
        LevenbergMarquardtOptimizer lm(B.graph, B.values, prm);
        result = lm.optimize();
        iters = lm.iterations(); inner = lm.getInnerIterations(); err = lm.error(); lam = lm.lambda();
      } else {
        GaussNewtonParams prm;
        prm.maxIterations = (int)optd("maxiter", prm.maxIterations);
        prm.relativeErrorTol = optd("reltol", prm.relativeErrorTol);
        prm.absoluteErrorTol = optd("abstol", prm.absoluteErrorTol);
        prm.setOrdering(ord);
        GaussNewtonOptimizer gn(B.graph, B.values, prm);
        result = gn.optimize();
        iters = gn.iterations(); err = gn.error();
      }
      double t1 = now();
      out.put_f64("final_error", {err});
      out.put_f64("final_lambda", {lam});
      out.put_i64("iterations", {iters});
      out.put_i64("inner_iterations", {inner});
      out.put_f64("seconds", {t1 - t0});
      dump_values(result, out, "final_");
      // LM log: rows [inner, time, error, lambda, outer]
      std::ifstream lf(log);
      std::vector<double> rows; std::string line; size_t nrows = 0;
      while (std::getline(lf, line)) {
        for (auto& ch : line) if (ch == ',') ch = ' ';
        std::istringstream is(line); double v; int n = 0;
        std::vector<double> r;
        while (is >> v) { r.push_back(v); ++n; }
        if (n == 5) { rows.insert(rows.end(), r.begin(), r.end()); ++nrows; }
      }
      out.put_f64("lm_log", rows, {nrows, 5});
      std::remove(log.c_str());
    } else if (mode == "time") {
      // CPU baseline: stage timings + LM inner iterations/s with the reference's own code.
      Ordering ord = make_ordering(P, B);
      int iters = (int)optd("iters", 3);
      double lambda = optd("lambda", 1e-5);
      int warm = (int)optd("warm", 1);
      double trial_error = 0, lin0 = 0, lin1 = 0;
      std::vector<double> tl, te, tb, tt;
      for (int it = 0; it < iters + warm; ++it) {
        double a = now();
        auto gfg = B.graph.linearize(B.values);
        double b = now();
        GaussianFactorGraph sys = damp(*gfg, B.values, lambda, optd("diag", 0) != 0);
        auto bt = sys.eliminateMultifrontal(ord);
        double c = now();
        VectorValues delta = bt->optimize();
        double d = now();
        double e0 = gfg->error(VectorValues::Zero(delta)), e1 = gfg->error(delta);
        Values nv = B.values.retract(delta);
        double en = B.graph.error(nv);
        double e = now();
        lin0 = e0; lin1 = e1; trial_error = en;
        if (it < warm) continue;  // warm-up
        tl.push_back(b - a); te.push_back(c - b); tb.push_back(d - c); tt.push_back(e - a);
      }
      out.put_f64("t_linearize", tl); out.put_f64("t_eliminate", te);
      out.put_f64("t_backsub", tb); out.put_f64("t_iter", tt);
      out.put_f64("trial_error", {trial_error}); out.put_f64("linear_error0", {lin0}); out.put_f64("linear_error1", {lin1});
    } else {
      throw std::runtime_error("unknown mode " + mode);
    }
    out.save(argv[2]);
  } catch (const std::exception& e) {
    std::cerr << "ref_tool: " << e.what() << "\n";
    return 1;
  }
  return 0;
}
