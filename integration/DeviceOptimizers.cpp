// integration/DeviceOptimizers.cpp -- see DeviceOptimizers.h.  Integration code: links the reference.
#include "DeviceOptimizers.h"

#include <gtsam/geometry/Cal3Bundler.h>
#include <gtsam/geometry/PinholeCamera.h>
#include <gtsam/geometry/Point3.h>
#include <gtsam/geometry/Pose2.h>
#include <gtsam/geometry/Pose3.h>
#include <gtsam/linear/JacobianFactor.h>
#include <gtsam/linear/NoiseModel.h>
#include <gtsam/inference/inferenceExceptions.h>
#include <gtsam/linear/linearExceptions.h>
#include <gtsam/nonlinear/PriorFactor.h>
#include <gtsam/nonlinear/internal/LevenbergMarquardtState.h>
#include <gtsam/slam/BetweenFactor.h>
#include <gtsam/slam/GeneralSFMFactor.h>

#include <algorithm>
#include <stdexcept>

namespace gtsam {

using device_detail::Bucket;
using device_detail::PackedGraph;
using device_detail::PackedValues;

namespace {

typedef PinholeCamera<Cal3Bundler> Camera;
typedef GeneralSFMFactor<Camera, Point3> SfmFactor;

const int kResidualDim[GTB_NUM_BUCKETS] = {3, 6, 2, 3, 6, 9, 3};
const int kMeasurement[GTB_NUM_BUCKETS] = {3, 12, 2, 3, 12, 17, 3};
const int kBucketKeys[GTB_NUM_BUCKETS] = {2, 2, 2, 1, 1, 1, 1};
const int kJacobianWidth[GTB_NUM_BUCKETS] = {7, 13, 13, 4, 7, 10, 4};
const int kFirstBlock[GTB_NUM_BUCKETS] = {3, 6, 9, 3, 6, 9, 3};
const int kTangent[GTB_NUM_VTYPES] = {3, 6, 9, 3};
const int kStateIn[GTB_NUM_VTYPES] = {3, 12, 17, 3};
const int kStateOut[GTB_NUM_VTYPES] = {4, 12, 17, 3};   // Pose2 comes back as (cos, sin, x, y)

void appendPose3(std::vector<double>& v, const Pose3& p) {
  const Matrix3 R = p.rotation().matrix();
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) v.push_back(R(r, c));
  v.push_back(p.x());
  v.push_back(p.y());
  v.push_back(p.z());
}
void appendCamera(std::vector<double>& v, const Camera& c) {
  appendPose3(v, c.pose());
  const Cal3Bundler& K = c.calibration();
  v.insert(v.end(), {K.fx(), K.k1(), K.k2(), K.px(), K.py()});
}
Pose3 pose3From(const double* p) {
  Matrix3 R;
  R << p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8];
  return Pose3(Rot3(R), Point3(p[9], p[10], p[11]));
}

// m-estimator of a Robust model -> (GTB_LOSS_*, parameter)  (linear/LossFunctions.h)
std::pair<int, double> lossOf(const noiseModel::mEstimator::Base::shared_ptr& est) {
  using namespace noiseModel::mEstimator;
  if (auto e = std::dynamic_pointer_cast<Fair>(est)) return {GTB_LOSS_FAIR, e->modelParameter()};
  if (auto e = std::dynamic_pointer_cast<Huber>(est)) return {GTB_LOSS_HUBER, e->modelParameter()};
  if (auto e = std::dynamic_pointer_cast<Cauchy>(est)) return {GTB_LOSS_CAUCHY, e->modelParameter()};
  if (auto e = std::dynamic_pointer_cast<Tukey>(est)) return {GTB_LOSS_TUKEY, e->modelParameter()};
  if (auto e = std::dynamic_pointer_cast<Welsch>(est)) return {GTB_LOSS_WELSCH, e->modelParameter()};
  if (auto e = std::dynamic_pointer_cast<GemanMcClure>(est)) return {GTB_LOSS_GEMANMCCLURE, e->modelParameter()};
  if (auto e = std::dynamic_pointer_cast<DCS>(est)) return {GTB_LOSS_DCS, e->modelParameter()};
  if (auto e = std::dynamic_pointer_cast<L2WithDeadZone>(est)) return {GTB_LOSS_L2WITHDEADZONE, e->modelParameter()};
  if (std::dynamic_pointer_cast<Null>(est)) return {GTB_LOSS_NONE, 0.0};
  throw std::invalid_argument("gtsam_b200: unsupported m-estimator (no device kernel, no CPU fallback)");
}

// One factor's noise model into the bucket's tables.  Unit / Isotropic / Diagonal / Gaussian go over as the user built
// them (the engine applies the reference's smart-factory rules itself); Constrained has no device path.
void appendNoise(Bucket& b, const SharedNoiseModel& model, int d) {
  SharedNoiseModel base = model;
  int loss = GTB_LOSS_NONE;
  double k = 0;
  if (auto rob = std::dynamic_pointer_cast<noiseModel::Robust>(model)) {
    std::tie(loss, k) = lossOf(rob->robust());
    base = rob->noise();
  }
  if (b.loss >= 0 && (b.loss != loss || b.lossK != k))
    throw std::invalid_argument("gtsam_b200: the factors of one type must share one m-estimator (gtb_set_robust is per bucket)");
  b.loss = loss;
  b.lossK = k;
  std::vector<double> nz(d * d, 0.0);
  int kind = GTB_NOISE_UNIT;
  if (!base || base->isUnit()) {
    kind = GTB_NOISE_UNIT;
  } else if (base->isConstrained()) {
    throw InvalidNoiseModel(d, d);   // QR path of the reference: not available on the device, rejected loudly
  } else if (auto iso = std::dynamic_pointer_cast<noiseModel::Isotropic>(base)) {
    kind = GTB_NOISE_ISO_SIGMA;
    nz[0] = iso->sigma();
  } else if (auto dg = std::dynamic_pointer_cast<noiseModel::Diagonal>(base)) {
    kind = GTB_NOISE_DIAG_SIGMAS;
    for (int i = 0; i < d; ++i) nz[i] = dg->sigmas()(i);
  } else if (auto g = std::dynamic_pointer_cast<noiseModel::Gaussian>(base)) {
    kind = GTB_NOISE_SQRT_INFORMATION;
    const Matrix R = g->R();
    if (R.rows() != d || R.cols() != d) throw InvalidNoiseModel(d, R.rows());
    for (int r = 0; r < d; ++r)
      for (int c = 0; c < d; ++c) nz[r * d + c] = R(r, c);
  } else {
    throw std::invalid_argument("gtsam_b200: unsupported noise model type");
  }
  b.kind.push_back(kind);
  b.noise.insert(b.noise.end(), nz.begin(), nz.end());
}

gtb_lm_params toParams(const LevenbergMarquardtParams& p) {
  gtb_lm_params q;
  gtb_lm_params_default(&q);
  q.max_iterations = (int32_t)p.maxIterations;
  q.relative_error_tol = p.relativeErrorTol;
  q.absolute_error_tol = p.absoluteErrorTol;
  q.error_tol = p.errorTol;
  q.lambda_initial = p.lambdaInitial;
  q.lambda_factor = p.lambdaFactor;
  q.lambda_upper_bound = p.lambdaUpperBound;
  q.lambda_lower_bound = p.lambdaLowerBound;
  q.min_model_fidelity = p.minModelFidelity;
  q.diagonal_damping = p.diagonalDamping ? 1 : 0;
  q.use_fixed_lambda_factor = p.useFixedLambdaFactor ? 1 : 0;
  q.min_diagonal = p.minDiagonal;
  q.max_diagonal = p.maxDiagonal;
  return q;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
void DeviceProblem::check(gtb_status st, Key nearby) const {
  if (st == GTB_OK) return;
  const std::string msg = eng_ ? gtb_last_error(eng_) : "gtsam_b200 error";
  switch (st) {
    case GTB_ERR_INDETERMINATE: throw IndeterminantLinearSystemException(nearby);   // linear/linearExceptions.h:94
    case GTB_ERR_KEY_MISSING: throw ValuesKeyDoesNotExist("gtsam_b200", nearby);
    case GTB_ERR_NOISE_MODEL: throw InvalidNoiseModel(0, 0);
    case GTB_ERR_INCONSISTENT: throw InconsistentEliminationRequested();
    case GTB_ERR_NO_DEVICE: throw std::runtime_error("gtsam_b200: no CUDA device (there is no CPU fallback): " + msg);
    default: throw std::runtime_error("gtsam_b200: " + msg);
  }
}

DeviceProblem::DeviceProblem(const NonlinearFactorGraph& graph, const Values& values, const Ordering& ordering, int device) {
  if (gtb_create(device, &eng_) != GTB_OK) throw std::runtime_error("gtsam_b200: cannot create an engine on this CUDA device");
  try {
    uploadValues(values);
    uploadGraph(graph);
    check(gtb_finalize_problem(eng_));
    std::vector<uint64_t> keys(ordering.begin(), ordering.end());
    check(gtb_set_ordering(eng_, (int64_t)keys.size(), keys.data()));   // symbolic Bayes tree, once
  } catch (...) {
    gtb_destroy(eng_);
    eng_ = nullptr;
    throw;
  }
}

DeviceProblem::~DeviceProblem() {
  if (eng_) gtb_destroy(eng_);
}

// Values -> one contiguous array per type (replaces the map lookups + dynamic_cast + copy of every Values::at,
// nonlinear/Values-inl.h:261-272)
PackedValues device_detail::packValues(const Values& v) {
  PackedValues out;
  for (const auto& kv : v) {
    int t;
    if (auto p = dynamic_cast<const GenericValue<Pose2>*>(&kv.value)) {
      t = GTB_POSE2;
      out.state[t].insert(out.state[t].end(), {p->value().x(), p->value().y(), p->value().theta()});
    } else if (auto p3 = dynamic_cast<const GenericValue<Pose3>*>(&kv.value)) {
      t = GTB_POSE3;
      appendPose3(out.state[t], p3->value());
    } else if (auto pc = dynamic_cast<const GenericValue<Camera>*>(&kv.value)) {
      t = GTB_CAMERA;
      appendCamera(out.state[t], pc->value());
    } else if (auto pp = dynamic_cast<const GenericValue<Point3>*>(&kv.value)) {
      t = GTB_POINT3;
      out.state[t].insert(out.state[t].end(), {pp->value().x(), pp->value().y(), pp->value().z()});
    } else {
      throw std::invalid_argument("gtsam_b200: unsupported Value type for key " + DefaultKeyFormatter(kv.key));
    }
    out.keys[t].push_back(kv.key);
  }
  return out;
}

void DeviceProblem::uploadValues(const Values& v) {
  const PackedValues pv = device_detail::packValues(v);
  std::vector<std::pair<Key, size_t>> all;
  for (int t = 0; t < GTB_NUM_VTYPES; ++t) {
    keys_[t] = pv.keys[t];
    for (Key k : pv.keys[t]) all.emplace_back(k, (size_t)kTangent[t]);
  }
  std::sort(all.begin(), all.end());
  for (const auto& a : all) {
    sortedKeys_.push_back(a.first);
    sortedDims_.push_back(a.second);
  }
  for (int t = 0; t < GTB_NUM_VTYPES; ++t)
    if (!keys_[t].empty()) {
      std::vector<uint64_t> k(keys_[t].begin(), keys_[t].end());
      check(gtb_set_variables(eng_, t, (int64_t)k.size(), k.data(), pv.state[t].data()));
    }
}

// NonlinearFactorGraph -> one table per factor type, discovered once by dynamic_cast (seam S5)
PackedGraph device_detail::packGraph(const NonlinearFactorGraph& g) {
  PackedGraph out;
  Bucket* b = out.bucket;
  for (const auto& f : g) {
    int bi = -1;
    if (!f) {
      out.slots.emplace_back(-1, 0);
      continue;
    }
    if (auto q = std::dynamic_pointer_cast<BetweenFactor<Pose2>>(f)) {
      bi = GTB_BETWEEN2;
      b[bi].keys.insert(b[bi].keys.end(), {q->key<1>(), q->key<2>()});
      b[bi].z.insert(b[bi].z.end(), {q->measured().x(), q->measured().y(), q->measured().theta()});
      appendNoise(b[bi], q->noiseModel(), 3);
    } else if (auto q3 = std::dynamic_pointer_cast<BetweenFactor<Pose3>>(f)) {
      bi = GTB_BETWEEN3;
      b[bi].keys.insert(b[bi].keys.end(), {q3->key<1>(), q3->key<2>()});
      appendPose3(b[bi].z, q3->measured());
      appendNoise(b[bi], q3->noiseModel(), 6);
    } else if (auto s = std::dynamic_pointer_cast<SfmFactor>(f)) {
      bi = GTB_SFM;
      b[bi].keys.insert(b[bi].keys.end(), {s->key<1>(), s->key<2>()});
      b[bi].z.insert(b[bi].z.end(), {s->measured().x(), s->measured().y()});
      appendNoise(b[bi], s->noiseModel(), 2);
    } else if (auto p2 = std::dynamic_pointer_cast<PriorFactor<Pose2>>(f)) {
      bi = GTB_PRIOR2;
      b[bi].keys.push_back(p2->key());
      b[bi].z.insert(b[bi].z.end(), {p2->prior().x(), p2->prior().y(), p2->prior().theta()});
      appendNoise(b[bi], p2->noiseModel(), 3);
    } else if (auto p3 = std::dynamic_pointer_cast<PriorFactor<Pose3>>(f)) {
      bi = GTB_PRIOR3;
      b[bi].keys.push_back(p3->key());
      appendPose3(b[bi].z, p3->prior());
      appendNoise(b[bi], p3->noiseModel(), 6);
    } else if (auto pc = std::dynamic_pointer_cast<PriorFactor<Camera>>(f)) {
      bi = GTB_PRIORCAM;
      b[bi].keys.push_back(pc->key());
      appendCamera(b[bi].z, pc->prior());
      appendNoise(b[bi], pc->noiseModel(), 9);
    } else if (auto pp = std::dynamic_pointer_cast<PriorFactor<Point3>>(f)) {
      bi = GTB_PRIORPT;
      b[bi].keys.push_back(pp->key());
      b[bi].z.insert(b[bi].z.end(), {pp->prior().x(), pp->prior().y(), pp->prior().z()});
      appendNoise(b[bi], pp->noiseModel(), 3);
    } else {
      throw std::invalid_argument("gtsam_b200: factor type without a device kernel (there is no CPU fallback)");
    }
    out.slots.emplace_back(bi, b[bi].n);
    b[bi].n++;
  }
  return out;
}

void DeviceProblem::uploadGraph(const NonlinearFactorGraph& g) {
  const PackedGraph pg = device_detail::packGraph(g);
  const Bucket* b = pg.bucket;
  std::vector<int64_t> order;
  graphSlots_.clear();
  for (const auto& s : pg.slots) {
    graphSlots_.push_back(Slot{s.first, s.second});
    if (s.first >= 0) {
      order.push_back(s.first);
      order.push_back((int64_t)s.second);
    }
  }
  bucketKeys_.assign(GTB_NUM_BUCKETS, {});
  for (int bi = 0; bi < GTB_NUM_BUCKETS; ++bi) {
    bucketCount_[bi] = b[bi].n;
    if (!b[bi].n) continue;
    bucketKeys_[bi].assign(b[bi].keys.begin(), b[bi].keys.end());
    check(gtb_set_factors(eng_, bi, (int64_t)b[bi].n, b[bi].keys.data(), b[bi].z.data(), b[bi].kind.data(), b[bi].noise.data()));
    if (b[bi].loss > 0) check(gtb_set_robust(eng_, bi, b[bi].loss, b[bi].lossK));
  }
  // the position of every factor in the graph: the clique layout and the summation order follow the graph order
  check(gtb_set_graph_order(eng_, (int64_t)order.size() / 2, order.data()));
}

Values device_detail::unpackValues(int t, const std::vector<Key>& keys, const double* rows) {
  Values out;
  for (size_t i = 0; i < keys.size(); ++i) {
    const double* p = rows + i * kStateOut[t];
    switch (t) {
      case GTB_POSE2: out.insert(keys[i], Pose2(Rot2::fromCosSin(p[0], p[1]), Point2(p[2], p[3]))); break;
      case GTB_POSE3: out.insert(keys[i], pose3From(p)); break;
      case GTB_CAMERA: out.insert(keys[i], Camera(pose3From(p), Cal3Bundler(p[12], p[13], p[14], p[15], p[16]))); break;
      default: out.insert(keys[i], Point3(p[0], p[1], p[2])); break;
    }
  }
  return out;
}

Values DeviceProblem::downloadValues(int which) const {
  Values out;
  for (int t = 0; t < GTB_NUM_VTYPES; ++t) {
    const size_t n = keys_[t].size();
    if (!n) continue;
    std::vector<double> s(n * kStateOut[t]);
    check(gtb_get_variables(eng_, t, which, s.data()));
    out.insert(device_detail::unpackValues(t, keys_[t], s.data()));
  }
  return out;
}

VectorValues DeviceProblem::downloadDelta() const {
  size_t n = 0;
  for (size_t d : sortedDims_) n += d;
  std::vector<double> d(n);
  check(gtb_get_delta(eng_, d.data()));
  VectorValues out;
  size_t off = 0;
  for (size_t i = 0; i < sortedKeys_.size(); ++i) {
    out.insert(sortedKeys_[i], Eigen::Map<const Vector>(d.data() + off, sortedDims_[i]));
    off += sortedDims_[i];
  }
  return out;
}

GaussianFactorGraph::shared_ptr DeviceProblem::downloadLinearGraph() const {
  std::vector<std::vector<double>> Ab(GTB_NUM_BUCKETS);
  for (int bi = 0; bi < GTB_NUM_BUCKETS; ++bi)
    if (bucketCount_[bi]) {
      Ab[bi].resize(bucketCount_[bi] * kResidualDim[bi] * kJacobianWidth[bi]);
      check(gtb_get_jacobians(eng_, bi, Ab[bi].data()));
    }
  auto out = std::make_shared<GaussianFactorGraph>();
  for (const Slot& s : graphSlots_) {
    if (s.bucket < 0) {
      out->push_back(GaussianFactor::shared_ptr());
      continue;
    }
    const int m = kResidualDim[s.bucket], w = kJacobianWidth[s.bucket], n0 = kFirstBlock[s.bucket];
    typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> RowMat;
    Eigen::Map<const RowMat> A(Ab[s.bucket].data() + s.index * m * w, m, w);
    const Key* k = bucketKeys_[s.bucket].data() + s.index * kBucketKeys[s.bucket];
    if (kBucketKeys[s.bucket] == 2)
      out->emplace_shared<JacobianFactor>(k[0], Matrix(A.leftCols(n0)), k[1], Matrix(A.middleCols(n0, w - 1 - n0)), Vector(A.col(w - 1)));
    else
      out->emplace_shared<JacobianFactor>(k[0], Matrix(A.leftCols(n0)), Vector(A.col(w - 1)));
  }
  return out;
}

// ---------------------------------------------------------------------------------------------------------------------
DeviceLevenbergMarquardtOptimizer::DeviceLevenbergMarquardtOptimizer(const NonlinearFactorGraph& graph, const Values& initialValues,
                                                                     const LevenbergMarquardtParams& params, int device)
    : LevenbergMarquardtOptimizer(graph, initialValues, params),   // computes params_.ordering once (COLAMD / METIS)
      dev_(std::make_shared<DeviceProblem>(graph, initialValues, *params_.ordering, device)) {}

DeviceLevenbergMarquardtOptimizer::DeviceLevenbergMarquardtOptimizer(const NonlinearFactorGraph& graph, const Values& initialValues,
                                                                     const Ordering& ordering, const LevenbergMarquardtParams& params,
                                                                     int device)
    : LevenbergMarquardtOptimizer(graph, initialValues, ordering, params),
      dev_(std::make_shared<DeviceProblem>(graph, initialValues, *params_.ordering, device)) {}

GaussianFactorGraph::shared_ptr DeviceLevenbergMarquardtOptimizer::iterate() {
  auto* st = static_cast<const internal::LevenbergMarquardtState*>(state_.get());
  const gtb_lm_params p = toParams(params_);
  double lambda = st->lambda, factor = st->currentFactor, err = st->error;
  int32_t inner = 0;
  dev_->check(gtb_lm_iterate(dev_->engine(), &p, &lambda, &factor, &err, &inner));
  // the iteration counter moves only when a step was accepted (LevenbergMarquardtState::decreaseLambda); a lambda search
  // that gave up leaves values and error unchanged
  const bool accepted = gtb_lm_last_accepted(dev_->engine()) != 0;
  Values next = accepted ? dev_->downloadValues(0) : st->values;
  state_.reset(new internal::LevenbergMarquardtState(std::move(next), err, lambda, factor, st->iterations + (accepted ? 1 : 0),
                                                      st->totalNumberInnerIterations + inner));
  return std::make_shared<GaussianFactorGraph>();
}

GaussianFactorGraph::shared_ptr DeviceLevenbergMarquardtOptimizer::linearize() const {
  dev_->check(gtb_linearize(dev_->engine()));
  return dev_->downloadLinearGraph();
}

VectorValues DeviceLevenbergMarquardtOptimizer::solveDamped(double lambda) const {
  uint64_t bad = 0;
  const gtb_status st = gtb_factorize(dev_->engine(), lambda, params_.diagonalDamping ? 1 : 0, params_.minDiagonal, params_.maxDiagonal, &bad);
  dev_->check(st, bad);
  dev_->check(gtb_backsolve(dev_->engine()));
  return dev_->downloadDelta();
}

VectorValues DeviceLevenbergMarquardtOptimizer::solve(const GaussianFactorGraph& gfg, const NonlinearOptimizerParams&) const {
  if (gfg.size() != graph_.size()) throw std::invalid_argument("DeviceLevenbergMarquardtOptimizer::solve: pass the graph linearize() returned");
  return solveDamped(0.0);
}

// ---------------------------------------------------------------------------------------------------------------------
DeviceGaussNewtonOptimizer::DeviceGaussNewtonOptimizer(const NonlinearFactorGraph& graph, const Values& initialValues,
                                                       const GaussNewtonParams& params, int device)
    : GaussNewtonOptimizer(graph, initialValues, params),
      dev_(std::make_shared<DeviceProblem>(graph, initialValues, *params_.ordering, device)) {}

GaussianFactorGraph::shared_ptr DeviceGaussNewtonOptimizer::iterate() {
  gtb_engine* e = dev_->engine();
  dev_->check(gtb_linearize(e));
  uint64_t bad = 0;
  dev_->check(gtb_factorize(e, 0.0, 0, 0.0, 0.0, &bad), bad);
  dev_->check(gtb_backsolve(e));
  dev_->check(gtb_retract(e));
  double err = 0;
  dev_->check(gtb_error(e, 1, &err));
  dev_->check(gtb_accept(e));
  Values next = dev_->downloadValues(0);
  state_.reset(new internal::NonlinearOptimizerState(std::move(next), err, state_->iterations + 1));
  return std::make_shared<GaussianFactorGraph>();
}

}  // namespace gtsam
