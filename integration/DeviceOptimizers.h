// integration/DeviceOptimizers.h -- the reference-side binding of libgtsam_b200.so: drop-in subclasses of GTSAM's own
// optimizers that keep the reference's signatures and run the iteration on the device through the C ABI
// (include/gtsam_b200.h).  INTEGRATION / TEST CODE: it links the reference by nature (oracle/_ref/libgtsam_ref.so in this
// repo, libgtsam in a GTSAM build); the numeric library itself stays free of it.
//
//   gtsam::LevenbergMarquardtOptimizer opt(graph, initial, params);          // before
//   gtsam::DeviceLevenbergMarquardtOptimizer opt(graph, initial, params);    // after: same params, same methods
//   Values result = opt.optimize();
//
// Seams bound: NonlinearOptimizer::iterate() (nonlinear/NonlinearOptimizer.h:136),
// LevenbergMarquardtOptimizer::linearize() (nonlinear/LevenbergMarquardtOptimizer.h:113), NonlinearOptimizer::solve()
// (nonlinear/NonlinearOptimizer.h:129-130).  The reference's own precedent for subclassing these is
// tests/testNonlinearOptimizer.cpp:507-551 (IterativeLM).  The ordering is the reference's own: the base-class
// constructors compute params_.ordering with Ordering::Create (COLAMD / METIS, LevenbergMarquardtParams.h:112-117,
// GaussNewtonOptimizer.cpp:70-75) and it is handed to gtb_set_ordering unchanged.
#pragma once
#include <gtsam/nonlinear/GaussNewtonOptimizer.h>
#include <gtsam/nonlinear/LevenbergMarquardtOptimizer.h>
#include <gtsam/nonlinear/NonlinearFactorGraph.h>
#include <gtsam/nonlinear/Values.h>

#include <memory>
#include <vector>

#include "../include/gtsam_b200.h"

namespace gtsam {

/// Packing of the reference's containers into the flat tables of the C ABI, shared by the batch optimizers and DeviceISAM2.
namespace device_detail {
struct PackedValues {
  std::vector<Key> keys[GTB_NUM_VTYPES];       // per type, in Values order
  std::vector<double> state[GTB_NUM_VTYPES];   // rows as gtb_set_variables takes them
};
struct Bucket {
  std::vector<uint64_t> keys;
  std::vector<double> z, noise;
  std::vector<int32_t> kind;
  int loss = -1;       // -1: no factor seen yet; GTB_LOSS_*
  double lossK = 0;
  size_t n = 0;
};
struct PackedGraph {
  Bucket bucket[GTB_NUM_BUCKETS];
  std::vector<std::pair<int, size_t>> slots;   // graph position -> (bucket, index); bucket < 0: null factor
};
PackedValues packValues(const Values& v);
PackedGraph packGraph(const NonlinearFactorGraph& g);
/// rows as gtb_get_variables returns them (POSE2: cos, sin, x, y) -> Values
Values unpackValues(int vtype, const std::vector<Key>& keys, const double* rows);
}  // namespace device_detail

/// NonlinearFactorGraph + Values packed into the engine's flat tables, and results unpacked into GTSAM objects.
class DeviceProblem {
 public:
  DeviceProblem(const NonlinearFactorGraph& graph, const Values& values, const Ordering& ordering, int device);
  ~DeviceProblem();
  DeviceProblem(const DeviceProblem&) = delete;
  DeviceProblem& operator=(const DeviceProblem&) = delete;

  gtb_engine* engine() const { return eng_; }
  /// Values of the engine's current (which = 0) or trial (which = 1) state, with the keys and types of the input.
  Values downloadValues(int which = 0) const;
  /// delta of the last back-substitution as VectorValues (one Vector per key, ascending keys).
  VectorValues downloadDelta() const;
  /// The whitened factors of the last gtb_linearize as a real GaussianFactorGraph of JacobianFactors, graph order.
  GaussianFactorGraph::shared_ptr downloadLinearGraph() const;
  /// Maps a status to the reference's exception types (linear/linearExceptions.h:94-134, nonlinear/Values.h).
  void check(gtb_status st, Key nearby = 0) const;

 private:
  struct Slot { int bucket; size_t index; };
  void uploadValues(const Values& v);
  void uploadGraph(const NonlinearFactorGraph& g);

  gtb_engine* eng_ = nullptr;
  std::vector<Key> keys_[GTB_NUM_VTYPES];      // per type, in upload order
  std::vector<Key> sortedKeys_;                // all keys ascending, with their tangent dimension
  std::vector<size_t> sortedDims_;
  std::vector<Slot> graphSlots_;               // graph position -> (bucket, index); bucket < 0: null factor
  std::vector<std::vector<Key>> bucketKeys_;   // per bucket: keys of every factor (flattened)
  size_t bucketCount_[GTB_NUM_BUCKETS] = {0, 0, 0, 0, 0, 0, 0};
};

/// LevenbergMarquardtOptimizer whose iterate() -- linearize, lambda search with damped multifrontal Cholesky,
/// back-substitution, error evaluations, retract -- runs on the device (gtb_lm_iterate).  optimize(), values(), error(),
/// iterations(), lambda(), getInnerIterations() and checkConvergence stay the reference's.
class DeviceLevenbergMarquardtOptimizer : public LevenbergMarquardtOptimizer {
 public:
  DeviceLevenbergMarquardtOptimizer(const NonlinearFactorGraph& graph, const Values& initialValues,
                                    const LevenbergMarquardtParams& params = LevenbergMarquardtParams(), int device = 0);
  DeviceLevenbergMarquardtOptimizer(const NonlinearFactorGraph& graph, const Values& initialValues, const Ordering& ordering,
                                    const LevenbergMarquardtParams& params = LevenbergMarquardtParams(), int device = 0);

  /// One outer iteration on the device.  Returns an EMPTY graph: defaultOptimize() ignores the return value; callers that
  /// want the linearised factors call linearize().
  GaussianFactorGraph::shared_ptr iterate() override;
  /// The whitened Jacobians of the current values, computed on the device and returned as JacobianFactors.
  GaussianFactorGraph::shared_ptr linearize() const override;
  /// Undamped multifrontal Cholesky solve of the device's current linearisation (call linearize() first).  `gfg` must be
  /// the graph linearize() returned: its contents are not read, the factors are already resident on the device.
  VectorValues solve(const GaussianFactorGraph& gfg, const NonlinearOptimizerParams& params) const override;
  /// The same with Levenberg-Marquardt damping (buildDampedSystem + solve of the reference's tryLambda).
  VectorValues solveDamped(double lambda) const;

  const DeviceProblem& device() const { return *dev_; }

 private:
  std::shared_ptr<DeviceProblem> dev_;
};

/// GaussNewtonOptimizer whose iterate() runs on the device (GaussNewtonOptimizer.cpp:44-66).
class DeviceGaussNewtonOptimizer : public GaussNewtonOptimizer {
 public:
  DeviceGaussNewtonOptimizer(const NonlinearFactorGraph& graph, const Values& initialValues,
                             const GaussNewtonParams& params = GaussNewtonParams(), int device = 0);
  GaussianFactorGraph::shared_ptr iterate() override;
  const DeviceProblem& device() const { return *dev_; }

 private:
  std::shared_ptr<DeviceProblem> dev_;
};

}  // namespace gtsam
