// integration/DeviceISAM2.h -- the reference-side binding of the ISAM2 layer of libgtsam_b200.so (include/gtsam_b200.h,
// gtb_isam2_*): a class with gtsam::ISAM2's update / calculateEstimate signatures (nonlinear/ISAM2.h:155-200, 233-246) that takes
// the reference's own NonlinearFactorGraph / Values / ISAM2Params objects.  INTEGRATION / TEST CODE, like DeviceOptimizers.h.
//
//   gtsam::ISAM2 isam(params);            ->   gtsam::DeviceISAM2 isam(params);
//   isam.update(newFactors, newTheta);         isam.update(newFactors, newTheta);
//   Values est = isam.calculateEstimate();     Values est = isam.calculateEstimate();
//
// Supported configuration: ISAM2GaussNewtonParams, a scalar relinearizeThreshold, enableRelinearization, relinearizeSkip.  The
// device layer always ends an update with the complete back-substitution, i.e. it behaves like wildfireThreshold = 0 (the setting
// of examples/ISAM2_City10000.cpp:66): with a positive threshold the reference's delta is the approximate one and the two
// estimates differ by up to that threshold.  Dogleg, per-type threshold maps, factor removal, constrained keys and marginal
// queries are not bound (std::invalid_argument / no method).
#pragma once
#include <gtsam/nonlinear/ISAM2.h>
#include <gtsam/nonlinear/NonlinearFactorGraph.h>
#include <gtsam/nonlinear/Values.h>

#include <vector>

#include "../include/gtsam_b200.h"

namespace gtsam {

class DeviceISAM2 {
 public:
  explicit DeviceISAM2(const ISAM2Params& params = ISAM2Params(), int device = 0);
  ~DeviceISAM2();
  DeviceISAM2(const DeviceISAM2&) = delete;
  DeviceISAM2& operator=(const DeviceISAM2&) = delete;

  /// ISAM2::update(newFactors, newTheta) (nonlinear/ISAM2.cpp:401-480).  Fills variablesRelinearized (the variables that moved
  /// their linearisation point), variablesReeliminated (all of them: the device re-eliminates the whole system) and cliques.
  ISAM2Result update(const NonlinearFactorGraph& newFactors = NonlinearFactorGraph(), const Values& newTheta = Values());
  /// ISAM2::calculateEstimate() (nonlinear/ISAM2.cpp:725-733): theta retracted by delta, for every variable.
  Values calculateEstimate() const;
  /// Elimination order for the variables known so far (optional; default: the library's nested dissection).  E.g.
  /// Ordering::Colamd(VariableIndex(graph)) of the whole graph every few hundred updates.
  void setOrdering(const Ordering& ordering);
  size_t size() const;   ///< number of variables

 private:
  void check(gtb_status st, Key nearby = 0) const;
  gtb_isam2* h_ = nullptr;
  std::vector<Key> keys_[GTB_NUM_VTYPES];   // per type, in the order the variables were added
};

}  // namespace gtsam
