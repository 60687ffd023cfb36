// integration/DeviceISAM2.cpp -- see DeviceISAM2.h.  Links the reference (Values, NonlinearFactorGraph, ISAM2Params / ISAM2Result
// types) and libgtsam_b200.so; packs what an update adds into the flat tables of gtb_isam2_add_variables / gtb_isam2_add_factors.
#include "DeviceISAM2.h"

#include <gtsam/linear/linearExceptions.h>

#include <stdexcept>
#include <string>
#include <variant>

#include "DeviceOptimizers.h"

namespace gtsam {

static const int kStateOut[GTB_NUM_VTYPES] = {4, 12, 17, 3};

void DeviceISAM2::check(gtb_status st, Key nearby) const {
  if (st == GTB_OK) return;
  const std::string msg = h_ ? gtb_isam2_last_error(h_) : "gtsam_b200 error";
  switch (st) {
    case GTB_ERR_INDETERMINATE: throw IndeterminantLinearSystemException(nearby);
    case GTB_ERR_KEY_MISSING: throw ValuesKeyDoesNotExist("gtsam_b200", nearby);
    case GTB_ERR_NO_DEVICE: throw std::runtime_error("gtsam_b200: no CUDA device (there is no CPU fallback): " + msg);
    case GTB_ERR_INVALID: throw std::invalid_argument("gtsam_b200: " + msg);
    default: throw std::runtime_error("gtsam_b200: " + msg);
  }
}

DeviceISAM2::DeviceISAM2(const ISAM2Params& params, int device) {
  if (!std::holds_alternative<ISAM2GaussNewtonParams>(params.optimizationParams))
    throw std::invalid_argument("gtsam_b200: DeviceISAM2 implements the Gauss-Newton variant (ISAM2GaussNewtonParams) only");
  const double* beta = std::get_if<double>(&params.relinearizeThreshold);
  if (!beta) throw std::invalid_argument("gtsam_b200: DeviceISAM2 takes a scalar relinearizeThreshold");
  // relinearisation switched off = a threshold no delta reaches
  const double threshold = params.enableRelinearization ? *beta : 1e300;
  if (gtb_isam2_create(device, threshold, (int32_t)std::max(1, params.relinearizeSkip), &h_) != GTB_OK)
    throw std::invalid_argument("gtsam_b200: bad ISAM2 parameters");
}

DeviceISAM2::~DeviceISAM2() {
  if (h_) gtb_isam2_destroy(h_);
}

ISAM2Result DeviceISAM2::update(const NonlinearFactorGraph& newFactors, const Values& newTheta) {
  const device_detail::PackedValues pv = device_detail::packValues(newTheta);
  for (int t = 0; t < GTB_NUM_VTYPES; ++t) {
    if (pv.keys[t].empty()) continue;
    std::vector<uint64_t> k(pv.keys[t].begin(), pv.keys[t].end());
    check(gtb_isam2_add_variables(h_, t, (int64_t)k.size(), k.data(), pv.state[t].data()), pv.keys[t].front());
    keys_[t].insert(keys_[t].end(), pv.keys[t].begin(), pv.keys[t].end());
  }
  const device_detail::PackedGraph pg = device_detail::packGraph(newFactors);
  for (int b = 0; b < GTB_NUM_BUCKETS; ++b) {
    const device_detail::Bucket& bk = pg.bucket[b];
    if (!bk.n) continue;
    if (bk.loss > 0) throw std::invalid_argument("gtsam_b200: DeviceISAM2 does not take robust noise models");
    check(gtb_isam2_add_factors(h_, b, (int64_t)bk.n, bk.keys.data(), bk.z.data(), bk.kind.data(), bk.noise.data()));
  }
  gtb_isam2_result r;
  uint64_t failed = 0;
  const gtb_status st = gtb_isam2_update(h_, &r, &failed);
  check(st, (Key)failed);
  ISAM2Result out;
  out.variablesRelinearized = (size_t)r.variables_relinearized;
  out.variablesReeliminated = (size_t)r.n_variables;
  out.factorsRecalculated = (size_t)r.n_factors;
  out.cliques = (size_t)r.n_cliques;
  return out;
}

Values DeviceISAM2::calculateEstimate() const {
  Values out;
  for (int t = 0; t < GTB_NUM_VTYPES; ++t) {
    const size_t n = keys_[t].size();
    if (!n) continue;
    std::vector<double> rows(n * kStateOut[t]);
    check(gtb_isam2_estimate(h_, t, rows.data()));
    out.insert(device_detail::unpackValues(t, keys_[t], rows.data()));
  }
  return out;
}

void DeviceISAM2::setOrdering(const Ordering& ordering) {
  std::vector<uint64_t> k(ordering.begin(), ordering.end());
  check(gtb_isam2_set_ordering(h_, (int64_t)k.size(), k.data()));
}

size_t DeviceISAM2::size() const {
  size_t n = 0;
  for (int t = 0; t < GTB_NUM_VTYPES; ++t) n += keys_[t].size();
  return n;
}

}  // namespace gtsam
