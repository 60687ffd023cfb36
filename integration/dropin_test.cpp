// integration/dropin_test.cpp -- TEST PROGRAM for the drop-in optimizers (DeviceOptimizers.h): runs the bodies of the
// reference's own examples twice -- stock optimizer and device optimizer, same graph, same initial values, ordering from
// the reference's own Ordering::Create (COLAMD or METIS) -- and compares iteration counts, final error and Values.
//
//   examples/Pose2SLAMExample_g2o.cpp:49-79   readG2o + prior on the first pose + GaussNewtonOptimizer (and LM here too)
//   examples/Pose3SLAMExample_g2o.cpp:42-55   readG2o(is3D) + prior + GaussNewtonOptimizer
//   examples/SFMExample_bal.cpp:44-81         SfmData::FromBalFile + GeneralSFMFactor<Cal3Bundler> + two priors + LM
//   tests/testNonlinearOptimizer.cpp:507-551  the subclassing precedent: linearize() / solve() used stage by stage
//
//   dropin_test <data dir>      (tests/golden/data: copies of the reference's examples/Data files)
// Prints one PASS / FAIL line per case; exit code = number of failures.  Built by oracle/Makefile (target dropin) against
// oracle/_ref/libgtsam_ref.so (the unmodified reference) and gtsam_petercdev_b200/libgtsam_b200.so; needs a CUDA device.
#include <gtsam/geometry/Cal3Bundler.h>
#include <gtsam/geometry/PinholeCamera.h>
#include <gtsam/inference/Symbol.h>
#include <gtsam/linear/linearExceptions.h>
#include <gtsam/nonlinear/PriorFactor.h>
#include <gtsam/sfm/SfmData.h>
#include <gtsam/slam/BetweenFactor.h>
#include <gtsam/slam/GeneralSFMFactor.h>
#include <gtsam/slam/dataset.h>

#include <cmath>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <iostream>

#include "DeviceISAM2.h"
#include "DeviceOptimizers.h"

#include <gtsam/nonlinear/internal/LevenbergMarquardtState.h>

using namespace gtsam;
using symbol_shorthand::P;
typedef PinholeCamera<Cal3Bundler> Camera;
typedef GeneralSFMFactor<Camera, Point3> SfmFactor;

static int g_fail = 0;
static void report(const std::string& name, bool ok, const std::string& detail) {
  std::printf("%s %s: %s\n", ok ? "PASS" : "FAIL", name.c_str(), detail.c_str());
  std::fflush(stdout);
  if (!ok) ++g_fail;
}

// largest difference between two Values in local coordinates, relative to 1 + |value|
static double valuesDiff(const Values& a, const Values& b) {
  double worst = 0;
  for (const auto& kv : a) {
    const Vector d = kv.value.localCoordinates_(b.at(kv.key));
    worst = std::max(worst, d.cwiseAbs().maxCoeff());
  }
  return worst;
}

template <class STOCK, class DEVICE, class PARAMS>
static void compareOptimizers(const std::string& name, const NonlinearFactorGraph& graph, const Values& initial, const PARAMS& params) {
  STOCK stock(graph, initial, params);
  const Values expected = stock.optimize();
  DEVICE dev(graph, initial, params);
  const Values actual = dev.optimize();
  const double rel = std::abs(dev.error() - stock.error()) / std::max(stock.error(), 1e-300);
  const double vd = valuesDiff(expected, actual);
  char buf[256];
  std::snprintf(buf, sizeof buf, "iterations %zu vs %zu, final error %.9g vs %.9g (rel %.2e), max |Values difference| %.2e", dev.iterations(),
                stock.iterations(), dev.error(), stock.error(), rel, vd);
  report(name, dev.iterations() == stock.iterations() && rel <= 1e-6 && vd <= 1e-6 && actual.size() == expected.size(), buf);
}

static NonlinearFactorGraph withPosePrior2(const NonlinearFactorGraph& graph, const Values& initial) {
  NonlinearFactorGraph g = graph;   // examples/Pose2SLAMExample_g2o.cpp:63-67
  auto priorModel = noiseModel::Diagonal::Variances(Vector3(1e-6, 1e-6, 1e-8));
  g.addPrior(initial.keys().front(), initial.at<Pose2>(initial.keys().front()), priorModel);
  return g;
}

int main(int argc, char** argv) {
  if (argc < 2) {
    std::fprintf(stderr, "usage: dropin_test <data dir>\n");
    return 2;
  }
  const std::string data = argv[1];
  try {
    // ---- Pose2SLAMExample_g2o: Gauss-Newton (the example) and Levenberg-Marquardt, COLAMD and METIS ----
    {
      NonlinearFactorGraph::shared_ptr graph;
      Values::shared_ptr initial;
      std::tie(graph, initial) = readG2o(data + "/noisyToyGraph.txt", false);
      const NonlinearFactorGraph g = withPosePrior2(*graph, *initial);
      GaussNewtonParams gn;
      gn.setVerbosity("SILENT");
      compareOptimizers<GaussNewtonOptimizer, DeviceGaussNewtonOptimizer>("Pose2SLAMExample_g2o noisyToyGraph GN colamd", g, *initial, gn);
      gn.orderingType = Ordering::METIS;
      compareOptimizers<GaussNewtonOptimizer, DeviceGaussNewtonOptimizer>("Pose2SLAMExample_g2o noisyToyGraph GN metis", g, *initial, gn);
      LevenbergMarquardtParams lm;
      compareOptimizers<LevenbergMarquardtOptimizer, DeviceLevenbergMarquardtOptimizer>("Pose2SLAMExample_g2o noisyToyGraph LM colamd", g, *initial, lm);
    }
    {
      NonlinearFactorGraph::shared_ptr graph;
      Values::shared_ptr initial;
      std::tie(graph, initial) = readG2o(data + "/pose2example.txt", false);
      const NonlinearFactorGraph g = withPosePrior2(*graph, *initial);
      GaussNewtonParams gn;
      compareOptimizers<GaussNewtonOptimizer, DeviceGaussNewtonOptimizer>("Pose2SLAMExample_g2o pose2example GN colamd", g, *initial, gn);
      LevenbergMarquardtParams lm;
      lm.orderingType = Ordering::METIS;
      compareOptimizers<LevenbergMarquardtOptimizer, DeviceLevenbergMarquardtOptimizer>("Pose2SLAMExample_g2o pose2example LM metis", g, *initial, lm);
    }
    // ---- Pose3SLAMExample_g2o ----
    {
      NonlinearFactorGraph::shared_ptr graph;
      Values::shared_ptr initial;
      std::tie(graph, initial) = readG2o(data + "/pose3example.txt", true);
      NonlinearFactorGraph g = *graph;   // examples/Pose3SLAMExample_g2o.cpp:42-48
      auto priorModel = noiseModel::Diagonal::Variances((Vector(6) << 1e-6, 1e-6, 1e-6, 1e-4, 1e-4, 1e-4).finished());
      Key first = 0;
      for (const auto key : initial->keys()) { first = key; break; }
      g.addPrior(first, Pose3(), priorModel);
      GaussNewtonParams gn;
      compareOptimizers<GaussNewtonOptimizer, DeviceGaussNewtonOptimizer>("Pose3SLAMExample_g2o pose3example GN colamd", g, *initial, gn);
      LevenbergMarquardtParams lm;
      compareOptimizers<LevenbergMarquardtOptimizer, DeviceLevenbergMarquardtOptimizer>("Pose3SLAMExample_g2o pose3example LM colamd", g, *initial, lm);
    }
    // ---- SFMExample_bal ----
    {
      const SfmData mydata = SfmData::FromBalFile(data + "/dubrovnik-3-7-pre.txt");
      auto noise = noiseModel::Isotropic::Sigma(2, 1.0);   // one pixel in u and v (becomes Unit under the smart rule)
      NonlinearFactorGraph graph;
      size_t j = 0;
      for (const SfmTrack& track : mydata.tracks) {
        for (const auto& m : track.measurements) graph.emplace_shared<SfmFactor>(m.second, noise, m.first, P(j));
        j += 1;
      }
      graph.addPrior(0, mydata.cameras[0], noiseModel::Isotropic::Sigma(9, 0.1));
      graph.addPrior(P(0), mydata.tracks[0].p, noiseModel::Isotropic::Sigma(3, 0.1));
      Values initial;
      size_t i = 0;
      j = 0;
      for (const SfmCamera& camera : mydata.cameras) initial.insert(i++, camera);
      for (const SfmTrack& track : mydata.tracks) initial.insert(P(j++), track.p);
      LevenbergMarquardtParams lm;
      compareOptimizers<LevenbergMarquardtOptimizer, DeviceLevenbergMarquardtOptimizer>("SFMExample_bal dubrovnik-3-7-pre LM colamd", graph, initial, lm);
      lm.orderingType = Ordering::METIS;
      compareOptimizers<LevenbergMarquardtOptimizer, DeviceLevenbergMarquardtOptimizer>("SFMExample_bal dubrovnik-3-7-pre LM metis", graph, initial, lm);
      LevenbergMarquardtParams ceres;
      LevenbergMarquardtParams::SetCeresDefaults(&ceres);
      compareOptimizers<LevenbergMarquardtOptimizer, DeviceLevenbergMarquardtOptimizer>("SFMExample_bal dubrovnik-3-7-pre LM ceres", graph, initial, ceres);

      // ---- the seams one by one (tests/testNonlinearOptimizer.cpp:507-551): linearize() and solve() ----
      LevenbergMarquardtParams p;
      DeviceLevenbergMarquardtOptimizer dev(graph, initial, p);
      const GaussianFactorGraph::shared_ptr ours = dev.linearize();
      const GaussianFactorGraph::shared_ptr ref = graph.linearize(initial);
      double worst = 0;
      bool shape = ours->size() == ref->size();
      for (size_t f = 0; shape && f < ref->size(); ++f) {
        const Matrix A = ref->at(f)->augmentedJacobian(), B = ours->at(f)->augmentedJacobian();
        shape = A.rows() == B.rows() && A.cols() == B.cols() && ref->at(f)->keys() == ours->at(f)->keys();
        if (shape) worst = std::max(worst, (A - B).cwiseAbs().maxCoeff() / std::max(1.0, A.cwiseAbs().maxCoeff()));
      }
      char buf[200];
      std::snprintf(buf, sizeof buf, "%zu factors, max |A_dev - A_ref| / max|A_ref| = %.2e", ref->size(), worst);
      report("linearize() returns the reference's JacobianFactors", shape && worst < 1e-12, buf);

      // damped solve of the same linearisation: reference = buildDampedSystem + eliminateMultifrontal + optimize
      const double lambda = 1e-4;
      const VectorValues dDev = dev.solveDamped(lambda);
      internal::LevenbergMarquardtState st(initial, 0.0, lambda, 10.0);
      const VectorValues dRef = st.buildDampedSystem(*ref).optimize(*dev.params().ordering, EliminatePreferCholesky);
      const double num = (dDev - dRef).norm(), den = dRef.norm();
      std::snprintf(buf, sizeof buf, "lambda = %g: |delta_dev - delta_ref| / |delta_ref| = %.2e", lambda, num / den);
      report("solveDamped() equals buildDampedSystem + eliminateMultifrontal", num / den < 1e-7, buf);

      // an underdetermined problem must surface as the reference's exception from solve()
      NonlinearFactorGraph noPrior;
      for (const auto& f : graph)
        if (std::dynamic_pointer_cast<SfmFactor>(f)) noPrior.push_back(f);
      DeviceLevenbergMarquardtOptimizer gauge(noPrior, initial, p);
      bool threw = false;
      try {
        const auto lin = gauge.linearize();
        gauge.solve(*lin, p);
      } catch (const IndeterminantLinearSystemException& e) {
        threw = true;
      }
      bool threwRef = false;
      try {
        noPrior.linearize(initial)->optimize(*gauge.params().ordering, EliminatePreferCholesky);
      } catch (const IndeterminantLinearSystemException& e) {
        threwRef = true;
      }
      report("solve() throws IndeterminantLinearSystemException where the reference does", threw == threwRef,
             std::string("device ") + (threw ? "threw" : "did not throw") + ", reference " + (threwRef ? "threw" : "did not throw"));
    }
    // ---- ISAM2: the schedule of examples/ISAM2_City10000.cpp:62-134 on the stock gtsam::ISAM2 and on DeviceISAM2 ----
    {
      std::ifstream in(data + "/T1_city10000_04_head1264.txt");
      ISAM2Params params;
      params.optimizationParams = ISAM2GaussNewtonParams(0.0);
      params.relinearizeThreshold = 0.01;
      params.relinearizeSkip = 1;
      ISAM2 stock(params);
      DeviceISAM2 dev(params);
      const auto priorModel = noiseModel::Diagonal::Sigmas(Vector3(1e-4, 1e-4, 1e-4));
      const auto odoModel = noiseModel::Diagonal::Sigmas(Vector3(1.0 / 30.0, 1.0 / 30.0, 1.0 / 100.0));
      const auto loopModel = noiseModel::Diagonal::Sigmas(Vector3(10.0, 10.0, 10.0));
      NonlinearFactorGraph fresh;
      Values freshValues;
      freshValues.insert(Symbol('x', 0), Pose2(0, 0, 0));
      fresh.addPrior<Pose2>(Symbol('x', 0), Pose2(0, 0, 0), priorModel);
      stock.update(fresh, freshValues);
      dev.update(fresh, freshValues);
      Values est = stock.calculateEstimate(), estDev = dev.calculateEstimate();
      double worst = 0;
      size_t updates = 0, relinStock = 0, relinDev = 0;
      std::string line;
      while (updates < 400 && std::getline(in, line)) {
        std::istringstream ls(line);
        std::string tag;
        long long from, to, d0, d1, nh;
        double x, y, th;
        if (!(ls >> tag >> from >> d0 >> to >> d1 >> nh >> x >> y >> th)) continue;
        const Pose2 z(x, y, th);
        fresh.resize(0);
        freshValues.clear();
        if (from == to - 1) {
          // both arms start the new pose from the STOCK estimate, so that they receive identical inputs
          freshValues.insert(Symbol('x', to), est.at<Pose2>(Symbol('x', from)) * z);
          fresh.emplace_shared<BetweenFactor<Pose2>>(Symbol('x', from), Symbol('x', to), z, odoModel);
        } else {
          fresh.emplace_shared<BetweenFactor<Pose2>>(Symbol('x', from), Symbol('x', to), z, loopModel);
        }
        stock.update(fresh, freshValues);
        const ISAM2Result r = dev.update(fresh, freshValues);
        relinDev += r.variablesRelinearized;
        est = stock.calculateEstimate();
        estDev = dev.calculateEstimate();
        ++updates;
        if (updates % 50 == 0 || updates == 400) worst = std::max(worst, valuesDiff(est, estDev));
      }
      char buf[200];
      std::snprintf(buf, sizeof buf, "%zu updates, %zu poses, max |estimate_dev - estimate_ref| = %.2e (bar 1e-9), %zu relinearisations on the device",
                    updates, dev.size(), worst, relinDev);
      report("DeviceISAM2 follows gtsam::ISAM2 update by update (City10000 schedule)", updates == 400 && est.size() == estDev.size() && worst < 1e-9, buf);
      (void)relinStock;
      // the same in 3D with full Gaussian information matrices: examples/Data/pose3example.txt fed edge by edge
      {
        const auto gv = readG2o(data + "/pose3example.txt", true);
        const NonlinearFactorGraph& g3 = *gv.first;
        const Values& init3 = *gv.second;
        ISAM2 stock3(params);
        DeviceISAM2 dev3(params);
        KeySet known;
        NonlinearFactorGraph f3;
        Values v3;
        const Key first = init3.keys().front();
        v3.insert(first, init3.at<Pose3>(first));
        f3.addPrior(first, init3.at<Pose3>(first), noiseModel::Diagonal::Variances((Vector(6) << 1e-6, 1e-6, 1e-6, 1e-4, 1e-4, 1e-4).finished()));
        known.insert(first);
        stock3.update(f3, v3);
        dev3.update(f3, v3);
        double worst3 = 0;
        size_t n3 = 0;
        for (const auto& f : g3) {
          f3.resize(0);
          v3.clear();
          for (Key k : f->keys())
            if (!known.count(k)) {
              v3.insert(k, init3.at<Pose3>(k));
              known.insert(k);
            }
          f3.push_back(f);
          stock3.update(f3, v3);
          dev3.update(f3, v3);
          worst3 = std::max(worst3, valuesDiff(stock3.calculateEstimate(), dev3.calculateEstimate()));
          ++n3;
        }
        // a few more updates without new factors: both keep relinearising towards the same optimum
        for (int it = 0; it < 3; ++it) {
          stock3.update();
          dev3.update();
          worst3 = std::max(worst3, valuesDiff(stock3.calculateEstimate(), dev3.calculateEstimate()));
        }
        std::snprintf(buf, sizeof buf, "%zu + 3 updates, %zu Pose3, max |estimate_dev - estimate_ref| = %.2e (bar 1e-9)", n3, dev3.size(), worst3);
        report("DeviceISAM2 follows gtsam::ISAM2 in 3D (pose3example.txt, Gaussian noise)", worst3 < 1e-9, buf);
      }
      // unsupported configurations are refused
      bool threw = false;
      try {
        ISAM2Params dl;
        dl.optimizationParams = ISAM2DoglegParams();
        DeviceISAM2 bad(dl);
      } catch (const std::invalid_argument&) {
        threw = true;
      }
      report("DeviceISAM2 refuses Dogleg parameters", threw, threw ? "std::invalid_argument" : "accepted");
    }
    // ---- unsupported input is rejected loudly (no CPU fallback) ----
    {
      NonlinearFactorGraph g;
      Values v;
      v.insert(0, Pose2(0, 0, 0));
      g.addPrior(0, Pose2(0, 0, 0), noiseModel::Constrained::All(3));
      bool threw = false;
      try {
        DeviceGaussNewtonOptimizer opt(g, v);
      } catch (const InvalidNoiseModel&) {
        threw = true;
      }
      report("constrained noise model is rejected with InvalidNoiseModel", threw, threw ? "thrown" : "not thrown");
    }
  } catch (const std::exception& e) {
    report("uncaught exception", false, e.what());
  }
  std::printf("%s: %d failure(s)\n", g_fail ? "FAILED" : "OK", g_fail);
  return g_fail;
}
