/* gtsam_b200.h -- C ABI of the B200-native batch nonlinear least-squares engine.
 *
 * This is the drop-in boundary.  GTSAM has no FFI or plugin
 * registry of its own; the seams a maintainer binds are the two virtuals
 *   LevenbergMarquardtOptimizer::linearize()            nonlinear/LevenbergMarquardtOptimizer.h:113
 *   NonlinearOptimizer::solve(gfg, params)              nonlinear/NonlinearOptimizer.h:129-130
 * (or iterate(), NonlinearOptimizer.h:136).  Every entry point below cites the reference code whose work it
 * replaces.  INTEGRATION.md shows the C++ subclass that calls them.
 *
 * Conventions: plain C types only; all pointers are caller-owned HOST buffers unless noted; device memory
 * is owned by the engine; one CUDA stream per engine; an engine is not thread-safe; every function returns
 * a gtb_status and never throws; gtb_last_error() gives the message for the last non-OK status.
 * There is NO CPU fallback: numeric entry points on an engine created without a device return
 * GTB_ERR_NO_DEVICE.
 */
#ifndef GTSAM_B200_H
#define GTSAM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gtb_engine gtb_engine;

typedef enum {
  GTB_OK = 0,
  GTB_ERR_INVALID = 1,        /* bad argument / call order (std::invalid_argument in the reference)        */
  GTB_ERR_NO_DEVICE = 2,      /* numeric call on a host-only engine, or CUDA unavailable                    */
  GTB_ERR_CUDA = 3,           /* CUDA runtime error                                                          */
  GTB_ERR_INDETERMINATE = 4,  /* IndeterminantLinearSystemException   linear/linearExceptions.h:94          */
  GTB_ERR_KEY_MISSING = 5,    /* ValuesKeyDoesNotExist                 nonlinear/Values.h                    */
  GTB_ERR_NOISE_MODEL = 6,    /* constrained / unsupported noise model (no QR path on device)               */
  GTB_ERR_INCONSISTENT = 7    /* InconsistentEliminationRequested     inference/inferenceExceptions.h       */
} gtb_status;

/* Variable types = the Values payloads on the hot path. State layout per row, float64:
 *   POSE2  (x, y, theta)                    -> kept on device as (cos, sin, x, y), geometry/Pose2.h, Rot2.h
 *   POSE3  (R row-major 9, t 3)             geometry/Pose3.h, Rot3M.cpp
 *   CAMERA (R 9, t 3, f, k1, k2, u0, v0)    PinholeCamera<Cal3Bundler>, geometry/PinholeCamera.h
 *   POINT3 (x, y, z)                                                                                       */
enum { GTB_POSE2 = 0, GTB_POSE3 = 1, GTB_CAMERA = 2, GTB_POINT3 = 3, GTB_NUM_VTYPES = 4 };

/* Factor buckets (one launch per bucket replaces the per-factor virtual call of
 * NonlinearFactorGraph::linearize, nonlinear/NonlinearFactorGraph.cpp:239-278):
 *   BETWEEN2  BetweenFactor<Pose2>   slam/BetweenFactor.h:111     keys 2, z = Pose2 (x,y,theta),  m = 3
 *   BETWEEN3  BetweenFactor<Pose3>                                keys 2, z = Pose3 (R9,t3),      m = 6
 *   SFM       GeneralSFMFactor<PinholeCamera<Cal3Bundler>,Point3> slam/GeneralSFMFactor.h:141
 *                                                                 keys 2 (camera, point), z = (u,v), m = 2
 *   PRIOR2 / PRIOR3 / PRIORCAM / PRIORPT   PriorFactor<T>  nonlinear/PriorFactor.h:98   keys 1            */
enum { GTB_BETWEEN2 = 0, GTB_BETWEEN3 = 1, GTB_SFM = 2, GTB_PRIOR2 = 3, GTB_PRIOR3 = 4, GTB_PRIORCAM = 5,
       GTB_PRIORPT = 6, GTB_NUM_BUCKETS = 7 };

/* Noise models as the user constructs them (linear/NoiseModel.h); the engine applies the reference's
 * smart-factory downgrade Gaussian -> Diagonal -> Isotropic -> Unit (linear/NoiseModel.cpp:84-132,284-315,
 * 625-634).  `noise` holds d*d doubles per factor (d = residual dimension), row-major, left-packed.       */
enum { GTB_NOISE_UNIT = 0, GTB_NOISE_ISO_SIGMA = 1, GTB_NOISE_DIAG_SIGMAS = 2, GTB_NOISE_INFORMATION = 3,
       GTB_NOISE_SQRT_INFORMATION = 4, GTB_NOISE_DIAG_VARIANCES = 5 };

/* ---- lifecycle ------------------------------------------------------------------------------------ */
/* device >= 0: CUDA device ordinal.  device < 0: host-only engine (problem set-up and symbolic analysis). */
gtb_status gtb_create(int device, gtb_engine** out);
void gtb_destroy(gtb_engine* e);
const char* gtb_last_error(const gtb_engine* e);
const char* gtb_version(void);

/* ---- problem upload (replaces Values / NonlinearFactorGraph containers) ----- */
/* Values::insert for n variables of one type.  Calling again with the same n re-uploads the state only
 * (the per-step host->device copy): the rows are converted into a page-locked staging buffer and copied from there, so
 * `state` may be reused as soon as the call returns -- EXCEPT when `state` itself is page-locked (cudaHostAlloc /
 * cudaHostRegister) and its rows need no conversion (every type but POSE2): then the copy is enqueued straight from
 * `state`, which must stay unchanged until the next call that synchronises (gtb_trial, gtb_error, gtb_get_*, ...).   */
gtb_status gtb_set_variables(gtb_engine* e, int vtype, int64_t n, const uint64_t* keys, const double* state);
/* NonlinearFactorGraph::push_back of n factors of one bucket; keys is n x (2|1).                         */
gtb_status gtb_set_factors(gtb_engine* e, int bucket, int64_t n, const uint64_t* keys, const double* z,
                           const int32_t* noise_kind, const double* noise);
/* Optional: position of every factor in the graph, rows of (bucket, index).  Default: bucket-major.      */
gtb_status gtb_set_graph_order(gtb_engine* e, int64_t n, const int64_t* bucket_index_pairs);
/* Finish set-up: key -> row tables, device upload.  Must precede everything below.                       */
gtb_status gtb_finalize_problem(gtb_engine* e);

/* ---- symbolic analysis, host, once per ordering -------------------------------- */
/* Takes the elimination ordering (params.ordering, nonlinear/NonlinearOptimizerParams.h:107; the reference
 * computes it once in the optimizer constructor, LevenbergMarquardtParams.h:112-117) and builds what
 * eliminateMultifrontal rebuilds on every solve (inference/EliminateableFactorGraph-inst.h:131-137):
 * VariableIndex -> EliminationTree (EliminationTree-inst.h:78-156) -> JunctionTree with the supernode
 * merge rule (JunctionTree-inst.h:65-121, ClusterTree-inst.h:46-96) -> clique layout in Scatter order
 * (linear/Scatter.cpp:39-73).  The result is bit-identical to the reference's Bayes-tree structure.      */
gtb_status gtb_set_ordering(gtb_engine* e, int64_t n, const uint64_t* keys);
/* Orderings computed inside the library: 0 = natural (sorted keys, Ordering::Natural), 1 = Schur (all POINT3 then all
 * CAMERA, timing/timeSFMBAL.h:74-83), 2 = nested dissection on the variable adjacency graph (csrc/ordering.cpp: level-
 * structure separators, written for pose graphs; NOT the permutation CCOLAMD or METIS would give -- a caller that wants
 * the reference's own ordering computes it with Ordering::Colamd / Ordering::Metis and passes it to gtb_set_ordering,
 * as integration/DeviceOptimizers.cpp does).                                                             */
gtb_status gtb_set_ordering_builtin(gtb_engine* e, int which);
gtb_status gtb_plan_sizes(const gtb_engine* e, int64_t* n_cliques, int64_t* n_clique_keys,
                          int64_t* n_conditional_doubles, int64_t* n_scalars);
/* Cliques in DFS post-order (children in stored order): frontal count, total key count, keys
 * (frontals then separator sorted by key), parent index or -1.                                           */
gtb_status gtb_plan_get(const gtb_engine* e, int64_t* nfrontal, int64_t* nkeys, uint64_t* keys,
                        int64_t* parent);
/* Plan statistics: out[0]=levels, [1]=max front dim, [2]=sum flops (POTRF+TRSM+SYRK), [3]=front arena
 * doubles, [4]=small-front cliques, [5]=large-front cliques, [6]=flops in large fronts, [7]=assembly bytes,
 * [8]=flops of the frontal trailing updates of large fronts, [9]=flops of their Schur updates, [10]=algorithmic
 * bytes of back-substitution, [11]=algorithmic bytes of small-front elimination, [12]=algorithmic bytes of
 * linearize (SURVEY 8d formula), [13]=the part of [11] in thin fronts, [14]=algorithmic bytes of point cliques
 * (factor rows read, three conditional rows written), [15]=the part of [11] in dense small fronts            */
gtb_status gtb_plan_stats(const gtb_engine* e, double* out16);

/* ---- per outer iteration ---------------------------------------------------------------------------- */
/* NonlinearFactorGraph::linearize(values): whitened [A|b] of every factor into per-bucket device buffers. */
gtb_status gtb_linearize(gtb_engine* e);
/* Row-major m x (sum n_j + 1) per factor, as JacobianFactor::augmentedJacobian (parity dump).            */
gtb_status gtb_get_jacobians(gtb_engine* e, int bucket, double* out);
/* GaussianFactorGraph::hessianDiagonal (linear/GaussianFactorGraph.cpp:279-287); out may be NULL, layout
 * = all variables concatenated in ascending key order.                                                   */
gtb_status gtb_hessian_diagonal(gtb_engine* e, double* out);

/* ---- per lambda trial --------------------------------------------------------------------------------- */
/* buildDampedSystem + eliminateMultifrontal(EliminatePreferCholesky) (LevenbergMarquardtState.h:125-156,
 * linear/HessianFactor.cpp:515-551, base/cholesky.cpp:108-159).  lambda <= 0: no damping (Gauss-Newton).
 * diagonal_damping != 0: lambda * clamp(diag(H), min_diagonal, max_diagonal).
 * Returns GTB_ERR_INDETERMINATE with *failed_key = first frontal key of the failing clique.              */
gtb_status gtb_factorize(gtb_engine* e, double lambda, int diagonal_damping, double min_diagonal,
                         double max_diagonal, uint64_t* failed_key);
/* GaussianBayesTree::optimize (linear/linearAlgorithms-inst.h:50-155).                                   */
gtb_status gtb_backsolve(gtb_engine* e);
/* One whole tryLambda worth of device work with ONE host synchronisation (LevenbergMarquardtOptimizer.cpp:121-270):
 * gtb_factorize(lambda ...) + gtb_backsolve + gtb_linear_errors + gtb_retract + gtb_error(1) enqueued back to back.
 * out3 = linear.error(zero), linear.error(delta), graph.error(retracted values).  GTB_ERR_INDETERMINATE + *failed_key as
 * gtb_factorize.  The trial values stay in the scratch copy (gtb_accept makes them current).                        */
gtb_status gtb_trial(gtb_engine* e, double lambda, int diagonal_damping, double min_diagonal, double max_diagonal,
                     double* out3, uint64_t* failed_key);
/* delta for all variables concatenated in ascending key order (VectorValues iteration order).            */
gtb_status gtb_get_delta(gtb_engine* e, double* out);
/* Conditionals [R S d], row-major F x (N+1) per clique, cliques in post-order, concatenated.             */
gtb_status gtb_get_conditionals(gtb_engine* e, double* out);
/* linear.error(zero), linear.error(delta) on the UNDAMPED system (LevenbergMarquardtOptimizer.cpp:170-171) */
gtb_status gtb_linear_errors(gtb_engine* e, double* error_zero, double* error_delta);
/* Values::retract(delta) into the scratch copy (nonlinear/Values.cpp:53-63,99).                          */
gtb_status gtb_retract(gtb_engine* e);
/* NonlinearFactorGraph::error (nonlinear/NonlinearFactorGraph.cpp:170-179); which: 0 current, 1 scratch. */
gtb_status gtb_error(gtb_engine* e, int which, double* error);
/* Accept the trial: scratch becomes current.                                                             */
gtb_status gtb_accept(gtb_engine* e);
/* Download variables (which: 0 current, 1 scratch); POSE2 rows come back as (cos, sin, x, y).            */
gtb_status gtb_get_variables(gtb_engine* e, int vtype, int which, double* state);

/* ---- optimizer drivers (host control flow identical to the reference) ------------------------------ */
typedef struct {
  int32_t max_iterations;         /* NonlinearOptimizerParams.h:44   default 100  */
  double relative_error_tol;      /* 1e-5 */
  double absolute_error_tol;      /* 1e-5 */
  double error_tol;               /* 0    */
  double lambda_initial;          /* LevenbergMarquardtParams.h:69-82: 1e-5 */
  double lambda_factor;           /* 10   */
  double lambda_upper_bound;      /* 1e5  */
  double lambda_lower_bound;      /* 0    */
  double min_model_fidelity;      /* 1e-3 */
  int32_t diagonal_damping;       /* 0    */
  int32_t use_fixed_lambda_factor;/* 1    */
  double min_diagonal;            /* 1e-6 */
  double max_diagonal;            /* 1e32 */
} gtb_lm_params;

typedef struct {
  double initial_error, final_error, final_lambda;
  int32_t iterations, inner_iterations;
  int32_t log_rows;               /* rows written to the trace (see gtb_lm_optimize) */
} gtb_lm_result;

void gtb_lm_params_default(gtb_lm_params* p);        /* LevenbergMarquardtParams()                     */
void gtb_lm_params_ceres(gtb_lm_params* p);          /* SetCeresDefaults, LevenbergMarquardtParams.h:85 */
/* LevenbergMarquardtOptimizer::optimize() (NonlinearOptimizer.cpp:62-117, LevenbergMarquardtOptimizer.cpp:
 * 121-304).  trace (may be NULL) receives up to trace_rows rows of [inner, 0, error, lambda, outer], the
 * reference's logFile columns (LevenbergMarquardtOptimizer.cpp:104-118; time column left 0).            */
gtb_status gtb_lm_optimize(gtb_engine* e, const gtb_lm_params* p, gtb_lm_result* result, double* trace,
                           int64_t trace_rows);
/* One LevenbergMarquardtOptimizer::iterate() (one linearize + lambda search).  State persists in the
 * engine between calls; lambda is in/out.                                                                */
gtb_status gtb_lm_iterate(gtb_engine* e, const gtb_lm_params* p, double* lambda_io, double* lambda_factor_io,
                          double* error_io, int32_t* inner_iterations);
/* 1 when the last gtb_lm_iterate ended with an accepted step (LevenbergMarquardtState::decreaseLambda: the reference's
 * iteration counter moves only then, nonlinear/internal/LevenbergMarquardtState.h:81-94), 0 when the lambda search gave up. */
int gtb_lm_last_accepted(const gtb_engine* e);
/* GaussNewtonOptimizer::optimize() (nonlinear/GaussNewtonOptimizer.cpp:44-66).                          */
gtb_status gtb_gn_optimize(gtb_engine* e, int32_t max_iterations, double relative_error_tol,
                           double absolute_error_tol, gtb_lm_result* result);

/* ---- robust noise models -------------------------------------------------------------
 * noiseModel::Robust::Create(mEstimator, base) with the default Block reweighting (linear/NoiseModel.cpp:709-745,
 * linear/LossFunctions.cpp): every factor of `bucket` keeps the base noise given to gtb_set_factors and gets the
 * m-estimator `loss` with parameter k (> 0): the whitened [A|b] is scaled by sqrt(w(|b|)) and the factor's error becomes
 * rho(|b|).  Call after gtb_set_factors for that bucket; GTB_LOSS_NONE removes it.  One estimator per bucket.       */
#define GTB_LOSS_NONE 0
#define GTB_LOSS_FAIR 1
#define GTB_LOSS_HUBER 2
#define GTB_LOSS_CAUCHY 3
#define GTB_LOSS_TUKEY 4
#define GTB_LOSS_WELSCH 5
#define GTB_LOSS_GEMANMCCLURE 6
#define GTB_LOSS_DCS 7
#define GTB_LOSS_L2WITHDEADZONE 8
gtb_status gtb_set_robust(gtb_engine* e, int bucket, int loss, double k);

/* ---- dataset ingestion straight to struct-of-arrays --------------------------------------
 * Replaces the file readers that sit in front of the path: readG2o / load2D / load3D (slam/dataset.cpp:621-633,
 * :505-570, :922-946) and SfmData::FromBalFile (sfm/SfmData.cpp:189-247) plus the graph construction loop of
 * examples/SFMExample_bal.cpp:52-63.  A dataset owns flat tables in exactly the layout gtb_set_variables /
 * gtb_set_factors take (variables in ascending key order, factors in file / track order); gtb_dataset_apply pushes
 * them into an engine (the caller still adds priors and calls gtb_finalize_problem).  Errors: GTB_ERR_INVALID with the
 * message in gtb_dataset_last_error() (missing file, malformed record, landmark records that have no device type). */
typedef struct gtb_dataset gtb_dataset;
/* readG2o(file, is3D, kernelFunctionType): kernel = GTB_LOSS_NONE | GTB_LOSS_HUBER (k = 1.345) | GTB_LOSS_TUKEY
 * (k = 4.6851), 2D only, as in the reference.  Also reads the TORO tags (VERTEX2 / EDGE2 / VERTEX3 / EDGE3).          */
gtb_status gtb_dataset_read_g2o(const char* path, int is3d, int kernel, gtb_dataset** out);
/* SfmData::FromBalFile + one GeneralSFMFactor<Cal3Bundler> per measurement, Unit noise, cameras keyed 0..n-1, points
 * P(j) (tests/testGeneralSFMFactorB.cpp:44-58).                                                                      */
gtb_status gtb_dataset_read_bal(const char* path, gtb_dataset** out);
void gtb_dataset_free(gtb_dataset* d);
const char* gtb_dataset_last_error(void);
int64_t gtb_dataset_num_variables(const gtb_dataset* d, int vtype);
const uint64_t* gtb_dataset_variable_keys(const gtb_dataset* d, int vtype);
const double* gtb_dataset_variable_state(const gtb_dataset* d, int vtype);
int64_t gtb_dataset_num_factors(const gtb_dataset* d, int bucket);
const uint64_t* gtb_dataset_factor_keys(const gtb_dataset* d, int bucket);
const double* gtb_dataset_factor_z(const gtb_dataset* d, int bucket);
const int32_t* gtb_dataset_factor_noise_kind(const gtb_dataset* d, int bucket);
const double* gtb_dataset_factor_noise(const gtb_dataset* d, int bucket);
int gtb_dataset_robust(const gtb_dataset* d, double* k);      /* m-estimator the file was read with (between factors) */
gtb_status gtb_dataset_apply(const gtb_dataset* d, gtb_engine* e);
/* writeG2o (slam/dataset.cpp:636-735): VERTEX_SE2 / VERTEX_SE3:QUAT for the poses, EDGE_SE2 / EDGE_SE3:QUAT for the
 * between factors; *_information holds R^T R of each factor's Gaussian model (d x d, row-major, GTSAM's (R, t) order
 * for Pose3; written in g2o's (t, R) order).  Default stream formatting, as the reference.                            */
gtb_status gtb_write_g2o(const char* path, int64_t n_pose2, const uint64_t* pose2_keys, const double* pose2,
                         int64_t n_pose3, const uint64_t* pose3_keys, const double* pose3,
                         int64_t n_between2, const uint64_t* between2_keys, const double* between2_z,
                         const double* between2_information,
                         int64_t n_between3, const uint64_t* between3_keys, const double* between3_z,
                         const double* between3_information);

/* ---- marginal covariances from the device Bayes tree ------------------------------------
 * Marginals(graph, values).marginalCovariance(key) (nonlinear/Marginals.cpp:35-150): the reference re-linearises,
 * eliminates into a Bayes tree and, per query, builds the marginal factor through shortcuts (inference/BayesTree-inst.h).
 * Here: gtb_linearize + gtb_factorize(lambda = 0) as usual, then gtb_marginals_compute sweeps the resident [R S d]
 * blocks once from the roots to the leaves (selective inversion), after which any number of d x d covariance blocks
 * can be read.  `out` receives the blocks back to back (dim(key)^2 doubles each, symmetric) in the order of `keys`.
 * Errors: GTB_ERR_KEY_MISSING for an unknown key; GTB_ERR_INVALID before a successful factorisation or with
 * several ranks.                                                                                                    */
gtb_status gtb_marginals_compute(gtb_engine* e);
gtb_status gtb_marginal_covariance(gtb_engine* e, int64_t n, const uint64_t* keys, double* out);

/* ---- ISAM2 update ---------------------------------------------------------------
 * gtsam::ISAM2::update(newFactors, newTheta) + calculateEstimate() (nonlinear/ISAM2.cpp:401-480, 725-784) for the
 * Gauss-Newton configuration of examples/ISAM2_City10000.cpp:62-68 (ISAM2GaussNewtonParams(0.0): wildfire threshold 0, so
 * every update ends in a complete back-substitution and delta is the exact solution of the current linear system).
 * What is kept of the reference: the split of the estimate into a linearisation point theta and a delta, fluid
 * relinearisation (a variable moves its linearisation point only when |delta_j|_inf >= relinearize_threshold, checked every
 * relinearize_skip updates, ISAM2-impl.h:341-351; the check and the masked retract run on the device:
 * gtb_relinearize_marked), factors linearised at theta (what the reference's cached linear factors amount to,
 * ISAM2.cpp:176-230), new variables entering with delta = 0.  With that the estimate after every update equals the
 * reference's (tests/test_isam2_gpu.py: 1e-9 on tests/golden/isam2__city10000_200loops.npz).
 * What is NOT kept: the reference re-eliminates only the top of the Bayes tree that the new factors and the relinearised
 * variables touch and re-uses the orphaned subtrees (ISAM2.cpp:117-175, 250-363); this layer re-plans and re-eliminates the
 * WHOLE system on the device at every update (batch engine underneath), so an update costs one symbolic analysis on the
 * host plus one multifrontal factorisation, not work proportional to the affected part.  DESIGN.md 8 says what an
 * incremental plan would take.
 * Usage: stage the new variables and factors (same row layouts as gtb_set_variables / gtb_set_factors; a key may be staged
 * once), then gtb_isam2_update.  Ordering: gtb_isam2_set_ordering gives an elimination order for the variables known so
 * far (the reference computes a constrained COLAMD per update, ISAM2.cpp:301-340; any order gives the same estimate);
 * variables that are not listed are eliminated last in the order they were added.  Without it: the built-in nested dissection (gtb_set_ordering_builtin(2)), recomputed every 8 updates (GTB_ISAM2_REORDER_EVERY), the variables added in between eliminated last. */
typedef struct gtb_isam2 gtb_isam2;
typedef struct {
  int64_t update_count;            /* ISAM2::update_count_                                            */
  int64_t variables_relinearized;  /* variables that moved their linearisation point in this update   */
  int64_t n_variables, n_factors;  /* totals after the update                                         */
  int64_t n_cliques;               /* cliques of the Bayes tree of this update                        */
  double plan_ms;                  /* host: engine set-up + symbolic analysis                         */
  double solve_ms;                 /* device: linearise + factorise + back-substitute + retract, with the sync */
} gtb_isam2_result;
gtb_status gtb_isam2_create(int device, double relinearize_threshold, int32_t relinearize_skip, gtb_isam2** out);
void gtb_isam2_destroy(gtb_isam2* s);
const char* gtb_isam2_last_error(const gtb_isam2* s);
gtb_status gtb_isam2_add_variables(gtb_isam2* s, int vtype, int64_t n, const uint64_t* keys, const double* state);
gtb_status gtb_isam2_add_factors(gtb_isam2* s, int bucket, int64_t n, const uint64_t* keys, const double* z,
                                 const int32_t* noise_kind, const double* noise);
gtb_status gtb_isam2_set_ordering(gtb_isam2* s, int64_t n, const uint64_t* keys);
/* Commits what was staged and runs the update.  GTB_ERR_INDETERMINATE + *failed_key as gtb_factorize (the staged
 * additions stay committed, as in the reference, whose update throws after pushing the factors).              */
gtb_status gtb_isam2_update(gtb_isam2* s, gtb_isam2_result* result, uint64_t* failed_key);
/* calculateEstimate(): theta (+) delta for every variable of one type, in the order the variables were added, device
 * row layout (gtb_get_variables).  n = number of variables of that type (gtb_isam2_count).                    */
gtb_status gtb_isam2_estimate(gtb_isam2* s, int vtype, double* state);
int64_t gtb_isam2_count(const gtb_isam2* s, int vtype);
/* The batch engine of the last update (linearised at theta, factorised, solved): gtb_get_delta, gtb_get_conditionals,
 * gtb_marginals_compute ... work on it.  Owned by the ISAM2 object, replaced by the next update.             */
gtb_engine* gtb_isam2_engine(gtb_isam2* s);
/* ISAM2 steps 4 and 6 on the device (ISAM2.cpp:456-468): every variable of `e` with |delta_j|_inf >= threshold takes its
 * retracted value as the new current value (Values::retractMasked).  Needs gtb_retract.  marks (may be NULL): one byte
 * per variable, types concatenated POSE2, POSE3, CAMERA, POINT3 in upload order.                              */
gtb_status gtb_relinearize_marked(gtb_engine* e, double threshold, uint8_t* marks, int64_t* n_marked);

/* ---- several GPUs: one engine per process / GPU ---------------------------------------------
 * The reference parallelises elimination over the children of a clique with TBB tasks
 * (inference/ClusterTree-inst.h:286-318).  Across GPUs the same tree is cut instead: every rank holds the whole
 * problem and the whole plan, eliminates the subtrees it owns, adds their Schur complements into its copy of the
 * fronts at the top of the tree, and the top fronts are summed onto rank 0 (fan-in), which factors them.
 * Back-substitution runs the other way: rank 0 solves the top, the solution is broadcast, every rank solves its
 * subtrees and the owned pieces of delta are summed so that each replica can retract.
 * The library does not link a communication library: the caller supplies the three collectives below
 * (ncclReduce / ncclBroadcast / ncclAllReduce on `stream` in a GTSAM build; torch.distributed in bench.py).
 * `buf` is a device pointer to `count` doubles; the call must be ordered on `stream` (a cudaStream_t) or complete
 * before it returns.  Return 0 on success.                                                               */
#define GTB_COMM_REDUCE_SUM_TO_0 0
#define GTB_COMM_BROADCAST_FROM_0 1
#define GTB_COMM_ALLREDUCE_SUM 2
#define GTB_COMM_ALLREDUCE_MIN 3
typedef int (*gtb_comm_fn)(void* user, int op, void* buf, int64_t count, void* stream);
/* Call after gtb_finalize_problem and before gtb_set_ordering*.  world = 1 restores the single-GPU schedule.
 * comm may be NULL on an engine without a device (host-side partition tests).                            */
gtb_status gtb_set_distributed(gtb_engine* e, int32_t rank, int32_t world, gtb_comm_fn comm, void* user);
/* Owner rank of every clique (post-order ids as in gtb_plan_get); -1 = top of the tree (fan-in, rank 0).
 * stats: out[0] = number of top cliques, out[1] = doubles reduced per factorisation, out[2] = cost share of the top
 * (model), out[3 + r] for r < world (at most 13) = cost share of rank r's subtrees.                            */
gtb_status gtb_partition_get(const gtb_engine* e, int32_t* owner, double* out16);
/* Host-only self-check of the static launch schedule that replaces the reference's task-parallel tree walk
 * (inference/ClusterTree-inst.h:286-318, base/treeTraversal-inst.h:156-168): every clique this rank owns is eliminated
 * exactly once and solved exactly once, every member of a launch class fits the class's kernel and shared memory, the
 * panel-step / Schur-step task lists, the inverse-block tasks and the camera-pair segments tile their buffers.  Works
 * on an engine without a device.  GTB_ERR_INCONSISTENT + gtb_last_error on a violated invariant.  out8 (may be NULL):
 * launch classes, panel-step launches, Schur-step launches, back-substitution launches, inverse blocks, camera-pair
 * segments, same-camera segments, panel + Schur tasks.                                                            */
gtb_status gtb_schedule_check(const gtb_engine* e, int64_t* out8);

/* ---- allocation-free entry point -------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * A variant of the nonlinear-error pass behind gtb_error (error_kernel + sum_partials_kernel): the error
 * sum over the n factors of ONE bucket, 0.5 * sum || whiten(h(x) - z) ||^2, whose per-CTA partial sums go to a fixed
 * file-scope __device__ array (256 KB per GPU that loads the library, held until exit) instead of engine memory.  The
 * call allocates nothing (no cudaMalloc / cudaFree / allocating library routine), only enqueues two kernels on `stream`
 * and returns without synchronising.  ALL pointers are caller-owned DEVICE pointers on `device`:
 *   state0 / state1   type arrays of key 0 / key 1 in the device layout (POSE2 rows are (cos, sin, x, y); others as above);
 *                     state1, row1 = NULL for the prior buckets
 *   row0 / row1       int32 row of each factor's key in its type array
 *   z                 measurements, device layout (BETWEEN2 / PRIOR2: (cos, sin, x, y); others as in gtb_set_factors)
 *   noise_kind, noise both NULL = every factor Unit; else RESOLVED kinds per factor (0 unit, 1 isotropic: noise[0] = 1/sigma,
 *                     2 diagonal: noise[0..d) = 1/sigma_i, 3 full: upper-triangular sqrt information, row-major) and d*d
 *                     doubles per factor
 *   error_out         one double
 * Limit: n <= GTB_NOALLOC_MAX_FACTORS, larger counts are rejected with GTB_ERR_INVALID (message in
 * gtb_noalloc_last_error()).  Every call shares the one scratch array, which keeps the last call's contents (nothing in
 * it needs to start at zero: each partial read is written by the same call first); calls that OVERLAP on the GPU
 * (different streams without a dependency, or several host threads) race on it -- order them.
 * gtb_noalloc_setup(device) resolves the array's device address (cudaGetSymbolAddress, once per device); the first call
 * on a device does it implicitly, so call it beforehand if the first use is under stream capture.                    */
#define GTB_NOALLOC_MAX_FACTORS (1 << 22)
gtb_status gtb_noalloc_setup(int device);
gtb_status gtb_bucket_error_noalloc(int device, int bucket, int64_t n, const double* state0, const double* state1,
                                    const int32_t* row0, const int32_t* row1, const double* z, const int8_t* noise_kind,
                                    const double* noise, double* error_out, void* stream);
const char* gtb_noalloc_last_error(void);

/* ---- instrumentation ------------------------------------------------------------------------------- */
/* Number of engine kernels launched since creation.                                                      */
int64_t gtb_kernel_launches(const gtb_engine* e);
/* Device time (ms, CUDA events on the engine stream) of the last call per stage:
 * out[0]=linearize, [1]=assemble+factorize, [2]=backsolve, [3]=linear errors, [4]=retract, [5]=error,
 * [6]=large-front factorisation share of [1], [7]=hessian diagonal                                       */
gtb_status gtb_stage_ms(gtb_engine* e, double* out8);
/* Per-stage event timing synchronises the stream after every stage; turn it off (on = 0) for throughput runs. */
gtb_status gtb_set_stage_timing(gtb_engine* e, int on);
/* Per-launch CUDA-event timing by kernel class (on the engine stream).  gtb_profile(e, 1) starts recording,
 * gtb_profile_read sums elapsed ms and launch counts per class since the last read (arrays of
 * gtb_profile_classes() entries) and clears the record.                                                   */
gtb_status gtb_profile(gtb_engine* e, int on);
int gtb_profile_classes(void);
const char* gtb_profile_class_name(int k);
gtb_status gtb_profile_read(gtb_engine* e, double* total_ms, int64_t* launches);
/* Every recorded launch in order: class id and elapsed ms; returns the number written (<= cap) and clears.  */
int64_t gtb_profile_read_raw(gtb_engine* e, int32_t* klass, float* ms, int64_t cap);
/* Raw stream handle (cudaStream_t) for callers that time with their own events.                          */
void* gtb_stream(gtb_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* GTSAM_B200_H */
