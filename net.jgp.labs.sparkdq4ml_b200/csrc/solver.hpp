// solver.hpp -- host-side (driver) half of LinearRegression.fit: what Spark runs on the driver JVM after / between
// the distributed aggregations.  No CUDA in here; everything is O(d) .. O(d^3) on vectors of d or d+1 doubles.
//
// Follows (algorithms, not code) org.apache.spark.ml.optim.{WeightedLeastSquares,NormalEquationSolver} and
// org.apache.spark.ml.regression.LinearRegression 2.4.4, and breeze.optimize.{LBFGS,OWLQN,BacktrackingLineSearch,
// StrongWolfeLineSearch,CachedDiffFunction} 0.13.2 -- the jars behind `lr.fit(df)` at
// DataQuality4MachineLearningApp.java:120-126 of the reference (pom.xml:13-32 pins the versions).
#pragma once

#include <cstdint>
#include <functional>
#include <string>
#include <vector>

namespace dq4ml {

using Vec = std::vector<double>;

struct LrParams {
  int max_iter = 100;
  double reg_param = 0.0;
  double elastic_net = 0.0;
  double tol = 1e-6;
  bool fit_intercept = true;
  bool standardization = true;
  int solver = 0;  // 0 auto, 1 normal, 2 l-bfgs
  int loss = 0;    // 0 squaredError, 1 huber
  double epsilon = 1.35;
};

struct LrModel {
  Vec coefficients;
  double intercept = 0.0;
  Vec objective_history;
  std::string solver_used;
  int data_passes = 0;
  int64_t num_rows = 0;
  // WeightedLeastSquaresModel.diagInvAtWA: diagonal of (A^T W A)^-1 in original units (features, then the intercept);
  // filled on the Cholesky path only, like Spark.  Empty = "No Std. Error of coefficients available".
  Vec diag_inv_atwa;
  double scale = 1.0;  // LinearRegressionModel.scale: the huber loss's sigma; 1.0 for squaredError
};

// Diagonal of the inverse of a packed upper SPD matrix (LAPACK dpptri, diagonal only).  Returns false if not SPD.
bool packed_spd_inverse_diag(Vec ap, int k, Vec* diag);

// Two-sided p-value of a Student-t statistic: 2 (1 - cdf_dof(|t|)), as LinearRegressionSummary.pValues.
double student_t_two_sided_p(double t, double dof);

struct SolveError {
  int code;  // dq4ml_status
  std::string what;
};

// Objective handed to the quasi-Newton drivers.  `x` is passed mutable on purpose: Spark's NormalEquationCostFun
// rewrites the intercept slot of the optimiser's own vector during evaluation, and the iterates depend on it.
using CostFn = std::function<double(Vec& x, Vec& grad)>;
// Per-coordinate L1 strength (Breeze: l1reg: Int => Double).  Empty function = plain L-BFGS.
using L1Fn = std::function<double(int)>;

struct QnResult {
  Vec x;
  Vec adjusted_values;  // one per optimiser state, initial state included
  int evaluations = 0;  // distinct points the objective was evaluated at
  std::string stop_reason;
};

// Breeze LBFGS(maxIter, m = 10, tol) when l1 is empty, else OWLQN(maxIter, 10, l1, tol).
QnResult quasi_newton_minimize(const CostFn& f, Vec x0, int max_iter, double tol, const L1Fn& l1, int memory = 10);

// Breeze LBFGSB(lowerBounds, upperBounds, maxIter, m = 10, tol): what LinearRegression runs for loss = huber.
QnResult lbfgsb_minimize(const CostFn& f, Vec x0, const Vec& lower, const Vec& upper, int max_iter, double tol, int memory = 10);

// LAPACK dppsv('U') equivalent on a packed upper, column-major SPD matrix; b is overwritten by the solution.
// Returns 0, or j+1 when the leading minor of order j+1 is not positive definite.
int packed_cholesky_solve(Vec& ap, Vec& b);

// WeightedLeastSquares.fit after its treeAggregate.  moments = [count,wSum,wwSum,bSum,bbSum,aSum[d],abSum[d],aaSum[tri]].
// Throws SolveError.
LrModel wls_fit_from_moments(const double* moments, int d, const LrParams& p);

// Summarizer read-out handed to the l-bfgs branch of LinearRegression.train.
struct ColumnStats {
  double count = 0;
  Vec mean, variance;  // d features then the label at index d; variance is the SAMPLE variance
};

// One evaluation of the least-squares objective on the data: fills loss_sum, weight_sum, grad_sum[d].
using LsqPassFn = std::function<void(const Vec& coef, const Vec& fstd, const Vec& fmean, double ystd, double ymean,
                                      double& loss_sum, double& weight_sum, Vec& grad_sum)>;

// LinearRegression.train, l-bfgs branch, after the summarizer pass.  Throws SolveError.
LrModel lbfgs_path_fit(const ColumnStats& st, int d, const LrParams& p, const LsqPassFn& pass);

// One evaluation of the huber objective on the data at params = [coefficients (d), intercept (if fitted), sigma]:
// fills loss_sum, weight_sum and grad_sum (same layout as params; HuberAggregator's raw sums).
using HuberPassFn = std::function<void(const Vec& params, const Vec& fstd, double& loss_sum, double& weight_sum, Vec& grad_sum)>;

// LinearRegression.train with loss = huber, after the summarizer pass.  Throws SolveError.
LrModel huber_path_fit(const ColumnStats& st, int d, const LrParams& p, const HuberPassFn& pass);

}  // namespace dq4ml
