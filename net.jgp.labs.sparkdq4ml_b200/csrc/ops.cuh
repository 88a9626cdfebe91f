// ops.cuh -- the per-row bodies plugged into stream_reduce_kernel.  Each one is the seqOp of a Spark ML 2.4.4
// aggregator reached from lr.fit / model.summary (DataQuality4MachineLearningApp.java:126,129,132-139), with w = 1.
//
// Contract with the kernel: x[j] and y arrive already zeroed for rows the DQ rules dropped (keep == false) and for
// j >= d; `c` holds Op::NCONST doubles the kernel preloaded from device memory.
#pragma once

#include "device_utils.cuh"

namespace dq4ml {

// WeightedLeastSquares.Aggregator.add: count, bSum, bbSum, aSum[d], abSum[d], aaSum (packed upper, column-major).
// wSum and wwSum equal count when every weight is 1, so they are not accumulated separately.
template <int DMAX>
struct WlsOp {
  static constexpr int TRI = DMAX * (DMAX + 1) / 2;
  static constexpr int NACC = 3 + 2 * DMAX + TRI;
  static constexpr int NCONST = 0;
  static constexpr bool HAS_OUT = false;
  __host__ __device__ static constexpr double identity(int) { return 0.0; }
  __host__ __device__ static constexpr int kind(int) { return kSum; }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double*, const double (&x)[DMAX], double y,
                                        bool keep, int) const {
    acc[0] += keep ? 1.0 : 0.0;
    acc[1] += y;
    acc[2] = fma(y, y, acc[2]);
#pragma unroll
    for (int j = 0; j < DMAX; ++j) {
      acc[3 + j] += x[j];
      acc[3 + DMAX + j] = fma(x[j], y, acc[3 + DMAX + j]);
    }
#pragma unroll
    for (int j = 0; j < DMAX; ++j)
#pragma unroll
      for (int i = 0; i <= j; ++i) acc[3 + 2 * DMAX + i + j * (j + 1) / 2] = fma(x[i], x[j], acc[3 + 2 * DMAX + i + j * (j + 1) / 2]);
    return 0.0;
  }
};

// LeastSquaresAggregator.add: diff = x.(c/sigma) - y/sigma_y + offset; lossSum += diff^2/2; gradSum_j += diff * x_j
// (the 1/sigma_j factor of Spark's per-row update is applied once on the host: w = 1 makes it a common factor).
template <int DMAX>
struct LsqOp {
  static constexpr int NACC = 2 + DMAX;
  static constexpr int NCONST = 0;
  static constexpr bool HAS_OUT = false;
  double eff[DMAX];
  double offset;
  double inv_ystd;
  __host__ __device__ static constexpr double identity(int) { return 0.0; }
  __host__ __device__ static constexpr int kind(int) { return kSum; }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double*, const double (&x)[DMAX], double y,
                                        bool keep, int) const {
    double dotp = 0.0;
#pragma unroll
    for (int j = 0; j < DMAX; ++j) dotp = fma(x[j], eff[j], dotp);
    double diff = dotp - y * inv_ystd + offset;
    diff = keep ? diff : 0.0;
    acc[0] = fma(0.5 * diff, diff, acc[0]);
    acc[1] += keep ? 1.0 : 0.0;
#pragma unroll
    for (int j = 0; j < DMAX; ++j) acc[2 + j] = fma(diff, x[j], acc[2 + j]);
    return 0.0;
  }
};

// MultivariateOnlineSummarizer, the part LinearRegression.train needs (mean and variance of every feature and of the
// label).  Spark updates a running mean per value (one division each); here each thread keeps shifted power sums
// S1 = sum(v - s), S2 = sum((v - s)^2) around a provisional centre s taken from the head of the column, which is as
// stable as Welford when s is within a few sigma of the mean and costs two FMAs.  Column DMAX is the label.
template <int DMAX>
struct SummCoreOp {
  static constexpr int NACC = 1 + 2 * (DMAX + 1);
  static constexpr int NCONST = DMAX + 1;  // the centres
  static constexpr bool HAS_OUT = false;
  __host__ __device__ static constexpr double identity(int) { return 0.0; }
  __host__ __device__ static constexpr int kind(int) { return kSum; }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double* c, const double (&x)[DMAX], double y,
                                        bool keep, int) const {
    acc[0] += keep ? 1.0 : 0.0;
#pragma unroll
    for (int j = 0; j < DMAX; ++j) {
      const double t = keep ? x[j] - c[j] : 0.0;
      acc[1 + 2 * j] += t;
      acc[2 + 2 * j] = fma(t, t, acc[2 + 2 * j]);
    }
    const double t = keep ? y - c[DMAX] : 0.0;
    acc[1 + 2 * DMAX] += t;
    acc[2 + 2 * DMAX] = fma(t, t, acc[2 + 2 * DMAX]);
    return 0.0;
  }
};

// Full MultivariateOnlineSummarizer read-out for a group of up to DMAX columns: shifted sums, min, max, numNonzeros,
// normL1, sum of squares.  Used by dq4ml_summarize only (the fit uses SummCoreOp).
template <int DMAX>
struct SummFullOp {
  static constexpr int PER = 7;
  static constexpr int NACC = 1 + PER * DMAX;
  static constexpr int NCONST = DMAX;
  static constexpr bool HAS_OUT = false;
  __host__ __device__ static constexpr int kind(int k) {
    return k == 0 ? kSum : ((k - 1) % PER == 2 ? kMin : ((k - 1) % PER == 3 ? kMax : kSum));
  }
  __host__ __device__ static constexpr double identity(int k) {
    return reduce_identity(kind(k));
  }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double* c, const double (&x)[DMAX], double,
                                        bool keep, int d) const {
    acc[0] += keep ? 1.0 : 0.0;
#pragma unroll
    for (int j = 0; j < DMAX; ++j) {
      const bool on = keep && j < d;
      const double v = x[j];
      const double t = on ? v - c[j] : 0.0;
      double* a = &acc[1 + PER * j];
      a[0] += t;
      a[1] = fma(t, t, a[1]);
      a[2] = combine(kMin, a[2], on ? v : __builtin_huge_val());
      a[3] = combine(kMax, a[3], on ? v : -__builtin_huge_val());
      a[4] += (on && v != 0.0) ? 1.0 : 0.0;
      a[5] += fabs(v);
      a[6] = fma(v, v, a[6]);
    }
    return 0.0;
  }
};

// Second half of the LeastSquaresAggregator for wide tables: gradSum_j += r * x_j over a group of columns, where the
// per-row scaled difference r was written by the wide row pass and arrives here as the "label" column.
template <int DMAX>
struct GradOp {
  static constexpr int NACC = DMAX;
  static constexpr int NCONST = 0;
  static constexpr bool HAS_OUT = false;
  __host__ __device__ static constexpr double identity(int) { return 0.0; }
  __host__ __device__ static constexpr int kind(int) { return kSum; }

  // This pass runs without the DQ columns: a row the rules dropped is recognised by its difference, which the row pass
  // wrote as exactly 0, and its features (possibly NaN/Inf garbage, 0 * NaN = NaN) are kept out of the sums.  A kept row
  // whose difference is exactly 0 contributes 0 either way.
  __device__ __forceinline__ double row(double (&acc)[NACC], const double*, const double (&x)[DMAX], double y, bool,
                                        int) const {
    const bool live = y != 0.0;
#pragma unroll
    for (int j = 0; j < DMAX; ++j) acc[j] = fma(y, live ? x[j] : 0.0, acc[j]);
    return 0.0;
  }
};

// ---- instance weights (LinearRegression.setWeightCol) ---------------------------------------------------------------
// The weighted ops take the weight as column 0 of the table and the features as columns 1..d, so that the weight rides
// the same vector loads as everything else and the unweighted kernels keep their register budget.  x[0] arrives as 0
// for dropped rows, which removes them from every weighted sum.

// MultivariateOnlineSummarizer.add(sample, weight): a zero-weight row is skipped entirely (not even counted); mean =
// sum(w v) / W and variance = sum(w (v - mean)^2) / (W - sum(w^2) / W), read out on the host from the shifted sums.
// acc: count, W, sum w^2, then (S1, S2) for feature columns 1..DMAX-1, then for the label.  Centres: slot j of the
// table, label at slot DMAX (slot 0, the weight's own head mean, is unused).
template <int DMAX>
struct SummCoreWOp {
  static constexpr int NACC = 3 + 2 * DMAX;
  static constexpr int NCONST = DMAX + 1;
  static constexpr bool HAS_OUT = false;
  __host__ __device__ static constexpr double identity(int) { return 0.0; }
  __host__ __device__ static constexpr int kind(int) { return kSum; }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double* c, const double (&x)[DMAX], double y,
                                        bool keep, int) const {
    const double w = x[0];
    const bool on = keep && w != 0.0;
    acc[0] += on ? 1.0 : 0.0;
    acc[1] += w >= 0.0 ? w : __longlong_as_double(0x7ff8000000000000LL);  // require(weight >= 0.0): a bad weight poisons W
    acc[2] = fma(w, w, acc[2]);
#pragma unroll
    for (int j = 1; j < DMAX; ++j) {
      const double t = on ? x[j] - c[j] : 0.0;
      const double wt = w * t;
      acc[1 + 2 * j] += wt;
      acc[2 + 2 * j] = fma(wt, t, acc[2 + 2 * j]);
    }
    const double t = on ? y - c[DMAX] : 0.0;
    const double wt = w * t;
    acc[1 + 2 * DMAX] += wt;
    acc[2 + 2 * DMAX] = fma(wt, t, acc[2 + 2 * DMAX]);
    return 0.0;
  }
};

// LeastSquaresAggregator.add with weights: lossSum += w diff^2 / 2, weightSum += w, gradSum_j += w diff x_j (eff[0] is
// 0 so the weight column stays out of the margin; accumulator 2 is unused).
template <int DMAX>
struct LsqWOp {
  static constexpr int NACC = 2 + DMAX;
  static constexpr int NCONST = 0;
  static constexpr bool HAS_OUT = false;
  double eff[DMAX];
  double offset;
  double inv_ystd;
  __host__ __device__ static constexpr double identity(int) { return 0.0; }
  __host__ __device__ static constexpr int kind(int) { return kSum; }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double*, const double (&x)[DMAX], double y,
                                        bool keep, int) const {
    double dotp = 0.0;
#pragma unroll
    for (int j = 1; j < DMAX; ++j) dotp = fma(x[j], eff[j], dotp);
    double diff = dotp - y * inv_ystd + offset;
    diff = keep ? diff : 0.0;
    const double wd = x[0] * diff;
    acc[0] = fma(0.5 * wd, diff, acc[0]);
    acc[1] += x[0];
#pragma unroll
    for (int j = 1; j < DMAX; ++j) acc[2 + j] = fma(wd, x[j], acc[2 + j]);
    return 0.0;
  }
};

// LinearRegressionSummary under a weight column: coefficientStandardErrors take rss = sum(w (label - prediction)^2) and
// devianceResiduals the extremes of sqrt(w) (label - prediction).  Table: column 0 = weight, column 1 = residual.
template <int DMAX>
struct WResidOp {
  static constexpr int NACC = 3;
  static constexpr int NCONST = 0;
  static constexpr bool HAS_OUT = false;
  __host__ __device__ static constexpr int kind(int k) { return k == 1 ? kMin : (k == 2 ? kMaxNan : kSum); }
  __host__ __device__ static constexpr double identity(int k) {
    return k == 1 ? __builtin_huge_val() : (k == 2 ? -__builtin_huge_val() : 0.0);
  }
  __device__ __forceinline__ double row(double (&acc)[NACC], const double*, const double (&x)[DMAX], double, bool keep,
                                        int) const {
    const double w = x[0], r = DMAX > 1 ? x[DMAX > 1 ? 1 : 0] : 0.0;
    acc[0] = fma(w * r, r, acc[0]);
    const double s = sqrt(w) * r;
    acc[1] = combine(kMin, acc[1], keep ? s : __builtin_huge_val());
    acc[2] = combine(kMaxNan, acc[2], keep ? s : -__builtin_huge_val());
    return 0.0;
  }
};

// ml.optim.aggregator.HuberAggregator.add (Spark 2.3+, LinearRegression.setLoss("huber")): the label is NOT standardised,
// features enter as value / featuresStd.  margin = x.(c / std) + intercept, r = label - margin; inside |r| <= sigma eps
// the row adds (sigma + r^2 / sigma) / 2 to the loss, -(r / sigma) x_j / std_j to the coefficient gradients, -(r / sigma)
// to the intercept's and (1 - (r / sigma)^2) / 2 to sigma's; outside it adds (sigma + 2 eps |r| - sigma eps^2) / 2,
// -+eps x_j / std_j, -+eps and (1 - eps^2) / 2.  Accumulators: loss, count, d/d intercept, d/d sigma, then the column
// sums of factor * x_j (the host applies the common 1 / std_j).
template <int DMAX>
struct HuberOp {
  static constexpr int NACC = 4 + DMAX;
  static constexpr int NCONST = 0;
  static constexpr bool HAS_OUT = false;
  double eff[DMAX];
  double intercept, sigma, epsilon;
  __host__ __device__ static constexpr double identity(int) { return 0.0; }
  __host__ __device__ static constexpr int kind(int) { return kSum; }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double*, const double (&x)[DMAX], double y,
                                        bool keep, int) const {
    double margin = intercept;
#pragma unroll
    for (int j = 0; j < DMAX; ++j) margin = fma(x[j], eff[j], margin);
    const double r = y - margin, rs = r / sigma;
    const bool inner = fabs(r) <= sigma * epsilon;
    const double loss = inner ? 0.5 * (sigma + r * rs) : 0.5 * (sigma + 2.0 * epsilon * fabs(r) - sigma * epsilon * epsilon);
    const double factor = keep ? (inner ? -rs : (r >= 0.0 ? -epsilon : epsilon)) : 0.0;
    const double dsig = inner ? 0.5 * (1.0 - rs * rs) : 0.5 * (1.0 - epsilon * epsilon);
    acc[0] += keep ? loss : 0.0;
    acc[1] += keep ? 1.0 : 0.0;
    acc[2] += factor;
    acc[3] += keep ? dsig : 0.0;
#pragma unroll
    for (int j = 0; j < DMAX; ++j) acc[4 + j] = fma(factor, x[j], acc[4 + j]);
    return 0.0;
  }
};

// The same with instance weights: column 0 of the table is the weight (eff[0] = 0), rows of weight 0 are skipped like
// HuberAggregator.add does; accumulator 1 is the weight sum.
template <int DMAX>
struct HuberWOp {
  static constexpr int NACC = 4 + DMAX;
  static constexpr int NCONST = 0;
  static constexpr bool HAS_OUT = false;
  double eff[DMAX];
  double intercept, sigma, epsilon;
  __host__ __device__ static constexpr double identity(int) { return 0.0; }
  __host__ __device__ static constexpr int kind(int) { return kSum; }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double*, const double (&x)[DMAX], double y,
                                        bool keep, int) const {
    const double w = x[0];
    const bool on = keep && w != 0.0;
    double margin = intercept;
#pragma unroll
    for (int j = 1; j < DMAX; ++j) margin = fma(x[j], eff[j], margin);
    const double r = y - margin, rs = r / sigma;
    const bool inner = fabs(r) <= sigma * epsilon;
    const double loss = inner ? 0.5 * (sigma + r * rs) : 0.5 * (sigma + 2.0 * epsilon * fabs(r) - sigma * epsilon * epsilon);
    const double factor = on ? w * (inner ? -rs : (r >= 0.0 ? -epsilon : epsilon)) : 0.0;
    const double dsig = inner ? 0.5 * (1.0 - rs * rs) : 0.5 * (1.0 - epsilon * epsilon);
    acc[0] += on ? w * loss : 0.0;
    acc[1] += keep ? (w >= 0.0 ? w : __longlong_as_double(0x7ff8000000000000LL)) : 0.0;  // require(weight >= 0.0)
    acc[2] += factor;
    acc[3] += on ? w * dsig : 0.0;
#pragma unroll
    for (int j = 1; j < DMAX; ++j) acc[4 + j] = fma(factor, on ? x[j] : 0.0, acc[4 + j]);
    return 0.0;
  }
};

// LinearRegressionModel.transform + RegressionMetrics / LinearRegressionSummary sums in one pass: yhat = x.beta + b;
// residual r = y - yhat; accumulators n, sum r^2, sum y, sum y^2, sum |r|, sum r, sum r*y, min r, max r (the last four
// feed explainedVariance and devianceResiduals).  Per-row output: prediction (mode 0) or residual (mode 1).
template <int DMAX>
struct EvalOp {
  static constexpr int NACC = 9;
  static constexpr int NCONST = 0;
  static constexpr bool HAS_OUT = true;
  double coef[DMAX];
  double intercept;
  int mode;
  __host__ __device__ static constexpr int kind(int k) { return k == 7 ? kMin : (k == 8 ? kMaxNan : kSum); }
  __host__ __device__ static constexpr double identity(int k) {
    return k == 7 ? __builtin_huge_val() : (k == 8 ? -__builtin_huge_val() : 0.0);
  }

  __device__ __forceinline__ double row(double (&acc)[NACC], const double*, const double (&x)[DMAX], double y,
                                        bool keep, int) const {
    double yhat = 0.0;
#pragma unroll
    for (int j = 0; j < DMAX; ++j) yhat = fma(x[j], coef[j], yhat);
    yhat += intercept;
    yhat = keep ? yhat : 0.0;
    const double res = y - yhat;  // y is already 0 for dropped rows
    acc[0] += keep ? 1.0 : 0.0;
    acc[1] = fma(res, res, acc[1]);
    acc[2] += y;
    acc[3] = fma(y, y, acc[3]);
    acc[4] += fabs(res);
    acc[5] += res;
    acc[6] = fma(res, y, acc[6]);
    acc[7] = combine(kMin, acc[7], keep ? res : __builtin_huge_val());
    acc[8] = combine(kMaxNan, acc[8], keep ? res : -__builtin_huge_val());
    return mode == 0 ? yhat : res;
  }
};

}  // namespace dq4ml
