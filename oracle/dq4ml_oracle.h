/*
 * dq4ml_oracle.h -- CPU oracle for the DQ + LinearRegression hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it, and only as the checker / the timed CPU
 * baseline.  The product (libdq4ml.so) never links or calls it.
 *
 * PARITY UNPINNED: the reference (frankyangdev/net.jgp.labs.sparkdq4ml) ships
 * no tests and no golden outputs, and its arithmetic lives in un-vendored jars
 * (spark-sql/spark-mllib 2.4.4, breeze 0.13.2 -- pom.xml:13-32).  This oracle
 * restates (a) the lab's own Java rule code and (b) the published Spark 2.4.4
 * algorithms reached from the lab's call sites; it is pinned only against the
 * three data files of the reference, closed-form lasso/OLS solutions and an
 * independent scikit-learn solve (tests/golden/make_golden.py).
 *
 * Citations "App:" = src/main/java/net/jgp/labs/sparkdq4ml/DataQuality4MachineLearningApp.java,
 * "MinSvc:" = .../dq/service/MinimumPriceDataQualityService.java,
 * "CorrSvc:" = .../dq/service/PriceCorrelationDataQualityService.java,
 * "MinUdf:"/"CorrUdf:" = .../dq/udf/{MinimumPrice,PriceCorrelation}DataQualityUdf.java.
 */
#ifndef DQ4ML_ORACLE_H
#define DQ4ML_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_I32 = 0, ORC_F64 = 1 };

typedef struct {
  const void* data; /* contiguous column of n values */
  int32_t dtype;    /* ORC_I32 or ORC_F64 */
} orc_column;

/* ---- DQ rules (MinSvc:5-13, CorrSvc:5-10, MinUdf:11-13, CorrUdf:11-16) ---- */
double orc_check_minimum_price(double price);
double orc_check_price_range(double price, int32_t guest);
/* Spark SQL "col > 0" on a double: NaN sorts above every number => true (App:78,90). */
int orc_spark_gt_zero(double v);
/* Columnar UDF forms.  valid==NULL means all valid.  Returns -1 if a null
 * price reaches minimumPriceRule (MinUdf:12 unboxes null -> NPE), else 0. */
int orc_udf_minimum_price(const double* price, const uint8_t* valid, int64_t n, double* out);
void orc_udf_price_correlation(const double* price, const uint8_t* pvalid, const int32_t* guest,
                               const uint8_t* gvalid, int64_t n, double* out);
/* App:68-90: rule 1 -> filter -> rule 2 -> filter, order preserving.  Returns survivors.
 * n_after_rule1 (nullable) receives the intermediate count. */
int64_t orc_dq_filter(const int32_t* guest, const double* price, int64_t n, int32_t* guest_out,
                      double* price_out, int64_t* n_after_rule1);
/* keep mask for one row under both rules */
int orc_dq_keep(int32_t guest, double price);

/* ---- CSV dialect of data/dataset-*.csv (App:52-55): records split on \r, \n or \r\n;
 * field 0 -> int32, field 1 -> double.  Returns #records or -1 on malformed input. */
int64_t orc_csv_parse(const char* buf, int64_t len, int32_t* guest, double* price, int64_t cap);

/* ---- the lab's whole d = 1 pipeline, one row at a time through the boxed UDF call chain (MinUdf:11-13 -> MinSvc:7-13,
 * App:77-78, CorrUdf:11-16 -> CorrSvc:5-10, App:89-90, App:101-113) into the WLS aggregator: the "row-at-a-time
 * scalar" CPU-baseline variant.  out[8] has the orc_wls_moments layout for d = 1 and equals it. */
void orc_lab_rowwise_moments(const int32_t* guest, const double* price, int64_t n, int threads, double* out);

/* ---- WeightedLeastSquares.Aggregator (Spark 2.4.4), w = 1 ----
 * out layout: [count, wSum, wwSum, bSum, bbSum, aSum[d], abSum[d], aaSum[d(d+1)/2]]
 * aaSum is packed upper, column-major (BLAS dspr 'U').
 * dq_guest/dq_price: if non-NULL rows failing orc_dq_keep are skipped (fused DQ).
 * threads<=1: one sequential left fold (Spark single-partition order);
 * threads>1 : OpenMP, one aggregator per contiguous partition, merged in partition order. */
void orc_wls_moments(const orc_column* features, int32_t d, const double* label, int64_t n,
                     const int32_t* dq_guest, const double* dq_price, int threads, double* out);

/* ---- MultivariateOnlineSummarizer x2 (features, label), w = 1 ----
 * out layout: [count, wSum, wwSum, then for j in 0..d (j==d is the label):
 *              mean_j, variance_j (sample), min_j, max_j, nnz_j, normL1_j, normL2_j] */
void orc_summarizer(const orc_column* features, int32_t d, const double* label, int64_t n,
                    const int32_t* dq_guest, const double* dq_price, int threads, double* out);

/* ---- LeastSquaresAggregator (Spark 2.4.4), one cost evaluation ----
 * coef: scaled-space coefficients c[d]; fstd/fmean: feature std/mean; out = [lossSum, weightSum, gradSum[d]] (raw sums) */
void orc_lsq_loss_grad(const orc_column* features, int32_t d, const double* label, int64_t n,
                       const int32_t* dq_guest, const double* dq_price, const double* coef,
                       const double* fstd, const double* fmean, double ystd, double ymean,
                       int fit_intercept, int threads, double* out);

/* ---- LinearRegressionModel.transform + RegressionMetrics ----
 * prediction (nullable) gets yhat per row (rows dropped by fused DQ get NaN-free 0 and are not counted);
 * out = [n, SSerr, sum_y, sum_yy, sum_abs_err] */
void orc_predict_metrics(const orc_column* features, int32_t d, const double* label, int64_t n,
                         const int32_t* dq_guest, const double* dq_price, const double* coef,
                         double intercept, int threads, double* prediction, double* out);

/* ---- synthetic generator: counter based, bit-identical to the device generator ---- */
typedef struct {
  uint64_t seed_data;  /* 20200601 */
  uint64_t seed_anom;  /* 7 */
  double p_low;        /* P(price <- U[3,15]) */
  double p_high;       /* P(price <- U[100,150] | guest < 14) is p_high */
  int32_t d;           /* number of features (guest + d-1 extra) */
} orc_synth_params;
/* rows [row0, row0+n): guest[n], price[n], extra[(d-1)][n] (column k at extra + (k-1)*extra_stride) */
void orc_synth_generate(const orc_synth_params* p, int64_t row0, int64_t n, int32_t* guest,
                        double* price, double* extra, int64_t extra_stride, int threads);

/* order-sensitive 128-bit digest of the packed survivors (guest,price) -- used for shard-of-shards checks */
void orc_rows_digest(const int32_t* guest, const double* price, int64_t n, uint64_t seed, uint64_t out[2]);

int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
