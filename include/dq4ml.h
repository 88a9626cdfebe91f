/*
 * dq4ml.h -- C ABI of libdq4ml.so: the B200-native (sm_100a) DQ + LinearRegression hot path of the
 * net.jgp.labs.sparkdq4ml lab.
 *
 * Every entry point below states the reference interface it replaces.  Citations are relative to the
 * reference repository root:
 *   App     = src/main/java/net/jgp/labs/sparkdq4ml/DataQuality4MachineLearningApp.java
 *   MinSvc  = src/main/java/net/jgp/labs/sparkdq4ml/dq/service/MinimumPriceDataQualityService.java
 *   CorrSvc = src/main/java/net/jgp/labs/sparkdq4ml/dq/service/PriceCorrelationDataQualityService.java
 *   MinUdf  = src/main/java/net/jgp/labs/sparkdq4ml/dq/udf/MinimumPriceDataQualityUdf.java
 *   CorrUdf = src/main/java/net/jgp/labs/sparkdq4ml/dq/udf/PriceCorrelationDataQualityUdf.java
 * The Spark classes the lab calls (spark-sql / spark-mllib 2.4.4, pom.xml:13-32) are named where an entry
 * point stands in for them.
 *
 * Conventions
 *   - plain C: pointers, sizes, POD structs; no exceptions cross the boundary.
 *   - every function returns DQ4ML_OK (0) or a negative dq4ml_status; dq4ml_last_error() gives the text
 *     (thread-local, valid until the next failing call on the same thread).
 *   - data pointers (columns, outputs marked "host or device") may be host or CUDA device memory; the library
 *     asks the driver which (cudaPointerGetAttributes).  Host buffers are copied into a per-context device staging
 *     arena first (pinned host memory makes that copy run at PCIe rate).  Device columns must be 32-byte aligned
 *     (the kernels use 256-bit loads); a row range of an aligned column stays aligned when it starts on a multiple of
 *     8 rows.
 *   - a dq4ml_ctx owns one CUDA device, one stream and its scratch buffers; use one ctx per thread (or lock).
 *     That stream is non-blocking (it does not synchronise with the legacy default stream): device inputs must be
 *     complete when an entry point is called, and every entry point returns after its own work has finished.
 *   - there is no CPU fallback: entry points that need a GPU fail with DQ4ML_ERR_NO_DEVICE without one.
 *     Entry points marked [host-only] never touch the GPU.
 */
#ifndef DQ4ML_H
#define DQ4ML_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DQ4ML_ABI_VERSION 2 /* 2: dq4ml_lr_params gained loss/epsilon; dq4ml_ctx_set_allreduce takes rank/world_size */

typedef enum {
  DQ4ML_OK = 0,
  DQ4ML_ERR_INVALID = -1,    /* bad argument (null pointer, negative size, unknown enum, misaligned device column) */
  DQ4ML_ERR_CUDA = -2,       /* a CUDA runtime call failed */
  DQ4ML_ERR_NULL_VALUE = -3, /* a null reached minimumPriceRule: the reference throws NullPointerException (MinUdf:12) */
  DQ4ML_ERR_NO_DEVICE = -4,  /* no usable CUDA device */
  DQ4ML_ERR_SINGULAR = -5,   /* Cholesky met a non-positive pivot and no fallback applies (solver = normal) */
  DQ4ML_ERR_COMM = -6,       /* NCCL / all-reduce callback failure */
  DQ4ML_ERR_UNSUPPORTED = -7,/* parameter combination outside what this build covers */
  DQ4ML_ERR_EMPTY = -8       /* no rows left to train on (Spark: "Training dataset is empty") */
} dq4ml_status;

typedef struct dq4ml_ctx dq4ml_ctx;
typedef struct dq4ml_lr_model dq4ml_lr_model;

typedef enum { DQ4ML_I32 = 0, DQ4ML_F64 = 1 } dq4ml_dtype;

/* One input column of the VectorAssembler (App:106-113): a contiguous array of n values. */
typedef struct {
  const void* data; /* host or device */
  int32_t dtype;    /* dq4ml_dtype; int32 columns are widened to double in registers (VectorAssembler's cast) */
} dq4ml_column;

/* VectorAssembler.setInputCols(...).setOutputCol("features") (App:110-113): the "features" vector is never
 * materialised; it is this descriptor table. */
typedef struct {
  const dq4ml_column* cols;
  int32_t n_cols;
} dq4ml_columns;

/* Thresholds of the two rules.  dq4ml_dq_params_default() gives the reference's constants:
 * min_price = 20 (MinSvc:5), min_guest = 14 and max_price = 90 (CorrSvc:6). */
typedef struct {
  double min_price;  /* rule 1 rejects price <  min_price                      (MinSvc:8)  */
  int32_t min_guest; /* rule 2 rejects guest <  min_guest && price > max_price  (CorrSvc:6) */
  double max_price;
} dq4ml_dq_params;

/* Fused-DQ descriptor: when passed to a training / evaluation entry point, rows that the two rules + filters
 * (App:68-90) would drop are skipped inside the same streaming pass, so no filtered copy is ever written. */
typedef struct {
  const int32_t* guest; /* host or device, n values */
  const double* price;  /* host or device, n values */
  dq4ml_dq_params params;
} dq4ml_dq_fused;

typedef enum { DQ4ML_SOLVER_AUTO = 0, DQ4ML_SOLVER_NORMAL = 1, DQ4ML_SOLVER_LBFGS = 2 } dq4ml_solver;
typedef enum { DQ4ML_LOSS_SQUARED_ERROR = 0, DQ4ML_LOSS_HUBER = 1 } dq4ml_loss;

/* org.apache.spark.ml.regression.LinearRegression params as the lab sets them (App:120-123) plus the Spark 2.4.4
 * defaults it leaves alone.  dq4ml_lr_params_default() = {100, 0.0, 0.0, 1e-6, 1, 1, AUTO, SQUARED_ERROR, 1.35}. */
typedef struct {
  int32_t max_iter;          /* setMaxIter        (App:121) */
  double reg_param;          /* setRegParam       (App:122) */
  double elastic_net_param;  /* setElasticNetParam(App:123) */
  double tol;                /* getTol            (App:145) */
  int32_t fit_intercept;
  int32_t standardization;
  int32_t solver;            /* dq4ml_solver */
  int32_t loss;              /* dq4ml_loss: setLoss("squaredError" | "huber"); the lab leaves the default (App:120-123) */
  double epsilon;            /* setEpsilon: the huber loss's shape parameter, > 1.0 (Spark default 1.35) */
} dq4ml_lr_params;

/* Sum-all-reduce of `count` doubles in place over all ranks; buf is HOST memory.  Return 0 on success. */
typedef int (*dq4ml_allreduce_fn)(void* user, double* buf, int64_t count);

/* ------------------------------------------------------------------------------------------------ context */

/* [host-only] Text of the last error on this thread. */
const char* dq4ml_last_error(void);
/* [host-only] ABI version of the loaded library (DQ4ML_ABI_VERSION). */
int dq4ml_abi_version(void);

/* Replaces SparkSession.builder().master("local[*]").getOrCreate() (App:38-41) as the owner of the execution
 * resources of the path: binds CUDA device `device`, creates the stream and scratch. */
int dq4ml_ctx_create(int device, dq4ml_ctx** out);
void dq4ml_ctx_destroy(dq4ml_ctx* ctx);
int dq4ml_ctx_synchronize(dq4ml_ctx* ctx);
/* The cudaStream_t every kernel of this ctx is launched on (for callers that time with CUDA events). */
void* dq4ml_ctx_stream(dq4ml_ctx* ctx);
/* Number of kernels this ctx has launched so far (bench.py's gpu_launches). */
int64_t dq4ml_ctx_launch_count(dq4ml_ctx* ctx);

/* Which kernel serves the normal-equation (WeightedLeastSquares.Aggregator) pass of lr.fit (App:126).  No counterpart
 * in the reference (Spark always accumulates in fp64 on the CPU).
 *   DQ4ML_GRAM_FP64   (default) fp64 throughout: streaming kernel for d <= 8, mma.sync DMMA Gram kernel above; 1e-9 parity.
 *   DQ4ML_GRAM_TF32X3 tcgen05 tensor cores with accumulators in TMEM, for 1 <= d <= 256: centred values split into two
 *                     TF32 parts (3 products per term), fp32 accumulation over spans of 1024 rows, fp64 across spans
 *                     and for every column sum.  Tolerance of the tier: 1e-5 relative on coefficients / intercept.
 *                     Wider tables fall back to the fp64 kernel. */
#define DQ4ML_GRAM_FP64 0
#define DQ4ML_GRAM_TF32X3 1
int dq4ml_ctx_set_gram_tier(dq4ml_ctx* ctx, int tier);

/* Measurement hook (no counterpart in the reference): when enabled, every aggregation-pass kernel (WLS moments,
 * summarizer, loss/gradient) is bracketed by CUDA events on the ctx stream; dq4ml_ctx_pass_timing returns the summed
 * kernel time and the number of passes since it was enabled. */
int dq4ml_ctx_enable_pass_timing(dq4ml_ctx* ctx, int enable);
int dq4ml_ctx_pass_timing(dq4ml_ctx* ctx, double* total_ms, int64_t* passes);
/* The individual kernel times behind that sum, in launch order (for an l-bfgs fit: the summarizer pass first, then one
 * loss/gradient pass per evaluation): copies up to cap values to ms (HOST) and returns how many passes were timed. */
int dq4ml_ctx_pass_times(dq4ml_ctx* ctx, double* ms, int32_t cap);

/* Row-sharded training across GPUs: replaces Spark's treeAggregate merge of partition aggregators
 * (RDD.treeAggregate inside WeightedLeastSquares.fit / RDDLossFunction.calculate, reached from App:126).
 * Either bind an NCCL communicator (device-side ncclAllReduce on the ctx stream) ... */
int dq4ml_nccl_unique_id(void* id128 /* 128 bytes, host */);
int dq4ml_ctx_init_nccl(dq4ml_ctx* ctx, const void* id128, int rank, int world_size);
/* ... or supply a host callback (e.g. gloo through the caller's runtime) together with this process's position among
 * the ranks: min / max statistics and the summarizer's centres travel through the sum in per-rank slots, so the library
 * has to know rank and world_size (0 <= rank < world_size).  fn == NULL unbinds. */
int dq4ml_ctx_set_allreduce(dq4ml_ctx* ctx, dq4ml_allreduce_fn fn, void* user, int rank, int world_size);
/* Exposed for tests: sum `count` doubles (host buffer) over the ranks bound to ctx; identity when unbound. */
int dq4ml_ctx_allreduce(dq4ml_ctx* ctx, double* buf, int64_t count);

/* Device / pinned memory helpers for callers without their own CUDA bindings (the Java drop-in). */
int dq4ml_device_alloc(dq4ml_ctx* ctx, int64_t bytes, void** out);
int dq4ml_device_free(dq4ml_ctx* ctx, void* p);
int dq4ml_host_alloc_pinned(int64_t bytes, void** out);
int dq4ml_host_free_pinned(void* p);
int dq4ml_memcpy(dq4ml_ctx* ctx, void* dst, const void* src, int64_t bytes); /* any direction, synchronous */

/* ------------------------------------------------------------------------------------------------ DQ rules */

/* [host-only] */
dq4ml_dq_params dq4ml_dq_params_default(void);

/* [host-only] Scalar forms, bit-for-bit the service methods:
 * MinimumPriceDataQualityService.checkMinimumPrice (MinSvc:7-13) and
 * PriceCorrelationDataQualityService.checkPriceRange (CorrSvc:5-10). */
double dq4ml_check_minimum_price(double price);
double dq4ml_check_price_range(double price, int32_t guest);

/* Columnar form of the UDF registered as "minimumPriceRule" (App:46-47), i.e. MinimumPriceDataQualityUdf.call
 * (MinUdf:11-13) applied to a whole column (App:68-69).  price_valid: optional byte-per-row validity (0 = null);
 * like the reference, a null price is an error (DQ4ML_ERR_NULL_VALUE; out is unspecified). */
int dq4ml_udf_minimumPriceRule(dq4ml_ctx* ctx, const double* price, const uint8_t* price_valid, int64_t n,
                               double* out, const dq4ml_dq_params* params /* NULL = defaults */);

/* Columnar form of the UDF registered as "priceCorrelationRule" (App:48-49): PriceCorrelationDataQualityUdf.call
 * (CorrUdf:11-16) over columns (App:86-87).  A null price or guest gives -1.0 (CorrUdf:12-14). */
int dq4ml_udf_priceCorrelationRule(dq4ml_ctx* ctx, const double* price, const int32_t* guest,
                                   const uint8_t* price_valid, const uint8_t* guest_valid, int64_t n, double* out,
                                   const dq4ml_dq_params* params);

/* The whole DQ section (App:68-90) in one pass: rule 1 -> WHERE price_no_min > 0 -> rule 2 ->
 * WHERE price_correct_correl > 0, projecting (guest, price).  Order preserving (stable compaction); Spark SQL
 * ordering for "> 0" (NaN survives).  guest_out/price_out need room for n rows.  n_out / n_after_rule1 are HOST
 * int64 (n_after_rule1 may be NULL).  In-place (guest_out == guest) is NOT allowed. */
int dq4ml_dq_filter(dq4ml_ctx* ctx, const int32_t* guest, const double* price, int64_t n,
                    const dq4ml_dq_params* params, int32_t* guest_out, double* price_out, int64_t* n_out,
                    int64_t* n_after_rule1);

/* Row-sharded DQ section: every rank compacts its own contiguous row shard with dq4ml_dq_filter and the output stays
 * sharded (Spark: each partition of the filtered Dataset, App:78,90, stays where it is).  This collective -- every rank
 * bound to the ctx must call it -- turns the local survivor counts into positions in the global, order-preserving
 * output: offset (HOST) = survivors on all lower ranks, total (HOST, nullable) = survivors on all ranks.  Identity
 * (offset 0, total n_out_local) when no ranks are bound. */
int dq4ml_dq_filter_global_offset(dq4ml_ctx* ctx, int64_t n_out_local, int64_t* offset, int64_t* total);

/* The same filter for callers that cannot let the library allocate.  No counterpart in the reference: it exists for
 * hosts where the GPU's memory is already owned by another allocator (an executor-wide memory pool, say), or where the
 * device-wide stall inside cudaMalloc/cudaFree is not tolerable.  It takes no dq4ml_ctx and calls no allocator of any
 * kind: the compaction kernel's look-back words and row counters live in a fixed array compiled into the library
 * (about 256 KB of every GPU that loads libdq4ml.so, held until the process exits), which caps one call at
 * DQ4ML_NOALLOC_MAX_ROWS rows -- larger inputs are refused with DQ4ML_ERR_UNSUPPORTED.
 *   - Call it with the device that owns the buffers current (cudaSetDevice); `stream` is a cudaStream_t (NULL = the
 *     default stream).  guest/price/guest_out/price_out must be DEVICE memory (outputs sized for n rows); n_out and
 *     n_after_rule1 (nullable) are HOST int64.
 *   - Every call copies the two row counts back to the host and waits for `stream` to drain before it returns
 *     (cudaMemcpyAsync + cudaStreamSynchronize), so it must never be issued on a stream that is being captured.
 *   - The first call on each device also looks up the array's device address (cudaGetSymbolAddress); make that first
 *     call before any stream capture starts anywhere in the process.
 *   - All calls on a device share the one array, and it keeps whatever the previous call left there; each call
 *     re-zeroes the part it is about to use, on `stream`, before its kernel runs.  Two calls whose GPU work overlaps
 *     (different streams or threads) would race on it: serialise them.
 * Results are identical to dq4ml_dq_filter's (same kernel). */
#define DQ4ML_NOALLOC_MAX_ROWS (1ll << 26) /* 67 108 864 rows per call */
int dq4ml_dq_filter_noalloc(void* stream, const int32_t* guest, const double* price, int64_t n,
                            const dq4ml_dq_params* params, int32_t* guest_out, double* price_out, int64_t* n_out,
                            int64_t* n_after_rule1);

/* CSV ingest for the lab's dialect: spark.read().format("csv").option("inferSchema","true").option("header","false")
 * .load("data/dataset-*.csv") (App:52-55) followed by the two renames (App:58-59).  Records end at CR, LF or CRLF
 * (the lab's files use bare CR), have no header, and hold `int,decimal`; empty records are skipped.  data (len bytes)
 * and the two outputs (room for cap records) may each be host or device memory; the decimal is converted exactly as
 * Double.parseDouble would (correctly rounded).  n_records (HOST) receives the record count; a malformed record or
 * cap < count is DQ4ML_ERR_INVALID. */
int dq4ml_csv_parse(dq4ml_ctx* ctx, const char* data, int64_t len, int32_t* guest_out, double* price_out, int64_t cap,
                    int64_t* n_records);

/* ------------------------------------------------------------------------------------------------ LinearRegression */

/* [host-only] */
dq4ml_lr_params dq4ml_lr_params_default(void);

/* Length of the WLS moment vector for d features: 5 + 2d + d(d+1)/2. [host-only] */
int64_t dq4ml_wls_moments_len(int32_t d);

/* One pass of WeightedLeastSquares.Aggregator.add/merge (Spark 2.4.4, reached from lr.fit, App:126) with w = 1:
 * out (HOST) = [count, wSum, wwSum, bSum, bbSum, aSum[d], abSum[d], aaSum[d(d+1)/2 packed upper, column-major]].
 * dq may be NULL.  When ranks are bound the result is already summed over ranks. */
int dq4ml_wls_moments(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                      const dq4ml_dq_fused* dq, double* out);

/* [host-only] WeightedLeastSquares.fit after the aggregation + NormalEquationSolver (Cholesky, or Breeze
 * OWL-QN / L-BFGS over the Gram): moments -> model. */
int dq4ml_wls_solve(const double* moments, int32_t d, const dq4ml_lr_params* params, dq4ml_lr_model** out);

/* Length of the summarizer output for d features: 3 + 7(d+1). [host-only] */
int64_t dq4ml_summary_len(int32_t d);

/* One pass of the two MultivariateOnlineSummarizers of LinearRegression.train (solver l-bfgs): out (HOST) =
 * [count, wSum, wwSum, then per column j = 0..d (j == d is the label):
 *  mean, sample variance, min, max, numNonzeros, normL1, normL2]. */
int dq4ml_summarize(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                    const dq4ml_dq_fused* dq, double* out);

/* One cost-function evaluation = one LeastSquaresAggregator treeAggregate (RDDLossFunction.calculate):
 * out (HOST) = [lossSum, weightSum, gradSum[d]] (raw sums, before the division by weightSum).
 * coef are the scaled-space coefficients the optimiser works on. */
int dq4ml_lsq_loss_grad(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                        const dq4ml_dq_fused* dq, const double* coef, const double* features_std,
                        const double* features_mean, double label_std, double label_mean, int fit_intercept,
                        double* out);

/* LinearRegression.fit(dataset) (App:126): solver selection, aggregation passes on the GPU, optimiser on the host.
 * features = the VectorAssembler table, label = the "label" column (App:101).  dq may be NULL.
 * loss = DQ4ML_LOSS_HUBER (the lab leaves Spark's default, squaredError, at App:120-123; SURVEY.md 8f lists huber as a
 * next parameter): summarizer pass for the feature scales, one HuberAggregator pass per evaluation, Breeze LBFGSB on the
 * host over [coefficients, intercept, sigma] with sigma > 0; needs elasticNetParam == 0, epsilon > 1, solver != normal
 * (as Spark requires).  Up to 32 streamed columns (features, plus the weight if any) one launch per evaluation; wider
 * tables a column-walking row pass plus gradient passes over column groups. */
int dq4ml_linreg_fit(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                     const dq4ml_lr_params* params, const dq4ml_dq_fused* dq, dq4ml_lr_model** out);

/* LinearRegression.setWeightCol(...).fit(dataset): the same fit with instance weights (the lab never sets weightCol,
 * App:111-126 leaves it at Spark's default; SURVEY.md 8f lists it as the next parameter).  weight: f64[n], host or
 * device, every value >= 0 (Spark: require(w >= 0.0)); NULL = dq4ml_linreg_fit.  Normal-equation branch: the Gram
 * kernel scales each row by sqrt(w) (any width).  l-bfgs branch: weighted summarizer (zero-weight rows are not counted,
 * variance denominator wSum - sum(w^2)/wSum) and weighted LeastSquaresAggregator; up to 31 feature columns the weight
 * rides as one more column of the streaming pass, wider tables take the grouped summarizer and the two-pass
 * loss/gradient (per-row weighted difference, then column groups). */
int dq4ml_linreg_fit_weighted(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, const double* weight,
                              int64_t n, const dq4ml_lr_params* params, const dq4ml_dq_fused* dq, dq4ml_lr_model** out);

/* [host-only] LinearRegressionModel / LinearRegressionTrainingSummary getters (App:132-146). */
int32_t dq4ml_linreg_model_num_features(const dq4ml_lr_model* m);
int dq4ml_linreg_model_coefficients(const dq4ml_lr_model* m, double* coef /* d */);
double dq4ml_linreg_model_intercept(const dq4ml_lr_model* m);             /* model.intercept()  App:141 */
int32_t dq4ml_linreg_model_total_iterations(const dq4ml_lr_model* m);     /* totalIterations()  App:134 */
int dq4ml_linreg_model_objective_history(const dq4ml_lr_model* m, double* hist, int32_t cap); /* App:135-136 */
int32_t dq4ml_linreg_model_data_passes(const dq4ml_lr_model* m);          /* aggregation passes the fit ran */
double dq4ml_linreg_model_scale(const dq4ml_lr_model* m);                 /* model.scale(): huber's sigma, else 1.0 */
int64_t dq4ml_linreg_model_num_rows(const dq4ml_lr_model* m);             /* rows that entered the fit */
const char* dq4ml_linreg_model_solver_used(const dq4ml_lr_model* m);
double dq4ml_linreg_predict1(const dq4ml_lr_model* m, const double* x /* d */); /* model.predict(Vector) App:150-151 */
/* [host-only] LinearRegressionTrainingSummary.coefficientStandardErrors / tValues / pValues (beyond what the lab
 * prints at App:132-139): k = numFeatures (+1 for the intercept, last) values each, from the diagonal of (A^T A)^-1 kept
 * by the normal-equation Cholesky solve, the residual sum of squares and the instance count of an evaluation
 * (metrics[1], metrics[0] of dq4ml_linreg_evaluate).  Any output may be NULL.  Returns k, or DQ4ML_ERR_UNSUPPORTED
 * for models fitted another way (Spark: "No Std. Error of coefficients available for this LinearRegressionModel"). */
int dq4ml_linreg_model_inference(const dq4ml_lr_model* m, double num_instances, double ss_err, int fit_intercept,
                                 double* std_err, double* t_values, double* p_values, int32_t cap);
void dq4ml_linreg_model_destroy(dq4ml_lr_model* m);
/* [host-only] build a model from known coefficients (LinearRegressionModel's constructor). */
int dq4ml_linreg_model_create(const double* coef, int32_t d, double intercept, dq4ml_lr_model** out);

/* LinearRegressionModel.transform(df) (App:129): prediction[i] = features_i . coefficients + intercept. */
int dq4ml_linreg_transform(dq4ml_ctx* ctx, const dq4ml_lr_model* m, const dq4ml_columns* features, int64_t n,
                           double* prediction /* host or device, n */);

/* LinearRegressionTrainingSummary.residuals()/rootMeanSquaredError()/r2() (App:137-139), one fused pass:
 * metrics (HOST, DQ4ML_METRICS_LEN = 16 doubles) = [n, SSerr, sum_y, sum_yy, sum_abs_err, rmse, r2, mse, mae,
 * explainedVariance, degreesOfFreedom, r2adj, min residual, max residual (devianceResiduals), sum_res, sum_res_y];
 * residuals (nullable, host or device,
 * n values, label - prediction; rows skipped by fused DQ get 0).  Summed over ranks when bound. */
#define DQ4ML_METRICS_LEN 16
int dq4ml_linreg_evaluate(dq4ml_ctx* ctx, const dq4ml_lr_model* m, const dq4ml_columns* features, const double* label,
                          int64_t n, const dq4ml_dq_fused* dq, int fit_intercept, double* metrics, double* residuals);

/* LinearRegressionSummary under a weight column (Spark 2.4.4 keeps RegressionMetrics unweighted; only these differ):
 * out3 (HOST) = [sum w r^2 (the rss of coefficientStandardErrors), min sqrt(w) r, max sqrt(w) r (devianceResiduals)]
 * over the residual column r that dq4ml_linreg_evaluate wrote (host or device) and the weight column.  Rows the fused
 * DQ rules skip do not take part.  Summed / folded over ranks when bound. */
int dq4ml_weighted_residual_stats(dq4ml_ctx* ctx, const double* residuals, const double* weight, int64_t n,
                                  const dq4ml_dq_fused* dq, double* out3);

/* ------------------------------------------------------------------------------------------------ synthetic data */

/* Counter-based generator of the benchmark data (no counterpart in the reference, which ships three CSVs only):
 * guest:int32[n], price:f64[n], extra features (d-1 columns, column k at extra + (k-1)*extra_stride), all DEVICE.
 * Bit-identical to oracle/dq4ml_oracle.c:orc_synth_generate for the same parameters. */
typedef struct {
  uint64_t seed_data;
  uint64_t seed_anom;
  double p_low;
  double p_high;
  int32_t d;
} dq4ml_synth_params;
int dq4ml_synth_generate(dq4ml_ctx* ctx, const dq4ml_synth_params* p, int64_t row0, int64_t n, int32_t* guest,
                         double* price, double* extra, int64_t extra_stride);

/* Order-sensitive 128-bit digest of (guest, price) rows (host or device), for bit-exactness checks at sizes where
 * copying the survivors back is impractical.  out: HOST uint64[2]. */
int dq4ml_rows_digest(dq4ml_ctx* ctx, const int32_t* guest, const double* price, int64_t n, uint64_t seed,
                      uint64_t* out);

#ifdef __cplusplus
}
#endif
#endif /* DQ4ML_H */
