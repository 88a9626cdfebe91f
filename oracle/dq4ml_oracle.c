/*
 * dq4ml_oracle.c -- CPU restatement of the reference's DQ + LinearRegression data passes.
 *
 * TEST INFRASTRUCTURE ONLY (see dq4ml_oracle.h).  PARITY UNPINNED by the reference's own
 * tests (it has none); pinned against its three CSV fixtures and closed forms instead.
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC  (no FMA contraction: Spark's f2j
 * BLAS does a separate multiply and add).
 */
#include "dq4ml_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------ DQ rules */

/* MinSvc:5-13  `if (price < MIN_PRICE) return -1; else return price;` with MIN_PRICE = int 20 */
double orc_check_minimum_price(double price) {
  if (price < 20) return -1;
  return price;
}

/* CorrSvc:5-10 `if (guest < 14 && price > 90) return -1; return price;` */
double orc_check_price_range(double price, int32_t guest) {
  if (guest < 14 && price > 90) return -1;
  return price;
}

/* App:78 / App:90  `WHERE col > 0` -- Spark SQL double ordering: NaN is larger than any value. */
int orc_spark_gt_zero(double v) { return (v != v) || (v > 0.0); }

/* MinUdf:11-13: call(Double t1) auto-unboxes -> a null input is a NullPointerException (job fails). */
int orc_udf_minimum_price(const double* price, const uint8_t* valid, int64_t n, double* out) {
  for (int64_t i = 0; i < n; ++i) {
    if (valid && !valid[i]) return -1;
    out[i] = orc_check_minimum_price(price[i]);
  }
  return 0;
}

/* CorrUdf:11-16: null price or null guest -> -1.0 */
void orc_udf_price_correlation(const double* price, const uint8_t* pvalid, const int32_t* guest,
                               const uint8_t* gvalid, int64_t n, double* out) {
  for (int64_t i = 0; i < n; ++i) {
    if ((pvalid && !pvalid[i]) || (gvalid && !gvalid[i])) {
      out[i] = -1.0;
    } else {
      out[i] = orc_check_price_range(price[i], guest[i]);
    }
  }
}

int orc_dq_keep(int32_t guest, double price) {
  double r1 = orc_check_minimum_price(price); /* App:68-69 */
  if (!orc_spark_gt_zero(r1)) return 0;       /* App:77-78 */
  double r2 = orc_check_price_range(r1, guest); /* App:86-87 */
  return orc_spark_gt_zero(r2);               /* App:89-90 */
}

int64_t orc_dq_filter(const int32_t* guest, const double* price, int64_t n, int32_t* guest_out,
                      double* price_out, int64_t* n_after_rule1) {
  int64_t m = 0, m1 = 0;
  for (int64_t i = 0; i < n; ++i) {
    double r1 = orc_check_minimum_price(price[i]);
    if (!orc_spark_gt_zero(r1)) continue;
    ++m1;
    double r2 = orc_check_price_range(r1, guest[i]);
    if (!orc_spark_gt_zero(r2)) continue;
    guest_out[m] = guest[i];
    price_out[m] = r2;
    ++m;
  }
  if (n_after_rule1) *n_after_rule1 = m1;
  return m;
}

/* ------------------------------------------------------------------ CSV (App:52-55) */

int64_t orc_csv_parse(const char* buf, int64_t len, int32_t* guest, double* price, int64_t cap) {
  int64_t nrec = 0, pos = 0;
  char tmp[128];
  while (pos < len) {
    int64_t e = pos;
    while (e < len && buf[e] != '\r' && buf[e] != '\n') ++e;
    int64_t rl = e - pos;
    if (rl > 0) {
      if (rl >= (int64_t)sizeof(tmp) || nrec >= cap) return -1;
      memcpy(tmp, buf + pos, (size_t)rl);
      tmp[rl] = 0;
      char* comma = strchr(tmp, ',');
      if (!comma) return -1;
      *comma = 0;
      char* endp = NULL;
      long g = strtol(tmp, &endp, 10);
      if (*endp != 0) return -1;
      double p = strtod(comma + 1, &endp);
      if (*endp != 0) return -1;
      guest[nrec] = (int32_t)g;
      price[nrec] = p;
      ++nrec;
    }
    /* \r\n counts as one separator; an empty record is skipped like Hadoop's LineRecordReader + CSV source */
    if (e < len && buf[e] == '\r' && e + 1 < len && buf[e + 1] == '\n') ++e;
    pos = e + 1;
  }
  return nrec;
}

/* ------------------------------------------------------------------ helpers */

static inline double col_at(const orc_column* c, int64_t i) {
  return c->dtype == ORC_I32 ? (double)((const int32_t*)c->data)[i] : ((const double*)c->data)[i];
}

int orc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

static int pick_threads(int threads, int64_t n) {
  if (threads < 1) threads = 1;
  if ((int64_t)threads > n) threads = n > 0 ? (int)n : 1;
  return threads;
}

/* ------------------------------------------------------------------ WLS aggregator */

static void wls_fold(const orc_column* f, int32_t d, const double* label, int64_t lo, int64_t hi,
                     const int32_t* dqg, const double* dqp, double* out, double* x) {
  const int64_t tri = (int64_t)d * (d + 1) / 2;
  double* aSum = out + 5;
  double* abSum = aSum + d;
  double* aaSum = abSum + d;
  for (int64_t i = lo; i < hi; ++i) {
    if (dqg && !orc_dq_keep(dqg[i], dqp[i])) continue;
    const double w = 1.0;
    const double l = label[i];
    for (int32_t j = 0; j < d; ++j) x[j] = col_at(&f[j], i);
    out[0] += 1.0;      /* count */
    out[1] += w;        /* wSum  */
    out[2] += w * w;    /* wwSum */
    out[3] += w * l;    /* bSum  */
    out[4] += w * l * l; /* bbSum: (w*l)*l */
    const double wl = w * l;
    for (int32_t j = 0; j < d; ++j) aSum[j] += w * x[j];   /* BLAS.axpy(w, f, aSum)   */
    for (int32_t j = 0; j < d; ++j) abSum[j] += wl * x[j]; /* BLAS.axpy(w*l, f, abSum) */
    int64_t p = 0;                                         /* BLAS.spr(w, f, aaSum): dspr 'U' */
    for (int32_t j = 0; j < d; ++j) {
      const double t = w * x[j];
      for (int32_t k = 0; k <= j; ++k) aaSum[p++] += x[k] * t;
    }
    (void)tri;
  }
}

void orc_wls_moments(const orc_column* features, int32_t d, const double* label, int64_t n,
                     const int32_t* dq_guest, const double* dq_price, int threads, double* out) {
  const int64_t len = 5 + 2 * (int64_t)d + (int64_t)d * (d + 1) / 2;
  threads = pick_threads(threads, n);
  memset(out, 0, (size_t)len * sizeof(double));
  if (threads == 1) {
    double* x = (double*)malloc(sizeof(double) * (size_t)(d > 0 ? d : 1));
    wls_fold(features, d, label, 0, n, dq_guest, dq_price, out, x);
    free(x);
    return;
  }
  double* parts = (double*)calloc((size_t)len * threads, sizeof(double));
#pragma omp parallel num_threads(threads)
  {
#ifdef _OPENMP
    int t = omp_get_thread_num();
    int T = omp_get_num_threads();
#else
    int t = 0, T = 1;
#endif
    double* x = (double*)malloc(sizeof(double) * (size_t)(d > 0 ? d : 1));
    double* acc = (double*)calloc((size_t)len + 16, sizeof(double)); /* thread-private: no false sharing */
    int64_t lo = n * t / T, hi = n * (t + 1) / T;
    wls_fold(features, d, label, lo, hi, dq_guest, dq_price, acc, x);
    memcpy(parts + (int64_t)t * len, acc, sizeof(double) * (size_t)len);
    free(acc);
    free(x);
  }
  for (int t = 0; t < threads; ++t) /* Aggregator.merge = element-wise sum */
    for (int64_t k = 0; k < len; ++k) out[k] += parts[(int64_t)t * len + k];
  free(parts);
}

/* ------------------------------------------------------------------ row-at-a-time variant of the lab's pipeline
 * The columnar fold above is the strong CPU baseline.  This one keeps the shape Spark executes for each partition of
 * the lab's job: one row at a time through the boxed UDF call chain
 *   MinimumPriceDataQualityUdf.call (MinUdf:11-13) -> MinimumPriceDataQualityService.checkMinimumPrice (MinSvc:7-13)
 *   -> filter price_no_min > 0 (App:77-78)
 *   -> PriceCorrelationDataQualityUdf.call (CorrUdf:11-16, null guard) -> checkPriceRange (CorrSvc:5-10)
 *   -> filter price_correct_correl > 0 (App:89-90)
 *   -> label alias + VectorAssembler (App:101-113: a dense vector per row) -> Aggregator.add
 * with every stage an out-of-line call on boxed (pointer) values, nothing fused or vectorised by the compiler.
 * Results equal orc_wls_moments with the DQ columns given (same adds in the same order per partition). */
typedef struct { double v; int is_null; } boxed_double;
typedef struct { int32_t v; int is_null; } boxed_int;

static __attribute__((noinline)) int rw_udf_min_price(const boxed_double* t1, boxed_double* out) {
  if (t1->is_null) return -1; /* NullPointerException on unboxing (MinUdf:12) */
  out->v = orc_check_minimum_price(t1->v);
  out->is_null = 0;
  return 0;
}
static __attribute__((noinline)) void rw_udf_price_corr(const boxed_double* price, const boxed_int* guest, boxed_double* out) {
  out->is_null = 0;
  if (price->is_null || guest->is_null) { out->v = -1.0; return; } /* CorrUdf:12-14 */
  out->v = orc_check_price_range(price->v, guest->v);
}
static __attribute__((noinline)) int rw_filter_gt_zero(const boxed_double* c) { return !c->is_null && orc_spark_gt_zero(c->v); }
static __attribute__((noinline)) void rw_assemble(const boxed_int* guest, double* vec) { vec[0] = (double)guest->v; } /* App:106-113 */
static __attribute__((noinline)) void rw_aggregator_add(double* out, int32_t d, const double* x, double l, double w) {
  double* aSum = out + 5;
  double* abSum = aSum + d;
  double* aaSum = abSum + d;
  out[0] += 1.0;
  out[1] += w;
  out[2] += w * w;
  out[3] += w * l;
  out[4] += w * l * l;
  const double wl = w * l;
  for (int32_t j = 0; j < d; ++j) aSum[j] += w * x[j];
  for (int32_t j = 0; j < d; ++j) abSum[j] += wl * x[j];
  int64_t p = 0;
  for (int32_t j = 0; j < d; ++j) {
    const double t = w * x[j];
    for (int32_t k = 0; k <= j; ++k) aaSum[p++] += x[k] * t;
  }
}

typedef int (*rw_udf1_fn)(const boxed_double*, boxed_double*);
typedef void (*rw_udf2_fn)(const boxed_double*, const boxed_int*, boxed_double*);

static void rw_partition(const int32_t* guest, const double* price, int64_t lo, int64_t hi, double* out,
                         rw_udf1_fn volatile udf1, rw_udf2_fn volatile udf2) {
  double vec[1];
  for (int64_t i = lo; i < hi; ++i) {
    boxed_double p0 = {price[i], 0}, p1, p2;
    boxed_int g = {guest[i], 0};
    if (udf1(&p0, &p1) != 0) continue;      /* callUDF("minimumPriceRule", price)            App:69 */
    if (!rw_filter_gt_zero(&p1)) continue;  /* WHERE price_no_min > 0                         App:77-78 */
    udf2(&p1, &g, &p2);                     /* callUDF("priceCorrelationRule", price, guest)  App:87 */
    if (!rw_filter_gt_zero(&p2)) continue;  /* WHERE price_correct_correl > 0                 App:89-90 */
    rw_assemble(&g, vec);
    rw_aggregator_add(out, 1, vec, p2.v, 1.0);
  }
}

void orc_lab_rowwise_moments(const int32_t* guest, const double* price, int64_t n, int threads, double* out /* 8 */) {
  const int64_t len = 8;
  threads = pick_threads(threads, n);
  memset(out, 0, (size_t)len * sizeof(double));
  double* parts = (double*)calloc((size_t)(len + 16) * threads, sizeof(double));
#pragma omp parallel num_threads(threads)
  {
#ifdef _OPENMP
    int t = omp_get_thread_num();
    int T = omp_get_num_threads();
#else
    int t = 0, T = 1;
#endif
    double acc[8 + 16];
    memset(acc, 0, sizeof(acc));
    rw_partition(guest, price, n * t / T, n * (t + 1) / T, acc, rw_udf_min_price, rw_udf_price_corr);
    memcpy(parts + (int64_t)t * (len + 16), acc, sizeof(double) * (size_t)len);
  }
  for (int t = 0; t < threads; ++t)
    for (int64_t k = 0; k < len; ++k) out[k] += parts[(int64_t)t * (len + 16) + k];
  free(parts);
}

/* ------------------------------------------------------------------ MultivariateOnlineSummarizer */

typedef struct {
  int64_t cnt;
  double wsum, wwsum;
  double *mean, *m2n, *m2, *l1, *mn, *mx, *ws, *nnz; /* each [d+1] */
} summ_t;

static void summ_init(summ_t* s, int32_t m) {
  s->cnt = 0;
  s->wsum = s->wwsum = 0;
  double* b = (double*)calloc((size_t)m * 8, sizeof(double));
  s->mean = b; s->m2n = b + m; s->m2 = b + 2 * m; s->l1 = b + 3 * m;
  s->mn = b + 4 * m; s->mx = b + 5 * m; s->ws = b + 6 * m; s->nnz = b + 7 * m;
  for (int32_t j = 0; j < m; ++j) { s->mn[j] = INFINITY; s->mx[j] = -INFINITY; } /* Double.MaxValue in Spark; same effect */
}

static inline void summ_add1(summ_t* s, int32_t j, double value, double weight) {
  if (value != 0.0) {
    if (s->mx[j] < value) s->mx[j] = value;
    if (s->mn[j] > value) s->mn[j] = value;
    const double prev = s->mean[j];
    const double diff = value - prev;
    s->mean[j] = prev + weight * diff / (s->ws[j] + weight);
    s->m2n[j] += weight * (value - s->mean[j]) * diff;
    s->m2[j] += weight * value * value;
    s->l1[j] += weight * fabs(value);
    s->ws[j] += weight;
    s->nnz[j] += 1.0;
  }
}

static void summ_merge(summ_t* a, const summ_t* b, int32_t m) {
  if (a->wsum != 0.0 && b->wsum != 0.0) {
    a->cnt += b->cnt; a->wsum += b->wsum; a->wwsum += b->wwsum;
    for (int32_t i = 0; i < m; ++i) {
      const double tn = a->ws[i], on = b->ws[i], tot = tn + on;
      if (tot != 0.0) {
        const double dm = b->mean[i] - a->mean[i];
        a->mean[i] += dm * on / tot;
        a->m2n[i] += b->m2n[i] + dm * dm * tn * on / tot;
        a->m2[i] += b->m2[i];
        a->l1[i] += b->l1[i];
        a->mx[i] = fmax(a->mx[i], b->mx[i]);
        a->mn[i] = fmin(a->mn[i], b->mn[i]);
      }
      a->ws[i] = tot;
      a->nnz[i] += b->nnz[i];
    }
  } else if (a->wsum == 0.0 && b->wsum != 0.0) {
    a->cnt = b->cnt; a->wsum = b->wsum; a->wwsum = b->wwsum;
    memcpy(a->mean, b->mean, sizeof(double) * (size_t)m * 8);
  }
}

static void summ_fold(summ_t* s, const orc_column* f, int32_t d, const double* label, int64_t lo,
                      int64_t hi, const int32_t* dqg, const double* dqp) {
  for (int64_t i = lo; i < hi; ++i) {
    if (dqg && !orc_dq_keep(dqg[i], dqp[i])) continue;
    for (int32_t j = 0; j < d; ++j) summ_add1(s, j, col_at(&f[j], i), 1.0);
    summ_add1(s, d, label[i], 1.0);
    s->wsum += 1.0; s->wwsum += 1.0; s->cnt += 1;
  }
}

void orc_summarizer(const orc_column* features, int32_t d, const double* label, int64_t n,
                    const int32_t* dq_guest, const double* dq_price, int threads, double* out) {
  const int32_t m = d + 1;
  threads = pick_threads(threads, n);
  summ_t* parts = (summ_t*)malloc(sizeof(summ_t) * (size_t)threads);
  for (int t = 0; t < threads; ++t) summ_init(&parts[t], m);
#pragma omp parallel num_threads(threads)
  {
#ifdef _OPENMP
    int t = omp_get_thread_num(), T = omp_get_num_threads();
#else
    int t = 0, T = 1;
#endif
    summ_fold(&parts[t], features, d, label, n * t / T, n * (t + 1) / T, dq_guest, dq_price);
  }
  for (int t = 1; t < threads; ++t) summ_merge(&parts[0], &parts[t], m);
  summ_t* s = &parts[0];
  out[0] = (double)s->cnt; out[1] = s->wsum; out[2] = s->wwsum;
  const double denom = s->wsum - (s->wsum != 0.0 ? s->wwsum / s->wsum : 0.0);
  for (int32_t j = 0; j < m; ++j) {
    double* o = out + 3 + 7 * j;
    const double rmean = s->wsum != 0.0 ? s->mean[j] * (s->ws[j] / s->wsum) : 0.0;
    double var = 0.0;
    if (denom > 0.0) {
      const double dm = s->mean[j];
      var = fmax((s->m2n[j] + dm * dm * s->ws[j] * (s->wsum - s->ws[j]) / s->wsum) / denom, 0.0);
    }
    /* min/max: zeros were skipped; Spark folds them back in when nnz < count */
    double mn = s->mn[j], mx = s->mx[j];
    if (s->nnz[j] < (double)s->cnt) { if (mn > 0.0) mn = 0.0; if (mx < 0.0) mx = 0.0; }
    if (s->cnt == 0) { mn = 0.0; mx = 0.0; }
    o[0] = rmean; o[1] = var; o[2] = mn; o[3] = mx; o[4] = s->nnz[j]; o[5] = s->l1[j]; o[6] = sqrt(s->m2[j]);
  }
  for (int t = 0; t < threads; ++t) free(parts[t].mean);
  free(parts);
}

/* ------------------------------------------------------------------ LeastSquaresAggregator */

void orc_lsq_loss_grad(const orc_column* features, int32_t d, const double* label, int64_t n,
                       const int32_t* dq_guest, const double* dq_price, const double* coef,
                       const double* fstd, const double* fmean, double ystd, double ymean,
                       int fit_intercept, int threads, double* out) {
  /* effectiveCoefAndOffset */
  double* eff = (double*)malloc(sizeof(double) * (size_t)(d > 0 ? d : 1));
  double sum = 0.0;
  for (int32_t j = 0; j < d; ++j) {
    if (fstd[j] != 0.0) { eff[j] = coef[j] / fstd[j]; sum += eff[j] * fmean[j]; }
    else eff[j] = 0.0;
  }
  const double offset = fit_intercept ? ymean / ystd - sum : 0.0;
  const int64_t len = 2 + d;
  threads = pick_threads(threads, n);
  double* parts = (double*)calloc((size_t)len * threads, sizeof(double));
#pragma omp parallel num_threads(threads)
  {
#ifdef _OPENMP
    int t = omp_get_thread_num(), T = omp_get_num_threads();
#else
    int t = 0, T = 1;
#endif
    double* o = (double*)calloc((size_t)len + 16, sizeof(double)); /* thread-private: no false sharing */
    double* x = (double*)malloc(sizeof(double) * (size_t)(d > 0 ? d : 1));
    for (int64_t i = n * t / T; i < n * (t + 1) / T; ++i) {
      if (dq_guest && !orc_dq_keep(dq_guest[i], dq_price[i])) continue;
      const double weight = 1.0;
      double dot = 0.0;
      for (int32_t j = 0; j < d; ++j) { x[j] = col_at(&features[j], i); dot += x[j] * eff[j]; }
      const double diff = dot - label[i] / ystd + offset;
      if (diff != 0) {
        for (int32_t j = 0; j < d; ++j)
          if (fstd[j] != 0.0 && x[j] != 0.0) o[2 + j] += weight * diff * x[j] / fstd[j];
        o[0] += weight * diff * diff / 2.0;
      }
      o[1] += weight;
    }
    memcpy(parts + (int64_t)t * len, o, sizeof(double) * (size_t)len);
    free(o);
    free(x);
  }
  for (int64_t k = 0; k < len; ++k) out[k] = 0.0;
  for (int t = 0; t < threads; ++t)
    for (int64_t k = 0; k < len; ++k) out[k] += parts[(int64_t)t * len + k];
  free(parts);
  free(eff);
}

/* ------------------------------------------------------------------ predict + RegressionMetrics */

void orc_predict_metrics(const orc_column* features, int32_t d, const double* label, int64_t n,
                         const int32_t* dq_guest, const double* dq_price, const double* coef,
                         double intercept, int threads, double* prediction, double* out) {
  threads = pick_threads(threads, n);
  double* parts = (double*)calloc((size_t)5 * threads, sizeof(double));
#pragma omp parallel num_threads(threads)
  {
#ifdef _OPENMP
    int t = omp_get_thread_num(), T = omp_get_num_threads();
#else
    int t = 0, T = 1;
#endif
    double o[16] = {0}; /* thread-private: no false sharing */
    for (int64_t i = n * t / T; i < n * (t + 1) / T; ++i) {
      if (dq_guest && !orc_dq_keep(dq_guest[i], dq_price[i])) {
        if (prediction) prediction[i] = 0.0;
        continue;
      }
      double yhat = 0.0; /* BLAS.dot(features, coefficients) + intercept */
      for (int32_t j = 0; j < d; ++j) yhat += col_at(&features[j], i) * coef[j];
      yhat += intercept;
      if (prediction) prediction[i] = yhat;
      if (label) {
        const double r = label[i] - yhat;
        o[0] += 1.0; o[1] += r * r; o[2] += label[i]; o[3] += label[i] * label[i]; o[4] += fabs(r);
      }
    }
    memcpy(parts + 5 * t, o, sizeof(double) * 5);
  }
  for (int k = 0; k < 5; ++k) out[k] = 0.0;
  for (int t = 0; t < threads; ++t)
    for (int k = 0; k < 5; ++k) out[k] += parts[5 * t + k];
  free(parts);
}

/* ------------------------------------------------------------------ synthetic generator */

#define ORC_PHI 0x9E3779B97F4A7C15ULL

static inline uint64_t mix64(uint64_t z) {
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL;
  z ^= z >> 27; z *= 0x94D049BB133111EBULL;
  z ^= z >> 31;
  return z;
}
static inline uint64_t row_key(uint64_t seed, uint64_t r) { return mix64(seed + (r + 1) * ORC_PHI); }
static inline uint64_t stream(uint64_t key, uint64_t k) { return mix64(key + k * ORC_PHI); }
static inline double u01(uint64_t h) { return (double)(h >> 11) * 0x1.0p-53; }

static inline double synth_scale(int32_t k) { return (double)(1 + (k & 3)); }
static inline double synth_gamma(int32_t k) { return (double)((k * 7) % 5 - 2) * 0.5; }

void orc_synth_generate(const orc_synth_params* p, int64_t row0, int64_t n, int32_t* guest,
                        double* price, double* extra, int64_t extra_stride, int threads) {
  const double thr1 = p->p_low, thr2 = p->p_low + p->p_high;
  threads = pick_threads(threads, n);
#pragma omp parallel for num_threads(threads) schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    const uint64_t r = (uint64_t)(row0 + i);
    const uint64_t key = row_key(p->seed_data, r);
    const int32_t g = 1 + (int32_t)(((stream(key, 1) >> 32) * 35ULL) >> 32);
    double s = u01(stream(key, 2)) + u01(stream(key, 3));
    s = s + u01(stream(key, 4));
    s = s + u01(stream(key, 5));
    const double noise = (s - 2.0) * 2.6;
    double pr = (double)(5 * g + 22) + noise;
    for (int32_t k = 1; k < p->d; ++k) {
      const double xk = u01(stream(key, 8 + (uint64_t)k)) * synth_scale(k);
      if (extra) extra[(int64_t)(k - 1) * extra_stride + i] = xk;
      pr = pr + synth_gamma(k) * xk;
    }
    const uint64_t akey = row_key(p->seed_anom, r);
    const double ua = u01(mix64(akey));
    const double ub = u01(stream(akey, 1));
    if (ua < thr1) pr = 3.0 + 12.0 * ub;
    else if (ua < thr2 && g < 14) pr = 100.0 + 50.0 * ub;
    guest[i] = g;
    price[i] = pr;
  }
}

/* order-sensitive digest: sum over rows of a hash keyed by the output index */
void orc_rows_digest(const int32_t* guest, const double* price, int64_t n, uint64_t seed, uint64_t out[2]) {
  uint64_t a = 0, b = 0;
#pragma omp parallel for reduction(+ : a, b) schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    uint64_t bits;
    memcpy(&bits, &price[i], 8);
    uint64_t h = mix64((seed + ((uint64_t)i + 1) * ORC_PHI) ^ bits);
    h = mix64(h + (uint64_t)(uint32_t)guest[i] * 0xD1B54A32D192ED03ULL);
    a += h;
    b += mix64(h ^ ((uint64_t)i * 0x2545F4914F6CDD1DULL));
  }
  out[0] = a; out[1] = b;
}
