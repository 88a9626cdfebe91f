package net.jgp.labs.sparkdq4ml.b200;

import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;

/**
 * LinearRegressionModel plus the LinearRegressionTrainingSummary getters the lab prints
 * (DataQuality4MachineLearningApp.java:129-151): transform, summary().totalIterations() / objectiveHistory() /
 * residuals() / rootMeanSquaredError() / r2(), intercept(), getRegParam(), getTol(), predict(Vector), coefficients().
 * NOT COMPILED in the repository's build image.
 */
public final class B200LinearRegressionModel implements AutoCloseable {
  private final Dq4ml q;
  private final B200LinearRegression estimator;
  private final MemorySegment handle;
  // the training frame, kept for summary() like Spark keeps the training summary on the model
  private final MemorySegment[] featureCols;
  private final int[] dtypes;
  private final MemorySegment label, dqFused;
  private final long n;
  private double[] metrics; // filled by the first summary getter

  B200LinearRegressionModel(Dq4ml q, B200LinearRegression estimator, MemorySegment handle, MemorySegment[] featureCols, int[] dtypes,
      MemorySegment label, long n, MemorySegment dqFused) {
    this.q = q;
    this.estimator = estimator;
    this.handle = handle;
    this.featureCols = featureCols;
    this.dtypes = dtypes;
    this.label = label;
    this.n = n;
    this.dqFused = dqFused;
  }

  public int numFeatures() throws Throwable { return (int) Dq4ml.MODEL_NUM_FEATURES.invokeExact(handle); }

  public double[] coefficients() throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      int d = numFeatures();
      MemorySegment buf = a.allocate(JAVA_DOUBLE, Math.max(d, 1));
      Dq4ml.check((int) Dq4ml.MODEL_COEF.invokeExact(handle, buf));
      return buf.asSlice(0, 8L * d).toArray(JAVA_DOUBLE);
    }
  }

  public double intercept() throws Throwable { return (double) Dq4ml.MODEL_INTERCEPT.invokeExact(handle); } // App:141
  public double scale() throws Throwable { return (double) Dq4ml.MODEL_SCALE.invokeExact(handle); }
  public double getRegParam() { return estimator.getRegParam(); } // App:143
  public double getTol() { return estimator.getTol(); }           // App:145

  /** model.predict(Vectors.dense(40.0)) (App:150-151). */
  public double predict(double[] features) throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      MemorySegment x = a.allocateFrom(JAVA_DOUBLE, features);
      return (double) Dq4ml.MODEL_PREDICT1.invokeExact(handle, x);
    }
  }

  /** model.transform(df) (App:129): prediction[i] = features_i . coefficients + intercept, written to `prediction`. */
  public void transform(MemorySegment[] cols, int[] types, long rows, MemorySegment prediction) throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      Dq4ml.check((int) Dq4ml.TRANSFORM.invokeExact(q.ctx, handle, Dq4ml.columns(a, cols, types), rows, prediction));
    }
  }

  // ---- LinearRegressionTrainingSummary (App:132-139)
  public int totalIterations() throws Throwable { return (int) Dq4ml.MODEL_ITERATIONS.invokeExact(handle); }

  public double[] objectiveHistory() throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      int len = totalIterations();
      MemorySegment buf = a.allocate(JAVA_DOUBLE, Math.max(len, 1));
      int got = (int) Dq4ml.MODEL_HISTORY.invokeExact(handle, buf, len);
      return buf.asSlice(0, 8L * Math.min(got, len)).toArray(JAVA_DOUBLE);
    }
  }

  /** One fused pass over the training frame: every metric, and the residual column when `residualsOut` is not NULL. */
  private double[] evaluate(MemorySegment residualsOut) throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      MemorySegment met = a.allocate(JAVA_DOUBLE, Dq4ml.METRICS_LEN);
      Dq4ml.check((int) Dq4ml.EVALUATE.invokeExact(q.ctx, handle, Dq4ml.columns(a, featureCols, dtypes), label, n, dqFused,
          estimator.getFitIntercept() ? 1 : 0, met, residualsOut));
      return met.toArray(JAVA_DOUBLE);
    }
  }

  private double[] metrics() throws Throwable {
    if (metrics == null) metrics = evaluate(MemorySegment.NULL);
    return metrics;
  }

  /** summary().residuals() (App:137): label - prediction into `out` (n doubles, host or device). */
  public void residuals(MemorySegment out) throws Throwable { metrics = evaluate(out); }
  public long numInstances() throws Throwable { return (long) metrics()[0]; }
  public double rootMeanSquaredError() throws Throwable { return metrics()[5]; } // App:138
  public double r2() throws Throwable { return metrics()[6]; }                   // App:139
  public double meanSquaredError() throws Throwable { return metrics()[7]; }
  public double meanAbsoluteError() throws Throwable { return metrics()[8]; }
  public double explainedVariance() throws Throwable { return metrics()[9]; }
  public long degreesOfFreedom() throws Throwable { return (long) metrics()[10]; }
  public double r2adj() throws Throwable { return metrics()[11]; }
  public double[] devianceResiduals() throws Throwable { return new double[] {metrics()[12], metrics()[13]}; }

  /** coefficientStandardErrors / tValues / pValues (intercept last): normal-equation Cholesky fits only, like Spark. */
  public double[][] inference() throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      int k = numFeatures() + (estimator.getFitIntercept() ? 1 : 0);
      MemorySegment se = a.allocate(JAVA_DOUBLE, k), tv = a.allocate(JAVA_DOUBLE, k), pv = a.allocate(JAVA_DOUBLE, k);
      double[] m = metrics();
      Dq4ml.check((int) Dq4ml.MODEL_INFERENCE.invokeExact(handle, m[0], m[1], estimator.getFitIntercept() ? 1 : 0, se, tv, pv, k));
      return new double[][] {se.toArray(JAVA_DOUBLE), tv.toArray(JAVA_DOUBLE), pv.toArray(JAVA_DOUBLE)};
    }
  }

  @Override
  public void close() {
    try {
      Dq4ml.MODEL_DESTROY.invokeExact(handle);
    } catch (Throwable t) {
      throw new RuntimeException(t);
    }
  }
}
