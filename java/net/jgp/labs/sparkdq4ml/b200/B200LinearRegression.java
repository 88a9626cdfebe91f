package net.jgp.labs.sparkdq4ml.b200;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;

/**
 * The estimator the lab builds at DataQuality4MachineLearningApp.java:120-123
 * (new LinearRegression().setMaxIter(40).setRegParam(1).setElasticNetParam(1)) with Spark 2.4.4's setters, defaults and
 * semantics; fit() = lr.fit(df) at line 126. NOT COMPILED in the repository's build image.
 *
 * Columns are passed as memory segments (host off-heap or CUDA device memory); the feature "vector" is the table of
 * column references VectorAssembler would have packed (lines 106-113 of the lab).
 */
public class B200LinearRegression {
  private int maxIter = 100;
  private double regParam = 0.0, elasticNetParam = 0.0, tol = 1e-6, epsilon = 1.35;
  private boolean fitIntercept = true, standardization = true;
  private int solver = Dq4ml.SOLVER_AUTO, loss = Dq4ml.LOSS_SQUARED_ERROR;

  public B200LinearRegression setMaxIter(int v) { maxIter = v; return this; }
  public B200LinearRegression setRegParam(double v) { regParam = v; return this; }
  public B200LinearRegression setElasticNetParam(double v) { elasticNetParam = v; return this; }
  public B200LinearRegression setTol(double v) { tol = v; return this; }
  public B200LinearRegression setFitIntercept(boolean v) { fitIntercept = v; return this; }
  public B200LinearRegression setStandardization(boolean v) { standardization = v; return this; }
  public B200LinearRegression setEpsilon(double v) { epsilon = v; return this; }

  public B200LinearRegression setSolver(String v) {
    switch (v) {
      case "auto": solver = Dq4ml.SOLVER_AUTO; break;
      case "normal": solver = Dq4ml.SOLVER_NORMAL; break;
      case "l-bfgs": solver = Dq4ml.SOLVER_LBFGS; break;
      default: throw new IllegalArgumentException("solver must be auto, normal or l-bfgs");
    }
    return this;
  }

  public B200LinearRegression setLoss(String v) {
    switch (v) {
      case "squaredError": loss = Dq4ml.LOSS_SQUARED_ERROR; break;
      case "huber": loss = Dq4ml.LOSS_HUBER; break;
      default: throw new IllegalArgumentException("loss must be squaredError or huber");
    }
    return this;
  }

  public int getMaxIter() { return maxIter; }
  public double getRegParam() { return regParam; }
  public double getElasticNetParam() { return elasticNetParam; }
  public double getTol() { return tol; }
  boolean getFitIntercept() { return fitIntercept; }

  /**
   * lr.fit(df). featureCols/dtypes: the VectorAssembler inputs in order (Dq4ml.I32 / Dq4ml.F64); label: the "label"
   * column (App:101); weight: MemorySegment.NULL unless setWeightCol was used; dqFused: Dq4ml.dqFused(...) to apply the
   * two DQ rules inside the fit's own passes, or MemorySegment.NULL for an already filtered table.
   */
  public B200LinearRegressionModel fit(Dq4ml q, MemorySegment[] featureCols, int[] dtypes, MemorySegment label, MemorySegment weight,
      long n, MemorySegment dqFused) throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      MemorySegment tab = Dq4ml.columns(a, featureCols, dtypes);
      MemorySegment p = a.allocate(Dq4ml.LR_PARAMS);
      p.set(JAVA_INT, 0, maxIter);
      p.set(JAVA_DOUBLE, 8, regParam);
      p.set(JAVA_DOUBLE, 16, elasticNetParam);
      p.set(JAVA_DOUBLE, 24, tol);
      p.set(JAVA_INT, 32, fitIntercept ? 1 : 0);
      p.set(JAVA_INT, 36, standardization ? 1 : 0);
      p.set(JAVA_INT, 40, solver);
      p.set(JAVA_INT, 44, loss);
      p.set(JAVA_DOUBLE, 48, epsilon);
      MemorySegment out = a.allocate(ADDRESS);
      Dq4ml.check((int) Dq4ml.LINREG_FIT.invokeExact(q.ctx, tab, label, weight, n, p, dqFused, out));
      return new B200LinearRegressionModel(q, this, out.get(ADDRESS, 0), featureCols, dtypes, label, n, dqFused);
    }
  }
}
