package net.jgp.labs.sparkdq4ml.dq.udf;

import java.lang.foreign.MemorySegment;

import org.apache.spark.sql.api.java.UDF1;

import net.jgp.labs.sparkdq4ml.b200.Dq4ml;

/**
 * Drop-in for the lab's class of the same name (reference: dq/udf/MinimumPriceDataQualityUdf.java:7-14): same fully
 * qualified name, same serialVersionUID (so jobs serialised against the lab's jar still deserialise), same
 * UDF1&lt;Double, Double&gt; row contract, registered as "minimumPriceRule" (DataQuality4MachineLearningApp.java:46-47).
 * The rule itself now lives in libdq4ml.so. NOT COMPILED in the repository's build image.
 */
public class MinimumPriceDataQualityUdf implements UDF1<Double, Double> {

  private static final long serialVersionUID = -201966159201746851L;

  /** Row at a time, as Spark's ScalaUDF calls it. Unboxing a null price throws NullPointerException, like the lab. */
  @Override
  public Double call(Double price) throws Exception {
    double p = price; // NPE on null, the reference's behaviour (MinimumPriceDataQualityUdf.java:12)
    try {
      return (double) Dq4ml.CHECK_MIN_HANDLE().invokeExact(p);
    } catch (Exception e) {
      throw e;
    } catch (Throwable t) {
      throw new RuntimeException(t);
    }
  }

  /**
   * A whole column at once (host off-heap or device memory): what a columnar operator calls instead of n boxed
   * invocations. valid may be MemorySegment.NULL (no nulls); a null price is an error, as in the row form.
   */
  public static void callColumns(Dq4ml q, MemorySegment price, MemorySegment valid, long n, MemorySegment out) throws Throwable {
    Dq4ml.udfMinimumPriceRule(q, price, valid, n, out);
  }
}
