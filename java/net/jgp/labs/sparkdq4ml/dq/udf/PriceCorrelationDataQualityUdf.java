package net.jgp.labs.sparkdq4ml.dq.udf;

import java.lang.foreign.MemorySegment;

import org.apache.spark.sql.api.java.UDF2;

import net.jgp.labs.sparkdq4ml.b200.Dq4ml;

/**
 * Drop-in for the lab's class of the same name (reference: dq/udf/PriceCorrelationDataQualityUdf.java:7-18): same
 * fully qualified name and serialVersionUID, UDF2&lt;Double, Integer, Double&gt;, registered as "priceCorrelationRule"
 * (DataQuality4MachineLearningApp.java:48-49). A null price or guest gives -1.0 without reaching the rule, as in the
 * lab (lines 12-14 there). NOT COMPILED in the repository's build image.
 */
public class PriceCorrelationDataQualityUdf implements UDF2<Double, Integer, Double> {

  private static final long serialVersionUID = 4949954702581973224L;

  @Override
  public Double call(Double price, Integer guest) throws Exception {
    if (price == null || guest == null) {
      return -1.0;
    }
    try {
      return (double) Dq4ml.CHECK_RANGE_HANDLE().invokeExact((double) price, (int) guest);
    } catch (Exception e) {
      throw e;
    } catch (Throwable t) {
      throw new RuntimeException(t);
    }
  }

  /** Columnar form: byte-per-row validity columns (0 = null, MemorySegment.NULL = all valid) give -1.0 rows. */
  public static void callColumns(Dq4ml q, MemorySegment price, MemorySegment guest, MemorySegment priceValid,
      MemorySegment guestValid, long n, MemorySegment out) throws Throwable {
    Dq4ml.udfPriceCorrelationRule(q, price, guest, priceValid, guestValid, n, out);
  }
}
