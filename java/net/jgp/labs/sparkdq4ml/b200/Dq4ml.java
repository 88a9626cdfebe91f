package net.jgp.labs.sparkdq4ml.b200;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemoryLayout;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.StructLayout;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;

/**
 * Panama binding of include/dq4ml.h (ABI version 2). NOT COMPILED in the repository's build image (no JDK there).
 *
 * One instance owns one dq4ml_ctx = one CUDA device + stream + scratch; a ctx is not thread-safe, so use one per
 * executor task thread. Data pointers handed to the library may be host (off-heap) or CUDA device memory.
 */
public final class Dq4ml implements AutoCloseable {
  public static final int ABI_VERSION = 2;
  public static final int I32 = 0, F64 = 1; // dq4ml_dtype
  public static final int SOLVER_AUTO = 0, SOLVER_NORMAL = 1, SOLVER_LBFGS = 2;
  public static final int LOSS_SQUARED_ERROR = 0, LOSS_HUBER = 1;
  public static final int METRICS_LEN = 16;

  private static final Linker LINKER = Linker.nativeLinker();
  private static final SymbolLookup LIB = SymbolLookup.libraryLookup("libdq4ml.so", Arena.global());

  private static MethodHandle h(String name, FunctionDescriptor d) {
    return LINKER.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), d);
  }

  // ---- context
  static final MethodHandle ABI = h("dq4ml_abi_version", FunctionDescriptor.of(JAVA_INT));
  static final MethodHandle LAST_ERROR = h("dq4ml_last_error", FunctionDescriptor.of(ADDRESS));
  static final MethodHandle CTX_CREATE = h("dq4ml_ctx_create", FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS));
  static final MethodHandle CTX_DESTROY = h("dq4ml_ctx_destroy", FunctionDescriptor.ofVoid(ADDRESS));
  static final MethodHandle NCCL_ID = h("dq4ml_nccl_unique_id", FunctionDescriptor.of(JAVA_INT, ADDRESS));
  static final MethodHandle INIT_NCCL = h("dq4ml_ctx_init_nccl", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT));
  static final MethodHandle DEVICE_ALLOC = h("dq4ml_device_alloc", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS));
  static final MethodHandle DEVICE_FREE = h("dq4ml_device_free", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
  static final MethodHandle MEMCPY = h("dq4ml_memcpy", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG));
  // ---- DQ rules
  static final MethodHandle CHECK_MIN = h("dq4ml_check_minimum_price", FunctionDescriptor.of(JAVA_DOUBLE, JAVA_DOUBLE));
  static final MethodHandle CHECK_RANGE = h("dq4ml_check_price_range", FunctionDescriptor.of(JAVA_DOUBLE, JAVA_DOUBLE, JAVA_INT));
  static final MethodHandle UDF_MIN = h("dq4ml_udf_minimumPriceRule",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
  static final MethodHandle UDF_CORR = h("dq4ml_udf_priceCorrelationRule",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
  static final MethodHandle DQ_FILTER = h("dq4ml_dq_filter",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
  static final MethodHandle DQ_FILTER_OFFSET = h("dq4ml_dq_filter_global_offset",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
  static final MethodHandle CSV_PARSE = h("dq4ml_csv_parse",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
  // ---- LinearRegression
  static final MethodHandle LINREG_FIT = h("dq4ml_linreg_fit_weighted",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS));
  static final MethodHandle MODEL_NUM_FEATURES = h("dq4ml_linreg_model_num_features", FunctionDescriptor.of(JAVA_INT, ADDRESS));
  static final MethodHandle MODEL_COEF = h("dq4ml_linreg_model_coefficients", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
  static final MethodHandle MODEL_INTERCEPT = h("dq4ml_linreg_model_intercept", FunctionDescriptor.of(JAVA_DOUBLE, ADDRESS));
  static final MethodHandle MODEL_SCALE = h("dq4ml_linreg_model_scale", FunctionDescriptor.of(JAVA_DOUBLE, ADDRESS));
  static final MethodHandle MODEL_ITERATIONS = h("dq4ml_linreg_model_total_iterations", FunctionDescriptor.of(JAVA_INT, ADDRESS));
  static final MethodHandle MODEL_HISTORY = h("dq4ml_linreg_model_objective_history", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT));
  static final MethodHandle MODEL_PREDICT1 = h("dq4ml_linreg_predict1", FunctionDescriptor.of(JAVA_DOUBLE, ADDRESS, ADDRESS));
  static final MethodHandle MODEL_INFERENCE = h("dq4ml_linreg_model_inference",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_DOUBLE, JAVA_DOUBLE, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT));
  static final MethodHandle MODEL_DESTROY = h("dq4ml_linreg_model_destroy", FunctionDescriptor.ofVoid(ADDRESS));
  static final MethodHandle TRANSFORM = h("dq4ml_linreg_transform", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
  static final MethodHandle EVALUATE = h("dq4ml_linreg_evaluate",
      FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, JAVA_INT, ADDRESS, ADDRESS));

  // ---- struct layouts, exactly as in the header
  /** struct dq4ml_column { const void* data; int32_t dtype; } */
  static final StructLayout COLUMN = MemoryLayout.structLayout(ADDRESS.withName("data"), JAVA_INT.withName("dtype"), MemoryLayout.paddingLayout(4));
  /** struct dq4ml_columns { const dq4ml_column* cols; int32_t n_cols; } */
  static final StructLayout COLUMNS = MemoryLayout.structLayout(ADDRESS.withName("cols"), JAVA_INT.withName("n_cols"), MemoryLayout.paddingLayout(4));
  /** struct dq4ml_dq_params { double min_price; int32_t min_guest; double max_price; } */
  static final StructLayout DQ_PARAMS = MemoryLayout.structLayout(JAVA_DOUBLE.withName("min_price"), JAVA_INT.withName("min_guest"),
      MemoryLayout.paddingLayout(4), JAVA_DOUBLE.withName("max_price"));
  /** struct dq4ml_dq_fused { const int32_t* guest; const double* price; dq4ml_dq_params params; } */
  static final StructLayout DQ_FUSED = MemoryLayout.structLayout(ADDRESS.withName("guest"), ADDRESS.withName("price"), DQ_PARAMS.withName("params"));
  /** struct dq4ml_lr_params: offsets 0 max_iter, 8 reg_param, 16 elastic_net_param, 24 tol, 32 fit_intercept,
   *  36 standardization, 40 solver, 44 loss, 48 epsilon; 56 bytes. */
  static final StructLayout LR_PARAMS = MemoryLayout.structLayout(
      JAVA_INT.withName("max_iter"), MemoryLayout.paddingLayout(4), JAVA_DOUBLE.withName("reg_param"),
      JAVA_DOUBLE.withName("elastic_net_param"), JAVA_DOUBLE.withName("tol"), JAVA_INT.withName("fit_intercept"),
      JAVA_INT.withName("standardization"), JAVA_INT.withName("solver"), JAVA_INT.withName("loss"), JAVA_DOUBLE.withName("epsilon"));

  final MemorySegment ctx;

  public Dq4ml(int device) throws Throwable {
    if ((int) ABI.invokeExact() != ABI_VERSION) throw new UnsatisfiedLinkError("libdq4ml.so ABI version mismatch");
    try (Arena a = Arena.ofConfined()) {
      MemorySegment out = a.allocate(ADDRESS);
      check((int) CTX_CREATE.invokeExact(device, out));
      ctx = out.get(ADDRESS, 0);
    }
  }

  /** Negative status -> exception with the library's message; DQ4ML_ERR_NULL_VALUE (-3) is the lab's NullPointerException. */
  static void check(int rc) throws Throwable {
    if (rc >= 0) return;
    String msg = ((MemorySegment) LAST_ERROR.invokeExact()).reinterpret(1024).getString(0);
    if (rc == -3) throw new NullPointerException(msg);
    if (rc == -1) throw new IllegalArgumentException(msg);
    throw new org.apache.spark.SparkException("dq4ml status " + rc + ": " + msg);
  }

  /** Row-sharded executors: rank 0 creates the id, ships the 128 bytes (e.g. a Spark broadcast), every rank binds. */
  public static byte[] ncclUniqueId() throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      MemorySegment id = a.allocate(128);
      check((int) NCCL_ID.invokeExact(id));
      return id.toArray(java.lang.foreign.ValueLayout.JAVA_BYTE);
    }
  }

  public void bindRanks(byte[] id128, int rank, int worldSize) throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      MemorySegment id = a.allocate(128);
      id.copyFrom(MemorySegment.ofArray(id128));
      check((int) INIT_NCCL.invokeExact(ctx, id, rank, worldSize));
    }
  }

  /** The whole DQ section (App:68-90) in one pass; returns {survivors, rows after rule 1}. Buffers: host or device. */
  public long[] dqFilter(MemorySegment guest, MemorySegment price, long n, MemorySegment guestOut, MemorySegment priceOut) throws Throwable {
    try (Arena a = Arena.ofConfined()) {
      MemorySegment nOut = a.allocate(JAVA_LONG), n1 = a.allocate(JAVA_LONG);
      check((int) DQ_FILTER.invokeExact(ctx, guest, price, n, MemorySegment.NULL, guestOut, priceOut, nOut, n1));
      return new long[] {nOut.get(JAVA_LONG, 0), n1.get(JAVA_LONG, 0)};
    }
  }

  /** Fused-DQ descriptor for fit / evaluate: the filtered copy of App:78,90 is never written. */
  public static MemorySegment dqFused(Arena a, MemorySegment guest, MemorySegment price) {
    MemorySegment f = a.allocate(DQ_FUSED);
    f.set(ADDRESS, 0, guest);
    f.set(ADDRESS, 8, price);
    f.set(JAVA_DOUBLE, 16, 20.0); // MinimumPriceDataQualityService.MIN_PRICE
    f.set(JAVA_INT, 24, 14);      // PriceCorrelationDataQualityService.checkPriceRange
    f.set(JAVA_DOUBLE, 32, 90.0);
    return f;
  }

  /** VectorAssembler (App:106-113): the "features" column is a table of column references, never a packed copy. */
  public static MemorySegment columns(Arena a, MemorySegment[] cols, int[] dtypes) {
    MemorySegment arr = a.allocate(COLUMN, cols.length);
    for (int j = 0; j < cols.length; j++) {
      arr.set(ADDRESS, j * COLUMN.byteSize(), cols[j]);
      arr.set(JAVA_INT, j * COLUMN.byteSize() + 8, dtypes[j]);
    }
    MemorySegment tab = a.allocate(COLUMNS);
    tab.set(ADDRESS, 0, arr);
    tab.set(JAVA_INT, 8, cols.length);
    return tab;
  }

  // ---- accessors for the classes that must live in the lab's own packages (the UDFs keep their FQCN)
  public static MethodHandle CHECK_MIN_HANDLE() { return CHECK_MIN; }
  public static MethodHandle CHECK_RANGE_HANDLE() { return CHECK_RANGE; }

  public static void udfMinimumPriceRule(Dq4ml q, MemorySegment price, MemorySegment valid, long n, MemorySegment out) throws Throwable {
    check((int) UDF_MIN.invokeExact(q.ctx, price, valid, n, out, MemorySegment.NULL));
  }

  public static void udfPriceCorrelationRule(Dq4ml q, MemorySegment price, MemorySegment guest, MemorySegment priceValid,
      MemorySegment guestValid, long n, MemorySegment out) throws Throwable {
    check((int) UDF_CORR.invokeExact(q.ctx, price, guest, priceValid, guestValid, n, out, MemorySegment.NULL));
  }

  @Override
  public void close() {
    try {
      CTX_DESTROY.invokeExact(ctx);
    } catch (Throwable t) {
      throw new RuntimeException(t);
    }
  }
}
