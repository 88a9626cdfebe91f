// api.cpp -- the extern "C" boundary declared in include/dq4ml.h: argument checking, host/device buffer resolution,
// kernel launches, the cross-rank sum, and the hand-off to the host solvers.  There is deliberately no CPU code path
// for the data passes: without a CUDA device these entry points fail with DQ4ML_ERR_NO_DEVICE.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <new>
#include <vector>

#include "context.hpp"
#include "gram.hpp"
#include "solver.hpp"
#include "wide.hpp"

using namespace dq4ml;

struct dq4ml_lr_model {
  LrModel m;
};

namespace dq4ml {

static thread_local std::string g_error;

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_error = buf;
}
const char* get_error() { return g_error.c_str(); }

void collect_pass_time(dq4ml_ctx* ctx) {  // call after the stream has been synchronised
  if (!ctx->pass_event_pending) return;
  ctx->pass_event_pending = false;
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, ctx->pass_ev0, ctx->pass_ev1) == cudaSuccess) {
    ctx->pass_ms_total += ms;
    ++ctx->pass_count;
    if (ctx->pass_ms_list.size() < 65536) ctx->pass_ms_list.push_back(ms);
  }
}

namespace {

int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) return DQ4ML_ERR_NO_DEVICE;
  return DQ4ML_ERR_CUDA;
}
#define CU(call)                                       \
  do {                                                 \
    cudaError_t e__ = (call);                          \
    if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
  } while (0)

bool is_device_ptr(const void* p) {
  cudaPointerAttributes at;
  const cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

// ---- staging arena -------------------------------------------------------------------------------------------------
struct CallScope {  // resets the arena bookkeeping at the start of every entry point
  explicit CallScope(dq4ml_ctx* c) : ctx(c) {
    cudaSetDevice(c->device);
    c->arena_used = 0;
    c->staged.clear();
  }
  dq4ml_ctx* ctx;
};

int arena_reserve(dq4ml_ctx* ctx, size_t total) {
  if (total <= ctx->arena_cap) return DQ4ML_OK;
  CU(cudaStreamSynchronize(ctx->stream));
  if (ctx->arena != nullptr) CU(cudaFree(ctx->arena));
  ctx->arena = nullptr;
  ctx->arena_cap = 0;
  const size_t cap = (total + (64u << 20)) & ~static_cast<size_t>((1u << 20) - 1);
  CU(cudaMalloc(reinterpret_cast<void**>(&ctx->arena), cap));
  ctx->arena_cap = cap;
  return DQ4ML_OK;
}

size_t round256(size_t b) { return (b + 255) & ~static_cast<size_t>(255); }

void* arena_take(dq4ml_ctx* ctx, size_t bytes) {
  void* p = ctx->arena + ctx->arena_used;
  ctx->arena_used += round256(bytes);
  return p;
}

// Plans which of `ptrs` (with byte sizes) live on the host, reserves arena space once, then uploads them.
struct Resolver {
  dq4ml_ctx* ctx;
  struct Item {
    const void** slot;
    size_t bytes;
  };
  std::vector<Item> inputs;
  size_t extra_bytes = 0;
  explicit Resolver(dq4ml_ctx* c) : ctx(c) {}
  void input(const void** slot, size_t bytes) {
    if (*slot != nullptr) inputs.push_back({slot, bytes});
  }
  void reserve_output(size_t bytes) { extra_bytes += round256(bytes); }
  int run() {
    size_t need = extra_bytes;
    std::vector<char> on_host(inputs.size(), 0);
    std::unordered_map<const void*, size_t> seen;
    for (size_t i = 0; i < inputs.size(); ++i) {
      const void* p = *inputs[i].slot;
      if (is_device_ptr(p)) {
        if ((reinterpret_cast<uintptr_t>(p) & 31u) != 0) {  // 256-bit loads
          set_error("device column %p is not 32-byte aligned", p);
          return DQ4ML_ERR_INVALID;
        }
        continue;
      }
      on_host[i] = 1;
      auto it = seen.find(p);
      if (it == seen.end()) {
        seen[p] = inputs[i].bytes;
        need += round256(inputs[i].bytes);
      } else if (inputs[i].bytes > it->second) {
        need += round256(inputs[i].bytes);  // conservative
        it->second = inputs[i].bytes;
      }
    }
    const int rc = arena_reserve(ctx, need);
    if (rc != DQ4ML_OK) return rc;
    for (size_t i = 0; i < inputs.size(); ++i) {
      if (!on_host[i]) continue;
      const void* hp = *inputs[i].slot;
      auto it = ctx->staged.find(hp);
      if (it == ctx->staged.end()) {
        void* dp = arena_take(ctx, inputs[i].bytes);
        CU(cudaMemcpyAsync(dp, hp, inputs[i].bytes, cudaMemcpyHostToDevice, ctx->stream));
        ctx->staged[hp] = dp;
        *inputs[i].slot = dp;
      } else {
        *inputs[i].slot = it->second;  // same host column referenced twice (e.g. price as label and as DQ column)
      }
    }
    return DQ4ML_OK;
  }
};

struct OutBuf {  // an output that may live on the host: kernels write `dev`, finish() copies back
  void* user = nullptr;
  void* dev = nullptr;
  bool staged = false;
};
OutBuf make_out(dq4ml_ctx* ctx, void* user, size_t bytes) {
  OutBuf o;
  o.user = user;
  if (user == nullptr) return o;
  if (is_device_ptr(user)) {
    o.dev = user;
  } else {
    o.dev = arena_take(ctx, bytes);
    o.staged = true;
  }
  return o;
}

int check_ctx(const dq4ml_ctx* ctx) {
  if (ctx == nullptr) {
    set_error("null ctx");
    return DQ4ML_ERR_INVALID;
  }
  return DQ4ML_OK;
}

// ---- how a streaming pass hands its result vector to the host ---------------------------------------------------------
// arm_epilogue() before the launch decides the route, finish_result() after it collects the vector into ctx->h_result:
//   (a) epilogue, one rank or host-callback ranks: the last CTA writes mapped host memory, the host polls a word;
//   (b) epilogue, NCCL-bound ranks with peer mailboxes: the same, after the in-kernel exchange over NVLink
//       (fold in rank order, or gather of all ranks' vectors);
//   (c) NCCL-bound ranks without mailboxes (IPC unavailable, DQ4ML_XRANK=nccl): ncclAllReduce + copy + synchronise.
int host_allreduce(dq4ml_ctx* ctx, double* buf, int64_t count);

struct PassOpts {
  bool reduce = true;    // sum (fold) over the bound ranks
  bool gather = false;   // instead of folding, give the host every rank's vector (world x len doubles)
  int extra_consts = 0;  // device constants appended to the vector (the summarizer's centres)
  bool timed = true;     // counts as an aggregation pass for dq4ml_ctx_enable_pass_timing
};

void arm_epilogue(dq4ml_ctx* ctx, const PassOpts& o) {
  XrEpilogue& ep = ctx->scratch.ep;
  ep = XrEpilogue{};
  ep.extra_consts = o.extra_consts;
  ctx->ep_armed = false;
  ctx->ep_world = 1;
  static const bool disabled = [] {  // A/B switch: DQ4ML_EPILOGUE=off keeps copy + stream synchronisation on every pass
    const char* env = std::getenv("DQ4ML_EPILOGUE");
    return env != nullptr && std::strcmp(env, "off") == 0;
  }();
  if (disabled) return;
  const bool cross = (o.reduce || o.gather) && ctx->nccl != nullptr && ctx->world > 1;
  if (cross) {
    if (!nccl_mailboxes(ctx, &ep)) return;  // route (c): the kernel leaves the vector in scratch.result
    ep.gather = o.gather ? 1 : 0;
    ep.xr_seq = ++ctx->xr_seq;
    ctx->ep_world = ctx->world;
  }
  ep.host_out = ctx->h_words_dev;
  ep.host_seq = ++ctx->host_seq;
  ctx->ep_armed = true;
}

// Polls the tagged word pairs the epilogue writes (epilogue.cuh) and decodes `count` doubles into ctx->h_result.
int wait_host_words(dq4ml_ctx* ctx, int64_t count) {
  const unsigned tag = static_cast<unsigned>(ctx->scratch.ep.host_seq);
  volatile unsigned long long* w = ctx->h_words;
  unsigned spins = 0;
  for (int64_t k = 0; k < count; ++k) {
    for (;;) {
      const unsigned long long a = w[2 * k], b = w[2 * k + 1];
      if (static_cast<unsigned>(a >> 32) == tag && static_cast<unsigned>(b >> 32) == tag) {
        const unsigned long long bits = (a & 0xffffffffull) | (b << 32);
        std::memcpy(&ctx->h_result[k], &bits, sizeof(double));
        break;
      }
      if (static_cast<unsigned>(w[0] >> 32) == kXrErrorTag) {
        cudaStreamSynchronize(ctx->stream);
        set_error("cross-rank epilogue timed out waiting for a peer's result vector");
        return DQ4ML_ERR_COMM;
      }
      if ((++spins & 4095u) == 0u) {  // a failed launch / kernel never writes the words: look at the stream now and then
        const cudaError_t e = cudaStreamQuery(ctx->stream);
        if (e == cudaSuccess) {
          const unsigned long long a2 = w[2 * k], b2 = w[2 * k + 1];
          if (static_cast<unsigned>(a2 >> 32) == tag && static_cast<unsigned>(b2 >> 32) == tag) continue;  // just landed
          set_error("the pass finished without publishing its result");
          return DQ4ML_ERR_CUDA;
        }
        if (e != cudaErrorNotReady) return cuda_fail(e, "streaming pass");
      }
#if defined(__x86_64__) || defined(__i386__)
      __builtin_ia32_pause();
#endif
    }
  }
  return DQ4ML_OK;
}

// Collects `count` doubles of the pass just launched into ctx->h_result (summed over ranks when the pass was armed with
// reduce; world x count doubles, rank-major, when it was armed with gather).  *gathered_world (nullable) receives the
// number of rank vectors now in h_result for a gather pass.
int finish_result(dq4ml_ctx* ctx, int64_t count, const PassOpts& o = PassOpts{}) {
  const bool multi = ctx->nccl != nullptr || ctx->allreduce_cb != nullptr;
  const int world = multi && ctx->world > 0 ? ctx->world : 1;
  if (static_cast<size_t>(count) * (o.gather ? world : 1) > kResultDoubles) {
    set_error("result vector too long (%lld)", static_cast<long long>(count));
    return DQ4ML_ERR_UNSUPPORTED;
  }
  bool have_all_ranks = false;  // h_result already holds the cross-rank fold / gather
  if (ctx->ep_armed) {
    ctx->ep_armed = false;
    const int rc = wait_host_words(ctx, count * (o.gather ? ctx->ep_world : 1));
    if (ctx->pass_event_pending) cudaEventSynchronize(ctx->pass_ev1);
    collect_pass_time(ctx);
    if (rc != DQ4ML_OK) return rc;
    have_all_ranks = ctx->ep_world > 1;
  } else {
    if (o.reduce && !o.gather && ctx->nccl != nullptr) {
      const int rc = nccl_allreduce_device(ctx, ctx->scratch.result, count);
      if (rc != DQ4ML_OK) return rc;
      have_all_ranks = true;
    }
    CU(cudaMemcpyAsync(ctx->h_result, ctx->scratch.result, sizeof(double) * count, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    collect_pass_time(ctx);
  }
  if (have_all_ranks || world == 1 || !(o.reduce || o.gather)) return DQ4ML_OK;
  if (o.gather) {  // every rank's vector through a zero-padded sum (host callback, or NCCL without mailboxes)
    std::vector<double> all(static_cast<size_t>(count) * world, 0.0);
    std::copy(ctx->h_result, ctx->h_result + count, all.begin() + static_cast<size_t>(ctx->rank) * count);
    const int rc = host_allreduce(ctx, all.data(), static_cast<int64_t>(all.size()));
    if (rc != DQ4ML_OK) return rc;
    std::copy(all.begin(), all.end(), ctx->h_result);
    return DQ4ML_OK;
  }
  if (ctx->allreduce_cb != nullptr && ctx->allreduce_cb(ctx->allreduce_user, ctx->h_result, count) != 0) {
    set_error("all-reduce callback failed");
    return DQ4ML_ERR_COMM;
  }
  return DQ4ML_OK;
}

int host_allreduce(dq4ml_ctx* ctx, double* buf, int64_t count) {
  if (count <= 0) return DQ4ML_OK;
  if (ctx->nccl != nullptr) {
    if (static_cast<size_t>(count) > kResultDoubles) {
      set_error("all-reduce of %lld doubles exceeds the staging buffer", static_cast<long long>(count));
      return DQ4ML_ERR_UNSUPPORTED;
    }
    CU(cudaMemcpyAsync(ctx->scratch.result, buf, sizeof(double) * count, cudaMemcpyHostToDevice, ctx->stream));
    const int rc = nccl_allreduce_device(ctx, ctx->scratch.result, count);
    if (rc != DQ4ML_OK) return rc;
    CU(cudaMemcpyAsync(buf, ctx->scratch.result, sizeof(double) * count, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
  } else if (ctx->allreduce_cb != nullptr) {
    if (ctx->allreduce_cb(ctx->allreduce_user, buf, count) != 0) {
      set_error("all-reduce callback failed");
      return DQ4ML_ERR_COMM;
    }
  }
  return DQ4ML_OK;
}

bool multi_rank(const dq4ml_ctx* ctx) { return ctx->nccl != nullptr || ctx->allreduce_cb != nullptr; }

// ---- column table plumbing -----------------------------------------------------------------------------------------
struct PassInputs {
  std::vector<const void*> ptr;  // every feature column (any width)
  std::vector<int> dtype;
  DeviceCols cols;               // the same table in the fixed-size form the streaming kernels take (width <= kMaxCols)
  const double* label = nullptr;
  const double* weight = nullptr;  // instance weights (weightCol); nullptr = every weight 1
  DeviceCols wcols;                // [weight, features...] for the weighted streaming ops (width + 1 <= kMaxCols)
  DeviceDq dq;
  int width() const { return static_cast<int>(ptr.size()); }
};

int gather_inputs(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                  const dq4ml_dq_fused* dq, PassInputs* out, size_t out_bytes, int max_cols,
                  const double* weight = nullptr) {
  if (features == nullptr || n < 0 || features->n_cols < 0 || (features->n_cols > 0 && features->cols == nullptr)) {
    set_error("bad feature table / row count");
    return DQ4ML_ERR_INVALID;
  }
  if (features->n_cols > max_cols) {
    set_error("%d feature columns: this entry point handles at most %d", features->n_cols, max_cols);
    return DQ4ML_ERR_UNSUPPORTED;
  }
  const int d = features->n_cols;
  out->ptr.assign(d, nullptr);
  out->dtype.assign(d, DQ4ML_F64);
  for (int j = 0; j < d; ++j) {
    const dq4ml_column& c = features->cols[j];
    if ((c.data == nullptr && n > 0) || (c.dtype != DQ4ML_I32 && c.dtype != DQ4ML_F64)) {
      set_error("feature column %d: null data or unknown dtype %d", j, c.dtype);
      return DQ4ML_ERR_INVALID;
    }
    out->ptr[j] = c.data;
    out->dtype[j] = c.dtype;
  }
  out->label = label;
  out->weight = weight;
  if (dq != nullptr) {
    if ((dq->guest == nullptr || dq->price == nullptr) && n > 0) {
      set_error("fused DQ needs both guest and price columns");
      return DQ4ML_ERR_INVALID;
    }
    out->dq.guest = dq->guest;
    out->dq.price = dq->price;
    out->dq.params = dq->params;
  }
  int rc = DQ4ML_OK;
  if (n == 0) {
    // an empty shard still takes part in the pass (and in the cross-rank sum): give the kernels valid, unread pointers
    const void* dummy = ctx->scratch.consts;
    for (int j = 0; j < d; ++j) out->ptr[j] = dummy;
    out->label = static_cast<const double*>(dummy);
    if (weight != nullptr) out->weight = static_cast<const double*>(dummy);
    if (dq != nullptr) {
      out->dq.guest = static_cast<const int32_t*>(dummy);
      out->dq.price = static_cast<const double*>(dummy);
    }
    rc = arena_reserve(ctx, out_bytes);
  } else {
    Resolver rs(ctx);
    for (int j = 0; j < d; ++j) rs.input(&out->ptr[j], static_cast<size_t>(n) * (out->dtype[j] == DQ4ML_I32 ? 4 : 8));
    rs.input(reinterpret_cast<const void**>(&out->label), static_cast<size_t>(n) * 8);
    rs.input(reinterpret_cast<const void**>(&out->weight), static_cast<size_t>(n) * 8);
    rs.input(reinterpret_cast<const void**>(&out->dq.guest), static_cast<size_t>(n) * 4);
    rs.input(reinterpret_cast<const void**>(&out->dq.price), static_cast<size_t>(n) * 8);
    rs.reserve_output(out_bytes);
    rc = rs.run();
  }
  if (rc != DQ4ML_OK) return rc;
  out->cols.n = d <= kMaxCols ? d : 0;
  for (int j = 0; j < out->cols.n; ++j) {
    out->cols.ptr[j] = out->ptr[j];
    out->cols.dtype[j] = out->dtype[j];
  }
  out->wcols.n = 0;
  if (out->weight != nullptr && d + 1 <= kMaxCols) {
    out->wcols.n = d + 1;
    out->wcols.ptr[0] = out->weight;
    out->wcols.dtype[0] = DQ4ML_F64;
    for (int j = 0; j < d; ++j) {
      out->wcols.ptr[1 + j] = out->ptr[j];
      out->wcols.dtype[1 + j] = out->dtype[j];
    }
  }
  return DQ4ML_OK;
}

// ---- the aggregation passes (device inputs) -------------------------------------------------------------------------

// One streaming pass: arms the epilogue, brackets the kernel with CUDA events on the ctx stream when pass timing is on.
cudaError_t run_stream_pass(dq4ml_ctx* ctx, StreamOpKind op, const DeviceCols& cols, const double* label, const DeviceDq& dq,
                            int64_t n, const void* host_consts, double* row_out, const PassOpts& o = PassOpts{}) {
  arm_epilogue(ctx, o);
  const bool timed = ctx->pass_timing && o.timed;
  if (timed) cudaEventRecord(ctx->pass_ev0, ctx->stream);
  const cudaError_t e = launch_stream_pass(op, cols, label, dq, n, host_consts, row_out, ctx->scratch, ctx->stream,
                                           &ctx->launches);
  if (timed) {
    cudaEventRecord(ctx->pass_ev1, ctx->stream);
    ctx->pass_event_pending = true;
  }
  if (e != cudaSuccess) ctx->ep_armed = false;
  return e;
}

// WLS moments of a small-d table via the streaming kernel; result in `out` with the public layout.
int wls_moments_small(dq4ml_ctx* ctx, const PassInputs& in, int64_t n, double* out) {
  const int d = in.cols.n;
  const int b = stream_bucket(d > 0 ? d : 1);
  const int nacc = stream_nacc(kOpWls, b);
  cudaError_t e = run_stream_pass(ctx, kOpWls, in.cols, in.label, in.dq, n, nullptr, nullptr);
  if (e != cudaSuccess) return cuda_fail(e, "wls moments pass");
  const int rc = finish_result(ctx, nacc);
  if (rc != DQ4ML_OK) return rc;
  const double* r = ctx->h_result;
  out[0] = r[0];
  out[1] = r[0];
  out[2] = r[0];
  out[3] = r[1];
  out[4] = r[2];
  for (int j = 0; j < d; ++j) {
    out[5 + j] = r[3 + j];
    out[5 + d + j] = r[3 + b + j];
  }
  const size_t tri = static_cast<size_t>(d) * (d + 1) / 2;
  for (size_t q = 0; q < tri; ++q) out[5 + 2 * d + q] = r[3 + 2 * b + q];  // packed index is independent of the order
  return DQ4ML_OK;
}

int wls_moments_any(dq4ml_ctx* ctx, const PassInputs& in, int64_t n, double* out) {
  if (in.label == nullptr) {
    set_error("label column is required");
    return DQ4ML_ERR_INVALID;
  }
  // The choice between the three kernels must not depend on this rank's row count: the streaming kernel sums a
  // bucket-padded vector across ranks and the two Gram kernels the public layout, so every rank has to take the same
  // branch.  A rank whose shard is empty under the tf32 tier runs the fp64 Gram kernel over zero rows, which adds the
  // same `len` doubles (all zero) to the cross-rank sum.
  const bool tf32_tier = ctx->gram_tier == DQ4ML_GRAM_TF32X3 && gram_tf32_supported(in.width());
  if (in.width() <= 8 && in.weight == nullptr && !tf32_tier) return wls_moments_small(ctx, in, n, out);
  const int rc = (tf32_tier && n > 0) ? gram_tf32_wls_moments(ctx, in.ptr, in.dtype, in.label, in.weight, in.dq, n, out)
                                      : gram_wls_moments(ctx, in.ptr, in.dtype, in.label, in.weight, in.dq, n, out);
  // WeightedLeastSquares.Aggregator.add: require(w >= 0.0).  A negative weight turns sqrt(w) into NaN in wSum.
  if (rc == DQ4ML_OK && in.weight != nullptr && !(out[1] == out[1])) {
    set_error("instance weight must be >= 0 and not NaN");
    return DQ4ML_ERR_INVALID;
  }
  return rc;
}

// ---- merging summarizers across ranks ---------------------------------------------------------------------------------
// Every rank accumulates shifted power sums S1 = sum w (v - c), S2 = sum w (v - c)^2 around ITS OWN provisional centre c
// (the head of its shard of the column), and publishes c next to them.  The host turns each rank's triple into
// (W, mean, M2 = sum w (v - mean)^2) and merges the ranks in rank order with the pairwise update of Chan et al. -- the
// same thing MultivariateOnlineSummarizer.merge does to the partition summarizers in Spark.  No rank needs to know
// another rank's centre before the pass, so the pass costs one launch and one hand-over, like every other pass.
struct Moments {
  double w = 0.0, mean = 0.0, m2 = 0.0;
  void merge(double wb, double centre, double s1, double s2) {
    if (!(wb > 0.0)) return;
    const double mean_b = centre + s1 / wb, m2_b = s2 - s1 * s1 / wb;
    if (!(w > 0.0)) {
      w = wb;
      mean = mean_b;
      m2 = m2_b;
      return;
    }
    const double tot = w + wb, delta = mean_b - mean;
    mean += delta * wb / tot;
    m2 += m2_b + delta * delta * w * wb / tot;
    w = tot;
  }
};

// Launches the provisional-centre kernel and the summarizer pass `op` over one column group and leaves every rank's
// vector [accumulators (nacc), centres (b + 1)] in ctx->h_result, rank-major.  Returns the number of rank vectors.
int summarizer_pass(dq4ml_ctx* ctx, StreamOpKind op, const DeviceCols& cols, const double* label, const DeviceDq& dq,
                    int64_t n, int b, int n_centres, bool timed, int* world_out, int* len_out) {
  cudaError_t e = launch_column_heads(cols, label, b, n, ctx->scratch, ctx->stream, &ctx->launches);
  if (e != cudaSuccess) return cuda_fail(e, "column heads");
  PassOpts o;
  o.gather = true;
  o.extra_consts = n_centres;
  o.timed = timed;
  e = run_stream_pass(ctx, op, cols, label, dq, n, nullptr, nullptr, o);
  if (e != cudaSuccess) return cuda_fail(e, "summarizer pass");
  const int len = stream_nacc(op, b) + n_centres;
  const int rc = finish_result(ctx, len, o);
  if (rc != DQ4ML_OK) return rc;
  *world_out = multi_rank(ctx) && ctx->world > 0 ? ctx->world : 1;
  *len_out = len;
  return DQ4ML_OK;
}

// Weighted summarizer for the l-bfgs branch: count (rows with w != 0), weighted mean and the unbiased weighted variance
// MultivariateOnlineSummarizer.variance returns, sum(w (v - mean)^2) / (W - sum(w^2) / W).  The weight rides as column 0
// of the streamed table, so wide tables go through in groups of kMaxCols - 1 feature columns (the label with the first).
int summarize_core_weighted(dq4ml_ctx* ctx, const PassInputs& in, int64_t n, ColumnStats* st) {
  const int d = in.width();
  st->mean.assign(d + 1, 0.0);
  st->variance.assign(d + 1, 0.0);
  double count = 0.0, w_sum = 0.0, ww_sum = 0.0;
  std::vector<Moments> mom(d + 1);
  constexpr int kGroup = kMaxCols - 1;
  for (int c0 = 0; c0 == 0 || c0 < d; c0 += kGroup) {
    DeviceCols g;
    const int gd = std::min(kGroup, d - c0);
    g.n = gd + 1;
    g.ptr[0] = in.weight;
    g.dtype[0] = DQ4ML_F64;
    for (int j = 0; j < gd; ++j) {
      g.ptr[1 + j] = in.ptr[c0 + j];
      g.dtype[1 + j] = in.dtype[c0 + j];
    }
    const bool with_label = c0 == 0;
    const int b = stream_bucket(g.n);
    int world = 1, len = 0;
    const int rc = summarizer_pass(ctx, kOpSummCoreW, g, with_label ? in.label : nullptr, in.dq, n, b, b + 1, true, &world, &len);
    if (rc != DQ4ML_OK) return rc;
    const int nacc = stream_nacc(kOpSummCoreW, b);
    for (int r = 0; r < world; ++r) {
      const double* v = ctx->h_result + static_cast<size_t>(r) * len;
      const double* c = v + nacc;  // centres: slot j of the table (slot 0 is the weight itself), the label at slot b
      if (with_label) {
        count += v[0];
        w_sum += v[1];
        ww_sum += v[2];
        mom[d].merge(v[1], c[b], v[1 + 2 * b], v[2 + 2 * b]);
      }
      for (int j = 0; j < gd; ++j) mom[c0 + j].merge(v[1], c[1 + j], v[1 + 2 * (1 + j)], v[2 + 2 * (1 + j)]);
    }
  }
  if (!(w_sum == w_sum) || w_sum < 0.0) {
    set_error("instance weight must be >= 0 and not NaN");
    return DQ4ML_ERR_INVALID;
  }
  st->count = count;
  const double denom = w_sum > 0.0 ? w_sum - ww_sum / w_sum : 0.0;
  for (int j = 0; j <= d; ++j) {
    if (w_sum > 0.0) {
      st->mean[j] = mom[j].mean;
      st->variance[j] = denom > 0.0 ? std::max(mom[j].m2 / denom, 0.0) : 0.0;
    }
  }
  return DQ4ML_OK;
}

// Summarizer, core statistics (count, mean, sample variance per column; label last) for the l-bfgs branch.
// One summarizer launch over a group of <= kMaxCols feature columns (+ the label when non-null): fills count and the
// mean / sample variance of the group's columns (group column c -> out index first + c; label -> out index label_out).
int summarize_group(dq4ml_ctx* ctx, const DeviceCols& cols, const double* label, const DeviceDq& dq, int64_t n, int first,
                    int label_out, ColumnStats* st) {
  const int d = cols.n;
  const int b = stream_bucket(d > 0 ? d : 1);
  int world = 1, len = 0;
  const int rc = summarizer_pass(ctx, kOpSummCore, cols, label, dq, n, b, b + 1, true, &world, &len);
  if (rc != DQ4ML_OK) return rc;
  const int nacc = stream_nacc(kOpSummCore, b);
  double cnt = 0.0;
  std::vector<Moments> mom(d + 1);
  for (int r = 0; r < world; ++r) {
    const double* v = ctx->h_result + static_cast<size_t>(r) * len;
    const double* c = v + nacc;
    cnt += v[0];
    for (int j = 0; j < d; ++j) mom[j].merge(v[0], c[j], v[1 + 2 * j], v[2 + 2 * j]);
    mom[d].merge(v[0], c[b], v[1 + 2 * b], v[2 + 2 * b]);
  }
  st->count = cnt;
  auto read_out = [&](int slot, int out) {
    if (cnt > 0) {
      st->mean[out] = mom[slot].mean;
      st->variance[out] = cnt > 1 ? std::max(mom[slot].m2 / (cnt - 1.0), 0.0) : 0.0;
    }
  };
  for (int c = 0; c < d; ++c) read_out(c, first + c);
  if (label != nullptr && label_out >= 0) read_out(d, label_out);
  return DQ4ML_OK;
}

// Summarizer, core statistics (count, mean, sample variance per column; label last) for the l-bfgs branch.  Tables
// wider than kMaxCols go through in column groups (each column is still read exactly once; the DQ columns once per group).
int summarize_core(dq4ml_ctx* ctx, const PassInputs& in, int64_t n, ColumnStats* st) {
  const int d = in.width();
  st->mean.assign(d + 1, 0.0);
  st->variance.assign(d + 1, 0.0);
  if (d <= kMaxCols) return summarize_group(ctx, in.cols, in.label, in.dq, n, 0, d, st);
  for (int c0 = 0; c0 < d; c0 += kMaxCols) {
    DeviceCols g;
    g.n = std::min(kMaxCols, d - c0);
    for (int j = 0; j < g.n; ++j) {
      g.ptr[j] = in.ptr[c0 + j];
      g.dtype[j] = in.dtype[c0 + j];
    }
    const bool with_label = c0 == 0;
    const int rc = summarize_group(ctx, g, with_label ? in.label : nullptr, in.dq, n, c0, with_label ? d : -1, st);
    if (rc != DQ4ML_OK) return rc;
  }
  return DQ4ML_OK;
}

int lsq_pass(dq4ml_ctx* ctx, const PassInputs& in, int64_t n, const double* coef, const double* fstd,
             const double* fmean, double ystd, double ymean, bool fit_intercept, double* loss_sum, double* weight_sum,
             double* grad_sum, double* diff_scratch = nullptr) {
  const int d = in.width();
  if (in.weight != nullptr && in.wcols.n > 0) {  // weight in column 0, features in 1..d (d <= kMaxCols - 1)
    int rc = DQ4ML_OK;
    const int b = stream_bucket(d + 1);
    LsqConsts c{};
    double s = 0.0;
    for (int j = 0; j < d; ++j) {
      c.eff[1 + j] = fstd[j] != 0.0 ? coef[j] / fstd[j] : 0.0;
      s += c.eff[1 + j] * fmean[j];
    }
    c.offset = fit_intercept ? ymean / ystd - s : 0.0;
    c.inv_ystd = 1.0 / ystd;
    cudaError_t e = run_stream_pass(ctx, kOpLsqW, in.wcols, in.label, in.dq, n, &c, nullptr);
    if (e != cudaSuccess) return cuda_fail(e, "weighted loss/gradient pass");
    rc = finish_result(ctx, stream_nacc(kOpLsqW, b));
    if (rc != DQ4ML_OK) return rc;
    const double* r = ctx->h_result;
    *loss_sum = r[0];
    *weight_sum = r[1];
    for (int j = 0; j < d; ++j) grad_sum[j] = fstd[j] != 0.0 ? r[3 + j] / fstd[j] : 0.0;
    return DQ4ML_OK;
  }
  if (d > kMaxCols || in.weight != nullptr) {
    // Wide table (or a weighted one too wide for the weight to ride as column 0): pass 1 writes the scaled difference
    // r = y/yStd - (x.eff + offset) per row -- times the instance weight when there is one -- and its sums, pass 2 takes
    // gradSum_j = -sum (w) r x_j over column groups.  The table is read twice; Spark reads it once, but keeps it in host RAM.
    if (diff_scratch == nullptr) {
      set_error("internal: no scratch column for the wide loss/gradient pass");
      return DQ4ML_ERR_INVALID;
    }
    std::vector<double> eff(d);
    double s = 0.0;
    for (int j = 0; j < d; ++j) {
      eff[j] = fstd[j] != 0.0 ? coef[j] / fstd[j] : 0.0;
      s += eff[j] * fmean[j];
    }
    const double offset = fit_intercept ? ymean / ystd - s : 0.0;
    WideSums ws;
    int rc = wide_rows_pass(ctx, in.ptr, in.dtype, eff.data(), offset, 1.0 / ystd, in.label, in.dq, n, 1, diff_scratch, false, &ws,
                            in.weight);
    if (rc != DQ4ML_OK) return rc;
    std::vector<double> all(2 + d, 0.0);
    all[0] = 0.5 * ws.ss_res;
    all[1] = ws.n;
    // groups of 16 columns: the 16-column streaming kernel runs at the HBM roofline (4 rows per thread, 256-bit loads),
    // the 32-column one at 0.78 of it, which outweighs reading the 8-byte difference column twice as often
    constexpr int kGradGroup = 16;  // measured at 10^8 rows x 64 columns: 19.0 ms per evaluation against 19.8 ms with groups of 32
    for (int c0 = 0; c0 < d; c0 += kGradGroup) {
      DeviceCols g;
      g.n = std::min(kGradGroup, d - c0);
      for (int j = 0; j < g.n; ++j) {
        g.ptr[j] = in.ptr[c0 + j];
        g.dtype[j] = in.dtype[c0 + j];
      }
      PassOpts local;
      local.reduce = false;  // the groups are summed over ranks together, below
      cudaError_t e = run_stream_pass(ctx, kOpGrad, g, diff_scratch, DeviceDq{}, n, nullptr, nullptr, local);
      if (e != cudaSuccess) return cuda_fail(e, "gradient pass");
      rc = finish_result(ctx, stream_nacc(kOpGrad, stream_bucket(g.n)), local);
      if (rc != DQ4ML_OK) return rc;
      for (int j = 0; j < g.n; ++j) all[2 + c0 + j] = -ctx->h_result[j];
    }
    if (multi_rank(ctx)) {
      rc = host_allreduce(ctx, all.data(), static_cast<int64_t>(all.size()));
      if (rc != DQ4ML_OK) return rc;
    }
    *loss_sum = all[0];
    *weight_sum = all[1];
    for (int j = 0; j < d; ++j) grad_sum[j] = fstd[j] != 0.0 ? all[2 + j] / fstd[j] : 0.0;
    return DQ4ML_OK;
  }
  const int b = stream_bucket(d > 0 ? d : 1);
  LsqConsts c{};
  double s = 0.0;
  for (int j = 0; j < d; ++j) {
    if (fstd[j] != 0.0) {
      c.eff[j] = coef[j] / fstd[j];
      s += c.eff[j] * fmean[j];
    } else {
      c.eff[j] = 0.0;
    }
  }
  c.offset = fit_intercept ? ymean / ystd - s : 0.0;
  c.inv_ystd = 1.0 / ystd;
  cudaError_t e = run_stream_pass(ctx, kOpLsq, in.cols, in.label, in.dq, n, &c, nullptr);
  if (e != cudaSuccess) return cuda_fail(e, "loss/gradient pass");
  const int rc = finish_result(ctx, stream_nacc(kOpLsq, b));
  if (rc != DQ4ML_OK) return rc;
  const double* r = ctx->h_result;
  *loss_sum = r[0];
  *weight_sum = r[1];
  for (int j = 0; j < d; ++j) grad_sum[j] = fstd[j] != 0.0 ? r[2 + j] / fstd[j] : 0.0;
  return DQ4ML_OK;
}

// One HuberAggregator evaluation at params = [coefficients (d), intercept (if fitted), sigma]: raw sums in the same
// layout.  The weight, when there is one, rides as column 0 of the streamed table (so d <= kMaxCols - 1 then).
int huber_pass(dq4ml_ctx* ctx, const PassInputs& in, int64_t n, const double* params, const double* fstd, bool fit_intercept,
               double epsilon, double* loss_sum, double* weight_sum, double* grad_sum, double* diff_scratch) {
  const int d = in.width();
  const bool weighted = in.weight != nullptr;
  if (d + (weighted ? 1 : 0) > kMaxCols) {
    // Wide table: pass 1 walks the columns per row, writes the factor that multiplies x_j / std_j in the gradient and
    // sums the loss, the weight and the intercept / sigma derivatives; pass 2 takes sum factor x_j over column groups.
    if (diff_scratch == nullptr) {
      set_error("internal: no scratch column for the wide huber pass");
      return DQ4ML_ERR_INVALID;
    }
    std::vector<double> eff(d);
    for (int j = 0; j < d; ++j) eff[j] = fstd[j] != 0.0 ? params[j] / fstd[j] : 0.0;
    const double sigma = params[d + (fit_intercept ? 1 : 0)];
    WideSums ws;
    int rc = wide_rows_pass(ctx, in.ptr, in.dtype, eff.data(), fit_intercept ? params[d] : 0.0, 1.0, in.label, in.dq, n, 2,
                            diff_scratch, false, &ws, in.weight, sigma, epsilon);
    if (rc != DQ4ML_OK) return rc;
    std::vector<double> all(4 + d, 0.0);
    all[0] = ws.ss_res;   // loss sum
    all[1] = ws.n;        // weight sum
    all[2] = ws.sum_res;  // d / d intercept
    all[3] = ws.sum_y;    // d / d sigma
    constexpr int kGradGroup = 16;
    for (int c0 = 0; c0 < d; c0 += kGradGroup) {
      DeviceCols g;
      g.n = std::min(kGradGroup, d - c0);
      for (int j = 0; j < g.n; ++j) {
        g.ptr[j] = in.ptr[c0 + j];
        g.dtype[j] = in.dtype[c0 + j];
      }
      PassOpts local;
      local.reduce = false;
      cudaError_t e = run_stream_pass(ctx, kOpGrad, g, diff_scratch, DeviceDq{}, n, nullptr, nullptr, local);
      if (e != cudaSuccess) return cuda_fail(e, "huber gradient pass");
      rc = finish_result(ctx, stream_nacc(kOpGrad, stream_bucket(g.n)), local);
      if (rc != DQ4ML_OK) return rc;
      for (int j = 0; j < g.n; ++j) all[4 + c0 + j] = ctx->h_result[j];
    }
    if (multi_rank(ctx)) {
      rc = host_allreduce(ctx, all.data(), static_cast<int64_t>(all.size()));
      if (rc != DQ4ML_OK) return rc;
    }
    *loss_sum = all[0];
    *weight_sum = all[1];
    for (int j = 0; j < d; ++j) grad_sum[j] = fstd[j] != 0.0 ? all[4 + j] / fstd[j] : 0.0;
    int k = d;
    if (fit_intercept) grad_sum[k++] = all[2];
    grad_sum[k] = all[3];
    return DQ4ML_OK;
  }
  const int off = weighted ? 1 : 0;
  const int b = stream_bucket(d + off > 0 ? d + off : 1);
  HuberConsts c{};
  for (int j = 0; j < d; ++j) c.eff[off + j] = fstd[j] != 0.0 ? params[j] / fstd[j] : 0.0;
  c.intercept = fit_intercept ? params[d] : 0.0;
  c.sigma = params[d + (fit_intercept ? 1 : 0)];
  c.epsilon = epsilon;
  cudaError_t e = run_stream_pass(ctx, weighted ? kOpHuberW : kOpHuber, weighted ? in.wcols : in.cols, in.label, in.dq, n, &c, nullptr);
  if (e != cudaSuccess) return cuda_fail(e, "huber loss/gradient pass");
  const int rc = finish_result(ctx, stream_nacc(kOpHuber, b));
  if (rc != DQ4ML_OK) return rc;
  const double* r = ctx->h_result;
  *loss_sum = r[0];
  *weight_sum = r[1];
  for (int j = 0; j < d; ++j) grad_sum[j] = fstd[j] != 0.0 ? r[4 + off + j] / fstd[j] : 0.0;
  int k = d;
  if (fit_intercept) grad_sum[k++] = r[2];
  grad_sum[k] = r[3];
  return DQ4ML_OK;
}

// Cross-rank reduction of the nine evaluation accumulators: seven sums, then min and max of the residuals (which travel
// as a zero-padded sum so that the only collective needed is the sum all-reduce).
int reduce_eval_sums(dq4ml_ctx* ctx, double* s) {
  if (!multi_rank(ctx)) return DQ4ML_OK;
  const int world = ctx->world > 0 ? ctx->world : 1;
  std::vector<double> v(7 + 2 * static_cast<size_t>(world), 0.0);
  std::copy(s, s + 7, v.begin());
  v[7 + 2 * ctx->rank] = s[7];
  v[7 + 2 * ctx->rank + 1] = s[8];
  // +-inf (a rank without surviving rows) cannot ride a sum next to zeros of other ranks: send finite stand-ins
  const double big = 1.7e308;
  if (v[7 + 2 * ctx->rank] > big) v[7 + 2 * ctx->rank] = big;            // a NaN stays a NaN (and survives the sum)
  if (v[7 + 2 * ctx->rank + 1] < -big) v[7 + 2 * ctx->rank + 1] = -big;
  const int rc = host_allreduce(ctx, v.data(), static_cast<int64_t>(v.size()));
  if (rc != DQ4ML_OK) return rc;
  std::copy(v.begin(), v.begin() + 7, s);
  // min skips NaN, max is NaN as soon as one rank's is (SQL ordering: NaN is the greatest double)
  double mn = big, mx = -big;
  for (int r = 0; r < world; ++r) {
    const double a = v[7 + 2 * r], b = v[7 + 2 * r + 1];
    if (a == a && a < mn) mn = a;
    if (b != b || mx != mx) mx = std::numeric_limits<double>::quiet_NaN();
    else if (b > mx) mx = b;
  }
  s[7] = mn;
  s[8] = mx;
  return DQ4ML_OK;
}

LrParams to_params(const dq4ml_lr_params* p) {
  LrParams q;
  if (p != nullptr) {
    q.max_iter = p->max_iter;
    q.reg_param = p->reg_param;
    q.elastic_net = p->elastic_net_param;
    q.tol = p->tol;
    q.fit_intercept = p->fit_intercept != 0;
    q.standardization = p->standardization != 0;
    q.solver = p->solver;
    q.loss = p->loss;
    q.epsilon = p->epsilon;
  }
  return q;
}

int check_params(const LrParams& p) {
  if (p.max_iter < 0 || p.reg_param < 0.0 || p.elastic_net < 0.0 || p.elastic_net > 1.0 || p.tol < 0.0 || p.solver < 0 ||
      p.solver > 2) {
    set_error("LinearRegression params out of range (maxIter >= 0, regParam >= 0, 0 <= elasticNetParam <= 1, tol >= 0)");
    return DQ4ML_ERR_INVALID;
  }
  if (p.loss != DQ4ML_LOSS_SQUARED_ERROR && p.loss != DQ4ML_LOSS_HUBER) {
    set_error("unknown loss %d", p.loss);
    return DQ4ML_ERR_INVALID;
  }
  if (p.loss == DQ4ML_LOSS_HUBER) {
    // LinearRegression.train: require(solver != normal) and require(elasticNetParam == 0.0) for huber; epsilon > 1.0
    if (p.solver == DQ4ML_SOLVER_NORMAL || p.elastic_net != 0.0 || !(p.epsilon > 1.0)) {
      set_error("huber loss needs epsilon > 1, elasticNetParam == 0 (L2 only) and a solver other than 'normal'");
      return DQ4ML_ERR_INVALID;
    }
  }
  return DQ4ML_OK;
}

}  // namespace
}  // namespace dq4ml

// =====================================================================================================================
extern "C" {

const char* dq4ml_last_error(void) { return get_error(); }
int dq4ml_abi_version(void) { return DQ4ML_ABI_VERSION; }

int dq4ml_ctx_create(int device, dq4ml_ctx** out) {
  if (out == nullptr) {
    set_error("null out");
    return DQ4ML_ERR_INVALID;
  }
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    set_error("no CUDA device available (%s); this library has no CPU fallback", e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return DQ4ML_ERR_NO_DEVICE;
  }
  if (device < 0 || device >= count) {
    set_error("device %d out of range (have %d)", device, count);
    return DQ4ML_ERR_INVALID;
  }
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    set_error("device %d is sm_%d%d; libdq4ml is built for sm_100a (B200) only", device, prop.major, prop.minor);
    return DQ4ML_ERR_NO_DEVICE;
  }
  auto* ctx = new (std::nothrow) dq4ml_ctx;
  if (ctx == nullptr) {
    set_error("out of memory");
    return DQ4ML_ERR_INVALID;
  }
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  auto fail = [&](cudaError_t err, const char* what) {
    const int rc = cuda_fail(err, what);
    dq4ml_ctx_destroy(ctx);
    return rc;
  };
  if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(e, "stream");
  const size_t result_b = kResultDoubles * sizeof(double), consts_b = 4096, ticket_b = 256, counters_b = 256, err_b = 256;
  const size_t total = kPartialsBytes + result_b + consts_b + ticket_b + counters_b + err_b;
  if ((e = cudaMalloc(reinterpret_cast<void**>(&ctx->scratch_base), total)) != cudaSuccess) return fail(e, "scratch");
  if ((e = cudaMemsetAsync(ctx->scratch_base, 0, total, ctx->stream)) != cudaSuccess) return fail(e, "scratch memset");
  char* p = ctx->scratch_base;
  ctx->scratch.partials = reinterpret_cast<double*>(p);
  ctx->scratch.partials_bytes = kPartialsBytes;
  p += kPartialsBytes;
  ctx->scratch.result = reinterpret_cast<double*>(p);
  p += result_b;
  ctx->scratch.consts = reinterpret_cast<double*>(p);
  p += consts_b;
  ctx->scratch.ticket = reinterpret_cast<unsigned int*>(p);
  p += ticket_b;
  ctx->d_counters = reinterpret_cast<unsigned long long*>(p);
  p += counters_b;
  ctx->d_err = reinterpret_cast<int*>(p);
  ctx->scratch.sm_count = ctx->sm_count;
  if ((e = cudaMallocHost(reinterpret_cast<void**>(&ctx->h_result), result_b)) != cudaSuccess) return fail(e, "pinned result");
  // the epilogue's hand-over buffer: pinned AND mapped, so the last CTA of a pass can write it directly
  void* words = nullptr;
  if ((e = cudaHostAlloc(&words, kEpilogueValues * 16, cudaHostAllocMapped | cudaHostAllocPortable)) != cudaSuccess)
    return fail(e, "mapped result words");
  std::memset(words, 0, kEpilogueValues * 16);
  ctx->h_words = static_cast<volatile unsigned long long*>(words);
  if ((e = cudaHostGetDevicePointer(reinterpret_cast<void**>(&ctx->h_words_dev), words, 0)) != cudaSuccess) return fail(e, "mapped result words");
  if ((e = cudaEventCreate(&ctx->pass_ev0)) != cudaSuccess) return fail(e, "event");
  if ((e = cudaEventCreate(&ctx->pass_ev1)) != cudaSuccess) return fail(e, "event");
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return fail(e, "sync");
  *out = ctx;
  return DQ4ML_OK;
}

void dq4ml_ctx_destroy(dq4ml_ctx* ctx) {
  if (ctx == nullptr) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream != nullptr) cudaStreamSynchronize(ctx->stream);
  nccl_unbind(ctx);
  if (ctx->arena != nullptr) cudaFree(ctx->arena);
  if (ctx->gram_scratch != nullptr) cudaFree(ctx->gram_scratch);
  if (ctx->gram_host != nullptr) cudaFreeHost(ctx->gram_host);
  if (ctx->d_tile_state != nullptr) cudaFree(ctx->d_tile_state);
  if (ctx->scratch_base != nullptr) cudaFree(ctx->scratch_base);
  if (ctx->h_result != nullptr) cudaFreeHost(ctx->h_result);
  if (ctx->h_words != nullptr) cudaFreeHost(const_cast<unsigned long long*>(ctx->h_words));
  if (ctx->pass_ev0 != nullptr) cudaEventDestroy(ctx->pass_ev0);
  if (ctx->pass_ev1 != nullptr) cudaEventDestroy(ctx->pass_ev1);
  if (ctx->stream != nullptr) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int dq4ml_ctx_synchronize(dq4ml_ctx* ctx) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  CU(cudaStreamSynchronize(ctx->stream));
  return DQ4ML_OK;
}

void* dq4ml_ctx_stream(dq4ml_ctx* ctx) { return ctx != nullptr ? ctx->stream : nullptr; }
int64_t dq4ml_ctx_launch_count(dq4ml_ctx* ctx) { return ctx != nullptr ? ctx->launches : 0; }

int dq4ml_ctx_set_gram_tier(dq4ml_ctx* ctx, int tier) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (tier != DQ4ML_GRAM_FP64 && tier != DQ4ML_GRAM_TF32X3) {
    set_error("unknown Gram tier %d", tier);
    return DQ4ML_ERR_INVALID;
  }
  ctx->gram_tier = tier;
  return DQ4ML_OK;
}

int dq4ml_ctx_enable_pass_timing(dq4ml_ctx* ctx, int enable) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  ctx->pass_timing = enable != 0;
  ctx->pass_event_pending = false;
  ctx->pass_ms_total = 0.0;
  ctx->pass_count = 0;
  ctx->pass_ms_list.clear();
  return DQ4ML_OK;
}

int dq4ml_ctx_pass_timing(dq4ml_ctx* ctx, double* total_ms, int64_t* passes) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (total_ms != nullptr) *total_ms = ctx->pass_ms_total;
  if (passes != nullptr) *passes = ctx->pass_count;
  return DQ4ML_OK;
}

int dq4ml_ctx_pass_times(dq4ml_ctx* ctx, double* ms, int32_t cap) {
  if (check_ctx(ctx) != DQ4ML_OK || (ms == nullptr && cap > 0) || cap < 0) {
    set_error("pass_times: bad arguments");
    return DQ4ML_ERR_INVALID;
  }
  const int32_t len = static_cast<int32_t>(ctx->pass_ms_list.size());
  for (int32_t i = 0; i < len && i < cap; ++i) ms[i] = ctx->pass_ms_list[i];
  return len;
}

int dq4ml_nccl_unique_id(void* id128) {
  if (id128 == nullptr) {
    set_error("null id buffer");
    return DQ4ML_ERR_INVALID;
  }
  return nccl_unique_id(id128);
}

int dq4ml_ctx_init_nccl(dq4ml_ctx* ctx, const void* id128, int rank, int world_size) {
  if (check_ctx(ctx) != DQ4ML_OK || id128 == nullptr) {
    set_error("null ctx or id");
    return DQ4ML_ERR_INVALID;
  }
  return nccl_bind(ctx, id128, rank, world_size);
}

int dq4ml_ctx_set_allreduce(dq4ml_ctx* ctx, dq4ml_allreduce_fn fn, void* user, int rank, int world_size) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (fn == nullptr) {  // unbind
    ctx->allreduce_cb = nullptr;
    ctx->allreduce_user = nullptr;
    if (ctx->nccl == nullptr) {
      ctx->rank = 0;
      ctx->world = 1;
    }
    return DQ4ML_OK;
  }
  // min / max and the summarizer centres are laid out per rank, so the position of this rank must be known
  if (world_size < 1 || rank < 0 || rank >= world_size) {
    set_error("set_allreduce: bad rank %d / world size %d", rank, world_size);
    return DQ4ML_ERR_INVALID;
  }
  if (ctx->nccl != nullptr) {
    set_error("set_allreduce: an NCCL communicator is already bound to this ctx");
    return DQ4ML_ERR_INVALID;
  }
  ctx->allreduce_cb = fn;
  ctx->allreduce_user = user;
  ctx->rank = rank;
  ctx->world = world_size;
  return DQ4ML_OK;
}

int dq4ml_ctx_allreduce(dq4ml_ctx* ctx, double* buf, int64_t count) {
  if (check_ctx(ctx) != DQ4ML_OK || (buf == nullptr && count > 0) || count < 0) {
    set_error("bad all-reduce arguments");
    return DQ4ML_ERR_INVALID;
  }
  cudaSetDevice(ctx->device);
  return host_allreduce(ctx, buf, count);
}

int dq4ml_device_alloc(dq4ml_ctx* ctx, int64_t bytes, void** out) {
  if (check_ctx(ctx) != DQ4ML_OK || out == nullptr || bytes < 0) {
    set_error("bad alloc arguments");
    return DQ4ML_ERR_INVALID;
  }
  CU(cudaSetDevice(ctx->device));
  CU(cudaMalloc(out, static_cast<size_t>(bytes > 0 ? bytes : 1)));
  return DQ4ML_OK;
}
int dq4ml_device_free(dq4ml_ctx* ctx, void* p) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  CU(cudaFree(p));
  return DQ4ML_OK;
}
int dq4ml_host_alloc_pinned(int64_t bytes, void** out) {
  if (out == nullptr || bytes < 0) {
    set_error("bad alloc arguments");
    return DQ4ML_ERR_INVALID;
  }
  CU(cudaMallocHost(out, static_cast<size_t>(bytes > 0 ? bytes : 1)));
  return DQ4ML_OK;
}
int dq4ml_host_free_pinned(void* p) {
  CU(cudaFreeHost(p));
  return DQ4ML_OK;
}
int dq4ml_memcpy(dq4ml_ctx* ctx, void* dst, const void* src, int64_t bytes) {
  if (check_ctx(ctx) != DQ4ML_OK || bytes < 0 || (bytes > 0 && (dst == nullptr || src == nullptr))) {
    set_error("bad memcpy arguments");
    return DQ4ML_ERR_INVALID;
  }
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemcpyAsync(dst, src, static_cast<size_t>(bytes), cudaMemcpyDefault, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DQ4ML_OK;
}

// ------------------------------------------------------------------------------------------------------------ DQ rules

dq4ml_dq_params dq4ml_dq_params_default(void) {
  dq4ml_dq_params p;
  p.min_price = 20.0;  // MinimumPriceDataQualityService.MIN_PRICE
  p.min_guest = 14;    // PriceCorrelationDataQualityService.checkPriceRange
  p.max_price = 90.0;
  return p;
}

double dq4ml_check_minimum_price(double price) {
  if (price < 20) return -1;
  return price;
}

double dq4ml_check_price_range(double price, int32_t guest) {
  if (guest < 14 && price > 90) return -1;
  return price;
}

int dq4ml_udf_minimumPriceRule(dq4ml_ctx* ctx, const double* price, const uint8_t* price_valid, int64_t n, double* out,
                               const dq4ml_dq_params* params) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (n < 0 || (n > 0 && (price == nullptr || out == nullptr))) {
    set_error("minimumPriceRule: null column or negative n");
    return DQ4ML_ERR_INVALID;
  }
  if (n == 0) return DQ4ML_OK;
  CallScope scope(ctx);
  const dq4ml_dq_params p = params != nullptr ? *params : dq4ml_dq_params_default();
  Resolver rs(ctx);
  rs.input(reinterpret_cast<const void**>(&price), static_cast<size_t>(n) * 8);
  rs.input(reinterpret_cast<const void**>(&price_valid), static_cast<size_t>(n));
  if (!is_device_ptr(out)) rs.reserve_output(static_cast<size_t>(n) * 8);
  int rc = rs.run();
  if (rc != DQ4ML_OK) return rc;
  OutBuf ob = make_out(ctx, out, static_cast<size_t>(n) * 8);
  CU(cudaMemsetAsync(ctx->d_err, 0, sizeof(int), ctx->stream));
  cudaError_t e = launch_udf_min_price(price, price_valid, n, static_cast<double*>(ob.dev), p, ctx->d_err, ctx->stream,
                                       &ctx->launches);
  if (e != cudaSuccess) return cuda_fail(e, "minimumPriceRule kernel");
  int flag = 0;
  CU(cudaMemcpyAsync(&flag, ctx->d_err, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if (ob.staged) CU(cudaMemcpyAsync(out, ob.dev, static_cast<size_t>(n) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  if (flag != 0) {
    set_error("minimumPriceRule: null price (the reference's MinimumPriceDataQualityUdf.call throws NullPointerException)");
    return DQ4ML_ERR_NULL_VALUE;
  }
  return DQ4ML_OK;
}

int dq4ml_udf_priceCorrelationRule(dq4ml_ctx* ctx, const double* price, const int32_t* guest, const uint8_t* price_valid,
                                   const uint8_t* guest_valid, int64_t n, double* out, const dq4ml_dq_params* params) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (n < 0 || (n > 0 && (price == nullptr || guest == nullptr || out == nullptr))) {
    set_error("priceCorrelationRule: null column or negative n");
    return DQ4ML_ERR_INVALID;
  }
  if (n == 0) return DQ4ML_OK;
  CallScope scope(ctx);
  const dq4ml_dq_params p = params != nullptr ? *params : dq4ml_dq_params_default();
  Resolver rs(ctx);
  rs.input(reinterpret_cast<const void**>(&price), static_cast<size_t>(n) * 8);
  rs.input(reinterpret_cast<const void**>(&guest), static_cast<size_t>(n) * 4);
  rs.input(reinterpret_cast<const void**>(&price_valid), static_cast<size_t>(n));
  rs.input(reinterpret_cast<const void**>(&guest_valid), static_cast<size_t>(n));
  if (!is_device_ptr(out)) rs.reserve_output(static_cast<size_t>(n) * 8);
  int rc = rs.run();
  if (rc != DQ4ML_OK) return rc;
  OutBuf ob = make_out(ctx, out, static_cast<size_t>(n) * 8);
  cudaError_t e = launch_udf_price_corr(price, guest, price_valid, guest_valid, n, static_cast<double*>(ob.dev), p,
                                        ctx->stream, &ctx->launches);
  if (e != cudaSuccess) return cuda_fail(e, "priceCorrelationRule kernel");
  if (ob.staged) CU(cudaMemcpyAsync(out, ob.dev, static_cast<size_t>(n) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DQ4ML_OK;
}

int dq4ml_dq_filter(dq4ml_ctx* ctx, const int32_t* guest, const double* price, int64_t n, const dq4ml_dq_params* params,
                    int32_t* guest_out, double* price_out, int64_t* n_out, int64_t* n_after_rule1) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (n < 0 || n_out == nullptr || (n > 0 && (guest == nullptr || price == nullptr || guest_out == nullptr || price_out == nullptr))) {
    set_error("dq_filter: null column / output or negative n");
    return DQ4ML_ERR_INVALID;
  }
  if (n > 0 && (static_cast<const void*>(guest) == static_cast<const void*>(guest_out) ||
                static_cast<const void*>(price) == static_cast<const void*>(price_out))) {
    set_error("dq_filter: in-place compaction is not supported");
    return DQ4ML_ERR_INVALID;
  }
  *n_out = 0;
  if (n_after_rule1 != nullptr) *n_after_rule1 = 0;
  if (n == 0) return DQ4ML_OK;
  CallScope scope(ctx);
  const dq4ml_dq_params p = params != nullptr ? *params : dq4ml_dq_params_default();
  const size_t need_state = dq_filter_state_bytes(n);
  if (need_state > ctx->tile_state_cap) {
    CU(cudaStreamSynchronize(ctx->stream));
    if (ctx->d_tile_state != nullptr) CU(cudaFree(ctx->d_tile_state));
    ctx->d_tile_state = nullptr;
    ctx->tile_state_cap = 0;
    CU(cudaMalloc(reinterpret_cast<void**>(&ctx->d_tile_state), need_state));
    ctx->tile_state_cap = need_state;
  }
  Resolver rs(ctx);
  rs.input(reinterpret_cast<const void**>(&guest), static_cast<size_t>(n) * 4);
  rs.input(reinterpret_cast<const void**>(&price), static_cast<size_t>(n) * 8);
  if (!is_device_ptr(guest_out)) rs.reserve_output(static_cast<size_t>(n) * 4);
  if (!is_device_ptr(price_out)) rs.reserve_output(static_cast<size_t>(n) * 8);
  int rc = rs.run();
  if (rc != DQ4ML_OK) return rc;
  OutBuf og = make_out(ctx, guest_out, static_cast<size_t>(n) * 4);
  OutBuf op = make_out(ctx, price_out, static_cast<size_t>(n) * 8);
  cudaError_t e = launch_dq_filter(guest, price, n, p, static_cast<int32_t*>(og.dev), static_cast<double*>(op.dev),
                                   ctx->d_tile_state, ctx->d_counters, ctx->stream, &ctx->launches);
  if (e != cudaSuccess) return cuda_fail(e, "dq_filter kernel");
  unsigned long long counts[2] = {0, 0};
  CU(cudaMemcpyAsync(counts, ctx->d_counters, sizeof(counts), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  *n_out = static_cast<int64_t>(counts[0]);
  if (n_after_rule1 != nullptr) *n_after_rule1 = static_cast<int64_t>(counts[1]);
  if (og.staged && counts[0] > 0) CU(cudaMemcpyAsync(guest_out, og.dev, counts[0] * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (op.staged && counts[0] > 0) CU(cudaMemcpyAsync(price_out, op.dev, counts[0] * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DQ4ML_OK;
}

int dq4ml_dq_filter_global_offset(dq4ml_ctx* ctx, int64_t n_out_local, int64_t* offset, int64_t* total) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (n_out_local < 0 || offset == nullptr) {
    set_error("dq_filter_global_offset: negative count or null output");
    return DQ4ML_ERR_INVALID;
  }
  cudaSetDevice(ctx->device);
  // an all-gather of the G survivor counts, carried by the one collective the ctx has: each rank fills its own slot of
  // a zeroed vector and the sum all-reduce assembles it (counts stay far below 2^53, so doubles carry them exactly)
  const int world = multi_rank(ctx) && ctx->world > 0 ? ctx->world : 1;
  const int rank = world > 1 ? ctx->rank : 0;
  std::vector<double> counts(static_cast<size_t>(world), 0.0);
  counts[rank] = static_cast<double>(n_out_local);
  const int rc = host_allreduce(ctx, counts.data(), world);
  if (rc != DQ4ML_OK) return rc;
  int64_t before = 0, all = 0;
  for (int r = 0; r < world; ++r) {
    const int64_t c = static_cast<int64_t>(counts[r]);
    if (r < rank) before += c;
    all += c;
  }
  *offset = before;
  if (total != nullptr) *total = all;
  return DQ4ML_OK;
}

// ------------------------------------------------------------------------------------------------------------ LinearRegression

dq4ml_lr_params dq4ml_lr_params_default(void) {
  dq4ml_lr_params p;
  p.max_iter = 100;
  p.reg_param = 0.0;
  p.elastic_net_param = 0.0;
  p.tol = 1e-6;
  p.fit_intercept = 1;
  p.standardization = 1;
  p.solver = DQ4ML_SOLVER_AUTO;
  p.loss = DQ4ML_LOSS_SQUARED_ERROR;
  p.epsilon = 1.35;
  return p;
}

int64_t dq4ml_wls_moments_len(int32_t d) { return d < 0 ? 0 : 5 + 2 * static_cast<int64_t>(d) + static_cast<int64_t>(d) * (d + 1) / 2; }
int64_t dq4ml_summary_len(int32_t d) { return d < 0 ? 0 : 3 + 7 * (static_cast<int64_t>(d) + 1); }

int dq4ml_wls_moments(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                      const dq4ml_dq_fused* dq, double* out) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (out == nullptr || (label == nullptr && n > 0)) {
    set_error("wls_moments: null label or output");
    return DQ4ML_ERR_INVALID;
  }
  CallScope scope(ctx);
  PassInputs in;
  int rc = gather_inputs(ctx, features, label, n, dq, &in, 0, 4096);
  if (rc != DQ4ML_OK) return rc;
  return wls_moments_any(ctx, in, n, out);
}

int dq4ml_wls_solve(const double* moments, int32_t d, const dq4ml_lr_params* params, dq4ml_lr_model** out) {
  if (moments == nullptr || out == nullptr || d < 0) {
    set_error("wls_solve: null moments / out or negative d");
    return DQ4ML_ERR_INVALID;
  }
  *out = nullptr;
  const LrParams p = to_params(params ? params : nullptr);
  dq4ml_lr_params defaults = dq4ml_lr_params_default();
  const LrParams q = params ? p : to_params(&defaults);
  if (check_params(q) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (q.loss == DQ4ML_LOSS_HUBER) {
    set_error("wls_solve: the normal-equation solver has no huber loss (Spark: huber requires solver l-bfgs)");
    return DQ4ML_ERR_INVALID;
  }
  try {
    auto m = std::make_unique<dq4ml_lr_model>();
    m->m = wls_fit_from_moments(moments, d, q);
    *out = m.release();
    return DQ4ML_OK;
  } catch (const SolveError& e) {
    set_error("%s", e.what.c_str());
    return e.code;
  } catch (const std::exception& e) {
    set_error("wls_solve: %s", e.what());
    return DQ4ML_ERR_INVALID;
  }
}

int dq4ml_summarize(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                    const dq4ml_dq_fused* dq, double* out) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (out == nullptr || (label == nullptr && n > 0)) {
    set_error("summarize: null label or output");
    return DQ4ML_ERR_INVALID;
  }
  CallScope scope(ctx);
  PassInputs in;
  int rc = gather_inputs(ctx, features, label, n, dq, &in, 0, 4096);
  if (rc != DQ4ML_OK) return rc;
  const int d = in.width();
  // every column (features, then the label) goes through the full summarizer in groups of four
  std::vector<const void*> ptrs;
  std::vector<int> dtypes;
  for (int j = 0; j < d; ++j) {
    ptrs.push_back(in.ptr[j]);
    dtypes.push_back(in.dtype[j]);
  }
  ptrs.push_back(in.label);
  dtypes.push_back(DQ4ML_F64);
  const int total = d + 1;
  double count = 0.0;
  for (int c0 = 0; c0 < total; c0 += 4) {
    DeviceCols g;
    g.n = std::min(4, total - c0);
    for (int j = 0; j < g.n; ++j) {
      g.ptr[j] = ptrs[c0 + j];
      g.dtype[j] = dtypes[c0 + j];
    }
    const int b = stream_bucket(g.n);
    int world = 1, len = 0;
    rc = summarizer_pass(ctx, kOpSummFull, g, nullptr, in.dq, n, b, b, false, &world, &len);
    if (rc != DQ4ML_OK) return rc;
    const int nacc = stream_nacc(kOpSummFull, b);
    // per column: shifted sums merged as moments; min / max folded (they skip NaNs like MultivariateOnlineSummarizer.add,
    // whose `currMin > value` is false on NaN); numNonzeros, normL1 and the sum of squares added
    count = 0.0;
    for (int r = 0; r < world; ++r) count += ctx->h_result[static_cast<size_t>(r) * len];
    for (int j = 0; j < g.n; ++j) {
      Moments m;
      double mn = std::numeric_limits<double>::infinity(), mx = -std::numeric_limits<double>::infinity();
      double nnz = 0.0, l1 = 0.0, sq = 0.0;
      for (int r = 0; r < world; ++r) {
        const double* v = ctx->h_result + static_cast<size_t>(r) * len;
        const double* a = v + 1 + 7 * j;
        m.merge(v[0], v[nacc + j], a[0], a[1]);
        if (a[2] < mn) mn = a[2];
        if (a[3] > mx) mx = a[3];
        nnz += a[4];
        l1 += a[5];
        sq += a[6];
      }
      double* o = out + 3 + 7 * (c0 + j);
      if (count > 0) {
        o[0] = m.mean;
        o[1] = count > 1 ? std::max(m.m2 / (count - 1.0), 0.0) : 0.0;
        o[2] = mn;
        o[3] = mx;
      } else {
        o[0] = o[1] = o[2] = o[3] = 0.0;
      }
      o[4] = nnz;
      o[5] = l1;
      o[6] = std::sqrt(sq);
    }
  }
  out[0] = out[1] = out[2] = count;
  return DQ4ML_OK;
}

int dq4ml_lsq_loss_grad(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                        const dq4ml_dq_fused* dq, const double* coef, const double* features_std,
                        const double* features_mean, double label_std, double label_mean, int fit_intercept, double* out) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (out == nullptr || (label == nullptr && n > 0) || features == nullptr ||
      (features->n_cols > 0 && (coef == nullptr || features_std == nullptr || features_mean == nullptr))) {
    set_error("lsq_loss_grad: null argument");
    return DQ4ML_ERR_INVALID;
  }
  CallScope scope(ctx);
  PassInputs in;
  const bool wide = features->n_cols > kMaxCols;
  int rc = gather_inputs(ctx, features, label, n, dq, &in, wide ? static_cast<size_t>(n) * 8 : 0, 4096);
  if (rc != DQ4ML_OK) return rc;
  double* diff = wide ? static_cast<double*>(arena_take(ctx, static_cast<size_t>(n) * 8)) : nullptr;
  return lsq_pass(ctx, in, n, coef, features_std, features_mean, label_std, label_mean, fit_intercept != 0, &out[0],
                  &out[1], out + 2, diff);
}

int dq4ml_linreg_fit(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, int64_t n,
                     const dq4ml_lr_params* params, const dq4ml_dq_fused* dq, dq4ml_lr_model** out) {
  return dq4ml_linreg_fit_weighted(ctx, features, label, nullptr, n, params, dq, out);
}

int dq4ml_linreg_fit_weighted(dq4ml_ctx* ctx, const dq4ml_columns* features, const double* label, const double* weight,
                              int64_t n, const dq4ml_lr_params* params, const dq4ml_dq_fused* dq, dq4ml_lr_model** out) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (out == nullptr || (label == nullptr && n > 0) || features == nullptr) {
    set_error("linreg_fit: null argument");
    return DQ4ML_ERR_INVALID;
  }
  *out = nullptr;
  dq4ml_lr_params defaults = dq4ml_lr_params_default();
  const LrParams p = to_params(params != nullptr ? params : &defaults);
  if (check_params(p) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  CallScope scope(ctx);
  const int d = features->n_cols;
  // LinearRegression.train: squared-error loss goes to WeightedLeastSquares when solver is normal, or auto with
  // numFeatures <= WeightedLeastSquares.MAX_NUM_FEATURES (4096); otherwise the summarizer + L-BFGS/OWL-QN branch.
  const bool huber = p.loss == DQ4ML_LOSS_HUBER;  // always the summarizer + aggregator branch
  const bool wls = !huber && (p.solver == DQ4ML_SOLVER_NORMAL || (p.solver == DQ4ML_SOLVER_AUTO && d <= 4096));
  PassInputs in;
  // tables too wide for one streaming launch (with the weight as an extra column, if any) need one scratch column for
  // the per-row scaled difference
  const bool wide_lbfgs = !wls && d + (weight != nullptr ? 1 : 0) > kMaxCols;
  int rc = gather_inputs(ctx, features, label, n, dq, &in, wide_lbfgs ? static_cast<size_t>(n) * 8 : 0, 1 << 20, weight);
  if (rc != DQ4ML_OK) return rc;
  double* diff = wide_lbfgs ? static_cast<double*>(arena_take(ctx, static_cast<size_t>(n) * 8)) : nullptr;
  try {
    auto model = std::make_unique<dq4ml_lr_model>();
    if (wls) {
      std::vector<double> mom(static_cast<size_t>(dq4ml_wls_moments_len(d)));
      rc = wls_moments_any(ctx, in, n, mom.data());
      if (rc != DQ4ML_OK) return rc;
      model->m = wls_fit_from_moments(mom.data(), d, p);
    } else if (huber) {
      ColumnStats st;
      rc = in.weight != nullptr ? summarize_core_weighted(ctx, in, n, &st) : summarize_core(ctx, in, n, &st);
      if (rc != DQ4ML_OK) return rc;
      HuberPassFn pass = [&](const Vec& par, const Vec& fstd, double& loss_sum, double& weight_sum, Vec& grad_sum) {
        grad_sum.assign(par.size(), 0.0);
        const int r = huber_pass(ctx, in, n, par.data(), fstd.data(), p.fit_intercept, p.epsilon, &loss_sum, &weight_sum,
                                 grad_sum.data(), diff);
        if (r != DQ4ML_OK) throw SolveError{r, get_error()};
      };
      model->m = huber_path_fit(st, d, p, pass);
    } else {
      ColumnStats st;
      rc = in.weight != nullptr ? summarize_core_weighted(ctx, in, n, &st) : summarize_core(ctx, in, n, &st);
      if (rc != DQ4ML_OK) return rc;
      int pass_rc = DQ4ML_OK;
      LsqPassFn pass = [&](const Vec& coef, const Vec& fstd, const Vec& fmean, double ystd, double ymean,
                           double& loss_sum, double& weight_sum, Vec& grad_sum) {
        grad_sum.assign(d, 0.0);
        const int r = lsq_pass(ctx, in, n, coef.data(), fstd.data(), fmean.data(), ystd, ymean, p.fit_intercept,
                               &loss_sum, &weight_sum, grad_sum.data(), diff);
        if (r != DQ4ML_OK) {
          pass_rc = r;
          throw SolveError{r, get_error()};
        }
      };
      model->m = lbfgs_path_fit(st, d, p, pass);
    }
    *out = model.release();
    return DQ4ML_OK;
  } catch (const SolveError& e) {
    set_error("%s", e.what.c_str());
    return e.code;
  } catch (const std::exception& e) {
    set_error("linreg_fit: %s", e.what());
    return DQ4ML_ERR_INVALID;
  }
}

int32_t dq4ml_linreg_model_num_features(const dq4ml_lr_model* m) { return m ? static_cast<int32_t>(m->m.coefficients.size()) : 0; }
int dq4ml_linreg_model_coefficients(const dq4ml_lr_model* m, double* coef) {
  if (m == nullptr || (coef == nullptr && !m->m.coefficients.empty())) {
    set_error("null model / buffer");
    return DQ4ML_ERR_INVALID;
  }
  std::copy(m->m.coefficients.begin(), m->m.coefficients.end(), coef);
  return DQ4ML_OK;
}
double dq4ml_linreg_model_intercept(const dq4ml_lr_model* m) { return m ? m->m.intercept : std::numeric_limits<double>::quiet_NaN(); }
int32_t dq4ml_linreg_model_total_iterations(const dq4ml_lr_model* m) {
  return m ? static_cast<int32_t>(m->m.objective_history.size()) : 0;  // Spark 2.4.4: totalIterations = objectiveHistory.length
}
int dq4ml_linreg_model_objective_history(const dq4ml_lr_model* m, double* hist, int32_t cap) {
  if (m == nullptr || (hist == nullptr && cap > 0)) {
    set_error("null model / buffer");
    return DQ4ML_ERR_INVALID;
  }
  const int32_t len = static_cast<int32_t>(m->m.objective_history.size());
  for (int32_t i = 0; i < len && i < cap; ++i) hist[i] = m->m.objective_history[i];
  return len;
}
int32_t dq4ml_linreg_model_data_passes(const dq4ml_lr_model* m) { return m ? m->m.data_passes : 0; }
double dq4ml_linreg_model_scale(const dq4ml_lr_model* m) { return m ? m->m.scale : std::numeric_limits<double>::quiet_NaN(); }
int64_t dq4ml_linreg_model_num_rows(const dq4ml_lr_model* m) { return m ? m->m.num_rows : 0; }
const char* dq4ml_linreg_model_solver_used(const dq4ml_lr_model* m) { return m ? m->m.solver_used.c_str() : ""; }

double dq4ml_linreg_predict1(const dq4ml_lr_model* m, const double* x) {
  if (m == nullptr || (x == nullptr && !m->m.coefficients.empty())) return std::numeric_limits<double>::quiet_NaN();
  double s = 0.0;  // BLAS.dot(features, coefficients) + intercept
  for (size_t j = 0; j < m->m.coefficients.size(); ++j) s += x[j] * m->m.coefficients[j];
  return s + m->m.intercept;
}

int dq4ml_linreg_model_inference(const dq4ml_lr_model* m, double num_instances, double ss_err, int fit_intercept,
                                 double* std_err, double* t_values, double* p_values, int32_t cap) {
  if (m == nullptr) {
    set_error("null model");
    return DQ4ML_ERR_INVALID;
  }
  const int k = static_cast<int>(m->m.diag_inv_atwa.size());
  const int d = static_cast<int>(m->m.coefficients.size());
  if (k == 0 || k != d + (fit_intercept ? 1 : 0)) {
    set_error("No Std. Error of coefficients available for this LinearRegressionModel (normal-equation Cholesky fits only)");
    return DQ4ML_ERR_UNSUPPORTED;
  }
  if (cap < k) {
    set_error("inference: room for %d values, need %d", cap, k);
    return DQ4ML_ERR_INVALID;
  }
  const double dof = num_instances - d - (fit_intercept ? 1.0 : 0.0);
  const double sigma2 = ss_err / dof;  // rss / degreesOfFreedom
  for (int i = 0; i < k; ++i) {
    const double se = std::sqrt(m->m.diag_inv_atwa[i] * sigma2);
    const double est = i < d ? m->m.coefficients[i] : m->m.intercept;
    const double t = est / se;
    if (std_err != nullptr) std_err[i] = se;
    if (t_values != nullptr) t_values[i] = t;
    if (p_values != nullptr) p_values[i] = student_t_two_sided_p(t, dof);
  }
  return k;
}

void dq4ml_linreg_model_destroy(dq4ml_lr_model* m) { delete m; }

int dq4ml_linreg_model_create(const double* coef, int32_t d, double intercept, dq4ml_lr_model** out) {
  if (out == nullptr || d < 0 || (d > 0 && coef == nullptr)) {
    set_error("model_create: bad arguments");
    return DQ4ML_ERR_INVALID;
  }
  auto* m = new dq4ml_lr_model;
  m->m.coefficients.assign(coef, coef + d);
  m->m.intercept = intercept;
  m->m.objective_history = {0.0};
  m->m.solver_used = "given";
  *out = m;
  return DQ4ML_OK;
}

static int eval_common(dq4ml_ctx* ctx, const dq4ml_lr_model* m, const dq4ml_columns* features, const double* label,
                       int64_t n, const dq4ml_dq_fused* dq, int mode, double* row_out_user, double* sums) {
  if (m == nullptr || features == nullptr) {
    set_error("null model / features");
    return DQ4ML_ERR_INVALID;
  }
  if (features->n_cols != static_cast<int>(m->m.coefficients.size())) {
    set_error("feature table has %d columns, model has %zu coefficients", features->n_cols, m->m.coefficients.size());
    return DQ4ML_ERR_INVALID;
  }
  CallScope scope(ctx);
  PassInputs in;
  int rc = gather_inputs(ctx, features, label, n, dq, &in, (row_out_user != nullptr && !is_device_ptr(row_out_user)) ? static_cast<size_t>(n) * 8 : 0, 4096);
  if (rc != DQ4ML_OK) return rc;
  OutBuf ob = make_out(ctx, row_out_user, static_cast<size_t>(n) * 8);
  if (in.width() > kMaxCols) {  // wide table: the column-walking row pass
    WideSums ws;
    rc = wide_rows_pass(ctx, in.ptr, in.dtype, m->m.coefficients.data(), m->m.intercept, 1.0, in.label, in.dq, n, mode,
                        static_cast<double*>(ob.dev), false, &ws);
    if (rc != DQ4ML_OK) return rc;
    if (ob.staged && n > 0) {
      CU(cudaMemcpyAsync(row_out_user, ob.dev, static_cast<size_t>(n) * 8, cudaMemcpyDeviceToHost, ctx->stream));
      CU(cudaStreamSynchronize(ctx->stream));
    }
    if (sums != nullptr) {
      const double loc[9] = {ws.n, ws.ss_res, ws.sum_y, ws.sum_yy, ws.sum_abs_res, ws.sum_res, ws.sum_res_y, ws.min_res, ws.max_res};
      std::copy(loc, loc + 9, sums);
      return reduce_eval_sums(ctx, sums);
    }
    return DQ4ML_OK;
  }
  EvalConsts c{};
  for (int j = 0; j < in.cols.n; ++j) c.coef[j] = m->m.coefficients[j];
  c.intercept = m->m.intercept;
  c.mode = mode;
  // with peer mailboxes the nine accumulators are folded over the ranks inside the kernel, each by its own kind (sums
  // add, the residual extremes take min / NaN-aware max); otherwise they come back per rank and are folded on the host
  PassOpts o;
  o.reduce = sums != nullptr && multi_rank(ctx) && ctx->nccl != nullptr && nccl_mailboxes(ctx, &ctx->scratch.ep);
  o.timed = false;
  cudaError_t e = run_stream_pass(ctx, kOpEval, in.cols, in.label, in.dq, n, &c, static_cast<double*>(ob.dev), o);
  if (e != cudaSuccess) return cuda_fail(e, "predict/evaluate pass");
  if (ob.staged && n > 0) CU(cudaMemcpyAsync(row_out_user, ob.dev, static_cast<size_t>(n) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  const bool folded = o.reduce;
  rc = finish_result(ctx, 9, o);
  if (rc != DQ4ML_OK) return rc;
  if (ob.staged && n > 0) CU(cudaStreamSynchronize(ctx->stream));  // the epilogue's hand-over does not wait for the copy
  if (sums != nullptr) {
    std::copy(ctx->h_result, ctx->h_result + 9, sums);
    return folded ? DQ4ML_OK : reduce_eval_sums(ctx, sums);
  }
  return DQ4ML_OK;
}

int dq4ml_linreg_transform(dq4ml_ctx* ctx, const dq4ml_lr_model* m, const dq4ml_columns* features, int64_t n,
                           double* prediction) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (prediction == nullptr && n > 0) {
    set_error("transform: null prediction buffer");
    return DQ4ML_ERR_INVALID;
  }
  // transform is per-rank: no cross-rank sum is wanted for the side statistics
  dq4ml::NcclBinding* saved_nccl = ctx->nccl;
  dq4ml_allreduce_fn saved_cb = ctx->allreduce_cb;
  ctx->nccl = nullptr;
  ctx->allreduce_cb = nullptr;
  const int rc = eval_common(ctx, m, features, nullptr, n, nullptr, 0, prediction, nullptr);
  ctx->nccl = saved_nccl;
  ctx->allreduce_cb = saved_cb;
  return rc;
}

int dq4ml_linreg_evaluate(dq4ml_ctx* ctx, const dq4ml_lr_model* m, const dq4ml_columns* features, const double* label,
                          int64_t n, const dq4ml_dq_fused* dq, int fit_intercept, double* metrics, double* residuals) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (metrics == nullptr || (label == nullptr && n > 0)) {
    set_error("evaluate: null label / metrics");
    return DQ4ML_ERR_INVALID;
  }
  double s[9];
  const int rc = eval_common(ctx, m, features, label, n, dq, 1, residuals, s);
  if (rc != DQ4ML_OK) return rc;
  const double cnt = s[0], ss_err = s[1], sy = s[2], syy = s[3], sabs = s[4], sr = s[5], sry = s[6];
  for (int k = 0; k < 5; ++k) metrics[k] = s[k];
  const double mse = cnt > 0 ? ss_err / cnt : 0.0;
  const double y_bar = cnt > 0 ? sy / cnt : 0.0;
  const double ss_tot = fit_intercept ? syy - sy * y_bar : syy;
  const double r2 = 1.0 - ss_err / ss_tot;
  metrics[5] = std::sqrt(mse);          // rootMeanSquaredError
  metrics[6] = r2;                      // r2
  metrics[7] = mse;                     // meanSquaredError
  metrics[8] = cnt > 0 ? sabs / cnt : 0.0;  // meanAbsoluteError
  // RegressionMetrics.explainedVariance = SSreg / n with SSreg = sum (prediction - mean(label))^2, prediction = y - r
  const double sum_pred = sy - sr, sum_pred2 = syy - 2.0 * sry + ss_err;
  metrics[9] = cnt > 0 ? (sum_pred2 - 2.0 * y_bar * sum_pred + cnt * y_bar * y_bar) / cnt : 0.0;
  // LinearRegressionSummary.degreesOfFreedom / r2adj
  const double k_int = fit_intercept ? 1.0 : 0.0;
  const double dof = cnt - static_cast<double>(m->m.coefficients.size()) - k_int;
  metrics[10] = dof;
  metrics[11] = 1.0 - (1.0 - r2) * (cnt - k_int) / dof;
  // devianceResiduals: SQL min / max of the residual column (NaN ordered greatest: max is NaN if any residual is, min
  // only if all are -- the kernel's NaN-skipping min is then still at its identity)
  metrics[12] = cnt > 0 ? ((s[8] != s[8] && !(s[7] < 1.7e308)) ? s[8] : s[7]) : 0.0;
  metrics[13] = cnt > 0 ? s[8] : 0.0;
  metrics[14] = sr;
  metrics[15] = sry;
  return DQ4ML_OK;
}

int dq4ml_weighted_residual_stats(dq4ml_ctx* ctx, const double* residuals, const double* weight, int64_t n,
                                  const dq4ml_dq_fused* dq, double* out3) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (out3 == nullptr || ((residuals == nullptr || weight == nullptr) && n > 0)) {
    set_error("weighted_residual_stats: null argument");
    return DQ4ML_ERR_INVALID;
  }
  CallScope scope(ctx);
  dq4ml_column col{residuals, DQ4ML_F64};
  dq4ml_columns table{&col, 1};
  PassInputs in;
  int rc = gather_inputs(ctx, &table, nullptr, n, dq, &in, 0, 1, weight);
  if (rc != DQ4ML_OK) return rc;
  PassOpts local;
  local.reduce = false;
  local.timed = false;
  cudaError_t e = run_stream_pass(ctx, kOpWResid, in.wcols, nullptr, in.dq, n, nullptr, nullptr, local);
  if (e != cudaSuccess) return cuda_fail(e, "weighted residual pass");
  rc = finish_result(ctx, 3, local);
  if (rc != DQ4ML_OK) return rc;
  // the cross-rank fold reuses the evaluate path: slots 0..6 are sums, 7 / 8 the min / max
  double s[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  s[0] = ctx->h_result[0];
  s[7] = ctx->h_result[1];
  s[8] = ctx->h_result[2];
  rc = reduce_eval_sums(ctx, s);
  if (rc != DQ4ML_OK) return rc;
  out3[0] = s[0];
  out3[1] = s[7];
  out3[2] = s[8];
  return DQ4ML_OK;
}

// ------------------------------------------------------------------------------------------------------------ synthetic data

int dq4ml_synth_generate(dq4ml_ctx* ctx, const dq4ml_synth_params* p, int64_t row0, int64_t n, int32_t* guest,
                         double* price, double* extra, int64_t extra_stride) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (p == nullptr || n < 0 || row0 < 0 || p->d < 1 || (n > 0 && (guest == nullptr || price == nullptr)) ||
      (p->d > 1 && extra != nullptr && extra_stride < n)) {
    set_error("synth_generate: bad arguments");
    return DQ4ML_ERR_INVALID;
  }
  if (n > 0 && (!is_device_ptr(guest) || !is_device_ptr(price) || (extra != nullptr && !is_device_ptr(extra)))) {
    set_error("synth_generate writes device memory only");
    return DQ4ML_ERR_INVALID;
  }
  CU(cudaSetDevice(ctx->device));
  cudaError_t e = launch_synth(*p, row0, n, guest, price, extra, extra_stride, ctx->stream, &ctx->launches);
  if (e != cudaSuccess) return cuda_fail(e, "synth kernel");
  CU(cudaStreamSynchronize(ctx->stream));
  return DQ4ML_OK;
}

int dq4ml_rows_digest(dq4ml_ctx* ctx, const int32_t* guest, const double* price, int64_t n, uint64_t seed, uint64_t* out) {
  if (check_ctx(ctx) != DQ4ML_OK) return DQ4ML_ERR_INVALID;
  if (out == nullptr || n < 0 || (n > 0 && (guest == nullptr || price == nullptr))) {
    set_error("rows_digest: bad arguments");
    return DQ4ML_ERR_INVALID;
  }
  CallScope scope(ctx);
  Resolver rs(ctx);
  rs.input(reinterpret_cast<const void**>(&guest), static_cast<size_t>(n) * 4);
  rs.input(reinterpret_cast<const void**>(&price), static_cast<size_t>(n) * 8);
  int rc = rs.run();
  if (rc != DQ4ML_OK) return rc;
  cudaError_t e = launch_rows_digest(guest, price, n, seed, ctx->d_counters, ctx->stream, &ctx->launches);
  if (e != cudaSuccess) return cuda_fail(e, "digest kernel");
  unsigned long long h[2];
  CU(cudaMemcpyAsync(h, ctx->d_counters, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  out[0] = h[0];
  out[1] = h[1];
  return DQ4ML_OK;
}

}  // extern "C"
