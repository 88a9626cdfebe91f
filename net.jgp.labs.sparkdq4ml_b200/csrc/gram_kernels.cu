// gram_kernels.cu -- K5: the WeightedLeastSquares aggregation for wide feature tables (d > 8): X^T X, X^T y, X^T 1,
// y^T y, y^T 1, 1^T 1 in one pass, fp64 throughout.
//
// Spark reaches this from lr.fit (DataQuality4MachineLearningApp.java:126) when solver is auto/normal:
// WeightedLeastSquares.Aggregator.add does BLAS.spr(w, x, aaSum) per row, i.e. a rank-1 update of the packed d x d
// Gram.  Here the same sums are a tiled Z^T Z with Z = [x_0 .. x_{d-1} | y | 1]:
//   * the d columns are cut into blocks of 64; one CTA owns one upper-triangular block pair (bi <= bj) and a slice of
//     the rows, and walks its rows in chunks of 32;
//   * a producer warp stages each chunk of the 64 (+64) columns it needs with one TMA bulk copy per column
//     (cp.async.bulk ... mbarrier::complete_tx, 256 B each: the columnar layout makes a 32-row chunk of a column one
//     contiguous run), two stages deep, full/empty mbarriers;
//   * four consumer warps first fix the chunk up in shared memory (int32 -> fp64 for integer columns, DQ predicate ->
//     rows the rules drop are zeroed, the ones column becomes the keep mask), then each owns a 32 x 32 output tile as
//     4 x 4 fragments of mma.sync.m8n8k4.f64 (DMMA) accumulated in registers; on diagonal pairs the fourth warp, whose
//     tile lies below the diagonal, computes the border columns (X^T y, X^T 1) and the y/1 corner instead;
//   * partial tiles go to scratch, and a second small kernel sums the row slices in a fixed order straight into the
//     WLS moment layout (deterministic, no atomics).
// fp64 DMMA and DFMA both peak at 37.2 TFLOP/s on B200 (tools/fp64_peak.cu), so this pass is bound by the fp64 pipe,
// not by HBM: 66 816 flop per row at d = 256 against 2 056 B.
#include <algorithm>
#include <vector>

#include "device_utils.cuh"
#include "gram.hpp"

namespace dq4ml {
namespace {

constexpr int GT = 64;          // CTA tile edge in columns
constexpr int GKC = 32;         // rows per pipeline stage
constexpr int GKP = 36;         // shared-memory pitch of one column chunk in doubles (32 + 4: conflict-free fragment loads)
constexpr int GSTAGES = 2;
constexpr int GCONSUMERS = 128; // 4 consumer warps; warp 4 is the producer
constexpr int GTHREADS = 160;
constexpr int GPART = GT * GT + GT * 8 + 64;  // doubles per CTA partial: tile, border (64 x 8), corner (8 x 8)

struct __align__(16) GramStage {
  double a[GT][GKP];       // column block bi
  double b[GT][GKP];       // column block bj (unused on diagonal pairs)
  double border[8][GKP];   // column 0 = y, column 1 = ones/keep mask, 2..7 stay zero
  double dqp[GKC];         // DQ price chunk
  int dqg[GKC];            // DQ guest chunk
};
constexpr size_t kGramSmem = GSTAGES * sizeof(GramStage) + 2 * GSTAGES * sizeof(unsigned long long) + 16 * sizeof(unsigned);

struct GramArgs {
  const void* const* col_ptr;  // device array [d]
  const unsigned char* col_i32;  // device array [d]: 1 = int32 column
  int d, nb, kinds, splits;
  const double* label;
  const int* dq_guest;
  const double* dq_price;
  DqThresholds t;
  int has_dq;
  long long n;
  double* partials;  // [splits * kinds][GPART]
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned ok;
  do {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (ok == 0u);
}
// TMA bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(GCONSUMERS) : "memory"); }
__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(GTHREADS) gram_tile_kernel(const GramArgs g) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  GramStage* st = reinterpret_cast<GramStage*>(smem_raw);
  unsigned long long* full = reinterpret_cast<unsigned long long*>(smem_raw + GSTAGES * sizeof(GramStage));
  unsigned long long* empty = full + GSTAGES;
  unsigned* s_mask = reinterpret_cast<unsigned*>(empty + GSTAGES);  // keep mask of the chunk in each stage

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kind = blockIdx.x % g.kinds, split = blockIdx.x / g.kinds;
  int bi = 0, bj = 0;
  {
    int k = kind;
    for (bi = 0;; ++bi) {
      const int row_len = g.nb - bi;
      if (k < row_len) { bj = bi + k; break; }
      k -= row_len;
    }
  }
  const bool diag = bi == bj;

  // zero every stage once: padding columns (>= d), border columns 2..7 and the unused rows stay zero for good
  for (int i = tid; i < static_cast<int>(GSTAGES * sizeof(GramStage) / 8); i += GTHREADS) reinterpret_cast<double*>(smem_raw)[i] = 0.0;
  if (tid == 0) {
    for (int s = 0; s < GSTAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], GCONSUMERS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  __syncthreads();

  const long long full_chunks = g.n / GKC;
  const long long my_iters = full_chunks > split ? (full_chunks - split + g.splits - 1) / g.splits : 0;

  if (warp == 4) {
    // ------------------------------------------------------------------ producer: one bulk copy per column chunk
    unsigned bytes = 0;
    for (int c = lane; c < GT; c += 32) {
      const int ca = bi * GT + c, cb = bj * GT + c;
      if (ca < g.d) bytes += g.col_i32[ca] ? GKC * 4u : GKC * 8u;
      if (!diag && cb < g.d) bytes += g.col_i32[cb] ? GKC * 4u : GKC * 8u;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) bytes += __shfl_xor_sync(0xffffffffu, bytes, off);
    if (diag) bytes += GKC * 8u;               // y
    if (g.has_dq) bytes += GKC * 4u + GKC * 8u;  // DQ guest + price
    for (long long it = 0; it < my_iters; ++it) {
      const int s = static_cast<int>(it % GSTAGES);
      const unsigned use = static_cast<unsigned>(it / GSTAGES);
      mbar_wait(&empty[s], (use & 1u) ^ 1u);
      const long long row = (split + it * g.splits) * GKC;
      if (lane == 0) mbar_expect_tx(&full[s], bytes);
      __syncwarp();
      for (int c = lane; c < GT; c += 32) {
        const int ca = bi * GT + c, cb = bj * GT + c;
        if (ca < g.d) {
          if (g.col_i32[ca]) bulk_g2s(reinterpret_cast<int*>(&st[s].a[c][0]) + GKC, static_cast<const int*>(g.col_ptr[ca]) + row, GKC * 4u, &full[s]);
          else bulk_g2s(&st[s].a[c][0], static_cast<const double*>(g.col_ptr[ca]) + row, GKC * 8u, &full[s]);
        }
        if (!diag && cb < g.d) {
          if (g.col_i32[cb]) bulk_g2s(reinterpret_cast<int*>(&st[s].b[c][0]) + GKC, static_cast<const int*>(g.col_ptr[cb]) + row, GKC * 4u, &full[s]);
          else bulk_g2s(&st[s].b[c][0], static_cast<const double*>(g.col_ptr[cb]) + row, GKC * 8u, &full[s]);
        }
      }
      if (lane == 0) {
        if (diag) bulk_g2s(&st[s].border[0][0], g.label + row, GKC * 8u, &full[s]);
        if (g.has_dq) {
          bulk_g2s(&st[s].dqg[0], g.dq_guest + row, GKC * 4u, &full[s]);
          bulk_g2s(&st[s].dqp[0], g.dq_price + row, GKC * 8u, &full[s]);
        }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumers
  const int wm = warp >> 1, wn = warp & 1;
  const bool border_warp = diag && warp == 2;  // tile (1,0) of a diagonal pair lies below the diagonal
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  double accb[8][2], corner[2] = {0.0, 0.0};
#pragma unroll
  for (int i = 0; i < 8; ++i) accb[i][0] = accb[i][1] = 0.0;

  // fix-up of one staged chunk + the MMAs over it.  rows_valid < GKC only for the tail chunk.
  auto consume = [&](GramStage& S, unsigned* mask_slot, bool convert, int rows_valid) {
    if (convert) {  // integer columns were staged as int32 in the upper half of their slot: widen in place
      for (int c = warp; c < GT; c += GCONSUMERS / 32) {
        const int ca = bi * GT + c, cb = bj * GT + c;
        if (ca < g.d && g.col_i32[ca]) {
          const int v = reinterpret_cast<const int*>(&S.a[c][0])[GKC + lane];
          __syncwarp();
          S.a[c][lane] = static_cast<double>(v);
        }
        if (!diag && cb < g.d && g.col_i32[cb]) {
          const int v = reinterpret_cast<const int*>(&S.b[c][0])[GKC + lane];
          __syncwarp();
          S.b[c][lane] = static_cast<double>(v);
        }
      }
    }
    if (warp == 0) {
      bool keep = lane < rows_valid;
      if (keep && g.has_dq) keep = dq_keep(S.dqg[lane], S.dqp[lane], g.t);
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (lane == 0) *mask_slot = m;
      if (diag) S.border[1][lane] = keep ? 1.0 : 0.0;
    }
    consumer_sync();
    const unsigned dropped = ~(*mask_slot);
    if (dropped != 0u) {  // rows the DQ rules (or the end of the data) remove contribute nothing: zero them in every column
      for (unsigned m = dropped; m != 0u; m &= m - 1u) {
        const int r = __ffs(m) - 1;
        if (tid < GT) S.a[tid][r] = 0.0;
        else if (!diag) S.b[tid - GT][r] = 0.0;
        if (diag && tid == GT) S.border[0][r] = 0.0;
      }
      consumer_sync();
    }
    const double(*A)[GKP] = S.a;
    const double(*B)[GKP] = diag ? S.a : S.b;
    const int fr = lane >> 2, fk = lane & 3;
    if (!border_warp) {
#pragma unroll
      for (int k0 = 0; k0 < GKC; k0 += 4) {
        double af[4], bf[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) af[i] = A[wm * 32 + i * 8 + fr][k0 + fk];
#pragma unroll
        for (int j = 0; j < 4; ++j) bf[j] = B[wn * 32 + j * 8 + fr][k0 + fk];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dmma884(acc[i][j], af[i], bf[j]);
      }
    } else {
#pragma unroll
      for (int k0 = 0; k0 < GKC; k0 += 4) {
        const double bb = S.border[fr][k0 + fk];
#pragma unroll
        for (int i = 0; i < 8; ++i) dmma884(accb[i], A[i * 8 + fr][k0 + fk], bb);
        if (bi == 0) dmma884(corner, bb, bb);
      }
    }
  };

  for (long long it = 0; it < my_iters; ++it) {
    const int s = static_cast<int>(it % GSTAGES);
    const unsigned use = static_cast<unsigned>(it / GSTAGES);
    mbar_wait(&full[s], use & 1u);
    consume(st[s], &s_mask[s], true, GKC);
    fence_proxy_async();  // our generic-proxy writes to this stage precede the producer's next bulk copy into it
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
  }

  // tail chunk (n % 32 rows): plain loads by the consumers of the CTA whose turn it is
  const int tail_rows = static_cast<int>(g.n - full_chunks * GKC);
  if (tail_rows > 0 && static_cast<int>(full_chunks % g.splits) == split) {
    consumer_sync();
    GramStage& S = st[0];  // every stage has been released by now
    const long long row0 = full_chunks * GKC;
    for (int idx = tid; idx < 2 * GT * GKC; idx += GCONSUMERS) {
      const int r = idx % GKC, c = (idx / GKC) % GT, which = idx / (GKC * GT);
      if (which == 1 && diag) continue;
      const int col = (which == 0 ? bi : bj) * GT + c;
      double v = 0.0;
      if (col < g.d && r < tail_rows)
        v = g.col_i32[col] ? static_cast<double>(static_cast<const int*>(g.col_ptr[col])[row0 + r])
                           : static_cast<const double*>(g.col_ptr[col])[row0 + r];
      (which == 0 ? S.a : S.b)[c][r] = v;
    }
    if (tid < GKC) {
      S.border[0][tid] = (diag && tid < tail_rows) ? g.label[row0 + tid] : 0.0;
      S.dqg[tid] = (g.has_dq && tid < tail_rows) ? g.dq_guest[row0 + tid] : 0;
      S.dqp[tid] = (g.has_dq && tid < tail_rows) ? g.dq_price[row0 + tid] : 0.0;
    }
    consumer_sync();
    consume(S, &s_mask[0], false, tail_rows);
  }

  // -------------------------------------------------------------------- partial tile -> scratch
  double* P = g.partials + static_cast<size_t>(blockIdx.x) * GPART;
  const int fr = lane >> 2, fc = (lane & 3) * 2;
  if (!border_warp) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = wm * 32 + i * 8 + fr, c = wn * 32 + j * 8 + fc;
        *reinterpret_cast<double2*>(&P[r * GT + c]) = make_double2(acc[i][j][0], acc[i][j][1]);
      }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      *reinterpret_cast<double2*>(&P[GT * GT + (i * 8 + fr) * 8 + fc]) = make_double2(accb[i][0], accb[i][1]);
    *reinterpret_cast<double2*>(&P[GT * GT + GT * 8 + fr * 8 + fc]) = make_double2(corner[0], corner[1]);
  }
}

// Sums the row slices in a fixed order and lays the result out as
// [count, wSum, wwSum, bSum, bbSum, aSum[d], abSum[d], aaSum packed upper column-major].
__global__ void __launch_bounds__(256) gram_reduce_kernel(const double* __restrict__ partials, int d, int nb, int kinds,
                                                          int splits, long long len, double* __restrict__ out) {
  const long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= len) return;
  int bi, bj, off;
  if (e < 5) {
    bi = bj = 0;
    off = GT * GT + GT * 8 + (e < 3 ? 1 * 8 + 1 : (e == 3 ? 0 * 8 + 1 : 0));  // corner: [1][1] = count, [0][1] = sum y, [0][0] = sum y^2
  } else if (e < 5 + 2LL * d) {
    const int j = static_cast<int>((e - 5) % d);
    const bool ab = (e - 5) >= d;
    bi = bj = j / GT;
    off = GT * GT + (j % GT) * 8 + (ab ? 0 : 1);  // border column 0 = x^T y, column 1 = x^T 1
  } else {
    const long long q = e - 5 - 2LL * d;
    long long j = static_cast<long long>((sqrt(8.0 * static_cast<double>(q) + 1.0) - 1.0) * 0.5);
    while (j * (j + 1) / 2 > q) --j;
    while ((j + 1) * (j + 2) / 2 <= q) ++j;
    const long long i = q - j * (j + 1) / 2;
    bi = static_cast<int>(i / GT);
    bj = static_cast<int>(j / GT);
    off = static_cast<int>(i % GT) * GT + static_cast<int>(j % GT);
  }
  const int kind = bi * nb - bi * (bi - 1) / 2 + (bj - bi);
  double s = 0.0;
  for (int sp = 0; sp < splits; ++sp) s += partials[(static_cast<size_t>(sp) * kinds + kind) * GPART + off];
  out[e] = s;
}

}  // namespace

int gram_wls_moments(dq4ml_ctx* ctx, const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes,
                     const double* label, const DeviceDq& dq, int64_t n, double* out) {
  const int d = static_cast<int>(col_ptrs.size());
  const int nb = (d + GT - 1) / GT;
  const int kinds = nb * (nb + 1) / 2;
  const int64_t len = 5 + 2LL * d + static_cast<int64_t>(d) * (d + 1) / 2;
  const int64_t full_chunks = n / GKC;
  int64_t splits = (2LL * ctx->sm_count + kinds - 1) / kinds;
  if (splits > full_chunks) splits = full_chunks;
  if (splits < 1) splits = 1;
  const size_t table_bytes = (static_cast<size_t>(d) * 9 + 255) & ~static_cast<size_t>(255);
  const size_t part_bytes = static_cast<size_t>(splits) * kinds * GPART * sizeof(double);
  const size_t out_bytes = static_cast<size_t>(len) * sizeof(double);
  const size_t need = table_bytes + part_bytes + out_bytes + 512;
  auto fail = [](cudaError_t e, const char* what) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return DQ4ML_ERR_CUDA;
  };
  cudaError_t e;
  if (need > ctx->gram_cap) {
    if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return fail(e, "sync");
    if (ctx->gram_scratch != nullptr) cudaFree(ctx->gram_scratch);
    ctx->gram_scratch = nullptr;
    ctx->gram_cap = 0;
    if ((e = cudaMalloc(reinterpret_cast<void**>(&ctx->gram_scratch), need)) != cudaSuccess) return fail(e, "gram scratch");
    ctx->gram_cap = need;
  }
  if (out_bytes > ctx->gram_host_cap) {
    if (ctx->gram_host != nullptr) cudaFreeHost(ctx->gram_host);
    ctx->gram_host = nullptr;
    ctx->gram_host_cap = 0;
    if ((e = cudaMallocHost(reinterpret_cast<void**>(&ctx->gram_host), out_bytes)) != cudaSuccess) return fail(e, "gram pinned result");
    ctx->gram_host_cap = out_bytes;
  }
  char* base = ctx->gram_scratch;
  const void** d_ptrs = reinterpret_cast<const void**>(base);
  unsigned char* d_i32 = reinterpret_cast<unsigned char*>(base + static_cast<size_t>(d) * 8);
  double* d_part = reinterpret_cast<double*>(base + table_bytes);
  double* d_out = reinterpret_cast<double*>(base + table_bytes + part_bytes);
  std::vector<unsigned char> flags(d);
  for (int j = 0; j < d; ++j) flags[j] = dtypes[j] == DQ4ML_I32 ? 1 : 0;
  if ((e = cudaMemcpyAsync(d_ptrs, col_ptrs.data(), static_cast<size_t>(d) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e, "column table upload");
  if ((e = cudaMemcpyAsync(d_i32, flags.data(), static_cast<size_t>(d), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e, "column table upload");
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return fail(e, "sync");  // the host vectors go out of scope

  GramArgs g;
  g.col_ptr = d_ptrs;
  g.col_i32 = d_i32;
  g.d = d;
  g.nb = nb;
  g.kinds = kinds;
  g.splits = static_cast<int>(splits);
  g.label = label;
  g.dq_guest = dq.guest;
  g.dq_price = dq.price;
  g.t.min_price = dq.params.min_price;
  g.t.max_price = dq.params.max_price;
  g.t.min_guest = dq.params.min_guest;
  g.has_dq = dq.guest != nullptr ? 1 : 0;
  g.n = n;
  g.partials = d_part;
  static bool attr_set[64] = {false};
  if (ctx->device < 64 && !attr_set[ctx->device]) {
    if ((e = cudaFuncSetAttribute(gram_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kGramSmem))) != cudaSuccess)
      return fail(e, "gram smem attribute");
    attr_set[ctx->device] = true;
  }
  if (ctx->pass_timing) cudaEventRecord(ctx->pass_ev0, ctx->stream);
  gram_tile_kernel<<<static_cast<unsigned>(splits * kinds), GTHREADS, kGramSmem, ctx->stream>>>(g);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "gram tile kernel");
  ++ctx->launches;
  gram_reduce_kernel<<<static_cast<unsigned>((len + 255) / 256), 256, 0, ctx->stream>>>(d_part, d, nb, kinds, static_cast<int>(splits), len, d_out);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "gram reduce kernel");
  ++ctx->launches;
  if (ctx->pass_timing) {
    cudaEventRecord(ctx->pass_ev1, ctx->stream);
    ctx->pass_event_pending = true;
  }
  if (ctx->nccl != nullptr) {
    const int rc = nccl_allreduce_device(ctx, d_out, len);
    if (rc != DQ4ML_OK) return rc;
  }
  if ((e = cudaMemcpyAsync(ctx->gram_host, d_out, out_bytes, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return fail(e, "gram D2H");
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return fail(e, "gram pass");
  if (ctx->pass_event_pending) {
    ctx->pass_event_pending = false;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ctx->pass_ev0, ctx->pass_ev1) == cudaSuccess) {
      ctx->pass_ms_total += ms;
      ++ctx->pass_count;
    }
  }
  if (ctx->nccl == nullptr && ctx->allreduce_cb != nullptr) {
    if (ctx->allreduce_cb(ctx->allreduce_user, ctx->gram_host, len) != 0) {
      set_error("all-reduce callback failed");
      return DQ4ML_ERR_COMM;
    }
  }
  std::copy(ctx->gram_host, ctx->gram_host + len, out);
  return DQ4ML_OK;
}

}  // namespace dq4ml
