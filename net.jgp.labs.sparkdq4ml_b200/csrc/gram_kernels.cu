// gram_kernels.cu -- K5: the WeightedLeastSquares aggregation for wide feature tables (d > 8): X^T X, X^T y, X^T 1,
// y^T y, y^T 1, 1^T 1 in one pass, fp64 throughout.
//
// Spark reaches this from lr.fit (DataQuality4MachineLearningApp.java:126) when solver is auto/normal:
// WeightedLeastSquares.Aggregator.add does BLAS.spr(w, x, aaSum) per row, i.e. a rank-1 update of the packed d x d
// Gram.  Here the same sums are a tiled Z^T Z with Z = [x_0 .. x_{d-1} | y | 1]:
//   * the d columns are cut into blocks of 64; a "pair" is one upper-triangular block pair (bi <= bj); the rows are cut
//     into chunks of 32 and dealt out to row slices (slice s of a pair takes chunks s, s + S, s + 2 S, ...);
//   * one CTA per SM = two independent halves ("sub-CTAs"), each 4 MMA warps + 2 producer warps with its own three
//     stages and mbarriers, working on adjacent slices of the same pair;
//   * the producer warps stage a chunk of the 64 (+64) columns with 16-byte cp.async copies (the columnar layout makes
//     a 32-row chunk of a column one contiguous 256-byte run) into an XOR-swizzled, padding-free layout, then fix it
//     up in shared memory (int32 -> fp64 for integer columns, DQ predicate -> rows the rules drop are zeroed, the ones
//     column becomes the keep mask, sqrt(w) scaling) and hand it over through a `full` mbarrier; the MMA warps hand
//     it back through an `empty` mbarrier.  No CTA-wide barrier inside the row loop;
//   * off the diagonal each MMA warp owns a 32 x 32 output tile as 4 x 4 fragments of mma.sync.m8n8k4.f64 (DMMA)
//     accumulated in registers; on the diagonal only the upper triangle is needed (36 of the 64 fragments, plus the 8
//     border fragments X^T y / X^T 1 and the y/1 corner), dealt out so that every warp has 11 (12 with the corner);
//   * partial tiles go to scratch, and a second small kernel sums the row slices in a fixed order straight into the
//     WLS moment layout (deterministic, no atomics).
// fp64 DMMA and DFMA both peak at 37.2 TFLOP/s on B200 (tools/fp64_peak.cu), so this pass is bound by the fp64 pipe,
// not by HBM: 66 816 flop per row at d = 256 against 2 056 B.  History of the kernel at 62.5 M x 256 (ms per pass,
// profiles/r2_gram_fp64_history.jsonl): round 1's kernel (4 warps staging, fixing up and multiplying behind two
// CTA-wide barriers per chunk, full 64 x 64 tile on the diagonal) 170; one producer warp with a branchy copy loop 333
// (the MMA warps waited for it 29 % of the time); lean copy loop 156; two producer warps 150; both halves of an SM on
// the same kind of pair, slices in proportion to the measured cost 136.  Round 1 had also tried one TMA bulk copy
// (cp.async.bulk) per column chunk: 130 copies of 256 B per stage, 23 % of the fp64 peak
// (profiles/r1_gram_tma_bulk_full.json); a column table of arbitrary pointers cannot be described by one tensor map.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "device_utils.cuh"
#include "gram.hpp"

namespace dq4ml {
namespace {

constexpr int GT = 64;          // CTA tile edge in columns
constexpr int GKC = 32;         // rows per pipeline stage: one column chunk is 256 B (fp64) or 128 B (int32)
constexpr int GSTAGES = 3;
constexpr int GPART = GT * GT + GT * 8 + 64;  // doubles per CTA partial: tile, border (64 x 8), corner (8 x 8)

// Shared-memory layout of a staged chunk.  Each column chunk is a dense 256-byte row of 16 x 16-byte pieces; the
// position of a piece is XOR-ed with twice the column's index mod 8, which makes the 8-byte fragment loads of
// mma.m8n8k4 (8 columns x 4 consecutive rows per warp) conflict-free without any padding: 3 stages of both panels fit
// twice (one per sub-CTA) in one SM's 227 KB.
struct __align__(16) GramStage {
  double a[GT][GKC];       // column block bi
  double b[GT][GKC];       // column block bj (unused on diagonal pairs)
  double border[8][GKC];   // column 0 = sqrt(w) y, 1 = sqrt(w) keep, 2 = w keep, 3 = keep (w = 1 unweighted); 4..7 stay zero
  // one copy per producer warp
  double wt[2][GKC];       // weight chunk, replaced by sqrt(w) * keep during the fix-up (weighted fits only)
  double dqp[2][GKC];      // DQ price chunk
  int dqg[2][GKC];         // DQ guest chunk
};
constexpr size_t kGramSmem = GSTAGES * sizeof(GramStage) + 2 * GT * sizeof(void*) + 2 * GT + 2 * GT + 64;

__device__ __forceinline__ int swz(int col, int k) {  // index (in doubles) of element k of column `col` inside its 32-double row
  return ((((k >> 1) ^ ((col & 7) << 1)) << 1) | (k & 1));
}

struct GramArgs {
  const void* const* col_ptr;  // device array [d]
  const unsigned char* col_i32;  // device array [d]: 1 = int32 column
  int d, nb, kinds;
  const double* label;
  const double* weight;  // instance weights (weightCol), nullptr = all 1
  const int* dq_guest;
  const double* dq_price;
  DqThresholds t;
  int has_dq;
  long long n;
  double* partials;  // [kinds][s_max][GPART]
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
// 16-byte asynchronous copy global -> shared (SASS: LDGSTS), bypassing L1
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

constexpr int GWS_THREADS = 192;  // 4 MMA warps + 2 producer warps
constexpr int GWS_MMA_WARPS = 4;
// Measured time per chunk of a sub-CTA next to one of its own kind (62.5 M x 256, DQ4ML_GRAM_DEBUG=1), on the scale
// where an off-diagonal pair costs its 16 fragments per warp: 2.26 us off the diagonal, 1.64 us on it (11 fragments per
// warp), 1.69 us for pair (0,0) (warp 3 also has the corner).
constexpr double kOffCost = 16.0, kDiagCost = 11.6, kDiag0Cost = 12.0;
constexpr size_t kGramWsSmem = (kGramSmem + 64 + 2 * GT * sizeof(void*) + 127) & ~static_cast<size_t>(127);  // one sub-CTA: + the stage mbarriers and the fp64-only pointer table

struct GramWsArgs {
  GramArgs g;
  int s_off, s_diag, s_diag0;     // row slices per off-diagonal pair, per diagonal pair, of pair (0,0) (it has the corner)
  int s_max;                      // partial slot stride: slot [pair][slice]
  unsigned long long* progress;   // [kinds] chunks issued per pair, zeroed before the launch; nullptr = no pacing
  int lead;                       // a producer does not run more than this many chunks ahead of the slowest pair
  int debug;                      // profile build only (DQ4ML_GRAM_PROFILE) + DQ4ML_GRAM_DEBUG=1: slice 0 of every pair prints its run time
};

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Waits for the phase with the given parity.  A pipeline bug would otherwise hang the GPU: after ~2^31 polls the CTA
// traps, which fails the launch with an error instead.
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned done = 0, polls = 0;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (!done && ++polls == 0x7fffffffu) __trap();
  } while (!done);
}
__device__ __forceinline__ unsigned long long ld_relaxed_gpu(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// The MMAs of warp WQ of a diagonal CTA over one staged chunk: fragment rows r0 = WQ and r1 = 7 - WQ of the upper
// triangle, the border fragment of both rows, and (warp 3 of pair (0,0)) the corner.  acc[] is used flat:
//   [0 .. 7-WQ]          (r0, r0 + t)      [8-WQ]      (r0, border)
//   [9-WQ .. 9]          (r1, r1 + t)      [10]        (r1, border)       [11]  corner
template <int WQ>
__device__ __forceinline__ void diag_chunk_mma(const GramStage& S, double (&acc)[16][2], int lane, bool corner) {
  constexpr int R0 = WQ, R1 = 7 - WQ;
  const int fr = lane >> 2, fk = lane & 3, sw = fr << 1;
#pragma unroll
  for (int k0 = 0; k0 < GKC; k0 += 4) {
    const int off = (((((k0 + fk) >> 1) ^ sw) << 1) | (fk & 1));
    double f[8];  // f[j]: fragment of column block j (A and B operand are the same data on a diagonal pair)
#pragma unroll
    for (int j = R0; j < 8; ++j) f[j] = S.a[j * 8 + fr][off];
    const double bb = S.border[fr][off];
#pragma unroll
    for (int j = R0; j < 8; ++j) dmma884(acc[j - R0], f[R0], f[j]);
    dmma884(acc[8 - WQ], f[R0], bb);
#pragma unroll
    for (int j = R1; j < 8; ++j) dmma884(acc[9 - WQ + (j - R1)], f[R1], f[j]);
    dmma884(acc[10], f[R1], bb);
    if (WQ == 3 && corner) dmma884(acc[11], bb, bb);
  }
}

template <int WQ>
__device__ __forceinline__ void diag_store(double* P, const double (&acc)[16][2], int lane, bool corner) {
  constexpr int R0 = WQ, R1 = 7 - WQ;
  const int fr = lane >> 2, fc = (lane & 3) * 2;
#pragma unroll
  for (int j = R0; j < 8; ++j)
    *reinterpret_cast<double2*>(&P[(R0 * 8 + fr) * GT + j * 8 + fc]) = make_double2(acc[j - R0][0], acc[j - R0][1]);
  *reinterpret_cast<double2*>(&P[GT * GT + (R0 * 8 + fr) * 8 + fc]) = make_double2(acc[8 - WQ][0], acc[8 - WQ][1]);
#pragma unroll
  for (int j = R1; j < 8; ++j)
    *reinterpret_cast<double2*>(&P[(R1 * 8 + fr) * GT + j * 8 + fc]) =
        make_double2(acc[9 - WQ + (j - R1)][0], acc[9 - WQ + (j - R1)][1]);
  *reinterpret_cast<double2*>(&P[GT * GT + (R1 * 8 + fr) * 8 + fc]) = make_double2(acc[10][0], acc[10][1]);
  if (WQ == 3 && corner) *reinterpret_cast<double2*>(&P[GT * GT + GT * 8 + fr * 8 + fc]) = make_double2(acc[11][0], acc[11][1]);
}

// dq_keep() on the bit pattern of the price.  The producer warps share their SM's fp64 pipe with the DMMAs, which keep
// it saturated: a DSETP of the producer waited for it for ~27 % of the producer's time (profiles/r2_gram_ws_full.json
// history in profiles/README.md), on the path that hands a chunk to the MMA warps.  Same truth table as dq_keep():
//   keep  <=>  !(price < min_price) && !(price <= 0) && !(guest < min_guest && price > max_price),  IEEE compares
// (false whenever an operand is NaN).  key() maps a non-NaN double to an unsigned integer with the same order; -0.0
// is folded into +0.0 first because the two compare equal.
struct DqKeys {
  unsigned long long min_price, max_price;  // keys of the thresholds
  bool min_nan, max_nan;
  int min_guest;
};
__host__ __device__ __forceinline__ bool f64_bits_nan(unsigned long long b) { return (b & 0x7fffffffffffffffULL) > 0x7ff0000000000000ULL; }
__host__ __device__ __forceinline__ unsigned long long f64_key(unsigned long long b) {
  if ((b << 1) == 0ULL) b = 0ULL;
  return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ DqKeys make_dq_keys(const DqThresholds& t) {
  DqKeys k;
  const unsigned long long bmin = static_cast<unsigned long long>(__double_as_longlong(t.min_price));
  const unsigned long long bmax = static_cast<unsigned long long>(__double_as_longlong(t.max_price));
  k.min_price = f64_key(bmin);
  k.max_price = f64_key(bmax);
  k.min_nan = f64_bits_nan(bmin);
  k.max_nan = f64_bits_nan(bmax);
  k.min_guest = t.min_guest;
  return k;
}
__device__ __forceinline__ bool dq_keep_bits(int guest, unsigned long long price_bits, const DqKeys& k) {
  if (f64_bits_nan(price_bits)) return true;  // every IEEE compare is false: the row passes both rules and both filters
  const unsigned long long key = f64_key(price_bits);
  const bool lt_min = !k.min_nan && key < k.min_price;
  const bool le_zero = key <= 0x8000000000000000ULL;
  const bool gt_max = !k.max_nan && key > k.max_price;
  return !lt_min && !le_zero && !(guest < k.min_guest && gt_max);
}

template <bool W>
__global__ void __launch_bounds__(2 * GWS_THREADS, 1) gram_ws_kernel(const GramWsArgs a) {
  const GramArgs& g = a.g;
  extern __shared__ __align__(128) unsigned char smem_all[];
  // one CTA per SM = two independent halves ("sub-CTAs": 4 MMA warps + 2 producer warps, own stages and mbarriers).
  // Two sub-CTAs share the SM's DMMA pipe, and a pair of different kinds does not finish together (measured: an
  // off-diagonal CTA next to a diagonal one ends at 117 ms, its neighbour at 139 ms, and the pipe is half idle in
  // between), so both halves of a CTA always work on the same block pair: adjacent row slices of it.
  const int sub = threadIdx.x / GWS_THREADS;
  unsigned char* smem_raw = smem_all + sub * kGramWsSmem;
  GramStage* st = reinterpret_cast<GramStage*>(smem_raw);
  const void** s_colptr = reinterpret_cast<const void**>(smem_raw + GSTAGES * sizeof(GramStage));  // [2*GT]: A slots then B slots
  unsigned char* s_isint = reinterpret_cast<unsigned char*>(s_colptr + 2 * GT);                    // [2*GT]
  unsigned char* s_intslots = s_isint + 2 * GT;                                                    // [2*GT] compact list of int32 slots
  int* s_meta = reinterpret_cast<int*>(s_intslots + 2 * GT);                                       // [0] = #int slots
  unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(s_meta + 16);                  // full[GSTAGES], empty[GSTAGES]
  const void** s_f64ptr = reinterpret_cast<const void**>(s_bar + 2 * GSTAGES + 2);                 // [2*GT]: s_colptr with the int32 slots nulled

  const int tid = threadIdx.x - sub * GWS_THREADS, warp = tid >> 5, lane = tid & 31;
#ifdef DQ4ML_GRAM_PROFILE
  const unsigned long long t_start = a.debug ? global_timer_ns() : 0ULL;
#endif
  // sub-CTA -> (pair, row slice): all slices of the off-diagonal pairs first, then those of the diagonal pairs, so that
  // the two halves of a CTA work on pairs of the same cost (at most one CTA straddles the two kinds)
  int kind = 0, split = static_cast<int>(blockIdx.x) * 2 + sub;
  bool active = false;
  for (int pass = 0; pass < 2 && !active; ++pass)
    for (int i = 0, k = 0; i < g.nb && !active; ++i)
      for (int j = i; j < g.nb; ++j, ++k) {
        if ((i == j) != (pass == 1)) continue;
        const int sk = i != j ? a.s_off : (i == 0 ? a.s_diag0 : a.s_diag);
        if (split < sk) { kind = k; active = true; break; }
        split -= sk;
      }
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
  const int my_splits = !diag ? a.s_off : (bi == 0 ? a.s_diag0 : a.s_diag);
  const int n_slots = diag ? GT : 2 * GT;

  for (int i = tid; i < static_cast<int>(GSTAGES * sizeof(GramStage) / 8); i += GWS_THREADS) reinterpret_cast<double*>(smem_raw)[i] = 0.0;
  if (tid < 2 * GT) {
    const int col = (tid < GT ? bi : bj) * GT + (tid & (GT - 1));
    const bool live = active && tid < n_slots && col < g.d;
    s_colptr[tid] = live ? g.col_ptr[col] : nullptr;
    s_isint[tid] = live ? g.col_i32[col] : 0;
    s_f64ptr[tid] = (live && g.col_i32[col] == 0) ? g.col_ptr[col] : nullptr;
  }
  if (tid == 0) {
    for (int s = 0; s < GSTAGES; ++s) {
      mbar_init(smem_u32(&s_bar[s]), 64);                        // full: every lane of both producer warps
      mbar_init(smem_u32(&s_bar[GSTAGES + s]), GWS_MMA_WARPS);   // empty: one lane of every MMA warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    int m = 0;
    for (int s = 0; s < n_slots; ++s)
      if (s_isint[s]) s_intslots[m++] = static_cast<unsigned char>(s);
    s_meta[0] = m;
  }
  __syncthreads();
  if (!active) return;  // the odd half of the last CTA when the slice count is odd (n < 64 rows)
  const int n_int = s_meta[0];

  const long long full_chunks = g.n / GKC;
  const long long my_iters = full_chunks > split ? (full_chunks - split + my_splits - 1) / my_splits : 0;
  const int tail_rows = static_cast<int>(g.n - full_chunks * GKC);
  const bool has_tail = tail_rows > 0 && static_cast<int>(full_chunks % my_splits) == split;
  const long long total = my_iters + (has_tail ? 1 : 0);  // the tail chunk (n % 32 rows) is one more, shorter, stage

  if (warp >= GWS_MMA_WARPS) {
    // ======================================================================================== producer warps
    // Producer p stages and fixes up half of the column slots (off the diagonal: p = 0 the A panel, p = 1 the B
    // panel); each keeps its own copy of the chunk's DQ columns and weights, so neither waits for the other's copies.
    const int pw = warp - GWS_MMA_WARPS;
    const int slot_lo = pw * (n_slots / 2), slot_hi = slot_lo + n_slots / 2;
    const int ch = lane & 15, half = lane >> 4;
    bool pacing = a.progress != nullptr && pw == 0;
    const DqKeys keys = make_dq_keys(g.t);
    // all of this warp's slots are live fp64 columns: the copy loop needs no per-slot test
    bool plain = true;
    for (int sl = slot_lo; sl < slot_hi; ++sl) plain = plain && s_f64ptr[sl] != nullptr;
    int my_int = 0;  // does this warp own an int32 slot at all
    for (int q = 0; q < n_int; ++q) my_int += (s_intslots[q] >= slot_lo && s_intslots[q] < slot_hi) ? 1 : 0;
    // destination of this lane's piece of slot s0 + 2u relative to slot s0 - half (a multiple of 16, so the swizzle
    // term depends on u alone): the column's 256-byte row, chunk position XOR-ed with twice the slot's index mod 8
    unsigned dst_off[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) dst_off[u] = static_cast<unsigned>((half + 2 * u) * 256 + ((ch ^ (((half + 2 * u) & 7) << 1)) << 4));
    // stage `it`: cp.async for a full chunk, plain loads for the tail
    auto fill = [&](long long it, int stage) {
      GramStage& S = st[stage];
      if (it < my_iters) {
        const long long row = (split + it * my_splits) * GKC;
        const unsigned sbase = smem_u32(&S.a[0][0]);
        // two column chunks (2 x 16 pieces of 16 bytes) per warp instruction; the column pointers of eight
        // instructions are fetched from shared memory first, so that the copies are not issued one shared-memory
        // latency apart (no "memory" clobber on the copies themselves: wait_group carries it).  int32 columns are
        // null in s_f64ptr and follow in their own short loop.
        const long long src_off = row * 8 + ch * 16;
        for (int s0 = slot_lo + half; s0 < slot_hi; s0 += 16) {
          const char* base[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) base[u] = static_cast<const char*>(s_f64ptr[s0 + 2 * u]);
          const unsigned dst0 = sbase + (s0 - half) * 256u;
          if (plain) {
#pragma unroll
            for (int u = 0; u < 8; ++u) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + dst_off[u]), "l"(base[u] + src_off));
          } else {
#pragma unroll
            for (int u = 0; u < 8; ++u)
              if (base[u] != nullptr) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + dst_off[u]), "l"(base[u] + src_off));
          }
        }
        if (my_int > 0 && ch < 8) {  // an int32 chunk is 128 B: 8 pieces, parked in the upper half of the row until widened
          for (int q = half; q < n_int; q += 2) {
            const int slot = s_intslots[q];
            if (slot >= slot_lo && slot < slot_hi)
              asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sbase + slot * 256u + 128u + ch * 16u),
                           "l"(static_cast<const char*>(s_colptr[slot]) + row * 4 + ch * 16));
          }
        }
        if (diag && pw == 0 && lane < 16) cp_async16(&S.border[0][0] + (lane << 1), g.label + row + lane * 2);  // y: border column 0 (no swizzle)
        if (g.has_dq) {
          if (lane >= 16) cp_async16(&S.dqp[pw][(lane - 16) * 2], g.dq_price + row + (lane - 16) * 2);
          if (lane < 8) cp_async16(&S.dqg[pw][lane * 4], g.dq_guest + row + lane * 4);
        }
        if (W && lane >= 16) cp_async16(&S.wt[pw][(lane - 16) * 2], g.weight + row + (lane - 16) * 2);
      } else {
        const long long row0 = full_chunks * GKC;
        for (int slot = slot_lo; slot < slot_hi; ++slot) {
          double v = 0.0;
          if (s_colptr[slot] != nullptr && lane < tail_rows)
            v = s_isint[slot] ? static_cast<double>(static_cast<const int*>(s_colptr[slot])[row0 + lane])
                              : static_cast<const double*>(s_colptr[slot])[row0 + lane];
          (slot < GT ? S.a[slot] : S.b[slot - GT])[swz(slot, lane)] = v;
        }
        if (pw == 0) S.border[0][swz(0, lane)] = (diag && lane < tail_rows) ? g.label[row0 + lane] : 0.0;
        S.dqg[pw][lane] = (g.has_dq && lane < tail_rows) ? g.dq_guest[row0 + lane] : 0;
        S.dqp[pw][lane] = (g.has_dq && lane < tail_rows) ? g.dq_price[row0 + lane] : 0.0;
        if (W) S.wt[pw][lane] = lane < tail_rows ? g.weight[row0 + lane] : 0.0;
      }
    };
    // everything the MMA warps must not see raw: int32 -> fp64, the DQ predicate (dropped rows are zeroed, the ones
    // column becomes the keep mask) and the sqrt(w) scaling.  lane = row of the chunk.
    auto fix_up = [&](long long it, int stage) {
      GramStage& S = st[stage];
      const bool is_tail = it >= my_iters;
      if (!is_tail && my_int > 0) {
        for (int q = 0; q < n_int; ++q) {
          const int slot = s_intslots[q];
          if (slot < slot_lo || slot >= slot_hi) continue;
          double* row = slot < GT ? S.a[slot] : S.b[slot - GT];
          const int v = reinterpret_cast<const int*>(row)[GKC + lane];
          __syncwarp();
          row[swz(slot, lane)] = static_cast<double>(v);
        }
        __syncwarp();
      }
      bool keep = lane < (is_tail ? tail_rows : GKC);
      if (keep && g.has_dq) keep = dq_keep_bits(S.dqg[pw][lane], reinterpret_cast<const unsigned long long*>(S.dqp[pw])[lane], keys);
      const unsigned dropped = ~__ballot_sync(0xffffffffu, keep);
      // WeightedLeastSquares.Aggregator: wSum += w, wwSum += w^2, bSum += w y, aSum += w x, aaSum += w x x^T.  Every row of
      // the chunk is scaled by sqrt(w) below, so Z^T Z carries the weights; count and wwSum ride in two more border columns.
      if constexpr (W) {
        const double w = keep ? S.wt[pw][lane] : 0.0;
        const double sw = sqrt(w);
        S.wt[pw][lane] = sw;
        if (diag && pw == 0) {
          S.border[1][swz(1, lane)] = sw;
          S.border[2][swz(2, lane)] = w;
          S.border[3][swz(3, lane)] = keep ? 1.0 : 0.0;
        }
      } else {
        if (diag && pw == 0) S.border[1][swz(1, lane)] = keep ? 1.0 : 0.0;  // count = wSum = wwSum: one border column is enough
      }
      for (unsigned m = dropped; m != 0u; m &= m - 1u) {
        const int r = __ffs(m) - 1;
        for (int slot = slot_lo + lane; slot < slot_hi; slot += 32) (slot < GT ? S.a[slot] : S.b[slot - GT])[swz(slot, r)] = 0.0;
        if (diag && pw == 0 && lane == 0) S.border[0][swz(0, r)] = 0.0;
      }
      if constexpr (W) {
        __syncwarp();
        // piece p of a slot holds rows 2 lch, 2 lch + 1 with lch = p ^ ((slot & 7) << 1)
        double* panel = &S.a[0][0];
#pragma unroll 4
        for (int slot = slot_lo + half; slot < slot_hi; slot += 2) {
          const int lch = ch ^ ((slot & 7) << 1);
          double2* q = reinterpret_cast<double2*>(panel + slot * GKC + 2 * ch);
          double2 v = *q;
          v.x *= S.wt[pw][2 * lch];
          v.y *= S.wt[pw][2 * lch + 1];
          *q = v;
        }
        if (diag && pw == 0) S.border[0][swz(0, lane)] *= S.wt[pw][lane];
      }
    };
    // Pacing (opt-in, DQ4ML_GRAM_PACE=<lead>): one counter per pair holds the chunks its CTAs have issued; a producer that
    // is more than `lead` chunks ahead of the slowest pair sleeps (bounded: after 200 us it stops pacing for good, a
    // stalled pair must not hang it).  It keeps the pairs on the same rows, so each block is read from DRAM once
    // instead of 3.3 times (16.6 GB against 54.8 GB at 8 M x 256), but the pass is bound by the DMMA pipe, not by
    // DRAM, and holding every pair to the slowest one costs more than the re-reads: 142.6 ms against 135.7 ms.
    unsigned long long seen = 0;  // the slowest pair's counter as this warp last read it (read one iteration early)
    auto read_progress = [&]() {
      unsigned long long lo = ~0ULL;
      for (int k = lane; k < g.kinds; k += 32) {
        const unsigned long long v = ld_relaxed_gpu(a.progress + k);
        lo = v < lo ? v : lo;
      }
      return lo;
    };
    auto warp_min = [&](unsigned long long lo) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, lo, o);
        lo = other < lo ? other : lo;
      }
      return lo;
    };
    auto pace = [&](long long it) {
      if (!pacing) return;
      if ((it & 3) == 3) seen = read_progress();  // consumed one iteration later: the load latency stays off the critical path
      if ((it & 3) != 0) return;
      if (lane == 0) atomicAdd(a.progress + kind, 4ULL);
      const long long my_chunk = split + it * my_splits;
      unsigned long long lo = warp_min(seen);
      unsigned long long t0 = 0;
      while (my_chunk > static_cast<long long>(lo) + a.lead) {
        const unsigned long long now = global_timer_ns();
        if (t0 == 0) t0 = now;
        if (now - t0 > 200000ULL) { pacing = false; break; }
        __nanosleep(500);
        lo = warp_min(read_progress());
      }
    };

    for (int pre = 0; pre < GSTAGES - 1; ++pre) {
      if (pre < total) fill(pre, pre);
      cp_async_commit();
    }
    int stage = 0, fill_stage = GSTAGES - 1;  // it % GSTAGES and (it + GSTAGES - 1) % GSTAGES, kept as counters
    unsigned fill_round = 0;                  // (it + GSTAGES - 1) / GSTAGES
    for (long long it = 0; it < total; ++it) {
      cp_async_wait<GSTAGES - 2>();  // this lane's copies of chunk `it` have landed ...
      __syncwarp();                  // ... and every lane's
      fix_up(it, stage);
      mbar_arrive(smem_u32(&s_bar[stage]));  // release: the 64 arrivals publish the chunk to the MMA warps
      const long long nxt = it + GSTAGES - 1;
      if (nxt < total) {
        if (fill_round > 0) mbar_wait(smem_u32(&s_bar[GSTAGES + fill_stage]), (fill_round - 1u) & 1u);  // chunk it-1 is consumed
        pace(nxt);
        fill(nxt, fill_stage);
      }
      cp_async_commit();
      stage = stage == GSTAGES - 1 ? 0 : stage + 1;
      if (fill_stage == GSTAGES - 1) { fill_stage = 0; ++fill_round; } else { ++fill_stage; }
    }
    cp_async_wait<0>();
    return;
  }

  // ========================================================================================== MMA warps
  double acc[16][2];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i][0] = acc[i][1] = 0.0;
  const int wm = warp >> 1, wn = warp & 1;
  const bool corner = bi == 0;
  int s = 0;
  unsigned phase = 0;
  for (long long it = 0; it < total; ++it) {
    mbar_wait(smem_u32(&s_bar[s]), phase);
    const GramStage& S = st[s];
    if (!diag) {
      const int fr = lane >> 2, fk = lane & 3, sw = fr << 1;
#pragma unroll
      for (int k0 = 0; k0 < GKC; k0 += 4) {
        const int off = (((((k0 + fk) >> 1) ^ sw) << 1) | (fk & 1));
        double af[4], bf[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) af[i] = S.a[wm * 32 + i * 8 + fr][off];
#pragma unroll
        for (int j = 0; j < 4; ++j) bf[j] = S.b[wn * 32 + j * 8 + fr][off];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dmma884(acc[i * 4 + j], af[i], bf[j]);
      }
    } else {
      switch (warp) {
        case 0: diag_chunk_mma<0>(S, acc, lane, corner); break;
        case 1: diag_chunk_mma<1>(S, acc, lane, corner); break;
        case 2: diag_chunk_mma<2>(S, acc, lane, corner); break;
        default: diag_chunk_mma<3>(S, acc, lane, corner); break;
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&s_bar[GSTAGES + s]));
    if (s == GSTAGES - 1) { s = 0; phase ^= 1u; } else { ++s; }
  }

#ifdef DQ4ML_GRAM_PROFILE  // make PTXAS_V=-DDQ4ML_GRAM_PROFILE: slice 0 of every pair prints its run time (how kOffCost : kDiagCost were measured)
  if (a.debug && split == 0 && tid == 0)
    printf("gram_ws pair %d (%d,%d) slices %d iters %lld: %.3f ms\n", kind, bi, bj, my_splits, total, (global_timer_ns() - t_start) * 1e-6);
#endif
  // -------------------------------------------------------------------- partial tile -> scratch slot [pair][slice]
  double* P = g.partials + (static_cast<size_t>(kind) * a.s_max + split) * GPART;
  if (!diag) {
    const int fr = lane >> 2, fc = (lane & 3) * 2;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = wm * 32 + i * 8 + fr, c = wn * 32 + j * 8 + fc;
        *reinterpret_cast<double2*>(&P[r * GT + c]) = make_double2(acc[i * 4 + j][0], acc[i * 4 + j][1]);
      }
  } else {
    switch (warp) {
      case 0: diag_store<0>(P, acc, lane, corner); break;
      case 1: diag_store<1>(P, acc, lane, corner); break;
      case 2: diag_store<2>(P, acc, lane, corner); break;
      default: diag_store<3>(P, acc, lane, corner); break;
    }
  }
}

// Sums the row slices in a fixed order and lays the result out as
// [count, wSum, wwSum, bSum, bbSum, aSum[d], abSum[d], aaSum packed upper column-major].
// Partial slots: [pair][slice] with stride s_max.
__global__ void __launch_bounds__(256) gram_reduce_kernel(const double* __restrict__ partials, int d, int nb, int s_off,
                                                          int s_diag, int s_diag0, int s_max, long long len, int weighted,
                                                          double* __restrict__ out) {
  const long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= len) return;
  int bi, bj, off;
  if (e < 5) {
    bi = bj = 0;
    // corner = [sqrt(w) y, sqrt(w), w, keep]^T [same]: [3][3] = count, [1][1] = wSum, [2][2] = wwSum, [0][1] = bSum, [0][0] = bbSum
    // (unweighted: only column 1 = keep is written, and [1][1] serves count, wSum and wwSum)
    const int diag3 = weighted ? (e == 0 ? 3 * 8 + 3 : (e == 1 ? 1 * 8 + 1 : 2 * 8 + 2)) : 1 * 8 + 1;
    off = GT * GT + GT * 8 + (e < 3 ? diag3 : (e == 3 ? 0 * 8 + 1 : 0));
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
  const int splits = bi != bj ? s_off : (bi == 0 ? s_diag0 : s_diag);
  double s = 0.0;
  for (int sp = 0; sp < splits; ++sp) s += partials[(static_cast<size_t>(kind) * s_max + sp) * GPART + off];
  out[e] = s;
}

}  // namespace

int gram_wls_moments(dq4ml_ctx* ctx, const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes,
                     const double* label, const double* weight, const DeviceDq& dq, int64_t n, double* out) {
  const int d = static_cast<int>(col_ptrs.size());
  const int nb = (d + GT - 1) / GT;
  const int kinds = nb * (nb + 1) / 2;
  const int64_t len = 5 + 2LL * d + static_cast<int64_t>(d) * (d + 1) / 2;
  const int64_t full_chunks = n / GKC;
  // One wave exactly: two sub-CTAs per SM.  The slices are dealt out in proportion to the measured cost of a chunk
  // (kOffCost : kDiagCost : kDiag0Cost), so that every pair finishes at about the same time.
  static const char* pace_env = std::getenv("DQ4ML_GRAM_PACE");
  const int lead = pace_env != nullptr ? std::atoi(pace_env) : 0;
  int s_off = 0, s_diag = 0, s_diag0 = 1;
  {
    const int n_off = kinds - nb, n_diag = nb - 1, cap = 2 * ctx->sm_count;
    double best = 1e300;
    for (int so = n_off > 0 ? 1 : 0; so * n_off < cap; ++so) {
      for (int sd = n_diag > 0 ? 1 : 0; so * n_off + sd * n_diag < cap; ++sd) {
        const int sd0 = cap - so * n_off - sd * n_diag;
        double t = kDiag0Cost / sd0;
        if (n_off > 0) t = std::max(t, kOffCost / so);
        if (n_diag > 0) t = std::max(t, kDiagCost / sd);
        if (t < best) { best = t; s_off = so; s_diag = sd; s_diag0 = sd0; }
        if (n_diag == 0) break;
      }
      if (n_off == 0) break;
    }
    static const char* slices_env = std::getenv("DQ4ML_GRAM_SLICES");  // "s_off,s_diag,s_diag0": tuning only
    if (slices_env != nullptr) std::sscanf(slices_env, "%d,%d,%d", &s_off, &s_diag, &s_diag0);
    const int most = static_cast<int>(std::max<int64_t>(std::min<int64_t>(full_chunks, cap), 1));
    s_off = n_off > 0 ? std::max(std::min(s_off, most), 1) : 0;
    s_diag = n_diag > 0 ? std::max(std::min(s_diag, most), 1) : 0;
    s_diag0 = std::max(std::min(s_diag0, most), 1);
  }
  const int s_max = std::max(s_diag0, std::max(s_diag, s_off));
  const int64_t ws_blocks = static_cast<int64_t>(s_off) * (kinds - nb) + static_cast<int64_t>(s_diag) * (nb - 1) + s_diag0;
  const int64_t splits = s_max;  // sizes the partial scratch: kinds * s_max slots
  const size_t table_bytes = (static_cast<size_t>(d) * 9 + 255) & ~static_cast<size_t>(255);
  const size_t part_bytes = static_cast<size_t>(splits) * kinds * GPART * sizeof(double);
  const size_t out_bytes = static_cast<size_t>(len) * sizeof(double);
  const size_t need = table_bytes + part_bytes + out_bytes + 512 + 256 + static_cast<size_t>(kinds) * 8;  // + the pacing counters
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
  unsigned long long* d_progress = reinterpret_cast<unsigned long long*>(base + ((table_bytes + part_bytes + out_bytes + 255) & ~static_cast<size_t>(255)));
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
  g.label = label;
  g.weight = weight;
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
    if ((e = cudaFuncSetAttribute(gram_ws_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(2 * kGramWsSmem))) != cudaSuccess)
      return fail(e, "gram smem attribute");
    if ((e = cudaFuncSetAttribute(gram_ws_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(2 * kGramWsSmem))) != cudaSuccess)
      return fail(e, "gram smem attribute");
    attr_set[ctx->device] = true;
  }
  GramWsArgs wa;
  wa.g = g;
  wa.s_off = s_off;
  wa.s_diag = s_diag;
  wa.s_diag0 = s_diag0;
  wa.s_max = s_max;
  wa.progress = (lead > 0 && kinds > 1) ? d_progress : nullptr;
  wa.lead = lead;
  static const char* debug_env = std::getenv("DQ4ML_GRAM_DEBUG");
  wa.debug = debug_env != nullptr && debug_env[0] == '1' ? 1 : 0;
  if (wa.progress != nullptr) {
    if ((e = cudaMemsetAsync(d_progress, 0, static_cast<size_t>(kinds) * sizeof(unsigned long long), ctx->stream)) != cudaSuccess)
      return fail(e, "gram pacing counters");
  }
  if (ctx->pass_timing) cudaEventRecord(ctx->pass_ev0, ctx->stream);
  const unsigned ws_grid = static_cast<unsigned>((ws_blocks + 1) / 2);
  if (weight != nullptr) gram_ws_kernel<true><<<ws_grid, 2 * GWS_THREADS, 2 * kGramWsSmem, ctx->stream>>>(wa);
  else gram_ws_kernel<false><<<ws_grid, 2 * GWS_THREADS, 2 * kGramWsSmem, ctx->stream>>>(wa);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "gram tile kernel");
  ++ctx->launches;
  gram_reduce_kernel<<<static_cast<unsigned>((len + 255) / 256), 256, 0, ctx->stream>>>(
      d_part, d, nb, s_off, s_diag, s_diag0, s_max, len,
      weight != nullptr ? 1 : 0, d_out);
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
  collect_pass_time(ctx);
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
