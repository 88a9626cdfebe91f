// gram_tf32_kernels.cu -- the tensor-core tier of the normal-equation pass: Z^T Z on tcgen05 (5th-gen tensor cores,
// accumulators in TMEM) with a 3xTF32 split, for 1 <= d <= 256 feature columns.  Opt-in (dq4ml_ctx_set_gram_tier).
//
// What it computes is the WeightedLeastSquares.Aggregator result of gram_kernels.cu (lr.fit, App:126, solver normal /
// auto) at a relaxed tolerance.  Every centred, weighted value z = sqrt(w) keep (x - c), c = mean of the column's first
// 256 values, is rounded to fp32 and split into hi (11 leading bits, exactly a tf32) and lo = z - hi; the tensor core
// accumulates hi.hi + hi.lo + lo.hi in fp32 in TMEM, which reproduces each product to ~2^-21.  Three things keep the
// fp32 accumulation -- which rounds towards zero, i.e. loses half an ulp of the running sum per step, always in the
// same direction for a sum of one sign -- from showing in the regression:
//   - the accumulators are drained into an fp64 tile every `drain_rows` rows (2048 by default);
//   - centring makes the off-diagonal sums covariances, which hover around zero unless two columns are correlated
//     (measured error of a covariance relative to sigma_i sigma_j: 3e-7 at 2048 rows, growing linearly with the span and
//     with the correlation);
//   - the sums that are of one sign or carry the signal never touch the tensor core: sum z^2 (the diagonal),
//     sum w (x - c) (which the centres multiply when the raw moments are rebuilt) and sum w (x - c)(y - c_y) are summed
//     over 8 rows in fp32 with round-to-nearest and across chunks in fp64 by the threads that convert the values; the
//     five scalars (count, sum w, sum w^2, sum w y, sum w y^2) are fp64 throughout.
// Measured against the fp64 kernel on the config-4 workload (62.5 M x 256, fused DQ): coefficients agree to 5e-7
// relative, the intercept to 1.4e-6.  Stated tolerance of the tier: 1e-5 (north_star: "1e-5 where fp32 accumulation is
// used and justified"); the fp64 DMMA kernel stays the default.
//
// Shape of the kernel: one persistent CTA per SM (all 512 TMEM columns, ~217 KB of shared memory), 18 warps:
//   - 16 loader / converter warps.  Rows are taken in chunks of 16.  A 4-stage cp.async ring stages the raw fp64 / int32
//     pieces of the columns; every warp fetches the 16 features it converts (8 neighbouring lanes fetch one 128-byte
//     line), so a __syncwarp is all that stands between the copies and their use (a CTA-wide barrier there measured
//     the same: the warps are paced by the mbarriers below, not by each other).  Two threads per feature convert 8
//     rows each and write the hi and lo operands in the canonical no-swizzle K-major UMMA
//     layout (8-row x 16-byte core matrices; LBO = 128 B between the K chunks, SBO = 512 B between 8-feature groups),
//     then fence.proxy.async and one mbarrier arrival per warp;
//   - the MMA warp: one thread waits for the operands of a chunk and issues, per 8 rows and per split term,
//     M128 x N256 (features 0..127 against all 256) and M128 x N128 (features 128..255 against themselves) of
//     kind::tf32; tcgen05.commit -> mbarrier frees the operand buffer (two buffers) and announces a finished span;
//   - the row warp: its own 16-deep cp.async ring for label / DQ price / DQ guest / weight, the keep mask of the DQ
//     rules, the per-row factors for the converters (eight buffers ahead), the five scalar sums;
//   - the drain: every warp reads its TMEM lane quadrant with tcgen05.ld.32x32b.x32 and adds it to the CTA's fp64 tile in
//     L2-resident scratch, stored column-major so that the 32 lanes read-modify-write 256 contiguous bytes; column
//     groups below the diagonal of a diagonal block are skipped.
// A small kernel pair folds the per-CTA tiles in a fixed order and rebuilds the public moment layout.
//
// Where the time goes (62.5 M x 256: 48.5 ms = 0.40 of the HBM roofline, tensor pipe 40 % busy, vs 170 ms for the fp64
// kernel): the converter warps.  Each executes ~200 instructions per chunk (conversion ~110 with the packed f32x2
// arithmetic, synchronisation and load issue ~90) as a mostly serial stream, four warps per scheduler; ablations
// (DQ4ML_TF32_EXPERIMENT) put loads alone at 30 ms, conversion alone at 32 ms, MMAs alone at 37 ms, and the bare
// synchronisation skeleton at 20 ms -- a lone warp pays ~8 cycles per dependent instruction, so the row warp's ~200
// instructions per chunk bound the skeleton.  History: 83 ms (one role per thread, 256 threads, row-major drain)
// -> 66 (MMA warp, column-major drain) -> 61 (512 converter threads, row warp) -> 50 (sums off the tensor core,
// own row ring) -> 48.5 (packed arithmetic).  Per chunk a converter warp spends ~390 cycles on load issue, ~380 waiting
// for the row factors / a free operand buffer, ~1700 in the conversion and ~270 (amortised) in the drain.  Tried and
// dropped: fp64 <-> fp32 conversion with integer instructions instead of F2F (55.7 ms: the phase follows its instruction
// count, not the conversion pipe); three operand buffers with three staging stages (same time); draining the CTAs at
// different chunks instead of all at once (same time: the drain is not an L2 burst problem).  Next: two chunks per row-warp iteration, a TMA tensor-map load path
// for uniformly strided tables (no per-thread load issue), 32-row chunks once the raw staging is gone.
#include <cuda.h>  // CUtensorMap and its enums only: the encoder is looked up at run time, libcuda is not linked

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "context.hpp"
#include "device_utils.cuh"
#include "gram.hpp"

namespace dq4ml {
namespace {

constexpr int TF = 256;                     // feature slots per CTA
constexpr int TKC = 16;                     // rows per chunk
constexpr int TSTAGES = 4;                  // staging stages (raw columns)
constexpr int TROWST = 16;                  // stages of the row warp's own ring (its loads are latency-bound: depth is what feeds it)
constexpr int TCONV = 512;                  // loader / converter threads: TCONV / TF per feature slot (256: 53 ms against 48.5)
constexpr int TFPW = 32 * TF / TCONV;       // features per converter warp
constexpr int TRPT = TKC * TF / TCONV;      // rows of a chunk per converter thread
constexpr int TTHREADS = TCONV + 64;        // + the MMA warp + the row warp (the TMA variant adds a producer warp)
constexpr int TOPS = 2;                     // operand stages (hi/lo tf32)
constexpr int TSCALE = 8;                   // row-factor buffers: how far the row warp may run ahead of the converters
constexpr int TSTRIDE = TKC * 8 + 16;       // bytes per feature in a staging stage: +16 keeps 128-bit reads conflict-free
constexpr int TLBO = 128;                   // bytes between the two 16-byte K chunks of one MMA (core matrix = 8 x 16 B)
constexpr int TSBO = (TKC / 4) * 128;       // bytes between groups of 8 feature rows
constexpr int TOPB = TF * TKC * 4;          // bytes of one operand buffer (hi or lo)
constexpr int TROWB = 512;                  // per row stage: y[16], dq price[16], weight[16], dq guest[16]
constexpr int TSCB = 160;                   // per row-factor buffer: float s[16], float s (y - c_y)[16], flag
constexpr int kStagingOff = 0;
constexpr int kOpsOff = kStagingOff + TSTAGES * TF * TSTRIDE;
constexpr int kRowsOff = kOpsOff + TOPS * 2 * TOPB;
constexpr int kScaleOff = kRowsOff + TROWST * TROWB;
constexpr int kBarOff = kScaleOff + TSCALE * TSCB;  // mbarriers + the TMEM base address
constexpr int kNumBars = 2 * TOPS + 2 * TSCALE + 1 + 2 * TSTAGES;  // free/ready, scale/sfree, drain, TMA full/empty
constexpr int kTf32Smem = kBarOff + 8 * (kNumBars + 1) + 1024;      // + slack to align the staging ring to 1024 bytes
// TMA staging (uniformly strided fp64 columns): one 2-D box of TKC rows x TF columns per stage, 128-byte swizzle.  It
// reuses the staging region: TSTAGES boxes of 32 KB, then the pieces of the (at most TFPW) columns that are not part of
// the strided table -- the lab's int32 guest column in front of the fp64 features -- in the cp.async layout.
constexpr int kTmaBoxBytes = TF * TKC * 8;                 // 32 768
constexpr int kTmaExtraOff = TSTAGES * kTmaBoxBytes;       // [TSTAGES][TFPW * TSTRIDE]
static_assert(kTmaExtraOff + TSTAGES * TFPW * TSTRIDE <= TSTAGES * TF * TSTRIDE, "TMA staging must fit the staging region");
// TMEM accumulator columns: [0, 256) = features 0..127 x features 0..255, [256, 384) = features 128..255 x features 128..255
constexpr int kColC = 256;
constexpr int SCOLS = 384;                  // fp64 tile columns per CTA (12 groups of 32), stored column-major
constexpr int kDiagOff = 128 * SCOLS;       // sum z^2 per feature
constexpr int kA1Off = kDiagOff + TF;       // sum w (x - c) per feature
constexpr int kAbOff = kA1Off + TF;         // sum w (x - c)(y - c_y) per feature
constexpr int kScalarOff = kAbOff + TF;     // count, sum w, sum w^2, sum w y, sum w y^2
constexpr int kCtaDoubles = kScalarOff + 8;

struct Tf32Args {
  const void* const* col_ptr;
  const unsigned char* col_i32;
  const double* centres;  // [d + 1]: features, then the label
  int d;
  const double* label;
  const double* weight;
  const int* dq_guest;
  const double* dq_price;
  DqThresholds t;
  int has_dq;
  long long n;
  int drain_chunks;
  int swap_offsets;  // debug: exchange LBO and SBO in the descriptors
  int experiment;    // ablation: bit 0 = converters skip the conversion, bit 1 = no MMAs are issued, bit 2 = no column loads, bit 3 = timers (profile build)
  double* partials;  // [grid][kCtaDoubles], zeroed by the launcher
  int tma_first;     // TMA variant: feature slots [tma_first, d) are columns 0.. of the tensor map; slots below it use cp.async
};

__device__ __forceinline__ unsigned s_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void cpa16(unsigned dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cpa_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cpa_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned ok = 0;
  for (unsigned spin = 0; ok == 0; ++spin) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (spin > (1u << 24)) __trap();  // a lost completion must not hang the GPU
  }
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// One TMA tile load: box (TKC rows, TF columns) of the column table at (row, col 0) -> shared memory, completion on `bar`.
// The tensor map covers this GPU's own buffers only.
__device__ __forceinline__ void tma_load_2d(unsigned dst, const CUtensorMap* tmap, int c0, int c1, unsigned bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(dst), "l"(reinterpret_cast<unsigned long long>(tmap)), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(unsigned tmem_d, unsigned long long da, unsigned long long db, unsigned idesc,
                                            unsigned accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread l receives lane (addr.lane + l), columns addr.col .. +31
__device__ __forceinline__ void tc_ld32(unsigned taddr, unsigned (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
      "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor, K-major, no swizzle: 8-row x 16-byte core matrices; `lbo` = byte distance between the
// two K chunks of one MMA, `sbo` = byte distance between 8-row groups (cute::UMMA::SmemDescriptor, version 1).
__device__ __forceinline__ unsigned long long umma_desc(unsigned saddr, unsigned lbo, unsigned sbo) {
  unsigned long long d = static_cast<unsigned long long>((saddr >> 4) & 0x3fffu);
  d |= static_cast<unsigned long long>((lbo >> 4) & 0x3fffu) << 16;
  d |= static_cast<unsigned long long>((sbo >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;
  return d;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D fp32, A/B tf32, both K-major, M = 128.
__host__ __device__ constexpr unsigned umma_idesc_tf32(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<unsigned>(n >> 3) << 17) | (static_cast<unsigned>(128 >> 4) << 24);
}

// One fp32 value -> the two tf32 operands (Dekker's split with 2^13 + 1, see split2 below): hi carries the 11 leading
// bits of f, rounded to nearest, so the tensor core -- which ignores the 13 low mantissa bits -- sees exactly hi;
// lo = f - hi is exact in fp32, and its own truncation by the tensor core is a 2^-21 relative effect of random sign.
// NaN / Inf stay NaN.
// packed fp32 pairs (FADD2 / FMUL2 / FFMA2 on sm_100): half the instruction count for the split and the three sums
using f32x2 = unsigned long long;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
// split_f32 on a pair.  The product with 2^13 + 1 is taken with scalar __fmul_rn: ptxas contracts a packed
// mul.rn.f32x2 followed by sub.rn.f32x2 into one FFMA2, which turns the split into hi = f, lo = 0.
__device__ __forceinline__ void split2(float f0, float f1, f32x2& h, f32x2& l) {
  const f32x2 f = pack2(f0, f1);
  const f32x2 t = pack2(__fmul_rn(f0, 8193.0f), __fmul_rn(f1, 8193.0f));
  h = sub2(t, sub2(t, f));
  l = sub2(f, h);
}

#ifdef DQ4ML_TF32_PROFILE
#define TICK(slot)                                   \
  do {                                               \
    if (prof) {                                      \
      const long long now_ = clock64();              \
      tacc[slot] += now_ - tlast;                    \
      tlast = now_;                                  \
    }                                                \
  } while (0)
#else
#define TICK(slot) \
  do {             \
  } while (0)
#endif

// does a warp of lane quadrant `quad` have to drain 32-column group cg?  Groups that lie entirely below the diagonal of
// a diagonal block are skipped.
__device__ __forceinline__ bool drain_needed(int cg, int quad) {
  if (cg < 4) return cg >= quad;
  if (cg < 8) return true;
  return cg - 8 >= quad;
}

template <bool TMA>
__global__ void __launch_bounds__(TTHREADS + (TMA ? 32 : 0), 1)
    gram_tf32_kernel(const Tf32Args g, const __grid_constant__ CUtensorMap tmap) {
#ifdef DQ4ML_TF32_PROFILE
  const bool prof = (g.experiment & 8) != 0 && blockIdx.x == 0 && (threadIdx.x & 31) == 0;
  long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long tlast = clock64();
#endif
  extern __shared__ __align__(128) unsigned char smem_raw[];
  // the 128-byte swizzle of the TMA boxes is a function of the shared-memory address: the ring starts on a 1024-byte line
  unsigned char* smem = smem_raw + ((1024u - (s_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int kThreads = TTHREADS + (TMA ? 32 : 0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned sbase = s_u32(smem);
  const unsigned bar_free0 = sbase + kBarOff;      // [TOPS] operand buffer o may be overwritten: arrived by tcgen05.commit
  const unsigned bar_ready0 = bar_free0 + 8 * TOPS;       // [TOPS] operand buffer o is written: one arrival per converter warp
  const unsigned bar_scale0 = bar_ready0 + 8 * TOPS;      // [TSCALE] row factors of a chunk are written: the row warp
  const unsigned bar_sfree0 = bar_scale0 + 8 * TSCALE;    // [TSCALE] row factors of a chunk have been used: one arrival per converter warp
  const unsigned bar_drain = bar_sfree0 + 8 * TSCALE;     // every MMA issued so far has completed
  const unsigned bar_full0 = bar_drain + 8;               // [TSTAGES] TMA box of stage s has landed (TMA variant)
  const unsigned bar_empty0 = bar_full0 + 8 * TSTAGES;    // [TSTAGES] every converter warp has read stage s
  unsigned* tmem_slot = reinterpret_cast<unsigned*>(smem + kBarOff + 8 * kNumBars);
  const int nb = g.d > 128 ? 2 : 1;

  // operand buffers start as zeros: feature slots >= d are never written again
  for (int i = tid; i < TOPS * 2 * TOPB / 16; i += kThreads) reinterpret_cast<int4*>(smem + kOpsOff)[i] = make_int4(0, 0, 0, 0);
  if (tid == 0) {
    for (int o = 0; o < TOPS; ++o) {
      mbar_init(bar_free0 + 8 * o, 1);
      mbar_init(bar_ready0 + 8 * o, TCONV / 32);
    }
    for (int b = 0; b < TSCALE; ++b) {
      mbar_init(bar_scale0 + 8 * b, 1);
      mbar_init(bar_sfree0 + 8 * b, TCONV / 32);
    }
    mbar_init(bar_drain, 1);
    for (int st = 0; st < TSTAGES; ++st) {
      mbar_init(bar_full0 + 8 * st, 1);
      mbar_init(bar_empty0 + 8 * st, TCONV / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {  // all 512 TMEM columns: this kernel is alone on its SM (shared memory sees to that)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32(tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the zeros above, for the tensor core's reads
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const unsigned tmem = *tmem_slot;

  const long long total_chunks = (g.n + TKC - 1) / TKC, full_chunks = g.n / TKC;
  const long long my_iters =
      total_chunks > blockIdx.x ? (total_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  double* P = g.partials + static_cast<size_t>(blockIdx.x) * kCtaDoubles;

  if (warp == TCONV / 32) {
    // ================================================= MMA warp: one thread feeds the tensor core ====================
    if (lane == 0) {
      const unsigned lbo = g.swap_offsets ? TSBO : TLBO, sbo = g.swap_offsets ? TLBO : TSBO;
      const unsigned idesc_row0 = umma_idesc_tf32(nb == 2 ? 256 : 128), idesc_blk = umma_idesc_tf32(128);
      int since_drain = 0;
      for (unsigned it = 0; it < my_iters; ++it) {
        const int o = static_cast<int>(it % TOPS);
        TICK(0);
        mbar_wait(bar_ready0 + 8 * o, static_cast<unsigned>((it / TOPS) & 1));
        TICK(1);
        tc_fence_after();
        const unsigned hi = sbase + kOpsOff + o * (2 * TOPB), lo = hi + TOPB;
#pragma unroll
        for (int ks = 0; ks < ((g.experiment & 2) ? 0 : TKC / 8); ++ks) {
          const unsigned koff = ks * 2 * TLBO;
#pragma unroll
          for (int term = 0; term < 3; ++term) {  // hi.hi, hi.lo, lo.hi
            const unsigned a = (term == 2 ? lo : hi) + koff, b = (term == 1 ? lo : hi) + koff;
            const unsigned acc = (since_drain == 0 && ks == 0 && term == 0) ? 0u : 1u;
            tc_mma_tf32(tmem, umma_desc(a, lbo, sbo), umma_desc(b, lbo, sbo), idesc_row0, acc);
            if (nb == 2)
              tc_mma_tf32(tmem + kColC, umma_desc(a + 16 * TSBO, lbo, sbo), umma_desc(b + 16 * TSBO, lbo, sbo), idesc_blk, acc);
          }
        }
        tc_commit(bar_free0 + 8 * o);
        if (++since_drain == g.drain_chunks || it == my_iters - 1) {
          tc_commit(bar_drain);
          since_drain = 0;
        }
        TICK(2);
      }
#ifdef DQ4ML_TF32_PROFILE
      if (prof) printf("mma thread: loop %lld wait_ready %lld issue+commit %lld (cycles, %lld chunks)\n", tacc[0], tacc[1], tacc[2], my_iters);
#endif
    }
  } else if (warp == TCONV / 32 + 1) {
    // ================================================= row warp ========================================================
    // Owns the per-row columns (label, DQ price / guest, weight): its own cp.async ring, the keep mask of the DQ rules,
    // the per-row factors s = sqrt(w) keep and s (y - c_y) the converters multiply with, and the five scalar sums.  It
    // runs up to TSCALE chunks ahead of the converters.
    auto issue_rows = [&](unsigned it) {
      const long long row = (blockIdx.x + it * static_cast<long long>(gridDim.x)) * TKC;
      unsigned char* rowb = smem + kRowsOff + static_cast<int>(it % TROWST) * TROWB;
      if (row + TKC <= g.n) {
        const unsigned rb = s_u32(rowb);
        if (lane < 8) cpa16(rb + lane * 16, g.label + row + lane * 2);
        else if (lane < 16) { if (g.has_dq) cpa16(rb + 128 + (lane - 8) * 16, g.dq_price + row + (lane - 8) * 2); }
        else if (lane < 24) { if (g.weight != nullptr) cpa16(rb + 256 + (lane - 16) * 16, g.weight + row + (lane - 16) * 2); }
        else if (lane < 28) { if (g.has_dq) cpa16(rb + 384 + (lane - 24) * 16, g.dq_guest + row + (lane - 24) * 4); }
      } else if (lane < TKC) {  // the ragged last chunk
        const bool in = row + lane < g.n;
        reinterpret_cast<double*>(rowb)[lane] = in ? g.label[row + lane] : 0.0;
        reinterpret_cast<double*>(rowb + 128)[lane] = (in && g.has_dq) ? g.dq_price[row + lane] : 0.0;
        reinterpret_cast<double*>(rowb + 256)[lane] = (in && g.weight != nullptr) ? g.weight[row + lane] : 0.0;
        reinterpret_cast<int*>(rowb + 384)[lane] = (in && g.has_dq) ? g.dq_guest[row + lane] : 0;
      }
    };
    for (int s = 0; s < TROWST - 1; ++s) {
      if (s < my_iters) issue_rows(s);
      cpa_commit();
    }
    double cnt = 0.0, wsum = 0.0, wwsum = 0.0, bsum = 0.0, bbsum = 0.0;
    const double ycentre = g.centres[g.d];
    for (unsigned it = 0; it < my_iters; ++it) {
      TICK(0);
      cpa_wait<TROWST - 2>();
      __syncwarp();
      if (it + TROWST - 1 < my_iters) issue_rows(it + TROWST - 1);
      cpa_commit();
      TICK(1);
      const unsigned char* rowb = smem + kRowsOff + static_cast<int>(it % TROWST) * TROWB;
      const int b = static_cast<int>(it % TSCALE);
      double s = 0.0, sy = 0.0;
      bool plain = true;
      if (lane < TKC) {
        const long long row = (blockIdx.x + it * static_cast<long long>(gridDim.x)) * TKC + lane;
        bool keep = row < g.n;
        if (keep && g.has_dq)
          keep = dq_keep(reinterpret_cast<const int*>(rowb + 384)[lane], reinterpret_cast<const double*>(rowb + 128)[lane], g.t);
        const double w = keep ? (g.weight != nullptr ? reinterpret_cast<const double*>(rowb + 256)[lane] : 1.0) : 0.0;
        s = g.weight != nullptr ? sqrt(w) : w;
        const double y = keep ? reinterpret_cast<const double*>(rowb)[lane] : 0.0;
        sy = keep ? s * (y - ycentre) : 0.0;
        plain = s == 1.0;
        cnt += keep ? 1.0 : 0.0;
        wsum += s * s;  // = w, and NaN for a negative weight (require(w >= 0))
        wwsum = fma(w, w, wwsum);
        bsum = fma(w, y, bsum);
        bbsum = fma(w * y, y, bbsum);
      }
      const bool all_plain = __all_sync(0xffffffffu, plain);
      if (it >= TSCALE) mbar_wait(bar_sfree0 + 8 * b, static_cast<unsigned>(((it / TSCALE) - 1) & 1));  // the converters are done with chunk it-3
      TICK(2);
      float* scale = reinterpret_cast<float*>(smem + kScaleOff + b * TSCB);
      if (lane < TKC) {
        scale[lane] = __double2float_rn(s);
        scale[TKC + lane] = __double2float_rn(sy);
      }
      if (lane == 0) reinterpret_cast<int*>(scale + 2 * TKC)[0] = all_plain ? 1 : 0;  // every factor is 1: the converters skip the scaling
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_scale0 + 8 * static_cast<unsigned>(b));
      TICK(3);
    }
    cpa_wait<0>();
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) {
      cnt += __shfl_down_sync(0xffffffffu, cnt, off);
      wsum += __shfl_down_sync(0xffffffffu, wsum, off);
      wwsum += __shfl_down_sync(0xffffffffu, wwsum, off);
      bsum += __shfl_down_sync(0xffffffffu, bsum, off);
      bbsum += __shfl_down_sync(0xffffffffu, bbsum, off);
    }
    if (lane == 0) {
      double* sc = P + kScalarOff;
      sc[0] = cnt; sc[1] = wsum; sc[2] = wwsum; sc[3] = bsum; sc[4] = bbsum;
    }
#ifdef DQ4ML_TF32_PROFILE
    if (prof) printf("row warp: loop %lld loads %lld compute+wait_sfree %lld write+arrive %lld\n", tacc[0], tacc[1], tacc[2], tacc[3]);
#endif
  } else if (TMA && warp == TCONV / 32 + 2) {
    // ================================================= TMA producer warp ===============================================
    // One thread keeps TSTAGES boxes of the strided fp64 table in flight: 16 rows x 256 columns = 32 KB per request,
    // no per-thread load issue in the converter warps.  Rows past the end of the table are zero-filled by the TMA unit.
    if (lane == 0) {
      for (unsigned it = 0; it < my_iters; ++it) {
        const int st = static_cast<int>(it % TSTAGES);
        if (it >= TSTAGES) mbar_wait(bar_empty0 + 8 * st, static_cast<unsigned>(((it / TSTAGES) - 1) & 1));
        const long long row = (blockIdx.x + it * static_cast<long long>(gridDim.x)) * TKC;
        mbar_arrive_expect_tx(bar_full0 + 8 * st, kTmaBoxBytes);
        if (!(g.experiment & 4)) tma_load_2d(sbase + kStagingOff + st * kTmaBoxBytes, &tmap, static_cast<int>(row), 0, bar_full0 + 8 * st);
        else asm volatile("mbarrier.complete_tx.shared::cta.b64 [%0], %1;" ::"r"(bar_full0 + 8 * st), "r"(static_cast<unsigned>(kTmaBoxBytes)) : "memory");
      }
    }
  } else {
    // ================================================= loader / converter warps ========================================
    // converter: two threads per feature, 8 rows of every chunk each (lanes 0..15 / 16..31 of a warp share 16 features)
    const int feat = warp * TFPW + (lane % TFPW), half = lane / TFPW;
    const bool active = feat < g.d;
    const bool own_int = active && g.col_i32[feat] != 0;
    const double centre = active ? g.centres[feat] : 0.0;
    // loader: every warp fetches exactly the 16 features it converts (so that a __syncwarp, not a CTA-wide barrier,
    // stands between the copies and their use, and the warps may drift apart by as much as the mbarriers allow):
    // 16-byte piece `ch` of the column slots 16 warp + (lane >> 3) + 4 i; 8 neighbouring lanes fetch one 128-byte line
    const int ch = lane & 7, slot0 = warp * TFPW + (lane >> 3);
    const char* src[TFPW / 4];
    unsigned live = 0, ints = 0;
#pragma unroll
    for (int i = 0; i < TFPW / 4; ++i) {
      const int slot = slot0 + 4 * i;
      const bool on = slot < g.d;
      const char* base = on ? static_cast<const char*>(g.col_ptr[slot]) : nullptr;
      const bool is_int = on && g.col_i32[slot] != 0;
      src[i] = base != nullptr ? base + ch * 16 : nullptr;
      if (base != nullptr && (!is_int || ch < 4) && (!TMA || slot < g.tma_first)) live |= 1u << i;  // TMA: the table's columns come as boxes
      if (is_int) ints |= 1u << i;
    }
    const unsigned piece_dst = static_cast<unsigned>(slot0 * TSTRIDE + ch * 16);
    // where the cp.async pieces of stage st start: the whole staging stage, or (TMA) the small region behind the boxes
    auto piece_stage = [&](int st) { return TMA ? kTmaExtraOff + st * (TFPW * TSTRIDE) : kStagingOff + st * (TF * TSTRIDE); };
    const bool by_tma = TMA && feat >= g.tma_first;  // this thread's feature arrives in the TMA box, row feat - tma_first

    auto issue_stage = [&](unsigned it) {
      const long long chunk = blockIdx.x + it * static_cast<long long>(gridDim.x);
      const long long row = chunk * TKC;
      const int st = static_cast<int>(it % TSTAGES);
      if (chunk >= full_chunks) {  // the ragged last chunk: guarded plain loads, same raw formats
        const int valid = static_cast<int>(g.n - row);
        if (half == 0 && active && !by_tma) {  // lanes 0..15 fill the rows of the warp's own 16 features
          unsigned char* dst = smem + piece_stage(st) + feat * TSTRIDE;
          for (int k = 0; k < TKC; ++k) {
            if (own_int) reinterpret_cast<int*>(dst)[k] = k < valid ? static_cast<const int*>(g.col_ptr[feat])[row + k] : 0;
            else reinterpret_cast<double*>(dst)[k] = k < valid ? static_cast<const double*>(g.col_ptr[feat])[row + k] : 0.0;
          }
        }
        return;
      }
      const unsigned stg = sbase + piece_stage(st) + piece_dst;
#pragma unroll
      for (int i = 0; i < TFPW / 4; ++i) {
        if (((live >> i) & 1u) && !(g.experiment & 4)) cpa16(stg + 4 * i * TSTRIDE, src[i] + (((ints >> i) & 1u) ? row * 4 : row * 8));
      }
    };

    // TMEM accumulators -> the CTA's fp64 tile, stored column-major (P[col * 128 + row]) so that the 32 lanes of a warp,
    // which hold 32 consecutive rows of one column, read-modify-write 256 contiguous bytes.
    auto drain = [&]() {
      tc_fence_after();
      const int quad = warp & 3, part = warp >> 2;
      const int groups = nb == 2 ? SCOLS / 32 : 4;
      const int row = quad * 32 + lane;
      for (int cg = part; cg < groups; cg += TCONV / 128) {
        if (!drain_needed(cg, quad)) continue;
        unsigned r[32];
        tc_ld32(tmem + (static_cast<unsigned>(quad * 32) << 16) + static_cast<unsigned>(cg * 32), r);
        double* dst = P + static_cast<size_t>(cg * 32) * 128 + row;
#pragma unroll
        for (int k = 0; k < 32; ++k) dst[k * 128] += static_cast<double>(__uint_as_float(r[k]));
      }
      tc_fence_before();
    };

    for (int s = 0; s < TSTAGES - 1; ++s) {
      if (s < my_iters) issue_stage(s);
      cpa_commit();
    }

    // sum z^2 (the diagonal of the Gram), sum s z = sum w (x - c) and sum z s (y - c_y) = sum w (x - c)(y - c_y) over this
    // thread's rows of its feature stay off the tensor core: they are sums of one sign, or get multiplied by the centres
    // afterwards, or carry the signal of the regression, and fp32 accumulation that rounds towards zero biases all three.
    // Per chunk they are summed over 8 rows in fp32 (round to nearest), across chunks in fp64.
    double zz = 0.0, a1 = 0.0, ab = 0.0;
    int since_drain = 0;
    unsigned drains = 0;
    const unsigned char* sp0 = smem + feat * TSTRIDE + half * (own_int ? TRPT * 4 : TRPT * 8);  // + piece_stage(st)
    // TMA box, 128-byte swizzle: row r = one feature's 16 values, its 16-byte piece c sits at r * 128 + ((c ^ (r & 7)) << 4)
    const int tma_row = feat - g.tma_first;
    unsigned char* op0 = smem + kOpsOff + (feat >> 3) * TSBO + (feat & 7) * 16 + half * (TRPT / 4) * TLBO;

    for (unsigned it = 0; it < my_iters; ++it) {
      TICK(0);
      cpa_wait<TSTAGES - 2>();
      __syncwarp();  // chunk `it` has landed for this warp's features; the stage chunk it+3 goes into was read by this warp only
      if (it + TSTAGES - 1 < my_iters) issue_stage(it + TSTAGES - 1);
      cpa_commit();
      TICK(1);

      const int st = static_cast<int>(it % TSTAGES);
      const int o = static_cast<int>(it % TOPS);
      const int b = static_cast<int>(it % TSCALE);
      if (TMA) mbar_wait(bar_full0 + 8 * st, static_cast<unsigned>((it / TSTAGES) & 1));        // the box of this chunk has landed
      mbar_wait(bar_scale0 + 8 * b, static_cast<unsigned>((it / TSCALE) & 1));                 // the row warp has written this chunk's factors
      if (it >= TOPS) mbar_wait(bar_free0 + 8 * o, static_cast<unsigned>(((it / TOPS) - 1) & 1));  // the MMAs of chunk it-2 have read this buffer
      TICK(2);

      if (active && !(g.experiment & 1)) {
        const float* scale = reinterpret_cast<const float*>(smem + kScaleOff + b * TSCB);
        const unsigned char* sp = sp0 + piece_stage(st);
        const unsigned char* box = smem + kStagingOff + st * kTmaBoxBytes + tma_row * 128;
        unsigned char* ohi = op0 + o * (2 * TOPB);
        unsigned char* olo = ohi + TOPB;
        const bool plain = reinterpret_cast<const int*>(scale + 2 * TKC)[0] != 0;
        f32x2 pzz = 0ull, pa1 = 0ull, pab = 0ull;  // (0.f, 0.f)
#pragma unroll
        for (int q = 0; q < TRPT / 4; ++q) {
          float z[4];
          if (own_int) {
            const int4 v = *reinterpret_cast<const int4*>(sp + q * 16);
            z[0] = __double2float_rn(static_cast<double>(v.x) - centre);
            z[1] = __double2float_rn(static_cast<double>(v.y) - centre);
            z[2] = __double2float_rn(static_cast<double>(v.z) - centre);
            z[3] = __double2float_rn(static_cast<double>(v.w) - centre);
          } else {
            double2 u, v;
            if (by_tma) {
              const int c = half * (TRPT / 2) + 2 * q, x = tma_row & 7;
              u = *reinterpret_cast<const double2*>(box + ((c ^ x) << 4));
              v = *reinterpret_cast<const double2*>(box + (((c + 1) ^ x) << 4));
            } else {
              u = *reinterpret_cast<const double2*>(sp + q * 32);
              v = *reinterpret_cast<const double2*>(sp + q * 32 + 16);
            }
            z[0] = __double2float_rn(u.x - centre);
            z[1] = __double2float_rn(u.y - centre);
            z[2] = __double2float_rn(v.x - centre);
            z[3] = __double2float_rn(v.y - centre);
          }
          const ulonglong2 yc = *reinterpret_cast<const ulonglong2*>(scale + TKC + half * TRPT + 4 * q);  // s (y - c_y) of the 4 rows
          f32x2 z01, z23;
          if (plain) {
            z01 = pack2(z[0], z[1]);
            z23 = pack2(z[2], z[3]);
            pa1 = add2(pa1, add2(z01, z23));
          } else {  // uniform over the CTA: a dropped row or a weight in this chunk
            const float4 sc = *reinterpret_cast<const float4*>(scale + half * TRPT + 4 * q);
            // dropped rows may hold NaN / Inf payload: they become exact zeros
            z[0] = sc.x != 0.f ? sc.x * z[0] : 0.f;
            z[1] = sc.y != 0.f ? sc.y * z[1] : 0.f;
            z[2] = sc.z != 0.f ? sc.z * z[2] : 0.f;
            z[3] = sc.w != 0.f ? sc.w * z[3] : 0.f;
            z01 = pack2(z[0], z[1]);
            z23 = pack2(z[2], z[3]);
            pa1 = fma2(pack2(sc.x, sc.y), z01, pa1);
            pa1 = fma2(pack2(sc.z, sc.w), z23, pa1);
          }
          pzz = fma2(z01, z01, pzz);
          pzz = fma2(z23, z23, pzz);
          pab = fma2(z01, yc.x, pab);
          pab = fma2(z23, yc.y, pab);
          f32x2 h01, l01, h23, l23;
          split2(z[0], z[1], h01, l01);
          split2(z[2], z[3], h23, l23);
          *reinterpret_cast<ulonglong2*>(ohi + q * TLBO) = make_ulonglong2(h01, h23);
          *reinterpret_cast<ulonglong2*>(olo + q * TLBO) = make_ulonglong2(l01, l23);
        }
        float e0, e1;
        unpack2(pzz, e0, e1);
        zz += static_cast<double>(e0 + e1);
        unpack2(pa1, e0, e1);
        a1 += static_cast<double>(e0 + e1);
        unpack2(pab, e0, e1);
        ab += static_cast<double>(e0 + e1);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the tensor core's reads
      __syncwarp();
      if (lane == 0) {  // one arrival per warp
        mbar_arrive(bar_ready0 + 8 * o);
        mbar_arrive(bar_sfree0 + 8 * b);
        if (TMA) mbar_arrive(bar_empty0 + 8 * st);  // this warp has read its rows of the box
      }
      TICK(3);

      if (++since_drain == g.drain_chunks || it == my_iters - 1) {
        mbar_wait(bar_drain, drains & 1u);
        ++drains;
        drain();
        since_drain = 0;
      }
      TICK(4);
    }
#ifdef DQ4ML_TF32_PROFILE
    if (prof && (warp == 0 || warp == 9)) printf("converter warp %d: loop %lld cpwait+barrier+issue %lld wait_scale+free %lld convert+arrive %lld drain %lld\n", warp, tacc[0], tacc[1], tacc[2], tacc[3], tacc[4]);
#endif
    cpa_wait<0>();
    if (active) {  // the two halves of a feature add up in either order to the same value
      atomicAdd(&P[kDiagOff + feat], zz);
      atomicAdd(&P[kA1Off + feat], a1);
      atomicAdd(&P[kAbOff + feat], ab);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

// centre of column j = mean of its first min(n, 256) values (0 when that is not finite): one warp per column
__global__ void tf32_centres_kernel(const void* const* col_ptr, const unsigned char* col_i32, const double* label, int d,
                                    long long n, double* centres) {
  const int j = blockIdx.x, lane = threadIdx.x;  // j == d: the label
  if (j > d) return;
  const long long m = n < 256 ? n : 256;
  double s = 0.0;
  for (long long r = lane; r < m; r += 32)
    s += j == d ? label[r]
                : (col_i32[j] ? static_cast<double>(static_cast<const int*>(col_ptr[j])[r]) : static_cast<const double*>(col_ptr[j])[r]);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
  if (lane == 0) {
    const double c = m > 0 ? s / static_cast<double>(m) : 0.0;
    centres[j] = (c - c == 0.0) ? c : 0.0;
  }
}

// where the tile keeps sum z_i z_j (i < j)
__device__ __forceinline__ size_t tile_off(long long i, long long j) {
  if (i < 128) return static_cast<size_t>(j) * 128 + i;             // columns [0, 256): column j, row i
  return static_cast<size_t>(kColC + (j - 128)) * 128 + (i - 128);  // columns [256, 384)
}

// scalars and the two border vectors, folded over the CTAs in a fixed order; tot = [5 scalars, sum w (x_j - c_j) per j]
__global__ void tf32_border_reduce_kernel(const double* __restrict__ partials, int grid, int d,
                                          const double* __restrict__ centres, double* __restrict__ tot,
                                          double* __restrict__ out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= 5 + 2 * d) return;
  auto fold = [&](size_t off) {
    double s = 0.0;
    for (int b = 0; b < grid; ++b) s += partials[static_cast<size_t>(b) * kCtaDoubles + off];
    return s;
  };
  if (e < 5) {
    const double v = fold(kScalarOff + e);
    tot[e] = v;
    out[e] = v;
  } else if (e < 5 + d) {
    const int j = e - 5;
    const double a = fold(kA1Off + j), w = fold(kScalarOff + 1);
    tot[5 + j] = a;
    out[e] = fma(centres[j], w, a);  // sum w x_j
  } else {
    // sum w x_j y = sum w (x_j - c_j)(y - c_y) + c_y sum w (x_j - c_j) + c_j sum w y
    const int j = e - 5 - d;
    const double v = fold(kAbOff + j), a = fold(kA1Off + j), by = fold(kScalarOff + 3);
    out[e] = v + centres[d] * a + centres[j] * by;
  }
}

// packed upper triangle: sum w x_i x_j = T_ij + c_j A'_i + c_i A'_j + c_i c_j W, with T the tensor-core tile (the exact
// fp64 sums on the diagonal) and A' the centred column sums
__global__ void tf32_pack_kernel(const double* __restrict__ partials, int grid, int d, const double* __restrict__ centres,
                                 const double* __restrict__ tot, double* __restrict__ out) {
  const long long q = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long tri = static_cast<long long>(d) * (d + 1) / 2;
  if (q >= tri) return;
  long long j = static_cast<long long>((sqrt(8.0 * static_cast<double>(q) + 1.0) - 1.0) * 0.5);
  while (j * (j + 1) / 2 > q) --j;
  while ((j + 1) * (j + 2) / 2 <= q) ++j;
  const long long i = q - j * (j + 1) / 2;  // i <= j
  const size_t off = i == j ? static_cast<size_t>(kDiagOff + i) : tile_off(i, j);
  double t = 0.0;
  for (int b = 0; b < grid; ++b) t += partials[static_cast<size_t>(b) * kCtaDoubles + off];
  const double ci = centres[i], cj = centres[j];
  out[5 + 2 * d + q] = t + cj * tot[5 + i] + ci * tot[5 + j] + ci * cj * tot[1];
}

}  // namespace

bool gram_tf32_supported(int d) { return d >= 1 && d <= TF; }

namespace {

// The TMA path needs the table's fp64 columns to be one uniformly strided 2-D array in THIS GPU's memory: columns
// [first, d) all fp64, 16-byte aligned, a constant multiple-of-16 byte distance apart, after at most TFPW other columns
// (the lab's int32 guest).  What dq4ml_synth_generate and any column-major allocation produce.  Returns false when the
// table does not have that shape (or the driver lacks the encoder): the per-thread cp.async path serves it then.
bool build_column_tensor_map(const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes, int64_t n,
                             CUtensorMap* tmap, int* first_out) {
  const int d = static_cast<int>(col_ptrs.size());
  int first = 0;
  while (first < d && dtypes[first] != DQ4ML_F64) ++first;
  if (first > TFPW || d - first < 2 || n < TKC) return false;
  for (int j = first; j < d; ++j)
    if (dtypes[j] != DQ4ML_F64) return false;
  const char* base = static_cast<const char*>(col_ptrs[first]);
  const long long stride = static_cast<const char*>(col_ptrs[first + 1]) - base;
  if (stride <= 0 || (stride & 15) != 0 || (reinterpret_cast<uintptr_t>(base) & 15u) != 0 || stride < n * 8) return false;
  for (int j = first + 1; j < d; ++j)
    if (static_cast<const char*>(col_ptrs[j]) - base != stride * (j - first)) return false;
  cudaPointerAttributes at;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaPointerGetAttributes(&at, base) != cudaSuccess ||
      at.type != cudaMemoryTypeDevice || at.device != dev)
    return false;  // only this GPU's own memory is ever described by the map
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      fn = nullptr;
    return reinterpret_cast<EncodeFn>(fn);
  }();
  if (encode == nullptr) return false;
  const cuuint64_t dims[2] = {static_cast<cuuint64_t>(n), static_cast<cuuint64_t>(d - first)};  // rows (contiguous), columns
  const cuuint64_t strides[1] = {static_cast<cuuint64_t>(stride)};
  const cuuint32_t box[2] = {TKC, TF};
  const cuuint32_t elem[2] = {1, 1};
  // L2 promotion: a box row is 128 bytes of one column; consecutive 128-byte pieces of a column belong to different CTAs
  static const CUtensorMapL2promotion promo = [] {
    const char* env = getenv("DQ4ML_TF32_TMA_L2");
    const int v = env != nullptr ? atoi(env) : 128;
    return v == 256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : (v == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE : CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
  }();
  const CUresult rc = encode(tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<char*>(base), dims, strides, box, elem,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) return false;
  *first_out = first;
  return true;
}

}  // namespace

int gram_tf32_wls_moments(dq4ml_ctx* ctx, const std::vector<const void*>& col_ptrs, const std::vector<int>& dtypes,
                          const double* label, const double* weight, const DeviceDq& dq, int64_t n, double* out) {
  const int d = static_cast<int>(col_ptrs.size());
  if (!gram_tf32_supported(d)) {
    set_error("the tf32 Gram tier handles 1..%d feature columns (got %d)", TF, d);
    return DQ4ML_ERR_UNSUPPORTED;
  }
  const int64_t len = 5 + 2LL * d + static_cast<int64_t>(d) * (d + 1) / 2;
  const int64_t chunks = (n + TKC - 1) / TKC;
  int grid = ctx->sm_count;
  if (grid > chunks) grid = static_cast<int>(chunks);
  if (grid < 1) grid = 1;
  const size_t table_bytes = (static_cast<size_t>(d) * 9 + 255) & ~static_cast<size_t>(255);
  const size_t centre_bytes = (TF + 8) * sizeof(double);  // features, then the label
  const size_t tot_bytes = (5 + TF + 3) * sizeof(double);
  const size_t part_bytes = static_cast<size_t>(grid) * kCtaDoubles * sizeof(double);
  const size_t out_bytes = static_cast<size_t>(len) * sizeof(double);
  const size_t need = table_bytes + centre_bytes + tot_bytes + part_bytes + out_bytes + 512;
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
  double* d_centres = reinterpret_cast<double*>(base + table_bytes);
  double* d_tot = reinterpret_cast<double*>(base + table_bytes + centre_bytes);
  double* d_part = reinterpret_cast<double*>(base + table_bytes + centre_bytes + tot_bytes);
  double* d_out = reinterpret_cast<double*>(base + table_bytes + centre_bytes + tot_bytes + part_bytes);
  std::vector<unsigned char> flags(d);
  for (int j = 0; j < d; ++j) flags[j] = dtypes[j] == DQ4ML_I32 ? 1 : 0;
  if ((e = cudaMemcpyAsync(d_ptrs, col_ptrs.data(), static_cast<size_t>(d) * 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e, "column table upload");
  if ((e = cudaMemcpyAsync(d_i32, flags.data(), static_cast<size_t>(d), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return fail(e, "column table upload");
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return fail(e, "sync");  // the host vectors go out of scope

  Tf32Args g;
  g.col_ptr = d_ptrs;
  g.col_i32 = d_i32;
  g.centres = d_centres;
  g.d = d;
  g.label = label;
  g.weight = weight;
  g.dq_guest = dq.guest;
  g.dq_price = dq.price;
  g.t.min_price = dq.params.min_price;
  g.t.max_price = dq.params.max_price;
  g.t.min_guest = dq.params.min_guest;
  g.has_dq = dq.guest != nullptr ? 1 : 0;
  g.n = n;
  int drain_rows = 2048;  // fp32 accumulation span; see the header comment
  if (const char* env = getenv("DQ4ML_TF32_DRAIN_ROWS")) { const int v = atoi(env); if (v >= TKC) drain_rows = v; }
  g.drain_chunks = drain_rows / TKC;
  g.swap_offsets = getenv("DQ4ML_TF32_SWAP_OFFSETS") != nullptr ? 1 : 0;
  g.experiment = getenv("DQ4ML_TF32_EXPERIMENT") != nullptr ? atoi(getenv("DQ4ML_TF32_EXPERIMENT")) : 0;
  g.partials = d_part;
  static bool attr_set[64] = {false};
  if (ctx->device < 64 && !attr_set[ctx->device]) {
    if ((e = cudaFuncSetAttribute(gram_tf32_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTf32Smem)) != cudaSuccess)
      return fail(e, "tf32 gram smem attribute");
    if ((e = cudaFuncSetAttribute(gram_tf32_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTf32Smem)) != cudaSuccess)
      return fail(e, "tf32 gram smem attribute");
    attr_set[ctx->device] = true;
  }
  // Staging: per-thread cp.async copies by default; DQ4ML_TF32_TMA=1 stages uniformly strided fp64 columns with TMA
  // tensor-map loads instead.  Both give bit-identical moments (tests/test_gpu_tf32.py); measured at 62.5 M x 256 the TMA
  // path is 3-5 % slower (50.4-51.0 ms against 48.5-49.3 ms, profiles/r2_config4_tf32_staging.jsonl): it removes the
  // ~390 cycles of load issue per chunk from the converter warps, but all 16 of them now have to release a stage before
  // its next box can be requested, and the pass is bound by their conversion arithmetic either way.
  CUtensorMap tmap;
  std::memset(&tmap, 0, sizeof(tmap));
  g.tma_first = 0;
  const char* tma_env = getenv("DQ4ML_TF32_TMA");  // read per call: the tests switch it
  const bool use_tma = tma_env != nullptr && atoi(tma_env) != 0 && build_column_tensor_map(col_ptrs, dtypes, n, &tmap, &g.tma_first);
  if (ctx->pass_timing) cudaEventRecord(ctx->pass_ev0, ctx->stream);
  tf32_centres_kernel<<<d + 1, 32, 0, ctx->stream>>>(d_ptrs, d_i32, label, d, n, d_centres);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "tf32 centres kernel");
  if ((e = cudaMemsetAsync(d_part, 0, part_bytes, ctx->stream)) != cudaSuccess) return fail(e, "tf32 tile clear");
  if (use_tma) gram_tf32_kernel<true><<<grid, TTHREADS + 32, kTf32Smem, ctx->stream>>>(g, tmap);
  else gram_tf32_kernel<false><<<grid, TTHREADS, kTf32Smem, ctx->stream>>>(g, tmap);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "tf32 gram kernel");
  tf32_border_reduce_kernel<<<(5 + 2 * d + 127) / 128, 128, 0, ctx->stream>>>(d_part, grid, d, d_centres, d_tot, d_out);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "tf32 border reduce kernel");
  const long long tri = static_cast<long long>(d) * (d + 1) / 2;
  tf32_pack_kernel<<<static_cast<unsigned>((tri + 255) / 256), 256, 0, ctx->stream>>>(d_part, grid, d, d_centres, d_tot, d_out);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "tf32 pack kernel");
  ctx->launches += 4;
  if (ctx->pass_timing) {
    cudaEventRecord(ctx->pass_ev1, ctx->stream);
    ctx->pass_event_pending = true;
  }
  if (ctx->nccl != nullptr) {
    const int rc = nccl_allreduce_device(ctx, d_out, len);
    if (rc != DQ4ML_OK) return rc;
  }
  if ((e = cudaMemcpyAsync(ctx->gram_host, d_out, out_bytes, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return fail(e, "gram D2H");
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return fail(e, "tf32 gram pass");
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
