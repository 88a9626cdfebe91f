// csv_kernels.cu -- ingest of the lab's CSV dialect straight into the two device columns.
//
// The lab loads `data/dataset-*.csv` with spark.read().format("csv").option("inferSchema","true")
// .option("header","false") (DataQuality4MachineLearningApp.java:52-55): two fields per record, `_c0` integer (guest),
// `_c1` decimal (price); the three files use a bare CR as record separator, which Hadoop's LineRecordReader accepts
// along with LF and CRLF.  Two kernels over 8 KB tiles of the byte buffer:
//   1. count the record starts of every tile (a start is a non-separator byte at offset 0 or after a separator);
//      a one-CTA kernel then turns the counts into exclusive offsets;
//   2. every thread owns 32 bytes, ranks the starts inside the tile (warp ballot-free: per-thread counts + block scan)
//      and parses each record that starts in its bytes: digits are accumulated in a 64-bit integer and, when the
//      decimal has at most 15 significant digits and a power of ten up to 10^22, converted with ONE correctly rounded
//      fp64 operation (digits * 10^k or digits / 10^k are exact operands) -- bit-identical to Double.parseDouble.
//      Anything longer is flagged and re-parsed on the host with strtod (also correctly rounded); the lab's files and
//      two-decimal prices never take that path.
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include "context.hpp"
#include "device_utils.cuh"

namespace dq4ml {
namespace {

constexpr int kCsvThreads = 256;
constexpr int kCsvBytesPerThread = 32;
constexpr int kCsvTile = kCsvThreads * kCsvBytesPerThread;  // 8192 bytes

__device__ __forceinline__ bool is_sep(char c) { return c == '\r' || c == '\n'; }
__device__ __forceinline__ bool is_digit(char c) { return c >= '0' && c <= '9'; }

__device__ __forceinline__ unsigned starts_in(const char* __restrict__ buf, long long len, long long lo, long long hi) {
  unsigned cnt = 0;
  char prev = lo > 0 ? buf[lo - 1] : '\n';
  for (long long i = lo; i < hi && i < len; ++i) {
    const char c = buf[i];
    cnt += (!is_sep(c) && is_sep(prev)) ? 1u : 0u;
    prev = c;
  }
  return cnt;
}

__global__ void __launch_bounds__(kCsvThreads) csv_count_kernel(const char* __restrict__ buf, long long len,
                                                                unsigned long long* __restrict__ tile_counts) {
  __shared__ double s_red[kCsvThreads / 32];
  const long long lo = static_cast<long long>(blockIdx.x) * kCsvTile + threadIdx.x * kCsvBytesPerThread;
  const unsigned c = starts_in(buf, len, lo, lo + kCsvBytesPerThread);
  const double tot = block_reduce(static_cast<double>(c), kSum, s_red);  // <= 8192: exact in fp64
  if (threadIdx.x == 0) tile_counts[blockIdx.x] = static_cast<unsigned long long>(tot);
}

// exclusive scan of the tile counts in place; total -> *total
__global__ void __launch_bounds__(1024) csv_scan_kernel(unsigned long long* __restrict__ counts, long long n_tiles,
                                                        unsigned long long* __restrict__ total) {
  __shared__ unsigned long long s_warp[32];
  __shared__ unsigned long long s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (long long base = 0; base < n_tiles; base += 1024) {
    const long long i = base + threadIdx.x;
    const unsigned long long v = i < n_tiles ? counts[i] : 0ull;
    unsigned long long incl = v;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      unsigned long long w = s_warp[lane], wi = w;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, wi, off);
        if (lane >= off) wi += t;
      }
      s_warp[lane] = wi - w;
    }
    __syncthreads();
    const unsigned long long carry = s_carry;
    if (i < n_tiles) counts[i] = carry + s_warp[warp] + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = carry + s_warp[warp] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = s_carry;
}

__constant__ double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                  1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

enum { kCsvOk = 0, kCsvSlow = 1, kCsvBad = 2 };

// Parses the record starting at `pos`.  Returns kCsvOk (outputs valid), kCsvSlow (integer valid, decimal needs the
// exact host path) or kCsvBad (malformed).
__device__ int parse_record(const char* __restrict__ buf, long long len, long long pos, int* guest, double* price) {
  long long i = pos;
  // ---- field 0: [+-]digits
  bool neg = false;
  if (i < len && (buf[i] == '-' || buf[i] == '+')) neg = buf[i++] == '-';
  long long iv = 0;
  int nd = 0;
  while (i < len && is_digit(buf[i])) {
    iv = iv * 10 + (buf[i] - '0');
    if (++nd > 10) return kCsvBad;
    ++i;
  }
  if (nd == 0 || i >= len || buf[i] != ',') return kCsvBad;
  iv = neg ? -iv : iv;
  if (iv > 2147483647LL || iv < -2147483648LL) return kCsvBad;
  *guest = static_cast<int>(iv);
  ++i;
  // ---- field 1: [+-]digits[.digits][(e|E)[+-]digits]
  neg = false;
  if (i < len && (buf[i] == '-' || buf[i] == '+')) neg = buf[i++] == '-';
  unsigned long long mant = 0;
  int sig = 0, any = 0, dec_exp = 0;
  bool slow = false;
  while (i < len && is_digit(buf[i])) {
    any = 1;
    if (sig < 19) { mant = mant * 10 + (buf[i] - '0'); sig += (mant != 0) ? 1 : 0; } else { slow = true; ++dec_exp; }
    ++i;
  }
  if (i < len && buf[i] == '.') {
    ++i;
    while (i < len && is_digit(buf[i])) {
      any = 1;
      if (sig < 19) { mant = mant * 10 + (buf[i] - '0'); sig += (mant != 0) ? 1 : 0; --dec_exp; } else { slow = true; }
      ++i;
    }
  }
  if (!any) return kCsvBad;
  if (i < len && (buf[i] == 'e' || buf[i] == 'E')) {
    ++i;
    bool eneg = false;
    if (i < len && (buf[i] == '-' || buf[i] == '+')) eneg = buf[i++] == '-';
    int ev = 0, ed = 0;
    while (i < len && is_digit(buf[i])) {
      if (ev < 100000) ev = ev * 10 + (buf[i] - '0');
      ++ed;
      ++i;
    }
    if (ed == 0) return kCsvBad;
    dec_exp += eneg ? -ev : ev;
  }
  if (i < len && !is_sep(buf[i])) return kCsvSlow;  // trailing characters (NaN, Infinity, 1.0d, ...): let the host decide
  if (slow || sig > 15 || dec_exp > 22 || dec_exp < -22) return kCsvSlow;
  double v = static_cast<double>(mant);  // < 10^15 < 2^53: exact
  v = dec_exp >= 0 ? __dmul_rn(v, kPow10[dec_exp]) : __ddiv_rn(v, kPow10[-dec_exp]);
  *price = neg ? -v : v;
  return kCsvOk;
}

__global__ void __launch_bounds__(kCsvThreads)
    csv_parse_kernel(const char* __restrict__ buf, long long len, const unsigned long long* __restrict__ tile_offsets,
                     long long cap, int* __restrict__ guest, double* __restrict__ price,
                     long long* __restrict__ slow_pos /* per record: byte offset if the host must re-parse, else -1 */,
                     unsigned long long* __restrict__ status /* [0] = 1 + index of first malformed record, 0 = none */) {
  __shared__ unsigned s_warp[kCsvThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long lo = static_cast<long long>(blockIdx.x) * kCsvTile + threadIdx.x * kCsvBytesPerThread;
  const unsigned cnt = starts_in(buf, len, lo, lo + kCsvBytesPerThread);
  unsigned incl = cnt;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const unsigned t = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  unsigned warp_base = 0;
  for (int w = 0; w < warp; ++w) warp_base += s_warp[w];
  long long rec = static_cast<long long>(tile_offsets[blockIdx.x]) + warp_base + incl - cnt;
  if (cnt == 0) return;
  char prev = lo > 0 ? buf[lo - 1] : '\n';
  for (long long i = lo; i < lo + kCsvBytesPerThread && i < len; ++i) {
    const char c = buf[i];
    if (!is_sep(c) && is_sep(prev)) {
      if (rec < cap) {
        int g = 0;
        double p = 0.0;
        const int st = parse_record(buf, len, i, &g, &p);
        guest[rec] = g;
        price[rec] = p;
        slow_pos[rec] = st == kCsvSlow ? i : -1;
        if (st == kCsvBad) atomicMin(status, static_cast<unsigned long long>(rec) + 1ull);
      }
      ++rec;
    }
    prev = c;
  }
}

int cuda_err(cudaError_t e, const char* what) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  return DQ4ML_ERR_CUDA;
}

}  // namespace
}  // namespace dq4ml

// Declared in include/dq4ml.h.
extern "C" int dq4ml_csv_parse(dq4ml_ctx* ctx, const char* data, int64_t len, int32_t* guest_out, double* price_out,
                               int64_t cap, int64_t* n_records) {
  using namespace dq4ml;
  if (ctx == nullptr || n_records == nullptr || len < 0 || cap < 0 || (len > 0 && data == nullptr) ||
      (cap > 0 && (guest_out == nullptr || price_out == nullptr))) {
    set_error("csv_parse: bad arguments");
    return DQ4ML_ERR_INVALID;
  }
  *n_records = 0;
  if (len == 0) return DQ4ML_OK;
  cudaError_t e;
  if ((e = cudaSetDevice(ctx->device)) != cudaSuccess) return cuda_err(e, "set device");
  cudaPointerAttributes at;
  auto on_device = [&](const void* p) {
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
  };
  const bool data_dev = on_device(data), g_dev = cap > 0 && on_device(guest_out), p_dev = cap > 0 && on_device(price_out);
  const long long n_tiles = (len + kCsvTile - 1) / kCsvTile;
  auto r256 = [](size_t b) { return (b + 255) & ~static_cast<size_t>(255); };
  const size_t need = r256(static_cast<size_t>(n_tiles) * 8) + 256 + r256(static_cast<size_t>(cap) * 8) +
                      (data_dev ? 0 : r256(static_cast<size_t>(len))) + (g_dev ? 0 : r256(static_cast<size_t>(cap) * 4)) +
                      (p_dev ? 0 : r256(static_cast<size_t>(cap) * 8)) + 1024;
  if (need > ctx->arena_cap) {
    if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return cuda_err(e, "sync");
    if (ctx->arena != nullptr) cudaFree(ctx->arena);
    ctx->arena = nullptr;
    ctx->arena_cap = 0;
    if ((e = cudaMalloc(reinterpret_cast<void**>(&ctx->arena), need + (16u << 20))) != cudaSuccess) return cuda_err(e, "csv staging");
    ctx->arena_cap = need + (16u << 20);
  }
  char* a = ctx->arena;
  auto take = [&](size_t b) { char* p = a; a += r256(b); return p; };
  auto* d_tiles = reinterpret_cast<unsigned long long*>(take(static_cast<size_t>(n_tiles) * 8));
  auto* d_status = reinterpret_cast<unsigned long long*>(take(256));  // [0] first bad record + 1 (init: max), [1] total
  auto* d_slow = reinterpret_cast<long long*>(take(static_cast<size_t>(cap) * 8));
  const char* d_data = data;
  if (!data_dev) {
    char* p = take(static_cast<size_t>(len));
    if ((e = cudaMemcpyAsync(p, data, static_cast<size_t>(len), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return cuda_err(e, "csv H2D");
    d_data = p;
  }
  int* d_guest = g_dev ? guest_out : reinterpret_cast<int*>(take(static_cast<size_t>(cap) * 4));
  double* d_price = p_dev ? price_out : reinterpret_cast<double*>(take(static_cast<size_t>(cap) * 8));
  const unsigned long long init[2] = {~0ull, 0ull};
  if ((e = cudaMemcpyAsync(d_status, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return cuda_err(e, "csv status");
  csv_count_kernel<<<static_cast<unsigned>(n_tiles), kCsvThreads, 0, ctx->stream>>>(d_data, len, d_tiles);
  csv_scan_kernel<<<1, 1024, 0, ctx->stream>>>(d_tiles, n_tiles, d_status + 1);
  csv_parse_kernel<<<static_cast<unsigned>(n_tiles), kCsvThreads, 0, ctx->stream>>>(d_data, len, d_tiles, cap, d_guest, d_price, d_slow, d_status);
  if ((e = cudaGetLastError()) != cudaSuccess) return cuda_err(e, "csv kernels");
  ctx->launches += 3;
  unsigned long long st[2] = {0, 0};
  if ((e = cudaMemcpyAsync(st, d_status, sizeof(st), cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return cuda_err(e, "csv D2H");
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return cuda_err(e, "csv parse");
  const int64_t total = static_cast<int64_t>(st[1]);
  *n_records = total;
  if (st[0] != ~0ull) {
    set_error("csv_parse: malformed record %llu (expected `int,decimal`)", st[0] - 1ull);
    return DQ4ML_ERR_INVALID;
  }
  if (total > cap) {
    set_error("csv_parse: %lld records but room for %lld", static_cast<long long>(total), static_cast<long long>(cap));
    return DQ4ML_ERR_INVALID;
  }
  // exact host path for the (rare) decimals outside the one-rounding fast path.  Every copy below goes through
  // ctx->stream (a non-blocking stream is not ordered against the legacy default stream) and is drained before returning.
  std::vector<long long> slow(static_cast<size_t>(total));
  if (total > 0) {
    if ((e = cudaMemcpyAsync(slow.data(), d_slow, static_cast<size_t>(total) * 8, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return cuda_err(e, "csv D2H");
    if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return cuda_err(e, "csv D2H");
  }
  std::vector<char> host_copy;
  std::vector<std::pair<int64_t, double>> fixups;
  for (int64_t r = 0; r < total; ++r) {
    if (slow[r] < 0) continue;
    if (host_copy.empty()) {
      host_copy.resize(static_cast<size_t>(len));
      if (data_dev) {
        if ((e = cudaMemcpyAsync(host_copy.data(), data, static_cast<size_t>(len), cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return cuda_err(e, "csv D2H");
        if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return cuda_err(e, "csv D2H");
      } else {
        std::memcpy(host_copy.data(), data, static_cast<size_t>(len));
      }
    }
    long long i = slow[r];
    while (i < len && host_copy[i] != ',') ++i;
    long long j = ++i;
    while (j < len && host_copy[j] != '\r' && host_copy[j] != '\n') ++j;
    const std::string field(host_copy.data() + i, static_cast<size_t>(j - i));
    char* endp = nullptr;
    const double v = std::strtod(field.c_str(), &endp);
    if (field.empty() || endp == nullptr || *endp != 0) {
      set_error("csv_parse: malformed record %lld (decimal field `%s`)", static_cast<long long>(r), field.c_str());
      return DQ4ML_ERR_INVALID;
    }
    fixups.emplace_back(r, v);
  }
  for (const auto& f : fixups)  // `fixups` outlives the copies: the stream is drained below
    if ((e = cudaMemcpyAsync(d_price + f.first, &f.second, 8, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return cuda_err(e, "csv fix-up");
  if (!g_dev && total > 0 && (e = cudaMemcpyAsync(guest_out, d_guest, static_cast<size_t>(total) * 4, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return cuda_err(e, "csv D2H");
  if (!p_dev && total > 0 && (e = cudaMemcpyAsync(price_out, d_price, static_cast<size_t>(total) * 8, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) return cuda_err(e, "csv D2H");
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return cuda_err(e, "csv drain");
  return DQ4ML_OK;
}
