// context.hpp -- dq4ml_ctx: one CUDA device, one stream, scratch, a grow-only staging arena for host inputs, and the
// (optional) binding to the other ranks.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/dq4ml.h"
#include "kernels.hpp"

namespace dq4ml {

void set_error(const char* fmt, ...);
const char* get_error();

struct NcclBinding;  // dist.cpp

}  // namespace dq4ml

struct dq4ml_ctx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  dq4ml::Scratch scratch{};
  char* scratch_base = nullptr;
  double* h_result = nullptr;  // pinned, kResultDoubles: where every pass's vector ends up for the host code
  volatile unsigned long long* h_words = nullptr;  // pinned + mapped: the tagged word pairs the epilogue of a streaming
                                                   // pass writes from the GPU (kEpilogueValues pairs, epilogue.cuh)
  unsigned long long* h_words_dev = nullptr;       // the same buffer as the device sees it
  unsigned long long host_seq = 0;  // passes published to the host so far
  unsigned long long xr_seq = 0;    // cross-rank (mailbox) passes so far: advances in lock step on every rank
  bool ep_armed = false;            // the pass just launched publishes through the epilogue
  int ep_world = 1;                 // ... folded / gathered over this many ranks inside the kernel
  unsigned long long* d_counters = nullptr;
  int* d_err = nullptr;
  unsigned long long* d_tile_state = nullptr;
  size_t tile_state_cap = 0;
  char* arena = nullptr;  // staging for host-resident inputs/outputs
  size_t arena_cap = 0, arena_used = 0;
  std::unordered_map<const void*, void*> staged;  // host pointer -> device copy, per call
  char* gram_scratch = nullptr;  // Gram pass (d > 8): column table, per-CTA partial tiles, packed result
  size_t gram_cap = 0;
  double* gram_host = nullptr;   // pinned copy of the packed result
  size_t gram_host_cap = 0;
  int gram_tier = DQ4ML_GRAM_FP64;  // which kernel serves the normal-equation pass (dq4ml_ctx_set_gram_tier)
  int64_t launches = 0;
  // optional CUDA-event timing of the aggregation-pass kernels (bench.py's roofline leg)
  bool pass_timing = false;
  bool pass_event_pending = false;
  cudaEvent_t pass_ev0 = nullptr, pass_ev1 = nullptr;
  double pass_ms_total = 0.0;
  int64_t pass_count = 0;
  std::vector<float> pass_ms_list;  // one entry per timed pass since timing was enabled (capped)
  // ranks
  dq4ml::NcclBinding* nccl = nullptr;
  int rank = 0, world = 1;
  dq4ml_allreduce_fn allreduce_cb = nullptr;
  void* allreduce_user = nullptr;
};

namespace dq4ml {

constexpr size_t kResultDoubles = 1 << 16;  // fits the packed Gram of d = 256 (33 411 doubles)
constexpr size_t kPartialsBytes = 8u << 20;
constexpr size_t kEpilogueValues = static_cast<size_t>(kXrMaxRanks) * kXrSlotDoubles;  // longest hand-over: a gather of every rank's vector

// api.cpp: after the ctx stream has been synchronised, folds the pending pass-timing event pair into the ctx counters
void collect_pass_time(dq4ml_ctx* ctx);

// dist.cpp
int nccl_unique_id(void* id128);
int nccl_bind(dq4ml_ctx* ctx, const void* id128, int rank, int world);
void nccl_unbind(dq4ml_ctx* ctx);
// In-stream sum all-reduce of `count` doubles at device pointer `buf` (no-op when no communicator is bound).
int nccl_allreduce_device(dq4ml_ctx* ctx, double* buf, int64_t count);
// Peer mailboxes of the bound ranks (epilogue.cuh): fills ep->peer / world / rank and returns true when every rank could
// map every other rank's mailbox at bind time; false = the per-pass vectors go through ncclAllReduce instead.
bool nccl_mailboxes(const dq4ml_ctx* ctx, XrEpilogue* ep);

}  // namespace dq4ml
