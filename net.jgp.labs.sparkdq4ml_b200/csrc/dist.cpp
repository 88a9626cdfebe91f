// dist.cpp -- binding a ctx to the other ranks of a row-sharded fit.
//
// Spark merges partition aggregators with treeAggregate inside the JVM; here every rank (one process per GPU) runs the
// same pass over its own row shard and the tiny result vector (moments / loss+gradient / Gram) is summed with one
// ncclAllReduce(ncclDouble, ncclSum) on the ctx stream over NVLink/NVSwitch.  NCCL is opened with dlopen so that the
// library has no link-time dependency on it: single-GPU callers (and the JVM drop-in) never need libnccl, and inside a
// PyTorch process the already-loaded libnccl.so.2 is reused.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "context.hpp"

namespace dq4ml {
namespace {

// Minimal mirror of the NCCL 2.x ABI used here (stable since 2.0).
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId {
  char internal[128];
};
constexpr int kNcclSuccess = 0;
constexpr int kNcclSum = 0;
constexpr int kNcclFloat64 = 8;
constexpr int kNcclInt8 = 0;

struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  std::string why;
};

NcclApi& api() {
  static NcclApi a;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {std::getenv("DQ4ML_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      if (nm == nullptr || *nm == 0) continue;
      a.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (a.handle != nullptr) break;
    }
    if (a.handle == nullptr) {
      a.why = std::string("cannot dlopen libnccl.so.2: ") + (dlerror() ? dlerror() : "?");
      return;
    }
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(a.handle, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(a.handle, "ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(a.handle, "ncclCommDestroy"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(a.handle, "ncclAllReduce"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(a.handle, "ncclAllGather"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(a.handle, "ncclGetErrorString"));
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce) {
      a.why = "libnccl is missing an expected symbol";
      a.handle = nullptr;
    }
  });
  return a;
}

const char* nccl_err(int rc) {
  NcclApi& a = api();
  return a.GetErrorString ? a.GetErrorString(rc) : "nccl error";
}

}  // namespace

struct NcclBinding {
  ncclComm_t comm = nullptr;
  // peer mailboxes of the in-kernel cross-rank epilogue (epilogue.cuh)
  unsigned char* mailbox = nullptr;         // ours (cudaMalloc)
  unsigned char* peer[kXrMaxRanks] = {};    // every rank's, as mapped here (peer[rank] == mailbox)
  bool mailboxes_ok = false;
};

namespace {

// two buffers x kXrMaxRanks slots x kXrSlotDoubles values, each value two tagged 8-byte words (epilogue.cuh)
constexpr size_t kMailboxBytes = 2 * static_cast<size_t>(kXrMaxRanks) * kXrSlotDoubles * 16;

// Allocates this rank's mailbox, exchanges CUDA IPC handles over the communicator, and maps every peer's mailbox.
// Collective.  Any failure on any rank leaves mailboxes_ok false on EVERY rank (they agree through an all-reduce), and
// the per-pass vectors keep going through ncclAllReduce.
void setup_mailboxes(dq4ml_ctx* ctx, NcclBinding* b, int rank, int world) {
  NcclApi& a = api();
  if (world > kXrMaxRanks || a.AllGather == nullptr) return;
  if (const char* env = std::getenv("DQ4ML_XRANK")) {
    if (std::strcmp(env, "nccl") == 0) return;  // A/B switch: keep the NCCL path
  }
  char* d_handles = nullptr;
  double* d_ok = nullptr;
  bool local_ok = cudaMalloc(reinterpret_cast<void**>(&b->mailbox), kMailboxBytes) == cudaSuccess &&
                  cudaMemsetAsync(b->mailbox, 0, kMailboxBytes, ctx->stream) == cudaSuccess &&
                  cudaMalloc(reinterpret_cast<void**>(&d_handles), sizeof(cudaIpcMemHandle_t) * world) == cudaSuccess &&
                  cudaMalloc(reinterpret_cast<void**>(&d_ok), sizeof(double)) == cudaSuccess;
  cudaIpcMemHandle_t mine;
  std::memset(&mine, 0, sizeof(mine));
  if (local_ok) local_ok = cudaIpcGetMemHandle(&mine, b->mailbox) == cudaSuccess;
  std::vector<cudaIpcMemHandle_t> all(world);
  // the collectives below must be entered by every rank, whatever happened locally
  if (d_handles != nullptr) {
    cudaMemcpyAsync(d_handles + sizeof(mine) * rank, &mine, sizeof(mine), cudaMemcpyHostToDevice, ctx->stream);
    if (a.AllGather(d_handles + sizeof(mine) * rank, d_handles, sizeof(mine), kNcclInt8, b->comm, ctx->stream) != kNcclSuccess) local_ok = false;
    cudaMemcpyAsync(all.data(), d_handles, sizeof(mine) * world, cudaMemcpyDeviceToHost, ctx->stream);
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) local_ok = false;
  }
  if (local_ok) {
    for (int r = 0; r < world && local_ok; ++r) {
      if (r == rank) {
        b->peer[r] = b->mailbox;
        continue;
      }
      void* p = nullptr;
      if (cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        local_ok = false;
      } else {
        b->peer[r] = static_cast<unsigned char*>(p);
      }
    }
  }
  // agreement: the mailboxes are used only if every rank mapped every peer (this all-reduce is also the barrier that
  // orders the zero-fill of every mailbox before anybody's first store into it)
  double ok = local_ok ? 1.0 : 0.0;
  if (d_ok != nullptr) {
    cudaMemcpyAsync(d_ok, &ok, sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
    a.AllReduce(d_ok, d_ok, 1, kNcclFloat64, kNcclSum, b->comm, ctx->stream);
    cudaMemcpyAsync(&ok, d_ok, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    cudaStreamSynchronize(ctx->stream);
  } else {
    ok = 0.0;  // (a rank that could not even allocate 8 bytes cannot take part; its peers will time out in NCCL)
  }
  b->mailboxes_ok = ok == static_cast<double>(world);
  if (d_handles != nullptr) cudaFree(d_handles);
  if (d_ok != nullptr) cudaFree(d_ok);
}

void release_mailboxes(dq4ml_ctx* ctx, NcclBinding* b) {
  for (int r = 0; r < kXrMaxRanks; ++r) {
    if (b->peer[r] != nullptr && b->peer[r] != b->mailbox) cudaIpcCloseMemHandle(b->peer[r]);
    b->peer[r] = nullptr;
  }
  if (b->mailbox != nullptr) cudaFree(b->mailbox);
  b->mailbox = nullptr;
  b->mailboxes_ok = false;
  (void)ctx;
}

}  // namespace

int nccl_unique_id(void* id128) {
  NcclApi& a = api();
  if (a.handle == nullptr) {
    set_error("%s", a.why.c_str());
    return DQ4ML_ERR_COMM;
  }
  ncclUniqueId id;
  const int rc = a.GetUniqueId(&id);
  if (rc != kNcclSuccess) {
    set_error("ncclGetUniqueId: %s", nccl_err(rc));
    return DQ4ML_ERR_COMM;
  }
  std::memcpy(id128, id.internal, sizeof(id.internal));
  return DQ4ML_OK;
}

int nccl_bind(dq4ml_ctx* ctx, const void* id128, int rank, int world) {
  NcclApi& a = api();
  if (a.handle == nullptr) {
    set_error("%s", a.why.c_str());
    return DQ4ML_ERR_COMM;
  }
  if (world < 1 || rank < 0 || rank >= world) {
    set_error("bad rank %d / world %d", rank, world);
    return DQ4ML_ERR_INVALID;
  }
  nccl_unbind(ctx);
  ncclUniqueId id;
  std::memcpy(id.internal, id128, sizeof(id.internal));
  cudaSetDevice(ctx->device);
  auto* b = new NcclBinding;
  const int rc = a.CommInitRank(&b->comm, world, id, rank);
  if (rc != kNcclSuccess) {
    delete b;
    set_error("ncclCommInitRank: %s", nccl_err(rc));
    return DQ4ML_ERR_COMM;
  }
  ctx->nccl = b;
  ctx->rank = rank;
  ctx->world = world;
  ctx->xr_seq = 0;
  if (world > 1) setup_mailboxes(ctx, b, rank, world);
  return DQ4ML_OK;
}

bool nccl_mailboxes(const dq4ml_ctx* ctx, XrEpilogue* ep) {
  if (ctx->nccl == nullptr || !ctx->nccl->mailboxes_ok) return false;
  ep->world = ctx->world;
  ep->rank = ctx->rank;
  for (int r = 0; r < kXrMaxRanks; ++r) ep->peer[r] = ctx->nccl->peer[r];
  return true;
}

void nccl_unbind(dq4ml_ctx* ctx) {
  if (ctx->nccl != nullptr) {
    cudaSetDevice(ctx->device);
    if (ctx->stream != nullptr) cudaStreamSynchronize(ctx->stream);
    release_mailboxes(ctx, ctx->nccl);
    api().CommDestroy(ctx->nccl->comm);
    delete ctx->nccl;
    ctx->nccl = nullptr;
  }
}

int nccl_allreduce_device(dq4ml_ctx* ctx, double* buf, int64_t count) {
  if (ctx->nccl == nullptr || count <= 0) return DQ4ML_OK;
  const int rc = api().AllReduce(buf, buf, static_cast<size_t>(count), kNcclFloat64, kNcclSum, ctx->nccl->comm, ctx->stream);
  if (rc != kNcclSuccess) {
    set_error("ncclAllReduce: %s", nccl_err(rc));
    return DQ4ML_ERR_COMM;
  }
  return DQ4ML_OK;
}

}  // namespace dq4ml
