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

struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
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
};

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
  return DQ4ML_OK;
}

void nccl_unbind(dq4ml_ctx* ctx) {
  if (ctx->nccl != nullptr) {
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
