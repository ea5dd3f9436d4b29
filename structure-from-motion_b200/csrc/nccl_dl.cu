// NCCL through dlopen: libsfmba has no link-time dependency on libnccl, so the library loads on a
// machine without NCCL and binds to the copy torch already mapped (soname libnccl.so.2) otherwise.
#include <dlfcn.h>
#include <string.h>

#include "ctx.h"

namespace {

struct NcclUniqueId { char internal[128]; };
typedef int (*fn_get_unique_id)(NcclUniqueId*);
typedef int (*fn_comm_init_rank)(void**, int, NcclUniqueId, int);
typedef int (*fn_comm_destroy)(void*);
typedef int (*fn_all_reduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef const char* (*fn_get_error_string)(int);
typedef int (*fn_all_gather)(const void*, void*, size_t, int, void*, cudaStream_t);

constexpr int NCCL_FLOAT64 = 8;   // ncclDataType_t ncclFloat64 / ncclDouble
constexpr int NCCL_INT8 = 0;      // ncclDataType_t ncclInt8 / ncclChar
constexpr int NCCL_SUM = 0;       // ncclRedOp_t ncclSum

void* open_nccl(std::string& err) {
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    void* h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) return h;
  }
  err = std::string("cannot dlopen libnccl.so.2: ") + dlerror();
  return nullptr;
}

}  // namespace

int sfm_nccl_unique_id(void* out128, std::string& err) {
  void* lib = open_nccl(err);
  if (!lib) return SFMBA_ERR_NCCL;
  auto f = (fn_get_unique_id)dlsym(lib, "ncclGetUniqueId");
  if (!f) {
    err = "ncclGetUniqueId not found";
    return SFMBA_ERR_NCCL;
  }
  NcclUniqueId id;
  memset(&id, 0, sizeof(id));
  const int rc = f(&id);
  if (rc != 0) {
    err = "ncclGetUniqueId failed";
    return SFMBA_ERR_NCCL;
  }
  memcpy(out128, &id, 128);
  return SFMBA_OK;
}

int sfm_nccl_init(sfmba_ctx* ctx, const void* unique_id) {
  std::string err;
  ctx->nccl_lib = open_nccl(err);
  if (!ctx->nccl_lib) return sfm_set_error(ctx, SFMBA_ERR_NCCL, err);
  auto f = (fn_comm_init_rank)dlsym(ctx->nccl_lib, "ncclCommInitRank");
  if (!f) return sfm_set_error(ctx, SFMBA_ERR_NCCL, "ncclCommInitRank not found");
  NcclUniqueId id;
  memcpy(&id, unique_id, 128);
  const int rc = f(&ctx->nccl_comm, ctx->world, id, ctx->rank);
  if (rc != 0) {
    auto es = (fn_get_error_string)dlsym(ctx->nccl_lib, "ncclGetErrorString");
    return sfm_set_error(ctx, SFMBA_ERR_NCCL, std::string("ncclCommInitRank: ") + (es ? es(rc) : "error"));
  }
  return SFMBA_OK;
}

int sfm_nccl_allreduce_f64(sfmba_ctx* ctx, double* buf, size_t count, cudaStream_t st) {
  if (ctx->world <= 1) return SFMBA_OK;
  static fn_all_reduce f = nullptr;
  if (!f) f = (fn_all_reduce)dlsym(ctx->nccl_lib, "ncclAllReduce");
  if (!f) return sfm_set_error(ctx, SFMBA_ERR_NCCL, "ncclAllReduce not found");
  const int rc = f(buf, buf, count, NCCL_FLOAT64, NCCL_SUM, ctx->nccl_comm, st);
  if (rc != 0) return sfm_set_error(ctx, SFMBA_ERR_NCCL, "ncclAllReduce failed");
  return SFMBA_OK;
}

int sfm_nccl_allgather_bytes(sfmba_ctx* ctx, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st) {
  static fn_all_gather f = nullptr;
  if (!f) f = (fn_all_gather)dlsym(ctx->nccl_lib, "ncclAllGather");
  if (!f) return sfm_set_error(ctx, SFMBA_ERR_NCCL, "ncclAllGather not found");
  const int rc = f(send, recv, bytes_per_rank, NCCL_INT8, ctx->nccl_comm, st);
  if (rc != 0) return sfm_set_error(ctx, SFMBA_ERR_NCCL, "ncclAllGather failed");
  return SFMBA_OK;
}

// all ranks have reached this point of `st` when the call returns
int sfm_nccl_barrier(sfmba_ctx* ctx, cudaStream_t st) {
  void* w = nullptr;
  int rc = sfm_check_cuda(ctx, ctx->reserve_misc(4096, &w), "cudaMalloc(barrier word)");
  if (rc != 0) return rc;
  double* word = reinterpret_cast<double*>(reinterpret_cast<char*>(w) + 2048);
  cudaMemsetAsync(word, 0, sizeof(double), st);
  if ((rc = sfm_nccl_allreduce_f64(ctx, word, 1, st)) != 0) return rc;
  return sfm_check_cuda(ctx, cudaStreamSynchronize(st), "nccl barrier");
}

void sfm_nccl_destroy(sfmba_ctx* ctx) {
  if (ctx->nccl_comm && ctx->nccl_lib) {
    auto f = (fn_comm_destroy)dlsym(ctx->nccl_lib, "ncclCommDestroy");
    if (f) f(ctx->nccl_comm);
  }
  ctx->nccl_comm = nullptr;
}
