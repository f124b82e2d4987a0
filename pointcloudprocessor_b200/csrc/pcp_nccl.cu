// pcp_nccl.cu — the one exchange step of the path: an NCCL all-reduce (sum, u32) of the per-point accumulators after
// the frames have been sharded across ranks.  libnccl is resolved with dlopen at first use so that the
// library loads on boxes without NCCL; integer sums are associative, so the result is bit-identical for any rank count.
#include <dlfcn.h>

#include <cstring>

#include "../../include/pcp/pcp.h"
#include "pcp_ctx.hpp"

namespace {

struct NcclId { char internal[128]; };
typedef void* NcclComm;
typedef int (*fn_get_unique_id)(NcclId*);
typedef int (*fn_comm_init_rank)(NcclComm*, int, NcclId, int);
typedef int (*fn_all_reduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*fn_comm_destroy)(NcclComm);
typedef const char* (*fn_error_string)(int);

struct NcclApi {
  void* handle = nullptr;
  fn_get_unique_id get_unique_id = nullptr;
  fn_comm_init_rank comm_init_rank = nullptr;
  fn_all_reduce all_reduce = nullptr;
  fn_comm_destroy comm_destroy = nullptr;
  fn_error_string error_string = nullptr;
  bool ok = false;
};

NcclApi& api() {
  static NcclApi a;
  if (a.handle) return a;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    a.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (a.handle) break;
  }
  if (!a.handle) return a;
  a.get_unique_id = (fn_get_unique_id)dlsym(a.handle, "ncclGetUniqueId");
  a.comm_init_rank = (fn_comm_init_rank)dlsym(a.handle, "ncclCommInitRank");
  a.all_reduce = (fn_all_reduce)dlsym(a.handle, "ncclAllReduce");
  a.comm_destroy = (fn_comm_destroy)dlsym(a.handle, "ncclCommDestroy");
  a.error_string = (fn_error_string)dlsym(a.handle, "ncclGetErrorString");
  a.ok = a.get_unique_id && a.comm_init_rank && a.all_reduce && a.comm_destroy;
  return a;
}

constexpr int kNcclUint32 = 3;  // ncclUint32
constexpr int kNcclSum = 0;     // ncclSum

}  // namespace

void nccl_teardown(pcp_ctx* ctx) {
  if (ctx->nccl_comm && api().ok) api().comm_destroy((NcclComm)ctx->nccl_comm);
  ctx->nccl_comm = nullptr;
}

extern "C" {

int pcp_nccl_unique_id(uint8_t id128[128]) {
  if (!id128 || !api().ok) return PCP_ENCCL;
  NcclId id;
  if (api().get_unique_id(&id) != 0) return PCP_ENCCL;
  std::memcpy(id128, id.internal, 128);
  return PCP_OK;
}

int pcp_nccl_init(pcp_ctx* ctx, const uint8_t id128[128], int rank, int nranks) {
  if (!ctx || !id128 || rank < 0 || rank >= nranks) return PCP_EINVAL;
  if (!api().ok) { ctx->fail("libnccl.so.2 not found"); return PCP_ENCCL; }
  if (cudaSetDevice(ctx->device) != cudaSuccess) { ctx->fail("cudaSetDevice failed"); return PCP_ECUDA; }
  nccl_teardown(ctx);
  NcclId id;
  std::memcpy(id.internal, id128, 128);
  NcclComm comm = nullptr;
  const int rc = api().comm_init_rank(&comm, nranks, id, rank);
  if (rc != 0) { ctx->fail("ncclCommInitRank: %s", api().error_string ? api().error_string(rc) : "error"); return PCP_ENCCL; }
  ctx->nccl_comm = comm;
  ctx->nccl_rank = rank;
  ctx->nccl_nranks = nranks;
  return PCP_OK;
}

int pcp_allreduce(pcp_ctx* ctx) {
  if (!ctx) return PCP_EINVAL;
  if (ctx->nccl_nranks <= 1 && !ctx->nccl_comm) return PCP_OK;  // single rank: nothing to exchange
  if (!ctx->nccl_comm) { ctx->fail("pcp_allreduce: call pcp_nccl_init first"); return PCP_ENCCL; }
  if (ctx->prm.accum_mode != PCP_ACCUM_SUM) { ctx->fail("pcp_allreduce: only PCP_ACCUM_SUM accumulators are all-reducible"); return PCP_EINVAL; }
  if (cudaSetDevice(ctx->device) != cudaSuccess) { ctx->fail("cudaSetDevice failed"); return PCP_ECUDA; }
  if (ctx->join() != cudaSuccess) { ctx->fail("pcp_allreduce: cudaStreamWaitEvent failed"); return PCP_ECUDA; }
  int rc = api().all_reduce(ctx->d_acc, ctx->d_acc, ctx->n * 4, kNcclUint32, kNcclSum, (NcclComm)ctx->nccl_comm, ctx->stream);
  if (rc == 0 && ctx->d_mask_acc) {
    // hits are summed; the per-point maximum is not a sum, so it is left rank-local (documented in DESIGN.md)
    rc = 0;
  }
  if (rc != 0) { ctx->fail("ncclAllReduce: %s", api().error_string ? api().error_string(rc) : "error"); return PCP_ENCCL; }
  ctx->launches++;
  return PCP_OK;
}

}  // extern "C"
