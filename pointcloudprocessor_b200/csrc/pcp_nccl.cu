// pcp_nccl.cu — the one exchange step of the path, after the frames have been sharded across ranks:
//   SUM accumulators   ncclAllReduce(u32, sum): integer sums are associative, bit-identical for any rank count;
//   mask accumulators  {hits, max} split into two planes, all-reduced with sum and max;
//   TOP5 slots         ncclAllGather of the ranks' sorted slot lists, block of points by block, and a merge kernel that keeps the
//                      top_m best of the union under the same total order (score descending, global frame number ascending) the
//                      insertion uses — so the merged lists equal the single-rank run's when frames carry their global numbers
//                      (pcp_set_frame_numbering).
// libnccl is resolved with dlopen at first use so that the library loads on boxes without NCCL.
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
typedef int (*fn_all_gather)(const void*, void*, size_t, int, NcclComm, cudaStream_t);
typedef int (*fn_comm_destroy)(NcclComm);
typedef const char* (*fn_error_string)(int);

struct NcclApi {
  void* handle = nullptr;
  fn_get_unique_id get_unique_id = nullptr;
  fn_comm_init_rank comm_init_rank = nullptr;
  fn_all_reduce all_reduce = nullptr;
  fn_all_gather all_gather = nullptr;
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
  a.all_gather = (fn_all_gather)dlsym(a.handle, "ncclAllGather");
  a.comm_destroy = (fn_comm_destroy)dlsym(a.handle, "ncclCommDestroy");
  a.error_string = (fn_error_string)dlsym(a.handle, "ncclGetErrorString");
  a.ok = a.get_unique_id && a.comm_init_rank && a.all_reduce && a.all_gather && a.comm_destroy;
  return a;
}

constexpr int kNcclUint32 = 3;  // ncclUint32
constexpr int kNcclSum = 0;     // ncclSum
constexpr int kNcclMax = 2;     // ncclMax

// ---- mask accumulators: {hits, max} per point <-> two planes -----------------------------------------------------------------
__global__ void mask_split_kernel(const uint2* __restrict__ acc, uint32_t* __restrict__ hits, uint32_t* __restrict__ mx, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint2 v = acc[i];
    hits[i] = v.x; mx[i] = v.y;
  }
}
__global__ void mask_join_kernel(uint2* __restrict__ acc, const uint32_t* __restrict__ hits, const uint32_t* __restrict__ mx, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) acc[i] = make_uint2(hits[i], mx[i]);
}

// ---- TOP5 slots ------------------------------------------------------------------------------------------------------
// staging layout of one rank's block of nb points: [array a in {score, rgb, frame}][slot k][point j]
__global__ void top_pack_kernel(const float* __restrict__ score, const uint32_t* __restrict__ rgbf, const uint32_t* __restrict__ frame, uint32_t n,
                                int top_m, uint32_t p0, uint32_t nb, uint32_t cap, uint32_t* __restrict__ out) {
  const uint32_t total = nb * (uint32_t)top_m;
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const uint32_t k = t / nb, j = t - k * nb;
    const size_t src = (size_t)k * n + p0 + j;
    out[((size_t)0 * top_m + k) * cap + j] = __float_as_uint(score[src]);
    out[((size_t)1 * top_m + k) * cap + j] = rgbf[src];
    out[((size_t)2 * top_m + k) * cap + j] = frame[src];
  }
}

constexpr int kMaxTopM = 16;  // pcp_create refuses more

__global__ void top_merge_kernel(const uint32_t* __restrict__ gathered, int nranks, float* __restrict__ score, uint32_t* __restrict__ rgbf,
                                 uint32_t* __restrict__ frame, uint32_t n, int top_m, uint32_t p0, uint32_t nb, uint32_t cap) {
  const size_t per_rank = (size_t)3 * top_m * cap;
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < nb; j += gridDim.x * blockDim.x) {
    float s[kMaxTopM];
    uint32_t c[kMaxTopM], f[kMaxTopM];
    int filled = 0;
    for (int r = 0; r < nranks; r++) {
      const uint32_t* g = gathered + (size_t)r * per_rank;
      for (int k = 0; k < top_m; k++) {
        const uint32_t kf = g[((size_t)2 * top_m + k) * cap + j];
        if (kf == 0xffffffffu) break;  // a rank's list is sorted: the empty slots are at its end
        const float ks = __uint_as_float(g[((size_t)0 * top_m + k) * cap + j]);
        const uint32_t kc = g[((size_t)1 * top_m + k) * cap + j];
        // insertion under the order of top_insert (pcp_kernels.cu): score descending, frame ascending
        int pos = filled;
        while (pos > 0 && (ks > s[pos - 1] || (ks == s[pos - 1] && kf < f[pos - 1]))) pos--;
        if (pos >= top_m) continue;
        const int last = filled < top_m ? filled : top_m - 1;
        for (int q = last; q > pos; q--) { s[q] = s[q - 1]; c[q] = c[q - 1]; f[q] = f[q - 1]; }
        s[pos] = ks; c[pos] = kc; f[pos] = kf;
        if (filled < top_m) filled++;
      }
    }
    for (int k = 0; k < top_m; k++) {
      const size_t o = (size_t)k * n + p0 + j;
      const bool have = k < filled;
      score[o] = have ? s[k] : 0.0f;
      rgbf[o] = have ? c[k] : 0u;
      frame[o] = have ? f[k] : 0xffffffffu;
    }
  }
}

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
  if (cudaSetDevice(ctx->device) != cudaSuccess) { ctx->fail("cudaSetDevice failed"); return PCP_ECUDA; }
  if (ctx->join() != cudaSuccess) { ctx->fail("pcp_allreduce: cudaStreamWaitEvent failed"); return PCP_ECUDA; }
  NcclComm comm = (NcclComm)ctx->nccl_comm;
  cudaStream_t st = ctx->stream;
  const uint32_t n = (uint32_t)ctx->n;
  const int grid = ctx->sm_count * 8;
  auto nccl_fail = [&](const char* what, int rc) { ctx->fail("%s: %s", what, api().error_string ? api().error_string(rc) : "error"); return PCP_ENCCL; };
  if (n == 0) return PCP_OK;

  if (ctx->prm.accum_mode == PCP_ACCUM_SUM) {
    const int rc = api().all_reduce(ctx->d_acc, ctx->d_acc, (size_t)n * 4, kNcclUint32, kNcclSum, comm, st);
    if (rc != 0) return nccl_fail("ncclAllReduce", rc);
    ctx->launches++;
  } else {
    // TOP5: all-gather the sorted slot lists block by block and merge.  Every rank calls this with the same cloud, hence the same
    // Morton permutation of the points: slot (k, i) means the same point everywhere.
    const int top_m = ctx->prm.top_m, G = ctx->nccl_nranks;
    uint32_t cap = 1u << 20;  // points per block: 3 * top_m * 4 MB sent, G times that received
    if (cap > n) cap = n;
    uint32_t* d_send = nullptr;
    uint32_t* d_recv = nullptr;
    const size_t per_rank = (size_t)3 * top_m * cap;
    if (cudaMalloc(&d_send, per_rank * sizeof(uint32_t)) != cudaSuccess || cudaMalloc(&d_recv, per_rank * sizeof(uint32_t) * G) != cudaSuccess) {
      if (d_send) cudaFree(d_send);
      ctx->fail("pcp_allreduce: out of device memory for the TOP5 exchange");
      return PCP_ENOMEM;
    }
    int rc = 0;
    for (uint32_t p0 = 0; p0 < n && rc == 0; p0 += cap) {
      const uint32_t nb = n - p0 < cap ? n - p0 : cap;
      top_pack_kernel<<<grid, 256, 0, st>>>(ctx->d_top_score, ctx->d_top_rgbf, ctx->d_top_frame, n, top_m, p0, nb, cap, d_send);
      rc = api().all_gather(d_send, d_recv, per_rank, kNcclUint32, comm, st);
      if (rc != 0) break;
      top_merge_kernel<<<grid, 256, 0, st>>>(d_recv, G, ctx->d_top_score, ctx->d_top_rgbf, ctx->d_top_frame, n, top_m, p0, nb, cap);
      ctx->launches += 3;
    }
    const cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(d_send);
    cudaFree(d_recv);
    if (rc != 0) return nccl_fail("ncclAllGather", rc);
    if (e != cudaSuccess) { ctx->fail("pcp_allreduce: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  }

  // mask accumulators exist only on ranks that were handed a mask: agree first (one u32, max), so that every rank makes the same calls
  uint32_t* d_flag = nullptr;
  if (cudaMalloc(&d_flag, sizeof(uint32_t)) != cudaSuccess) { ctx->fail("pcp_allreduce: cudaMalloc failed"); return PCP_ENOMEM; }
  uint32_t any_mask = ctx->d_mask_acc ? 1u : 0u;
  cudaMemcpyAsync(d_flag, &any_mask, sizeof(uint32_t), cudaMemcpyHostToDevice, st);
  {
    const int rc = api().all_reduce(d_flag, d_flag, 1, kNcclUint32, kNcclMax, comm, st);
    cudaMemcpyAsync(&any_mask, d_flag, sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    const cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(d_flag);
    if (rc != 0) return nccl_fail("ncclAllReduce (mask flag)", rc);
    if (e != cudaSuccess) { ctx->fail("pcp_allreduce: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  }
  if (any_mask && !ctx->d_mask_acc) {
    if (cudaMalloc(&ctx->d_mask_acc, sizeof(uint2) * (size_t)n) != cudaSuccess) { ctx->fail("pcp_allreduce: out of device memory for the mask accumulators"); return PCP_ENOMEM; }
    cudaMemsetAsync(ctx->d_mask_acc, 0, sizeof(uint2) * (size_t)n, st);
  }
  if (any_mask) {
    uint32_t* d_planes = nullptr;
    if (cudaMalloc(&d_planes, sizeof(uint32_t) * 2 * (size_t)n) != cudaSuccess) { ctx->fail("pcp_allreduce: out of device memory for the mask exchange"); return PCP_ENOMEM; }
    mask_split_kernel<<<grid, 256, 0, st>>>(ctx->d_mask_acc, d_planes, d_planes + n, n);
    int rc = api().all_reduce(d_planes, d_planes, n, kNcclUint32, kNcclSum, comm, st);
    if (rc == 0) rc = api().all_reduce(d_planes + n, d_planes + n, n, kNcclUint32, kNcclMax, comm, st);
    if (rc == 0) mask_join_kernel<<<grid, 256, 0, st>>>(ctx->d_mask_acc, d_planes, d_planes + n, n);
    const cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(d_planes);
    if (rc != 0) return nccl_fail("ncclAllReduce (mask)", rc);
    if (e != cudaSuccess) { ctx->fail("pcp_allreduce: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
    ctx->launches += 4;
  }
  return PCP_OK;
}

/* Global keyframe numbers for the frames this context submits: the next frame gets `next`, then next + stride, ...
 * (rank r of G with interleaved sharding: next = r, stride = G).  Restarts from `next` at pcp_set_cloud / pcp_reset_accumulators. */
int pcp_set_frame_numbering(pcp_ctx* ctx, int next, int stride) {
  if (!ctx || next < 0 || stride < 1) return PCP_EINVAL;
  ctx->frame_first = ctx->frame_index = next;
  ctx->frame_stride = stride;
  return PCP_OK;
}

}  // extern "C"
