// pcp_nccl.cu — the one exchange step of the path, after the frames have been sharded across ranks:
//   SUM accumulators   ncclAllReduce(u32, sum): integer sums are associative, bit-identical for any rank count;
//   mask accumulators  {hits, max} split into two planes, all-reduced with sum and max;
//   SUM + finalize     pcp_exchange_finalize*: ncclReduceScatter of the sums (every rank ends up owning the job's sums of 1/G of the
//                      points), finalize of that slice only, ncclAllGather of the packed colours + counts (8 B per point instead of
//                      the 16 B of the sums): 21 instead of 28 bytes per point cross NVLink and the finalize arithmetic is done once;
//   TOP5 slots         ncclAllGather of the ranks' sorted slot lists, block of points by block, and a merge kernel that keeps the
//                      top_m best of the union under the same total order (score descending, global frame number ascending) the
//                      insertion uses — so the merged lists equal the single-rank run's when frames carry their global numbers
//                      (pcp_set_frame_numbering).
// libnccl is resolved with dlopen at first use so that the library loads on boxes without NCCL.
#include <dlfcn.h>

#include <cstring>

#include "../../include/pcp/pcp.h"
#include "pcp_ctx.hpp"

using namespace pcp;

namespace {

struct NcclId { char internal[128]; };
typedef void* NcclComm;
typedef int (*fn_get_unique_id)(NcclId*);
typedef int (*fn_comm_init_rank)(NcclComm*, int, NcclId, int);
typedef int (*fn_all_reduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*fn_all_gather)(const void*, void*, size_t, int, NcclComm, cudaStream_t);
typedef int (*fn_reduce_scatter)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*fn_comm_destroy)(NcclComm);
typedef const char* (*fn_error_string)(int);

struct NcclApi {
  void* handle = nullptr;
  fn_get_unique_id get_unique_id = nullptr;
  fn_comm_init_rank comm_init_rank = nullptr;
  fn_all_reduce all_reduce = nullptr;
  fn_all_gather all_gather = nullptr;
  fn_reduce_scatter reduce_scatter = nullptr;
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
  a.reduce_scatter = (fn_reduce_scatter)dlsym(a.handle, "ncclReduceScatter");
  a.comm_destroy = (fn_comm_destroy)dlsym(a.handle, "ncclCommDestroy");
  a.error_string = (fn_error_string)dlsym(a.handle, "ncclGetErrorString");
  a.ok = a.get_unique_id && a.comm_init_rank && a.all_reduce && a.all_gather && a.reduce_scatter && a.comm_destroy;
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

// SUM: the rank's slice is packed after the reduce-scatter and everything is un-permuted after the all-gather by launch_pack_sum /
// launch_unpack_colours (pcp_kernels.cu), the same pair pcp_finalize runs on one GPU.

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

int nccl_all_gather_cloud(pcp_ctx* ctx, float4* d_cloud, size_t slice) {
  if (!ctx->nccl_comm || !api().ok) { ctx->fail("pcp_set_cloud_sharded: call pcp_nccl_init first"); return PCP_EINVAL; }
  const int rc = api().all_gather(d_cloud + (size_t)ctx->nccl_rank * slice, d_cloud, slice * 4, kNcclUint32, (NcclComm)ctx->nccl_comm, ctx->stream);
  if (rc != 0) { ctx->fail("ncclAllGather (cloud): %s", api().error_string ? api().error_string(rc) : "error"); return PCP_ENCCL; }
  return PCP_OK;
}

// mask accumulators {hits, max}: split into two planes, all-reduced with sum and max, joined (every rank makes the same calls)
static int exchange_masks(pcp_ctx* ctx) {
  NcclComm comm = (NcclComm)ctx->nccl_comm;
  cudaStream_t st = ctx->stream;
  const uint32_t n = (uint32_t)ctx->n;
  const int grid = ctx->sm_count * 8;
  auto nccl_fail = [&](const char* what, int rc) { ctx->fail("%s: %s", what, api().error_string ? api().error_string(rc) : "error"); return PCP_ENCCL; };
  // mask accumulators exist only on ranks that were handed a mask: agree first (one u32, max), so that every rank makes the same calls
  uint32_t* d_flag = nullptr;
  if (cudaMalloc(&d_flag, sizeof(uint32_t)) != cudaSuccess) { ctx->fail("pcp exchange: cudaMalloc failed"); return PCP_ENOMEM; }
  uint32_t any_mask = ctx->d_mask_acc ? 1u : 0u;
  cudaMemcpyAsync(d_flag, &any_mask, sizeof(uint32_t), cudaMemcpyHostToDevice, st);
  {
    const int rc = api().all_reduce(d_flag, d_flag, 1, kNcclUint32, kNcclMax, comm, st);
    cudaMemcpyAsync(&any_mask, d_flag, sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    const cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(d_flag);
    if (rc != 0) return nccl_fail("ncclAllReduce (mask flag)", rc);
    if (e != cudaSuccess) { ctx->fail("pcp exchange: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  }
  if (any_mask && !ctx->d_mask_acc) {
    if (cudaMalloc(&ctx->d_mask_acc, sizeof(uint2) * (size_t)n) != cudaSuccess) { ctx->fail("pcp exchange: out of device memory for the mask accumulators"); return PCP_ENOMEM; }
    cudaMemsetAsync(ctx->d_mask_acc, 0, sizeof(uint2) * (size_t)n, st);
  }
  if (any_mask) {
    uint32_t* d_planes = nullptr;
    if (cudaMalloc(&d_planes, sizeof(uint32_t) * 2 * (size_t)n) != cudaSuccess) { ctx->fail("pcp exchange: out of device memory for the mask exchange"); return PCP_ENOMEM; }
    mask_split_kernel<<<grid, 256, 0, st>>>(ctx->d_mask_acc, d_planes, d_planes + n, n);
    int rc = api().all_reduce(d_planes, d_planes, n, kNcclUint32, kNcclSum, comm, st);
    if (rc == 0) rc = api().all_reduce(d_planes + n, d_planes + n, n, kNcclUint32, kNcclMax, comm, st);
    if (rc == 0) mask_join_kernel<<<grid, 256, 0, st>>>(ctx->d_mask_acc, d_planes, d_planes + n, n);
    const cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(d_planes);
    if (rc != 0) return nccl_fail("ncclAllReduce (mask)", rc);
    if (e != cudaSuccess) { ctx->fail("pcp exchange: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
    ctx->launches += 4;
  }
  return PCP_OK;
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
  if (ctx->reduced) {  // a second exchange would add the other ranks' sums again (and duplicate TOP5 observations)
    ctx->fail("pcp_allreduce: the accumulators were already exchanged; call pcp_reset_accumulators before the next pass");
    return PCP_EINVAL;
  }
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
    // the plain sums / observation counts are kept in TOP5 mode too: exchange them as well, so that pcp_get_accumulators describes the job
    rc = api().all_reduce(ctx->d_acc, ctx->d_acc, (size_t)n * 4, kNcclUint32, kNcclSum, comm, st);
    if (rc != 0) return nccl_fail("ncclAllReduce", rc);
    ctx->launches++;
  }

  ctx->reduced = true;
  return exchange_masks(ctx);
}

/* Reduce-scatter + finalize of the rank's slice + all-gather of colours and counts (SUM accumulators); TOP5: pcp_allreduce + finalize.
 * Outputs are device pointers (either may be null).  Afterwards only this rank's slice of the sums is the job's: the accumulators
 * count as consumed until pcp_reset_accumulators / pcp_set_cloud. */
int pcp_exchange_finalize_device(pcp_ctx* ctx, uint8_t* d_rgb_out, uint32_t* d_count_out) {
  if (!ctx) return PCP_EINVAL;
  if (!ctx->d_acc) { ctx->fail("no cloud"); return PCP_EINVAL; }
  if (ctx->nccl_nranks <= 1 && !ctx->nccl_comm) return pcp_finalize_device(ctx, d_rgb_out, d_count_out);
  if (!ctx->nccl_comm) { ctx->fail("pcp_exchange_finalize: call pcp_nccl_init first"); return PCP_ENCCL; }
  if (ctx->prm.accum_mode != PCP_ACCUM_SUM) {
    const int rc = pcp_allreduce(ctx);
    return rc != PCP_OK ? rc : pcp_finalize_device(ctx, d_rgb_out, d_count_out);
  }
  if (ctx->reduced) {
    ctx->fail("pcp_exchange_finalize: the accumulators were already exchanged; call pcp_reset_accumulators before the next pass");
    return PCP_EINVAL;
  }
  if (cudaSetDevice(ctx->device) != cudaSuccess) { ctx->fail("cudaSetDevice failed"); return PCP_ECUDA; }
  if (ctx->join() != cudaSuccess) { ctx->fail("pcp_exchange_finalize: cudaStreamWaitEvent failed"); return PCP_ECUDA; }
  const uint32_t n = (uint32_t)ctx->n;
  if (n == 0) return PCP_OK;
  NcclComm comm = (NcclComm)ctx->nccl_comm;
  cudaStream_t st = ctx->stream;
  const uint32_t G = (uint32_t)ctx->nccl_nranks, r = (uint32_t)ctx->nccl_rank;
  const uint32_t slice = (n + G - 1) / G;  // points per rank; the accumulator array is padded to G * slice entries (pcp_ctx::kAccPad)
  if ((size_t)slice * G > ctx->n + pcp_ctx::kAccPad) { ctx->fail("pcp_exchange_finalize: more than %d ranks", (int)pcp_ctx::kAccPad); return PCP_EINVAL; }
  if (ctx->xchg_points < (size_t)slice * G) {
    if (ctx->d_xchg) { cudaFree(ctx->d_xchg); ctx->d_xchg = nullptr; }
    if (cudaMalloc(&ctx->d_xchg, sizeof(uint2) * (size_t)slice * G) != cudaSuccess) { ctx->fail("pcp_exchange_finalize: out of device memory"); return PCP_ENOMEM; }
    ctx->xchg_points = (size_t)slice * G;
  }
  auto nccl_fail = [&](const char* what, int rc) { ctx->fail("%s: %s", what, api().error_string ? api().error_string(rc) : "error"); return PCP_ENCCL; };
  uint32_t* acc_words = reinterpret_cast<uint32_t*>(ctx->d_acc);
  // optional phase timing (pcp_profile_enable): events after every phase, read with pcp_debug_exchange_times
  cudaEvent_t* xe = nullptr;
  if (ctx->profiling) {
    for (int i = 0; i < 5; i++)
      if (!ctx->xchg_ev[i] && cudaEventCreate(&ctx->xchg_ev[i]) != cudaSuccess) { ctx->fail("pcp_exchange_finalize: cudaEventCreate failed"); return PCP_ECUDA; }
    xe = ctx->xchg_ev;
    cudaEventRecord(xe[0], st);
  }
  int rc = api().reduce_scatter(acc_words, acc_words + (size_t)r * slice * 4, (size_t)slice * 4, kNcclUint32, kNcclSum, comm, st);  // in place
  if (xe) cudaEventRecord(xe[1], st);
  if (rc != 0) return nccl_fail("ncclReduceScatter", rc);
  const uint32_t j0 = r * slice, j1 = (r + 1) * slice < n ? (r + 1) * slice : n;
  launch_pack_sum(ctx->d_acc, j0, j1, ctx->d_xchg, ctx->prm.sample_mode == PCP_SAMPLE_BILINEAR, st);
  if (xe) cudaEventRecord(xe[2], st);
  rc = api().all_gather(ctx->d_xchg + (size_t)r * slice, ctx->d_xchg, (size_t)slice * 2, kNcclUint32, comm, st);  // in place
  if (xe) cudaEventRecord(xe[3], st);
  if (rc != 0) return nccl_fail("ncclAllGather", rc);
  launch_unpack_colours(ctx->d_xchg, ctx->d_perm ? ctx->d_inv : nullptr, n, d_rgb_out, d_count_out, st);
  if (xe) { cudaEventRecord(xe[4], st); ctx->xchg_timed = true; }
  ctx->launches += 4;
  ctx->reduced = true;
  ctx->acc_scattered = true;  // only this rank's slice of d_acc holds the job's sums now
  if (cudaGetLastError() != cudaSuccess) { ctx->fail("pcp_exchange_finalize: kernel launch failed"); return PCP_ECUDA; }
  return exchange_masks(ctx);
}

/* Debug: milliseconds of the last pcp_exchange_finalize[_device] call made while profiling was enabled (pcp_profile_enable):
 * {reduce-scatter, finalize of the slice, all-gather, unpack}.  Synchronises the context's stream. */
int pcp_debug_exchange_times(pcp_ctx* ctx, double ms4[4]) {
  if (!ctx || !ms4) return PCP_EINVAL;
  if (!ctx->xchg_timed) { ctx->fail("pcp_debug_exchange_times: no exchange was timed"); return PCP_EINVAL; }
  if (cudaSetDevice(ctx->device) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess) { ctx->fail("pcp_debug_exchange_times: synchronisation failed"); return PCP_ECUDA; }
  for (int i = 0; i < 4; i++) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ctx->xchg_ev[i], ctx->xchg_ev[i + 1]) != cudaSuccess) { ctx->fail("pcp_debug_exchange_times: cudaEventElapsedTime failed"); return PCP_ECUDA; }
    ms4[i] = ms;
  }
  return PCP_OK;
}

/* Same with host outputs (synchronises). */
int pcp_exchange_finalize(pcp_ctx* ctx, uint8_t* rgb_out, uint32_t* count_out) {
  if (!ctx) return PCP_EINVAL;
  if (!ctx->d_acc) { ctx->fail("no cloud"); return PCP_EINVAL; }
  const size_t n = ctx->n;
  if (n == 0) return PCP_OK;
  if (cudaSetDevice(ctx->device) != cudaSuccess) { ctx->fail("cudaSetDevice failed"); return PCP_ECUDA; }
  if (ctx->out_points < n) {
    if (ctx->d_out_rgb) { cudaFree(ctx->d_out_rgb); ctx->d_out_rgb = nullptr; }
    if (ctx->d_out_cnt) { cudaFree(ctx->d_out_cnt); ctx->d_out_cnt = nullptr; }
    if (cudaMalloc(&ctx->d_out_rgb, n * 3) != cudaSuccess || cudaMalloc(&ctx->d_out_cnt, n * sizeof(uint32_t)) != cudaSuccess) {
      ctx->fail("pcp_exchange_finalize: out of device memory");
      return PCP_ENOMEM;
    }
    ctx->out_points = n;
  }
  const int rc = pcp_exchange_finalize_device(ctx, rgb_out ? ctx->d_out_rgb : nullptr, count_out ? ctx->d_out_cnt : nullptr);
  if (rc != PCP_OK) return rc;
  cudaError_t e = cudaSuccess;
  if (rgb_out) e = cudaMemcpyAsync(rgb_out, ctx->d_out_rgb, n * 3, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && count_out) e = cudaMemcpyAsync(count_out, ctx->d_out_cnt, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) { ctx->fail("pcp_exchange_finalize: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
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
