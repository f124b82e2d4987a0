// pcp_capi.cu — the C ABI of include/pcp/pcp.h: context lifecycle, cloud ingest, per-frame submission, culling,
// finalize.  Host code only orchestrates; all arithmetic on points runs in the kernels of pcp_kernels.cu.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/pcp/pcp.h"
#include "pcp_ctx.hpp"

using namespace pcp;

static thread_local std::string g_create_error;

void pcp_ctx::fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  err = buf;
}

#define CU(call)                                                                             \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      ctx->fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return PCP_ECUDA;                                                                      \
    }                                                                                        \
  } while (0)

extern "C" {

void pcp_default_params(pcp_params* p) {
  if (!p) return;
  std::memset(p, 0, sizeof(*p));
  p->image_width = 4096;  // the reference camera (src/PointCloudProcessor.cpp:57-62, :525)
  p->image_height = 3000;
  p->cull_width = 4096;
  p->cull_height = 3000;
  p->downsample = 14;
  p->depth_tolerance = 0.05;
  p->enable_depth_buffer_culling = 1;
  p->accum_mode = PCP_ACCUM_SUM;
  p->sample_mode = PCP_SAMPLE_NEAREST;
  p->top_m = 5;
  p->frames_in_flight = 0;  // auto
  p->hsv_adjust = 0;
  p->saturation_scale = 1.0f;  // cpp:726-727
  p->brightness_scale = 1.0f;
  p->hsv_vector_block = 32;
  p->prebin = 1;  // Morton-sorted chunks with AABBs; prebin = 0 keeps the caller's order on the device
}

int pcp_abi_version(void) { return PCP_ABI_VERSION; }

const char* pcp_last_error(const pcp_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int pcp_create(pcp_ctx** out, int device, const pcp_camera* cam, const pcp_params* params) {
  if (!out || !cam || !params) { g_create_error = "pcp_create: null argument"; return PCP_EINVAL; }
  *out = nullptr;
  if (params->image_width <= 0 || params->image_height <= 0 || params->cull_width <= 0 || params->cull_height <= 0 ||
      params->downsample <= 0 || params->top_m < 1 || params->top_m > 16 ||
      (params->accum_mode != PCP_ACCUM_SUM && params->accum_mode != PCP_ACCUM_TOP5) ||
      (params->sample_mode != PCP_SAMPLE_NEAREST && params->sample_mode != PCP_SAMPLE_BILINEAR) || params->frames_in_flight < 0 ||
      params->frames_in_flight > pcp_ctx::kMaxFlight || params->hsv_vector_block < 0 ||
      (params->frames_per_launch != 0 && params->frames_per_launch != 1 && params->frames_per_launch != 2 && params->frames_per_launch != 4)) {
    g_create_error = "pcp_create: invalid params";
    return PCP_EINVAL;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_error = std::string("pcp_create: no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e);
    return PCP_ECUDA;
  }
  if (device < 0 || device >= ndev) { g_create_error = "pcp_create: bad device index"; return PCP_EINVAL; }
  pcp_ctx* ctx = new pcp_ctx();
  ctx->device = device;
  ctx->cam = *cam;
  ctx->prm = *params;
  ctx->Wc = params->cull_width / params->downsample;   // view_culling.cpp:64 integer division
  ctx->Hc = params->cull_height / params->downsample;
  auto bail = [&](const char* what, cudaError_t ce) {
    g_create_error = std::string("pcp_create: ") + what + ": " + cudaGetErrorString(ce);
    pcp_destroy(ctx);
    return PCP_ECUDA;
  };
  if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail("cudaGetDeviceProperties", e);
  ctx->sm_count = prop.multiProcessorCount;
  if ((e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
  ctx->stream = ctx->own_stream;
  if ((e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
  for (int i = 0; i < pcp_ctx::kSlots; i++) {
    if ((e = cudaEventCreateWithFlags(&ctx->ev_copied[i], cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
    if ((e = cudaEventCreateWithFlags(&ctx->ev_free[i], cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
  }
  ctx->flight_auto = params->frames_in_flight <= 0;  // decided per cloud in alloc_cloud_state
  ctx->flight = params->frames_in_flight > 0 ? params->frames_in_flight : 3;
  for (int i = 0; i < pcp_ctx::kMaxFlight; i++) {
    if ((e = cudaStreamCreateWithFlags(&ctx->fs[i].stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
    if ((e = cudaEventCreateWithFlags(&ctx->fs[i].ev_fork, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
    if ((e = cudaEventCreateWithFlags(&ctx->fs[i].ev_done, cudaEventDisableTiming)) != cudaSuccess) return bail("cudaEventCreate", e);
  }
  const size_t cells = (size_t)ctx->Wc * ctx->Hc;
  if ((e = cudaMalloc(&ctx->d_zgrid, sizeof(unsigned long long) * pcp_ctx::kMaxFlight * (cells ? cells : 1))) != cudaSuccess) return bail("cudaMalloc zgrid", e);
  if ((e = cudaMalloc(&ctx->d_zprev, sizeof(unsigned long long) * pcp_ctx::kMaxFlight * (cells ? cells : 1))) != cudaSuccess) return bail("cudaMalloc zprev", e);
  if ((e = cudaMalloc(&ctx->d_counters, 2 * pcp_ctx::kMaxFlight * sizeof(FrameCounters))) != cudaSuccess) return bail("cudaMalloc counters", e);
  // every grid empty (FLT_MAX, index 0xffffffff: no seed point before a slot's first frame), every counter block zero
  for (int i = 0; i < pcp_ctx::kMaxFlight; i++) {
    launch_zgrid_init(ctx->d_zgrid + (size_t)i * (cells ? cells : 1), (int)cells, ctx->d_counters + 2 * i, ctx->stream);
    launch_zgrid_init(ctx->d_zprev + (size_t)i * (cells ? cells : 1), (int)cells, ctx->d_counters + 2 * i + 1, ctx->stream);
  }
  if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) return bail("z-grid set-up", e);
  if ((e = cudaMallocHost(&ctx->h_counters, sizeof(FrameCounters))) != cudaSuccess) return bail("cudaMallocHost", e);
  if ((e = cudaMalloc(&ctx->d_total, sizeof(uint32_t))) != cudaSuccess) return bail("cudaMalloc", e);
  ctx->scan_grid = ctx->sm_count * scan_ctas_per_sm();
#ifndef PCP_RES_GRID_MULT
#define PCP_RES_GRID_MULT 4  // = the resident CTAs per SM of resolve_kernel: one persistent wave
#endif
  ctx->resolve_grid = ctx->sm_count * PCP_RES_GRID_MULT;
  ctx->prefilter = !(params->debug_flags & PCP_DEBUG_NO_PREFILTER);
  ctx->use_seed = !(params->debug_flags & PCP_DEBUG_NO_SEED);
  *out = ctx;
  return PCP_OK;
}

void pcp_destroy(pcp_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  ctx->drain();
  nccl_teardown(ctx);
  for (cudaEvent_t& e : ctx->xchg_ev) if (e) { cudaEventDestroy(e); e = nullptr; }
  for (int i = 0; i < pcp_ctx::kMaxFlight; i++) {
    if (ctx->fs[i].stream) cudaStreamDestroy(ctx->fs[i].stream);
    if (ctx->fs[i].ev_fork) cudaEventDestroy(ctx->fs[i].ev_fork);
    if (ctx->fs[i].ev_done) cudaEventDestroy(ctx->fs[i].ev_done);
  }
  ctx->free_cloud();
  for (int i = 0; i < pcp_ctx::kSlots; i++) {
    if (ctx->d_img[i]) cudaFree(ctx->d_img[i]);
    if (ctx->d_mask[i]) cudaFree(ctx->d_mask[i]);
    if (ctx->ev_copied[i]) cudaEventDestroy(ctx->ev_copied[i]);
    if (ctx->ev_free[i]) cudaEventDestroy(ctx->ev_free[i]);
  }
  if (ctx->d_zgrid) cudaFree(ctx->d_zgrid);
  if (ctx->d_zprev) cudaFree(ctx->d_zprev);
  if (ctx->d_counters) cudaFree(ctx->d_counters);
  if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
  if (ctx->d_total) cudaFree(ctx->d_total);
  if (ctx->d_tmp) cudaFree(ctx->d_tmp);
  for (cudaEvent_t e : ctx->prof_events) if (e) cudaEventDestroy(e);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  delete ctx;
}

int pcp_set_stream(pcp_ctx* ctx, void* cuda_stream) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  CU(ctx->drain());
  ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
  return PCP_OK;
}
void* pcp_get_stream(pcp_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int pcp_synchronize(pcp_ctx* ctx) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->copy_stream));
  CU(ctx->drain());
  return PCP_OK;
}

int pcp_host_alloc(void** out, size_t bytes) {
  if (!out) return PCP_EINVAL;
  return cudaMallocHost(out, bytes ? bytes : 1) == cudaSuccess ? PCP_OK : PCP_ENOMEM;
}
int pcp_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? PCP_OK : PCP_ECUDA; }

size_t pcp_num_points(const pcp_ctx* ctx) { return ctx ? ctx->n : 0; }
uint64_t pcp_launch_count(const pcp_ctx* ctx) { return ctx ? ctx->launches : 0; }

int pcp_pose_matrices(const double pose7[7], const double* T_opt16, float w2c16[16], float c2w16[16]) {
  if (!pose7 || !w2c16 || !c2w16) return PCP_EINVAL;
  host::M4f a, b;
  host::pose_matrices(pose7, T_opt16, a, b);
  std::memcpy(w2c16, a.a, sizeof(a.a));
  std::memcpy(c2w16, b.a, sizeof(b.a));
  return PCP_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// cloud
// ---------------------------------------------------------------------------------------------------------------------
// ctx->stream waits for the frames of EVERY slot: SUM frames on different slots are independent of each other (their resolve
// kernels are not chained), so the last submitted frame's event says nothing about the other slots.
cudaError_t pcp_ctx::join() {
  if (!last_done) return cudaSuccess;
  for (int i = 0; i < kMaxFlight; i++) {
    if (!fs[i].used) continue;
    const cudaError_t e = cudaStreamWaitEvent(stream, fs[i].ev_done, 0);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

void pcp_ctx::end_batch() {
  for (int i = 0; i < kMaxFlight; i++)
    if ((batch_pending >> i) & 1u) cudaEventRecord(fs[i].ev_done, fs[i].stream);
  batch_pending = 0;
  batch_forked = 0;
  in_batch = false;
}

cudaError_t pcp_ctx::drain() {
  cudaError_t e = cudaSuccess;
  for (int i = 0; i < kMaxFlight; i++)
    if (fs[i].stream) { const cudaError_t r = cudaStreamSynchronize(fs[i].stream); if (r != cudaSuccess) e = r; }
  if (stream) { const cudaError_t r = cudaStreamSynchronize(stream); if (r != cudaSuccess) e = r; }
  return e;
}

void pcp_ctx::free_cloud() {
  auto F = [](auto*& p) { if (p) { cudaFree(p); p = nullptr; } };
  F(d_cloud); F(d_perm); F(d_inv); F(d_inv_local); F(d_aabb); F(d_order); F(d_active); F(d_acc); F(d_mask_acc); F(d_top_score); F(d_top_rgbf); F(d_top_frame); F(d_records); F(d_vis_bits);
  F(d_pix); F(d_idx); F(d_scratch); F(d_out_rgb); F(d_out_cnt); F(d_xchg); F(d_aos);
  aos_bytes = 0;
  xchg_points = 0;
  inv_local = -1.0f;
  if (d_prebin_scratch) { cudaFree(d_prebin_scratch); d_prebin_scratch = nullptr; prebin_scratch_bytes = 0; }
  out_points = 0;
  n = 0;
}

// Sort the uploaded cloud into Morton chunks (pcp_prebin.cu).  Called after every upload, before the accumulators are reset.
// Frames in flight and frames per launch, when left to the library.  Small clouds are latency-bound (three dependent kernels per frame
// of a few us each) and their launches cost the submitting thread more than the GPU needs: twelve slots, four frames per launch
// (C1: 17.7 us per frame with 3 slots, 12.2 with 6).  Large clouds fill the GPU with three frames in flight.
static void choose_flight(pcp_ctx* ctx, size_t n) {
  const bool small = n <= 4000000;
  if (ctx->flight_auto) ctx->flight = small ? 12 : 3;
  ctx->group = ctx->prm.frames_per_launch > 0 ? ctx->prm.frames_per_launch : (small ? (ctx->flight >= 8 ? 4 : (ctx->flight >= 4 ? 2 : 1)) : 1);
  while (ctx->group > 1 && ctx->group > ctx->flight) ctx->group /= 2;
}

static int prebin_if_enabled(pcp_ctx* ctx, size_t n) {
  const bool on = ctx->prm.prebin != 0 && n > 0;
  if (!on) {
    if (ctx->d_perm) { cudaFree(ctx->d_perm); ctx->d_perm = nullptr; }
    if (ctx->d_inv) { cudaFree(ctx->d_inv); ctx->d_inv = nullptr; }
    if (ctx->d_inv_local) { cudaFree(ctx->d_inv_local); ctx->d_inv_local = nullptr; }
    if (ctx->d_aabb) { cudaFree(ctx->d_aabb); ctx->d_aabb = nullptr; }
    if (ctx->d_order) { cudaFree(ctx->d_order); ctx->d_order = nullptr; }
    if (ctx->d_active) { cudaFree(ctx->d_active); ctx->d_active = nullptr; }
    ctx->n_chunks = 0;
    return PCP_OK;
  }
  const uint32_t n_chunks = (uint32_t)((n + kChunk - 1) / kChunk);
  if (!ctx->d_perm) CU(cudaMalloc(&ctx->d_perm, sizeof(uint32_t) * n));
  if (!ctx->d_inv) CU(cudaMalloc(&ctx->d_inv, sizeof(uint32_t) * n));
  if (!ctx->d_aabb) CU(cudaMalloc(&ctx->d_aabb, sizeof(float4) * 2 * n_chunks));
  if (!ctx->d_order) CU(cudaMalloc(&ctx->d_order, sizeof(uint32_t) * n_chunks));
  if (!ctx->d_active) CU(cudaMalloc(&ctx->d_active, sizeof(uint4) * n_chunks * (size_t)ctx->flight));
  CU(prebin_cloud(ctx->d_cloud, (uint32_t)n, ctx->d_perm, ctx->d_aabb, ctx->d_order, &ctx->d_prebin_scratch, &ctx->prebin_scratch_bytes, ctx->sm_count,
                  ctx->stream));
  launch_invert_perm(ctx->d_perm, (uint32_t)n, ctx->d_inv, ctx->stream);
  if (!ctx->d_inv_local) CU(cudaMalloc(&ctx->d_inv_local, 2 * sizeof(uint32_t)));
  launch_inv_locality(ctx->d_inv, (uint32_t)n, ctx->d_inv_local, ctx->stream);
  ctx->inv_local = -1.0f;
  ctx->n_chunks = n_chunks;
  ctx->launches += 7;
  return PCP_OK;
}

static int alloc_cloud_state(pcp_ctx* ctx, size_t n) {
  const size_t n1 = n ? n : 1;
  CU(cudaMalloc(&ctx->d_acc, sizeof(uint4) * (n1 + pcp_ctx::kAccPad)));
  CU(cudaMalloc(&ctx->d_records, sizeof(Record) * (n1 + 64) * ctx->flight));
  if (ctx->prm.accum_mode == PCP_ACCUM_TOP5) {
    CU(cudaMalloc(&ctx->d_top_score, sizeof(float) * n1 * ctx->prm.top_m));
    CU(cudaMalloc(&ctx->d_top_rgbf, sizeof(uint32_t) * n1 * ctx->prm.top_m));
    CU(cudaMalloc(&ctx->d_top_frame, sizeof(uint32_t) * n1 * ctx->prm.top_m));
  }
  ctx->n = n;
  return pcp_reset_accumulators(ctx);
}

extern "C" {

/* The context's stream waits (asynchronously) for every frame submitted so far: work enqueued on it afterwards, e.g. an event
 * recorded by the caller, follows the frames' kernels. */
int pcp_join(pcp_ctx* ctx) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  CU(ctx->join());
  return PCP_OK;
}

int pcp_reset_accumulators(pcp_ctx* ctx) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  const size_t n1 = ctx->n ? ctx->n : 1;
  CU(ctx->join());  // frames in flight still update the accumulators
  if (ctx->d_acc) CU(cudaMemsetAsync(ctx->d_acc, 0, sizeof(uint4) * (n1 + pcp_ctx::kAccPad), ctx->stream));
  ctx->reduced = ctx->acc_scattered = false;
  if (ctx->d_mask_acc) CU(cudaMemsetAsync(ctx->d_mask_acc, 0, sizeof(uint2) * n1, ctx->stream));
  if (ctx->d_top_frame) {
    CU(cudaMemsetAsync(ctx->d_top_frame, 0xff, sizeof(uint32_t) * n1 * ctx->prm.top_m, ctx->stream));
    CU(cudaMemsetAsync(ctx->d_top_score, 0, sizeof(float) * n1 * ctx->prm.top_m, ctx->stream));
    CU(cudaMemsetAsync(ctx->d_top_rgbf, 0, sizeof(uint32_t) * n1 * ctx->prm.top_m, ctx->stream));
  }
  ctx->frame_index = ctx->frame_first;
  return PCP_OK;
}

// pcp_set_cloud and pcp_set_cloud_sharded.  `first` / `count`: the part of the cloud `points` holds (everything, or the rank's slice);
// slice > 0: all-gather the slices over NCCL afterwards (every rank's slice is `slice` points, the last one possibly shorter).
static int set_cloud_impl(pcp_ctx* ctx, const void* points, size_t n, size_t stride_bytes, size_t intensity_offset, size_t first, size_t count,
                          size_t slice) {
  CU(cudaSetDevice(ctx->device));
  const bool reuse = ctx->d_cloud && ctx->n == n && n > 0;  // same-size cloud: keep every allocation, only re-upload
  const size_t n1 = n ? n : 1;
  CU(ctx->join());  // frames in flight still read the cloud
  ctx->pix_valid = false;  // the kept pixel indices belong to a frame over the previous cloud
  if (!reuse) {
    CU(ctx->drain());
    ctx->free_cloud();
    CU(cudaMalloc(&ctx->d_cloud, sizeof(float4) * (n1 + pcp_ctx::kAccPad)));  // padded like the accumulators: nranks equal slices fit
  }
  if (count) {
    if (stride_bytes == 16 && intensity_offset == 12) {
      CU(cudaMemcpyAsync(ctx->d_cloud + first, points, sizeof(float4) * count, cudaMemcpyHostToDevice, ctx->stream));
    } else {
      // record layouts other than float4 (pcl::PointXYZI: 32 bytes, intensity at 16) are staged on the device and converted there; the
      // staging buffer is kept across uploads and nothing synchronises (the copy, the conversion and the pre-binning are stream-ordered)
      const size_t need = count * stride_bytes;
      if (ctx->aos_bytes < need) {
        if (ctx->d_aos) { CU(cudaStreamSynchronize(ctx->stream)); cudaFree(ctx->d_aos); ctx->d_aos = nullptr; ctx->aos_bytes = 0; }
        CU(cudaMalloc(&ctx->d_aos, need));
        ctx->aos_bytes = need;
      }
      CU(cudaMemcpyAsync(ctx->d_aos, points, need, cudaMemcpyHostToDevice, ctx->stream));
      launch_aos_to_soa(ctx->d_aos, count, stride_bytes, intensity_offset, ctx->d_cloud + first, ctx->stream);
      ctx->launches++;
    }
  }
  if (slice && n) {
    const int rc = nccl_all_gather_cloud(ctx, ctx->d_cloud, slice);
    if (rc != PCP_OK) return rc;
  }
  choose_flight(ctx, n);
  {
    const int rc = prebin_if_enabled(ctx, n);
    if (rc != PCP_OK) return rc;
  }
  if (reuse) return pcp_reset_accumulators(ctx);
  return alloc_cloud_state(ctx, n);
}

static int check_cloud_layout(pcp_ctx* ctx, bool points_missing, size_t n, size_t stride_bytes, size_t intensity_offset, const char* who) {
  if (points_missing || stride_bytes < 12 || (stride_bytes & 3) || (intensity_offset & 3) || intensity_offset + 4 > stride_bytes ||
      n >= 0xfffffff0ull) {
    ctx->fail("%s: invalid layout (n=%zu stride=%zu intensity_offset=%zu)", who, n, stride_bytes, intensity_offset);
    return PCP_EINVAL;
  }
  return PCP_OK;
}

int pcp_set_cloud(pcp_ctx* ctx, const void* points, size_t n, size_t stride_bytes, size_t intensity_offset) {
  if (!ctx) return PCP_EINVAL;
  const int rc = check_cloud_layout(ctx, !points && n, n, stride_bytes, intensity_offset, "pcp_set_cloud");
  if (rc != PCP_OK) return rc;
  return set_cloud_impl(ctx, points, n, stride_bytes, intensity_offset, 0, n, 0);
}

int pcp_cloud_slice(size_t n_total, int rank, int nranks, size_t* first, size_t* count) {
  if (nranks < 1 || rank < 0 || rank >= nranks || !first || !count) return PCP_EINVAL;
  const size_t slice = (n_total + (size_t)nranks - 1) / (size_t)nranks;
  const size_t f = (size_t)rank * slice < n_total ? (size_t)rank * slice : n_total;
  *first = f;
  *count = n_total - f < slice ? n_total - f : slice;
  return PCP_OK;
}

/* The map of a multi-GPU job crosses the host link ONCE in total: every rank uploads only its slice (pcp_cloud_slice) and the slices
 * are all-gathered over NVLink into every rank's full copy.  Collective: every rank of the communicator calls it with the same n_total. */
int pcp_set_cloud_sharded(pcp_ctx* ctx, const void* slice_points, size_t n_total, size_t stride_bytes, size_t intensity_offset) {
  if (!ctx) return PCP_EINVAL;
  if (!ctx->nccl_comm) { ctx->fail("pcp_set_cloud_sharded: call pcp_nccl_init first"); return PCP_EINVAL; }
  if (ctx->nccl_nranks > pcp_ctx::kAccPad) { ctx->fail("pcp_set_cloud_sharded: more than %d ranks", (int)pcp_ctx::kAccPad); return PCP_EINVAL; }
  size_t first = 0, count = 0;
  pcp_cloud_slice(n_total, ctx->nccl_rank, ctx->nccl_nranks, &first, &count);
  const int rc = check_cloud_layout(ctx, !slice_points && count, n_total, stride_bytes, intensity_offset, "pcp_set_cloud_sharded");
  if (rc != PCP_OK) return rc;
  const size_t slice = (n_total + (size_t)ctx->nccl_nranks - 1) / (size_t)ctx->nccl_nranks;
  return set_cloud_impl(ctx, slice_points, n_total, stride_bytes, intensity_offset, first, count, slice);
}

int pcp_set_cloud_device(pcp_ctx* ctx, const void* d_xyzi, size_t n) {
  if (!ctx || (!d_xyzi && n) || n >= 0xfffffff0ull) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  CU(ctx->drain());
  ctx->free_cloud();
  ctx->pix_valid = false;
  CU(cudaMalloc(&ctx->d_cloud, sizeof(float4) * ((n ? n : 1) + pcp_ctx::kAccPad)));
  if (n) CU(cudaMemcpyAsync(ctx->d_cloud, d_xyzi, sizeof(float4) * n, cudaMemcpyDeviceToDevice, ctx->stream));
  choose_flight(ctx, n);
  {
    const int rc = prebin_if_enabled(ctx, n);
    if (rc != PCP_OK) return rc;
  }
  return alloc_cloud_state(ctx, n);
}

}  // extern "C"
