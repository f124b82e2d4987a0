// pcp_frames.cu — per-frame submission (the keyframe loop body), ViewCulling::cull drop-in, accumulators, finalize.
#include <cmath>
#include <cstring>
#include <vector>

#include "../../include/pcp/pcp.h"
#include "pcp_ctx.hpp"

#ifdef PCP_HOST_TIMING  // tools/ variant only: where the submitting thread's time goes (printed at the end of every loop call)
#include <chrono>
#include <cstdio>
namespace { double g_ht[8]; uint64_t g_ht_frames; inline double ht_now() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); } }
#define HT(k, stmt) do { const double t0_ = ht_now(); stmt; g_ht[k] += ht_now() - t0_; } while (0)
#else
#define HT(k, stmt) stmt
#endif

using namespace pcp;

#define CU(call)                                                                             \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      ctx->fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return PCP_ECUDA;                                                                      \
    }                                                                                        \
  } while (0)

namespace {

void fill_xf(Xf34& x, const float m16[16]) {  // column-major 4x4 -> row-major 3x4
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 4; c++) x.r[r * 4 + c] = m16[c * 4 + r];
}
void fill_xd(Xd34& x, const double m16[16]) {
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 4; c++) x.r[r * 4 + c] = m16[c * 4 + r];
}

FrameParams base_params(const pcp_ctx* ctx) {
  FrameParams fp;
  std::memset(&fp, 0, sizeof(fp));
  const pcp_camera& c = ctx->cam;
  fp.cam = CamD{c.fx, c.fy, c.cx, c.cy, c.dist[0], c.dist[1], c.dist[2], c.dist[3], c.dist[4]};
  fp.tol = ctx->prm.depth_tolerance;
  fp.cullW = ctx->prm.cull_width;
  fp.cullH = ctx->prm.cull_height;
  fp.Wc = ctx->Wc;
  fp.Hc = ctx->Hc;
  fp.dsf = (float)ctx->prm.downsample;
  fp.imgW = ctx->prm.image_width;
  fp.imgH = ctx->prm.image_height;
  fp.depth_culling = ctx->prm.enable_depth_buffer_culling;
  // FP32 pre-filter constants: magnitudes are rounded away from zero so that the bound stays an upper bound
  auto up = [](double v) { float f = (float)std::fabs(v); return (double)f < std::fabs(v) ? std::nextafterf(f, INFINITY) : f; };
  PreF& q = fp.pre;
  q.fx = (float)c.fx; q.fy = (float)c.fy; q.cx = (float)c.cx; q.cy = (float)c.cy;
  q.k1 = (float)c.dist[0]; q.k2 = (float)c.dist[1]; q.p1 = (float)c.dist[2]; q.p2 = (float)c.dist[3]; q.k3 = (float)c.dist[4];
  q.ak1 = up(c.dist[0]); q.ak2 = up(c.dist[1]); q.ak3 = up(c.dist[4]);
  q.apt = up(3.0 * (std::fabs(c.dist[2]) + std::fabs(c.dist[3])) * (1.0 + 1e-6));
  q.afmax = up(std::fmax(std::fabs(c.fx), std::fabs(c.fy))); q.acmax = up(std::fmax(std::fabs(c.cx), std::fabs(c.cy)));
  const double ds = (double)ctx->prm.downsample;
  q.inv_ds = (float)(1.0 / ds); q.wc = (float)ctx->Wc; q.hc = (float)ctx->Hc; q.Wc = ctx->Wc;
  q.lo = -(float)ds;
  q.u_hi_cand = (float)(ds * ctx->prm.cull_width); q.v_hi_cand = (float)(ds * ctx->prm.cull_height);
  q.err = 64.0f * 5.9604644775390625e-08f;  // 64 * 2^-24
  q.slack = 1.0f / 32.0f;                     // pixels; also covers the 1-ulp fuzz of the float division by ds
  q.dscale = 1.0f - 8.0f * 5.9604644775390625e-08f;  // squared norm: 3 roundings + the final product < 5 eps
  {
    float t = (float)ctx->prm.depth_tolerance;
    if ((double)t < ctx->prm.depth_tolerance) t = std::nextafterf(t, INFINITY);
    q.tol_up = t >= 0.0f ? t : INFINITY;  // a negative tolerance disables the FP32 early-z (the squared compare needs lim >= 0)
  }
  fp.use_prefilter = ctx->prefilter ? 1 : 0;
  return fp;
}

int ensure_outputs(pcp_ctx* ctx, bool vis, bool pix) {
  const size_t n1 = ctx->n ? ctx->n : 1;
  if (vis && !ctx->d_vis_bits) CU(cudaMalloc(&ctx->d_vis_bits, sizeof(uint32_t) * ((n1 + 31) / 32)));
  if (pix && !ctx->d_pix) CU(cudaMalloc(&ctx->d_pix, sizeof(int32_t) * n1));
  return PCP_OK;
}

// Enqueue z-grid clear + classify/scan + resolve for one frame on its frame slot's stream (forked from ctx->stream).
// img_ready: event the resolve kernel must wait for (the image upload), or null; img_free: recorded after the resolve kernel.
// Arguments of one frame on slot `si`: the slot's z-grids, counter blocks (by parity), record buffer and chunk list.  Two z-grids and
// counter blocks per slot: this frame's pair was cleared by the slot's previous frame (or at set-up), the other pair holds that
// frame's final grid (the seed points) and is cleared by this frame's classify_kernel for the next one.
void fill_frame_args(pcp_ctx* ctx, int si, const FrameParams& fp, const uint8_t* d_bgr, size_t pitch, const uint8_t* d_mask, size_t mask_pitch,
                     bool want_vis, bool want_pix, bool accumulate, ScanArgs& sa, ResolveArgs& ra, int* counters_index) {
  const int cells = ctx->Wc * ctx->Hc;
  const size_t c1 = cells ? cells : 1;
  const int par = ctx->fs[si].parity;
  unsigned long long* zg = (par ? ctx->d_zprev : ctx->d_zgrid) + (size_t)si * c1;
  unsigned long long* zp = (par ? ctx->d_zgrid : ctx->d_zprev) + (size_t)si * c1;
  FrameCounters* ctr = ctx->d_counters + 2 * si + par;
  FrameCounters* ctr_next = ctx->d_counters + 2 * si + (par ^ 1);
  Record* recs = ctx->d_records + (size_t)si * ((ctx->n ? ctx->n : 1) + 64);
  const bool binned = ctx->d_perm != nullptr;
  if (binned) ctx->fs[si].parity ^= 1;
  *counters_index = 2 * si + par;
  std::memset(&sa, 0, sizeof(sa));
  sa.cloud = ctx->d_cloud;
  sa.perm = ctx->d_perm;
  sa.aabb = ctx->d_perm ? ctx->d_aabb : nullptr;
  sa.active = ctx->d_active ? ctx->d_active + (size_t)si * ctx->n_chunks : nullptr;
  sa.n_chunks = ctx->n_chunks;
  sa.order = ctx->d_order;
  sa.seed_blocks = binned ? ((cells > 16 ? cells : 16) + 255) / 256 : 0;
  sa.counters_next = ctr_next;
  sa.seed = (ctx->use_seed && fp.depth_culling) ? 1 : 0;
  sa.zgrid = zg;
  sa.zprev = zp;
  sa.cells = cells;
  sa.records = recs;
  sa.counters = ctr;
  sa.n = (uint32_t)ctx->n;
  std::memset(&ra, 0, sizeof(ra));
  ra.records = recs;
  ra.zgrid = zg;
  ra.counters = ctr;
  ra.perm = ctx->d_perm;
  ra.bgr = accumulate ? d_bgr : nullptr;
  ra.pitch = pitch;
  ra.mask = accumulate ? d_mask : nullptr;
  ra.mask_pitch = mask_pitch;
  ra.acc = ctx->d_acc;
  ra.mask_acc = ctx->d_mask_acc;
  ra.top_score = ctx->d_top_score;
  ra.top_rgbf = ctx->d_top_rgbf;
  ra.top_frame = ctx->d_top_frame;
  ra.vis_bits = want_vis ? ctx->d_vis_bits : nullptr;
  ra.pix_out = want_pix ? ctx->d_pix : nullptr;
  ra.n = (uint32_t)ctx->n;
  ra.top_m = ctx->prm.top_m;
  ra.accum_mode = ctx->prm.accum_mode;
  ra.sample_mode = ctx->prm.sample_mode;
}

// small clouds: fewer resolve CTAs (the records are a fraction of the points), SMs stay free for the frames on the other slots
int resolve_grid_for(const pcp_ctx* ctx) {
  int rgrid = ctx->resolve_grid;
  const int by_work = (int)(ctx->n / 4096);
  if (by_work < rgrid) rgrid = by_work > ctx->sm_count / 2 ? by_work : ctx->sm_count / 2;
  return rgrid;
}

// The resources of slot `si` (z-grids, lists, records) were last used by work on the stream of slot fs[si].owner: `st` waits for it.
int wait_for_owner(pcp_ctx* ctx, int si, cudaStream_t st) {
  const int o = ctx->fs[si].owner;
  if (o < 0 || ctx->fs[o].stream == st || !ctx->fs[o].used) return PCP_OK;
  if ((ctx->batch_pending >> o) & 1u) {  // its done-event was being saved up for the end of the loop call
    CU(cudaEventRecord(ctx->fs[o].ev_done, ctx->fs[o].stream));
    ctx->batch_pending &= ~(1u << o);
  }
  CU(cudaStreamWaitEvent(st, ctx->fs[o].ev_done, 0));
  return PCP_OK;
}

int run_frame(pcp_ctx* ctx, const FrameParams& fp, const uint8_t* d_bgr, size_t pitch, const uint8_t* d_mask, size_t mask_pitch,
              bool want_vis, bool want_pix, bool accumulate, bool count_candidates, cudaEvent_t img_ready = nullptr,
              cudaEvent_t img_free = nullptr) {
  if (!ctx->d_cloud) { ctx->fail("no cloud: call pcp_set_cloud first"); return PCP_EINVAL; }
  if (accumulate && ctx->acc_scattered) { ctx->fail("the accumulators were consumed by pcp_exchange_finalize: call pcp_reset_accumulators first"); return PCP_EINVAL; }
  int rc = ensure_outputs(ctx, want_vis, want_pix);
  if (rc != PCP_OK) return rc;
  if (d_mask && !ctx->d_mask_acc) {
    CU(cudaMalloc(&ctx->d_mask_acc, sizeof(uint2) * (ctx->n ? ctx->n : 1)));
    CU(cudaMemsetAsync(ctx->d_mask_acc, 0, sizeof(uint2) * (ctx->n ? ctx->n : 1), ctx->stream));
    ctx->batch_forked = 0;  // every slot has to see this memset
  }
  if (want_vis || want_pix) CU(ctx->join());  // the previous frame's resolve kernel may still write these shared buffers
  if (want_vis) CU(cudaMemsetAsync(ctx->d_vis_bits, 0, sizeof(uint32_t) * ((ctx->n + 31) / 32), ctx->stream));
  if (want_pix) CU(cudaMemsetAsync(ctx->d_pix, 0xff, sizeof(int32_t) * ctx->n, ctx->stream));
  // fork: everything enqueued on ctx->stream so far (cloud upload, accumulator reset, the memsets above) precedes this frame
  const int si = (int)(ctx->frames_run % (uint64_t)ctx->flight);
  pcp_ctx::FrameSlot& slot = ctx->fs[si];
  cudaStream_t st = slot.stream;
  const bool batch = ctx->in_batch && !want_vis && !want_pix && !ctx->profiling && !img_ready && !img_free;
  if (!batch || !((ctx->batch_forked >> si) & 1u)) {
    HT(0, CU(cudaEventRecord(slot.ev_fork, ctx->stream)); CU(cudaStreamWaitEvent(st, slot.ev_fork, 0)));
    if (batch) ctx->batch_forked |= 1u << si;
  }

  cudaEvent_t* pe = nullptr;
  if (ctx->profiling) {
    if (ctx->prof_used + 4 > ctx->prof_events.size()) {
      const size_t old_n = ctx->prof_events.size();
      ctx->prof_events.resize(old_n + 1024, nullptr);
      for (size_t i = old_n; i < ctx->prof_events.size(); i++) CU(cudaEventCreate(&ctx->prof_events[i]));
    }
    pe = &ctx->prof_events[ctx->prof_used];
    ctx->prof_used += 4;
  }
  {
    const int rc_o = wait_for_owner(ctx, si, st);
    if (rc_o != PCP_OK) return rc_o;
  }
  ScanArgs sa;
  ResolveArgs ra;
  int counters_index = 0;
  fill_frame_args(ctx, si, fp, d_bgr, pitch, d_mask, mask_pitch, want_vis, want_pix, accumulate, sa, ra, &counters_index);
  if (!ctx->d_perm) launch_zgrid_init(sa.zgrid, sa.cells, sa.counters, st);  // unbinned path: no classify_kernel, the grid is cleared per frame
  if (pe) CU(cudaEventRecord(pe[1], st));
  int scan_rc = 0;
  HT(1, scan_rc = launch_scan(sa, fp, ctx->scan_grid, ctx->sm_count, count_candidates, ctx->device, st));
  if (scan_rc != 0) { ctx->fail("pass 1: kernel launch failed: %s", cudaGetErrorString(cudaGetLastError())); return PCP_ECUDA; }
  if (pe) CU(cudaEventRecord(pe[2], st));
  // TOP5 slot lists are read-modify-write: those resolve kernels run in frame order.  SUM accumulators are updated with
  // atomics (integer sums commute): frames on different slots are fully independent.
  if (ctx->last_done && ctx->flight > 1 && ctx->prm.accum_mode != PCP_ACCUM_SUM) CU(cudaStreamWaitEvent(st, ctx->last_done, 0));
  if (img_ready) CU(cudaStreamWaitEvent(st, img_ready, 0));

  const int rgrid = resolve_grid_for(ctx);
  HT(2, launch_resolve(ra, fp, rgrid, st));
  if (pe) CU(cudaEventRecord(pe[3], st));
  if (batch && ctx->prm.accum_mode == PCP_ACCUM_SUM) ctx->batch_pending |= 1u << si;  // recorded once, by pcp_ctx::end_batch
  else HT(3, CU(cudaEventRecord(slot.ev_done, st)));
  slot.used = true;
  if (img_free) CU(cudaEventRecord(img_free, st));
  ctx->last_done = slot.ev_done;
  ctx->last_slot = si;
  ctx->last_counters = counters_index;
  ctx->pix_valid = want_pix && accumulate;
  ctx->mask_idx_valid = false;
  ctx->last_bgr = d_bgr; ctx->last_pitch = pitch; ctx->last_mask = d_mask; ctx->last_mask_pitch = mask_pitch; ctx->last_w2c = fp.w2c;
  slot.owner = si;
  ctx->frames_run++;
  ctx->launches += 3;  // classify + scan + resolve, or clear + scan + resolve (unbinned)
  CU(cudaGetLastError());
  return PCP_OK;
}

// k = 2 or 4 consecutive frames of a loop call in ONE classify, ONE scan and ONE resolve launch (blockIdx.y = frame), on the slots
// base .. base + k - 1 and the stream of slot `base`.  SUM mode, pre-binned cloud, no per-frame outputs.  For small clouds the three
// launches of a frame cost the submitting thread more (~9 us) than the GPU needs for the frame: see DESIGN 4.5.
int run_group(pcp_ctx* ctx, int k, const FrameParams* fps, const uint8_t* const* d_bgr, size_t pitch, const uint8_t* const* d_mask, size_t mask_pitch) {
  if (!ctx->d_cloud) { ctx->fail("no cloud: call pcp_set_cloud first"); return PCP_EINVAL; }
  if (ctx->acc_scattered) { ctx->fail("the accumulators were consumed by pcp_exchange_finalize: call pcp_reset_accumulators first"); return PCP_EINVAL; }
  bool any_mask = false;
  for (int j = 0; j < k; j++) any_mask = any_mask || (d_mask && d_mask[j]);
  if (any_mask && !ctx->d_mask_acc) {
    CU(cudaMalloc(&ctx->d_mask_acc, sizeof(uint2) * (ctx->n ? ctx->n : 1)));
    CU(cudaMemsetAsync(ctx->d_mask_acc, 0, sizeof(uint2) * (ctx->n ? ctx->n : 1), ctx->stream));
    ctx->batch_forked = 0;  // every slot has to see this memset
  }
  const int base = (int)(ctx->frames_run % (uint64_t)ctx->flight);
  pcp_ctx::FrameSlot& slot = ctx->fs[base];
  cudaStream_t st = slot.stream;
  if (!((ctx->batch_forked >> base) & 1u)) {
    CU(cudaEventRecord(slot.ev_fork, ctx->stream));
    CU(cudaStreamWaitEvent(st, slot.ev_fork, 0));
    ctx->batch_forked |= 1u << base;
  }
  ScanArgs sa[kGroupMax];
  ResolveArgs ra[kGroupMax];
  int counters_index = 0;
  for (int j = 0; j < k; j++) {
    const int rc = wait_for_owner(ctx, base + j, st);
    if (rc != PCP_OK) return rc;
    fill_frame_args(ctx, base + j, fps[j], d_bgr[j], pitch, d_mask ? d_mask[j] : nullptr, mask_pitch, false, false, true, sa[j], ra[j], &counters_index);
  }
  if (launch_scan_group(sa, fps, k, ctx->sm_count, false, ctx->device, st) != 0) { ctx->fail("pass 1: kernel launch failed: %s", cudaGetErrorString(cudaGetLastError())); return PCP_ECUDA; }
  if (launch_resolve_group(ra, fps, k, resolve_grid_for(ctx), st) != 0) { ctx->fail("resolve: kernel launch failed"); return PCP_ECUDA; }
  ctx->batch_pending |= 1u << base;  // the done-event of the group's stream is recorded once, by pcp_ctx::end_batch
  slot.used = true;
  for (int j = 0; j < k; j++) ctx->fs[base + j].owner = base;
  ctx->last_done = slot.ev_done;
  ctx->last_slot = base + k - 1;
  ctx->last_counters = counters_index;
  ctx->pix_valid = false;
  ctx->mask_idx_valid = false;
  ctx->frames_run += (uint64_t)k;
  ctx->launches += 3;
  CU(cudaGetLastError());
  return PCP_OK;
}

int fetch_outputs(pcp_ctx* ctx, pcp_frame_out* out) {
  CU(ctx->join());
  CU(cudaMemcpyAsync(ctx->h_counters, ctx->d_counters + ctx->last_counters, sizeof(FrameCounters), cudaMemcpyDeviceToHost, ctx->stream));
  if (out->visible_bits) CU(cudaMemcpyAsync(out->visible_bits, ctx->d_vis_bits, sizeof(uint32_t) * ((ctx->n + 31) / 32), cudaMemcpyDeviceToHost, ctx->stream));
  if (out->pixel_index) CU(cudaMemcpyAsync(out->pixel_index, ctx->d_pix, sizeof(int32_t) * ctx->n, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  out->n_candidates = (out->flags & PCP_OUT_SKIP_CANDIDATES) ? -1 : (int64_t)ctx->h_counters->n_candidates;
  out->n_visible = ctx->h_counters->n_visible;
  out->n_colored = ctx->h_counters->n_colored;
  return PCP_OK;
}

int frame_params_for_pose(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, FrameParams& fp) {
  if (!pose7) { ctx->fail("null pose"); return PCP_EINVAL; }
  fp = base_params(ctx);
  host::M4f w2c, c2w;
  host::pose_matrices(pose7, T_opt16, w2c, c2w);  // cpp:495-519
  fill_xf(fp.w2c, w2c.a);
  fp.use_w2c = 1;
  fp.use_T = 0;                                   // cull(cloudInCameraPose, Identity)  cpp:527
  fp.cam_pos[0] = pose7[0]; fp.cam_pos[1] = pose7[1]; fp.cam_pos[2] = pose7[2];  // hpp:208
  fp.frame_index = ctx->frame_index;
  return PCP_OK;
}

}  // namespace

static int submit_device(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, const uint8_t* d_bgr, size_t pitch,
                         const uint8_t* d_mask, size_t mask_pitch, pcp_frame_out* out, cudaEvent_t img_ready, cudaEvent_t img_free) {
  if (!ctx) return PCP_EINVAL;
  if (!d_bgr || pitch < (size_t)ctx->prm.image_width * 3 || (d_mask && mask_pitch < (size_t)ctx->prm.image_width)) {
    ctx->fail("pcp_submit_frame: bad image/pitch");
    return PCP_EINVAL;
  }
  CU(cudaSetDevice(ctx->device));
  FrameParams fp;
  int rc = PCP_OK;
  HT(4, rc = frame_params_for_pose(ctx, pose7, T_opt16, fp));
  if (rc != PCP_OK) return rc;
  HT(5, rc = run_frame(ctx, fp, d_bgr, pitch, d_mask, mask_pitch, out && out->visible_bits, out && (out->pixel_index || (out->flags & PCP_OUT_PIXELS_ON_DEVICE)), true,
                       out && !(out->flags & PCP_OUT_SKIP_CANDIDATES), img_ready, img_free));
  if (rc != PCP_OK) return rc;
  ctx->frame_index += ctx->frame_stride;
  if (out) return fetch_outputs(ctx, out);
  return PCP_OK;
}

// SUM finalize on ctx->stream.  With a pre-binned cloud the sums (sorted order) and the outputs (caller's order) are a permutation apart
// and the cost is the gather's locality (tools/bench_finalize.py):
//   * caller's order spatially coherent (a voxel-grid filter's output, a scan-ordered map): one kernel gathers the 16-byte sums, 69 us
//     at 1e7 points and 656 us at 1e8 = 0.60 / 0.64 of the HBM peak on 27 B per point (packing first: 102 / 1023 us);
//   * caller's order random (the synthetic clouds): every gather is its own DRAM sector and, past a few hundred MB, its own TLB
//     entry.  Up to 32 M points it pays to pack the sums to 8 bytes per point first (coalesced) and gather from that array, which is
//     mostly still in L2: 139 against 156 us at 1e7 points, 620 / 654 at 3e7, 1212 / 1161 at 5e7, 3141 / 2448 at 1e8.
// Which case a cloud is was measured when it was pre-binned (inv_locality_kernel).
constexpr size_t kPackedMaxBytes = 256u << 20;
static int finalize_sum_on_stream(pcp_ctx* ctx, uint8_t* d_rgb, uint32_t* d_cnt) {
  const uint32_t n = (uint32_t)ctx->n;
  const int fixed88 = ctx->prm.sample_mode == PCP_SAMPLE_BILINEAR;
  const int dbg = ctx->prm.debug_flags;
  if (ctx->d_perm && ctx->inv_local < 0.0f) {
    uint32_t h[2] = {0, 0};
    cudaError_t e = cudaMemcpyAsync(h, ctx->d_inv_local, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { ctx->fail("pcp_finalize: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
    ctx->inv_local = h[1] ? (float)h[0] / (float)h[1] : 1.0f;
  }
  const bool random_order = ctx->d_perm && ctx->inv_local < 0.5f;
  const bool two_pass = ctx->d_perm && !(dbg & PCP_DEBUG_FINALIZE_ONEPASS) &&
                        ((dbg & PCP_DEBUG_FINALIZE_TWOPASS) || (random_order && sizeof(uint2) * ctx->n <= kPackedMaxBytes));
  if (!two_pass) {
    launch_finalize_sum(ctx->d_acc, ctx->d_perm ? ctx->d_inv : nullptr, n, d_rgb, d_cnt, fixed88, ctx->stream);
    ctx->launches++;
    return PCP_OK;
  }
  if (ctx->xchg_points < ctx->n) {
    if (ctx->d_xchg) { cudaFree(ctx->d_xchg); ctx->d_xchg = nullptr; ctx->xchg_points = 0; }
    if (cudaMalloc(&ctx->d_xchg, sizeof(uint2) * ctx->n) != cudaSuccess) { ctx->fail("pcp_finalize: out of device memory (%zu bytes)", sizeof(uint2) * ctx->n); return PCP_ENOMEM; }
    ctx->xchg_points = ctx->n;
  }
  launch_pack_sum(ctx->d_acc, 0, n, ctx->d_xchg, fixed88, ctx->stream);
  launch_unpack_colours(ctx->d_xchg, ctx->d_inv, n, d_rgb, d_cnt, ctx->stream);
  ctx->launches += 2;
  return PCP_OK;
}

extern "C" {

int pcp_submit_frame_device(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, const uint8_t* d_bgr, size_t pitch,
                            const uint8_t* d_mask, size_t mask_pitch, pcp_frame_out* out) {
  return submit_device(ctx, pose7, T_opt16, d_bgr, pitch, d_mask, mask_pitch, out, nullptr, nullptr);
}

/* The keyframe loop (cpp:488-595) over device-resident images in one call: frame i uses poses7[7 i .. 7 i + 6] and the image
 * d_bgr[i] (mask d_mask[i] when d_mask is not null).  Same work and results as n_frames calls of pcp_submit_frame_device without
 * per-frame outputs; only the per-call host overhead is paid once. */
int pcp_submit_frames_device(pcp_ctx* ctx, size_t n_frames, const double* poses7, const double* T_opt16, const uint8_t* const* d_bgr, size_t pitch,
                             const uint8_t* const* d_mask, size_t mask_pitch) {
  if (!ctx) return PCP_EINVAL;
  if (n_frames && (!poses7 || !d_bgr)) { ctx->fail("pcp_submit_frames_device: null argument"); return PCP_EINVAL; }
  if (n_frames && (pitch < (size_t)ctx->prm.image_width * 3 || (d_mask && mask_pitch < (size_t)ctx->prm.image_width))) {
    ctx->fail("pcp_submit_frames_device: bad image/pitch");
    return PCP_EINVAL;
  }
  CU(cudaSetDevice(ctx->device));
  ctx->in_batch = true;
  ctx->batch_forked = 0;
  ctx->batch_pending = 0;
  const int group = (ctx->prm.accum_mode == PCP_ACCUM_SUM && ctx->d_perm && !ctx->profiling) ? ctx->group : 1;
  int rc = PCP_OK;
  size_t i = 0;
  while (i < n_frames && rc == PCP_OK) {
    const int base = (int)(ctx->frames_run % (uint64_t)ctx->flight);
    int k = 1;
    if (group >= 4 && base % 4 == 0 && base + 4 <= ctx->flight && n_frames - i >= 4) k = 4;
    else if (group >= 2 && base % 2 == 0 && base + 2 <= ctx->flight && n_frames - i >= 2) k = 2;
    if (k == 1) {
      rc = submit_device(ctx, poses7 + 7 * i, T_opt16, d_bgr[i], pitch, d_mask ? d_mask[i] : nullptr, mask_pitch, nullptr, nullptr, nullptr);
    } else {
      FrameParams fps[kGroupMax];
      const int frame_index0 = ctx->frame_index;
      for (int j = 0; j < k && rc == PCP_OK; j++) {
        if (!d_bgr[i + j]) { ctx->fail("pcp_submit_frames_device: null image"); rc = PCP_EINVAL; break; }
        rc = frame_params_for_pose(ctx, poses7 + 7 * (i + j), T_opt16, fps[j]);
        ctx->frame_index += ctx->frame_stride;
      }
      if (rc == PCP_OK) HT(5, rc = run_group(ctx, k, fps, d_bgr + i, pitch, d_mask ? d_mask + i : nullptr, mask_pitch));
      if (rc != PCP_OK) ctx->frame_index = frame_index0;  // none of the group's frames was submitted
    }
    i += (size_t)k;
  }
  HT(6, ctx->end_batch());  // also after an error: the frames submitted so far stay joinable
#ifdef PCP_HOST_TIMING
  g_ht_frames += n_frames;
  if (g_ht_frames >= 1000) {
    std::fprintf(stderr, "[host us per frame] fork %.2f | classify+scan launch %.2f | resolve launch %.2f | done event %.2f | pose->params %.2f | run_frame total %.2f | end_batch %.2f (%llu frames; sizeof FrameParams %zu ScanArgs %zu ResolveArgs %zu)\n",
                 g_ht[0] / g_ht_frames, g_ht[1] / g_ht_frames, g_ht[2] / g_ht_frames, g_ht[3] / g_ht_frames, g_ht[4] / g_ht_frames, g_ht[5] / g_ht_frames, g_ht[6] / g_ht_frames,
                 (unsigned long long)g_ht_frames, sizeof(FrameParams), sizeof(ScanArgs), sizeof(ResolveArgs));
    for (double& v : g_ht) v = 0;
    g_ht_frames = 0;
  }
#endif
  return rc;
}

int pcp_submit_frame(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, const uint8_t* bgr, size_t pitch,
                     const uint8_t* mask, size_t mask_pitch, pcp_frame_out* out) {
  if (!ctx) return PCP_EINVAL;
  const size_t W = ctx->prm.image_width, H = ctx->prm.image_height;
  if (!bgr || pitch < W * 3 || (mask && mask_pitch < W)) { ctx->fail("pcp_submit_frame: bad image/pitch"); return PCP_EINVAL; }
  CU(cudaSetDevice(ctx->device));
  const int s = ctx->slot;
  ctx->slot = (ctx->slot + 1) % pcp_ctx::kSlots;
  if (!ctx->d_img[s]) CU(cudaMalloc(&ctx->d_img[s], W * 3 * H));
  if (mask && !ctx->d_mask[s]) CU(cudaMalloc(&ctx->d_mask[s], W * H));
  // copy stream: wait until the kernels that last read this slot are done, then stage the image
  CU(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_free[s], 0));
  CU(cudaMemcpy2DAsync(ctx->d_img[s], W * 3, bgr, pitch, W * 3, H, cudaMemcpyHostToDevice, ctx->copy_stream));
  if (mask) CU(cudaMemcpy2DAsync(ctx->d_mask[s], W, mask, mask_pitch, W, H, cudaMemcpyHostToDevice, ctx->copy_stream));
  if (ctx->prm.hsv_adjust) {  // cpp:722-741, in place, on the copy stream (overlaps the scan of this frame)
    launch_hsv_roundtrip(ctx->d_img[s], W * 3, ctx->d_img[s], W * 3, (int)W, (int)H, ctx->prm.saturation_scale, ctx->prm.brightness_scale,
                         ctx->prm.hsv_vector_block, ctx->sm_count, ctx->copy_stream);
    ctx->launches++;
  }
  CU(cudaEventRecord(ctx->ev_copied[s], ctx->copy_stream));
  // only the resolve kernel reads the image: the scan of this frame does not wait for the upload
  return submit_device(ctx, pose7, T_opt16, ctx->d_img[s], W * 3, mask ? ctx->d_mask[s] : nullptr, W, out, ctx->ev_copied[s], ctx->ev_free[s]);
}

int pcp_hsv_roundtrip(pcp_ctx* ctx, const uint8_t* bgr_in, size_t pitch_in, uint8_t* bgr_out, size_t pitch_out, int width, int height,
                      float saturation_scale, float brightness_scale) {
  if (!ctx) return PCP_EINVAL;
  if (!bgr_in || !bgr_out || width <= 0 || height <= 0 || pitch_in < (size_t)width * 3 || pitch_out < (size_t)width * 3) {
    ctx->fail("pcp_hsv_roundtrip: bad image/pitch");
    return PCP_EINVAL;
  }
  CU(cudaSetDevice(ctx->device));
  uint8_t* d = nullptr;
  const size_t row = (size_t)width * 3;
  CU(cudaMalloc(&d, row * height));
  cudaError_t e = cudaMemcpy2DAsync(d, row, bgr_in, pitch_in, row, height, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) {
    launch_hsv_roundtrip(d, row, d, row, width, height, saturation_scale, brightness_scale, ctx->prm.hsv_vector_block, ctx->sm_count, ctx->stream);
    ctx->launches++;
    e = cudaMemcpy2DAsync(bgr_out, pitch_out, d, row, row, height, cudaMemcpyDeviceToHost, ctx->stream);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d);
  if (e != cudaSuccess) { ctx->fail("pcp_hsv_roundtrip: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  return PCP_OK;
}

static int cull_common(pcp_ctx* ctx, const FrameParams& fp, int32_t* idx_out, size_t* m_out) {
  if (!idx_out || !m_out) { ctx->fail("pcp_cull: null output"); return PCP_EINVAL; }
  int rc = run_frame(ctx, fp, nullptr, 0, nullptr, 0, true, false, false, false);
  if (rc != PCP_OK) return rc;
  CU(ctx->join());
  const size_t n1 = ctx->n ? ctx->n : 1;
  if (!ctx->d_idx) CU(cudaMalloc(&ctx->d_idx, sizeof(int32_t) * n1));
  if (!ctx->d_scratch) CU(cudaMalloc(&ctx->d_scratch, sizeof(uint32_t) * compact_scratch_words((uint32_t)ctx->n)));
  ctx->mask_idx_valid = false;  // d_idx is shared with pcp_frame_mask_cloud
  launch_compact_bits(ctx->d_vis_bits, (uint32_t)ctx->n, nullptr, ctx->d_scratch, ctx->d_idx, ctx->d_total, ctx->stream);
  ctx->launches += 3;
  uint32_t total = 0;
  CU(cudaMemcpyAsync(&total, ctx->d_total, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  if (total) CU(cudaMemcpyAsync(idx_out, ctx->d_idx, sizeof(int32_t) * total, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  *m_out = total;
  return PCP_OK;
}

int pcp_cull(pcp_ctx* ctx, const float* w2c16, const double* T16, int32_t* idx_out, size_t* m_out) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  FrameParams fp = base_params(ctx);
  if (w2c16) { fill_xf(fp.w2c, w2c16); fp.use_w2c = 1; }
  if (T16) { fill_xd(fp.T, T16); fp.use_T = 1; fp.use_prefilter = 0; }  // the gate is taken after the double T: exact path only
  return cull_common(ctx, fp, idx_out, m_out);
}

int pcp_cull_pose(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, int32_t* idx_out, size_t* m_out) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  FrameParams fp;
  const int rc = frame_params_for_pose(ctx, pose7, T_opt16, fp);
  if (rc != PCP_OK) return rc;
  return cull_common(ctx, fp, idx_out, m_out);
}

int pcp_splat_masks_pose(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, float* distance_out, float* xyz_out, int32_t* index_out,
                         size_t* n_filled_out) {
  if (!ctx) return PCP_EINVAL;
  if (!distance_out) { ctx->fail("pcp_splat_masks_pose: null distance_out"); return PCP_EINVAL; }
  CU(cudaSetDevice(ctx->device));
  FrameParams fp;
  int rc = frame_params_for_pose(ctx, pose7, T_opt16, fp);
  if (rc != PCP_OK) return rc;
  // the frame's culled cloud and its pixels: pass 1 + resolve without image (nothing is accumulated)
  rc = run_frame(ctx, fp, nullptr, 0, nullptr, 0, false, true, false, false);
  if (rc != PCP_OK) return rc;
  CU(ctx->join());
  const size_t npix = (size_t)fp.imgW * fp.imgH;
  // one allocation: winner keys | distance | xyz | index | counter
  const size_t off_dist = npix * 8, off_xyz = off_dist + npix * 4, off_idx = off_xyz + npix * 12, off_cnt = off_idx + npix * 4;
  uint8_t* d = nullptr;
  CU(cudaMalloc(&d, off_cnt + 16));
  cudaError_t e = cudaMemsetAsync(d, 0, off_dist, ctx->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(d + off_cnt, 0, 16, ctx->stream);
  if (e == cudaSuccess) {
    launch_splat(ctx->d_pix, ctx->d_perm, (uint32_t)ctx->n, ctx->d_cloud, fp, reinterpret_cast<unsigned long long*>(d),
                 reinterpret_cast<float*>(d + off_dist), xyz_out ? reinterpret_cast<float*>(d + off_xyz) : nullptr,
                 index_out ? reinterpret_cast<int32_t*>(d + off_idx) : nullptr, reinterpret_cast<unsigned int*>(d + off_cnt), ctx->sm_count,
                 ctx->stream);
    ctx->launches += ctx->n ? 2 : 1;
    e = cudaGetLastError();
  }
  unsigned int filled = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(distance_out, d + off_dist, npix * 4, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && xyz_out) e = cudaMemcpyAsync(xyz_out, d + off_xyz, npix * 12, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && index_out) e = cudaMemcpyAsync(index_out, d + off_idx, npix * 4, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(&filled, d + off_cnt, sizeof(filled), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d);
  if (e != cudaSuccess) { ctx->fail("pcp_splat_masks_pose: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  if (n_filled_out) *n_filled_out = filled;
  return PCP_OK;
}

int pcp_get_accumulators(pcp_ctx* ctx, uint32_t* acc4, uint32_t* mask_hits, uint32_t* mask_max) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  if (!ctx->d_acc) { ctx->fail("no cloud"); return PCP_EINVAL; }
  if (acc4 && ctx->acc_scattered) { ctx->fail("pcp_get_accumulators: the colour sums were consumed by pcp_exchange_finalize (the mask accumulators are still available)"); return PCP_EINVAL; }
  if (ctx->n == 0) return PCP_OK;
  CU(ctx->join());
  // per-point state lives in sorted order when the cloud is pre-binned: hand it back in the caller's order
  uint32_t* d_tmp = nullptr;
  if (ctx->d_perm) CU(cudaMalloc(&d_tmp, sizeof(uint4) * ctx->n));
  if (acc4) {
    const void* src = ctx->d_acc;
    if (ctx->d_perm) { launch_unpermute(reinterpret_cast<const uint32_t*>(ctx->d_acc), ctx->d_perm, (uint32_t)ctx->n, 4, d_tmp, ctx->stream); src = d_tmp; }
    CU(cudaMemcpyAsync(acc4, src, sizeof(uint4) * ctx->n, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CU(cudaStreamSynchronize(ctx->stream));
  if (mask_hits || mask_max) {
    std::vector<uint2> tmp(ctx->n);
    if (ctx->d_mask_acc) {
      const void* src = ctx->d_mask_acc;
      if (ctx->d_perm) { launch_unpermute(reinterpret_cast<const uint32_t*>(ctx->d_mask_acc), ctx->d_perm, (uint32_t)ctx->n, 2, d_tmp, ctx->stream); src = d_tmp; }
      CU(cudaMemcpyAsync(tmp.data(), src, sizeof(uint2) * ctx->n, cudaMemcpyDeviceToHost, ctx->stream));
      CU(cudaStreamSynchronize(ctx->stream));
    } else std::memset(tmp.data(), 0, sizeof(uint2) * ctx->n);
    for (size_t i = 0; i < ctx->n; i++) {
      if (mask_hits) mask_hits[i] = tmp[i].x;
      if (mask_max) mask_max[i] = tmp[i].y;
    }
  }
  if (d_tmp) cudaFree(d_tmp);
  return PCP_OK;
}

/* Debug: the 16 u32 device counters of the last submitted frame {records, candidates, visible, coloured, active chunks,
 * [PCP_STATS builds: chunks skipped whole, points kept by stage A0, points kept by stage A1], ...}. */
int pcp_debug_counters(pcp_ctx* ctx, uint32_t out16[16]) {
  if (!ctx || !out16) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  CU(ctx->drain());
  CU(cudaMemcpy(out16, ctx->d_counters + ctx->last_counters, 16 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  if (ctx->d_inv_local) CU(cudaMemcpy(out16 + 14, ctx->d_inv_local, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  return PCP_OK;
}

void* pcp_accumulators_device(pcp_ctx* ctx) { return ctx ? (void*)ctx->d_acc : nullptr; }

int pcp_profile_enable(pcp_ctx* ctx, int on) {
  if (!ctx) return PCP_EINVAL;
  ctx->profiling = on != 0;
  return PCP_OK;
}

int pcp_profile_read(pcp_ctx* ctx, double ms3[3], uint64_t* frames) {
  if (!ctx || !ms3) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  CU(ctx->drain());
  ms3[0] = ms3[1] = ms3[2] = 0.0;
  for (size_t i = 0; i + 4 <= ctx->prof_used; i += 4) {
    for (int k = 1; k < 3; k++) {  // interval 0 is unused (the z-grid clear precedes the first event)
      float ms = 0.f;
      CU(cudaEventElapsedTime(&ms, ctx->prof_events[i + k], ctx->prof_events[i + k + 1]));
      ms3[k] += ms;
    }
  }
  if (frames) *frames = ctx->prof_used / 4;
  ctx->prof_used = 0;
  return PCP_OK;
}

int pcp_finalize_device(pcp_ctx* ctx, uint8_t* d_rgb_out, uint32_t* d_count_out) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  if (!ctx->d_acc) { ctx->fail("no cloud"); return PCP_EINVAL; }
  if (ctx->acc_scattered) { ctx->fail("pcp_finalize: the accumulators were consumed by pcp_exchange_finalize"); return PCP_EINVAL; }
  if (ctx->n == 0) return PCP_OK;
  CU(ctx->join());
  if (ctx->prm.accum_mode == PCP_ACCUM_SUM) {
    const int rc = finalize_sum_on_stream(ctx, d_rgb_out, d_count_out);
    if (rc != PCP_OK) return rc;
  } else {
    launch_finalize_top(ctx->d_top_score, ctx->d_top_rgbf, ctx->d_top_frame, ctx->d_perm, (uint32_t)ctx->n, ctx->prm.top_m, d_rgb_out, d_count_out, ctx->stream);
    ctx->launches++;
  }
  CU(cudaGetLastError());
  return PCP_OK;
}

/* smoothColors + removePointsWithNoColor (cpp:604-631, hpp:238-252) with the compaction on the device: the coloured points only, as
 * pcl::PointXYZRGB payloads {x, y, z, b | g << 8 | r << 16 | 255 << 24} (16 bytes), in the caller's point order.
 * `*n_out` receives their number; at most `cap` records are written (call with points_out = NULL to learn the number first). */
int pcp_finalize_compact(pcp_ctx* ctx, void* points_out, size_t cap, size_t* n_out) {
  if (!ctx || !n_out) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  if (!ctx->d_acc) { ctx->fail("no cloud"); return PCP_EINVAL; }
  if (ctx->acc_scattered) { ctx->fail("pcp_finalize_compact: the accumulators were consumed by pcp_exchange_finalize"); return PCP_EINVAL; }
  const size_t n = ctx->n;
  *n_out = 0;
  if (n == 0) return PCP_OK;
  if (ctx->out_points < n) {
    if (ctx->d_out_rgb) { cudaFree(ctx->d_out_rgb); ctx->d_out_rgb = nullptr; }
    if (ctx->d_out_cnt) { cudaFree(ctx->d_out_cnt); ctx->d_out_cnt = nullptr; }
    CU(cudaMalloc(&ctx->d_out_rgb, n * 3));
    CU(cudaMalloc(&ctx->d_out_cnt, n * sizeof(uint32_t)));
    ctx->out_points = n;
  }
  int rc = pcp_finalize_device(ctx, ctx->d_out_rgb, nullptr);
  if (rc != PCP_OK) return rc;
  if (!ctx->d_vis_bits) CU(cudaMalloc(&ctx->d_vis_bits, sizeof(uint32_t) * ((n + 31) / 32)));
  if (!ctx->d_idx) CU(cudaMalloc(&ctx->d_idx, sizeof(int32_t) * n));
  if (!ctx->d_scratch) CU(cudaMalloc(&ctx->d_scratch, sizeof(uint32_t) * compact_scratch_words((uint32_t)n)));
  ctx->mask_idx_valid = false;  // d_idx is shared with pcp_frame_mask_cloud
  launch_coloured_bits(ctx->d_out_rgb, (uint32_t)n, ctx->d_vis_bits, ctx->stream);
  launch_compact_bits(ctx->d_vis_bits, (uint32_t)n, nullptr, ctx->d_scratch, ctx->d_idx, ctx->d_total, ctx->stream);
  uint32_t total = 0;
  CU(cudaMemcpyAsync(&total, ctx->d_total, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->launches += 4;
  *n_out = total;
  if (!points_out || total == 0) return PCP_OK;
  const uint32_t m = (uint32_t)(total < cap ? total : cap);
  if (m == 0) return PCP_OK;
  // the 16-byte records are staged in the (idle) count buffer when they fit, else in a temporary
  uint4* d_pts = nullptr;
  bool own = false;
  if ((size_t)m * sizeof(uint4) <= n * sizeof(uint32_t)) d_pts = reinterpret_cast<uint4*>(ctx->d_out_cnt);
  else { CU(cudaMalloc(&d_pts, (size_t)m * sizeof(uint4))); own = true; }
  launch_gather_coloured(ctx->d_idx, m, ctx->d_cloud, ctx->d_perm ? ctx->d_inv : nullptr, ctx->d_out_rgb, d_pts, ctx->stream);
  ctx->launches++;
  cudaError_t e = cudaMemcpyAsync(points_out, d_pts, (size_t)m * sizeof(uint4), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (own) cudaFree(d_pts);
  if (e != cudaSuccess) { ctx->fail("pcp_finalize_compact: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  return PCP_OK;
}

/* generateSegmentMap (cpp:769-816) + the per-frame part of cpp:533-551 for the frame just submitted, on the device: the points that
 * received a pixel, as 20-byte PointXYZRGBMask records {x, y, z in the camera frame, b, g, r, 255, uint16 mask, 2 bytes of padding}
 * in the map's order; mask == 255 paints the point red.  The frame must have been submitted alone (pcp_submit_frame[_device]) with a
 * mask and with its pixel indices kept (pcp_frame_out.pixel_index, or flags & PCP_OUT_PIXELS_ON_DEVICE), and no frame after it.
 * `*n_out` receives the number of records; at most `cap` are written (records_out = NULL: only count). */
int pcp_frame_mask_cloud(pcp_ctx* ctx, void* records_out, size_t cap, size_t* n_out) {
  if (!ctx || !n_out) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  *n_out = 0;
  if (!ctx->d_cloud || !ctx->pix_valid || !ctx->d_pix || !ctx->last_mask || !ctx->last_bgr) {
    ctx->fail("pcp_frame_mask_cloud: the last frame was not a single frame with a mask and kept pixel indices");
    return PCP_EINVAL;
  }
  const size_t n = ctx->n;
  if (n == 0) return PCP_OK;
  CU(ctx->join());
  uint32_t total = ctx->mask_idx_total;
  if (!ctx->mask_idx_valid) {  // the usual second call (count first, then the records) finds the indices already compacted
    if (!ctx->d_vis_bits) CU(cudaMalloc(&ctx->d_vis_bits, sizeof(uint32_t) * ((n + 31) / 32)));
    if (!ctx->d_idx) CU(cudaMalloc(&ctx->d_idx, sizeof(int32_t) * n));
    if (!ctx->d_scratch) CU(cudaMalloc(&ctx->d_scratch, sizeof(uint32_t) * compact_scratch_words((uint32_t)n)));
    launch_pixel_bits(ctx->d_pix, (uint32_t)n, ctx->d_vis_bits, ctx->stream);
    launch_compact_bits(ctx->d_vis_bits, (uint32_t)n, nullptr, ctx->d_scratch, ctx->d_idx, ctx->d_total, ctx->stream);
    CU(cudaMemcpyAsync(&total, ctx->d_total, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->launches += 4;
    ctx->mask_idx_total = total;
    ctx->mask_idx_valid = true;
  }
  *n_out = total;
  const uint32_t m = (uint32_t)(total < cap ? total : cap);
  if (!records_out || m == 0) return PCP_OK;
  uint32_t* d_rec = nullptr;
  CU(cudaMalloc(&d_rec, (size_t)m * 20));
  launch_gather_mask_cloud(ctx->d_idx, m, ctx->d_cloud, ctx->d_perm ? ctx->d_inv : nullptr, ctx->d_pix, ctx->last_bgr, ctx->last_pitch, ctx->last_mask,
                           ctx->last_mask_pitch, ctx->prm.image_width, ctx->last_w2c, d_rec, ctx->stream);
  ctx->launches++;
  cudaError_t e = cudaMemcpyAsync(records_out, d_rec, (size_t)m * 20, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d_rec);
  if (e != cudaSuccess) { ctx->fail("pcp_frame_mask_cloud: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  return PCP_OK;
}

int pcp_finalize(pcp_ctx* ctx, uint8_t* rgb_out, uint32_t* count_out) {
  if (!ctx) return PCP_EINVAL;
  CU(cudaSetDevice(ctx->device));
  if (!ctx->d_acc) { ctx->fail("no cloud"); return PCP_EINVAL; }
  const size_t n = ctx->n;
  if (ctx->acc_scattered) { ctx->fail("pcp_finalize: the accumulators were consumed by pcp_exchange_finalize"); return PCP_EINVAL; }
  if (n == 0) return PCP_OK;
  CU(ctx->join());
  if (ctx->out_points < n) {
    if (ctx->d_out_rgb) { cudaFree(ctx->d_out_rgb); ctx->d_out_rgb = nullptr; }
    if (ctx->d_out_cnt) { cudaFree(ctx->d_out_cnt); ctx->d_out_cnt = nullptr; }
    CU(cudaMalloc(&ctx->d_out_rgb, n * 3));
    CU(cudaMalloc(&ctx->d_out_cnt, n * sizeof(uint32_t)));
    ctx->out_points = n;
  }
  uint8_t* d_rgb = rgb_out ? ctx->d_out_rgb : nullptr;
  uint32_t* d_cnt = count_out ? ctx->d_out_cnt : nullptr;
  if (ctx->prm.accum_mode == PCP_ACCUM_SUM) {
    const int rc = finalize_sum_on_stream(ctx, d_rgb, d_cnt);
    if (rc != PCP_OK) return rc;
  } else {
    launch_finalize_top(ctx->d_top_score, ctx->d_top_rgbf, ctx->d_top_frame, ctx->d_perm, (uint32_t)n, ctx->prm.top_m, d_rgb, d_cnt, ctx->stream);
    ctx->launches++;
  }
  if (rgb_out) CU(cudaMemcpyAsync(rgb_out, d_rgb, n * 3, cudaMemcpyDeviceToHost, ctx->stream));
  if (count_out) CU(cudaMemcpyAsync(count_out, d_cnt, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

}  // extern "C"
