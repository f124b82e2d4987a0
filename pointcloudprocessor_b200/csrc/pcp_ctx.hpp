// pcp_ctx.hpp — the opaque context behind the C ABI (one per GPU).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/pcp/pcp.h"
#include "pcp_host_math.hpp"
#include "pcp_internal.cuh"

struct pcp_ctx {
  static constexpr int kSlots = 16;  // image staging slots (allocated on first use): uploads run up to 16 frames ahead of the kernels,
                                    // so the host link stays busy while the cloud is still being pre-binned

  int device = 0;
  int sm_count = 148;
  pcp_camera cam{};
  pcp_params prm{};
  int Wc = 0, Hc = 0;

  cudaStream_t own_stream = nullptr, stream = nullptr, copy_stream = nullptr;
  cudaEvent_t ev_copied[kSlots] = {}, ev_free[kSlots] = {};
  uint8_t* d_img[kSlots] = {};
  uint8_t* d_mask[kSlots] = {};
  size_t img_bytes = 0, mask_bytes = 0;
  int slot = 0;

  // cloud + per-point state (HBM layout: see DESIGN.md)
  size_t n = 0;
  float4* d_cloud = nullptr;      // N x 16 B  {x,y,z,intensity}
  uint8_t* d_aos = nullptr;       // staging of clouds handed over in a record layout other than float4 (kept across uploads)
  size_t aos_bytes = 0;
  uint32_t* d_perm = nullptr;     // N x 4 B   sorted position -> original index (pre-binned mode; null otherwise)
  uint32_t* d_inv = nullptr;      // N x 4 B   original index -> sorted position (pre-binned mode): finalize gathers through it
  float4* d_aabb = nullptr;       // 2 float4 per chunk of 64 sorted points, stored in visiting order (pre-binned mode)
  uint32_t* d_inv_local = nullptr; // {close pairs, pairs looked at} of inv_locality_kernel (pre-binned mode)
  float inv_local = -1.0f;        // their ratio once read back (pcp_finalize), -1 = not read yet
  uint32_t* d_order = nullptr;    // visit slot -> chunk id (pre-binned mode)
  uint4* d_active = nullptr;      // per-frame list of chunks to scan + their cell rectangle / distance bound (pre-binned mode)
  uint32_t n_chunks = 0;
  void* d_prebin_scratch = nullptr;  // sort keys / temporaries of the pre-binning, kept across uploads
  size_t prebin_scratch_bytes = 0;
  uint8_t* d_out_rgb = nullptr;      // staging of pcp_finalize's host outputs, kept across calls
  uint32_t* d_out_cnt = nullptr;
  size_t out_points = 0;
  static constexpr size_t kAccPad = 64;  // d_acc holds N + kAccPad entries: the reduce-scatter of pcp_exchange_finalize works on G equal slices
  uint4* d_acc = nullptr;         // N x 16 B  {sum_r,sum_g,sum_b,count}
  uint8_t* d_tmp = nullptr;       // staging for pcp_nid_create (image plane + bin table + non-float4 point records), kept until pcp_destroy
  size_t tmp_bytes = 0;
  uint2* d_xchg = nullptr;        // pcp_exchange_finalize: packed {rgb, count} of every point in sorted order (all-gather buffer)
  size_t xchg_points = 0;
  cudaEvent_t xchg_ev[5] = {};    // phase timing of pcp_exchange_finalize while profiling is enabled
  bool xchg_timed = false;
  bool reduced = false;           // the accumulators were exchanged since the last reset (a second exchange would double-count)
  bool acc_scattered = false;     // ... by pcp_exchange_finalize: only this rank's slice of d_acc is meaningful until the next reset
  uint2* d_mask_acc = nullptr;    // N x 8 B   {hits, max}; allocated at the first frame that carries a mask
  float* d_top_score = nullptr;   // top_m x N
  uint32_t* d_top_rgbf = nullptr;
  uint32_t* d_top_frame = nullptr;
  pcp::Record* d_records = nullptr;  // flight x (N + 64) x 32 B worst case (one region per frame slot)
  uint32_t* d_vis_bits = nullptr;    // ceil(N/32) words, lazily allocated
  int32_t* d_pix = nullptr;          // N, lazily allocated
  int32_t* d_idx = nullptr;          // N, lazily allocated (pcp_cull)
  uint32_t* d_scratch = nullptr;     // compaction tile sums

  // Frame slots: frame f runs on slot f % flight, each slot with its own stream, z-grid, counters, record region and
  // active-chunk list.  The scan of frame f+1 therefore overlaps the tail of the scan of frame f and its resolve kernel
  // (whose accumulator updates stay ordered: resolve f+1 waits for resolve f).  Work is forked from / joined to ctx->stream.
  static constexpr int kMaxFlight = 24;
  struct FrameSlot {
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_done = nullptr;
    int parity = 0;               // which of the slot's two z-grids / counter blocks the next frame uses
    bool used = false;            // ev_done has been recorded at least once (or is pending in a loop call)
    int owner = -1;               // slot whose STREAM last ran work on this slot's z-grids / lists / records (-1: none yet)
  };
  FrameSlot fs[kMaxFlight];
  int flight = 3;                 // frames in flight (pcp_params.frames_in_flight)
  bool flight_auto = false;       // frames_in_flight = 0: chosen per cloud (12 for <= 4e6 points, else 3)
  int group = 1;                  // frames per launch in pcp_submit_frames_device (1, 2 or 4; pcp_params.frames_per_launch, 0 = per cloud)
  uint64_t frames_run = 0;        // frames submitted since creation (slot = frames_run % flight)
  int last_slot = 0;              // slot of the last submitted frame
  int last_counters = 0;          // index of its counter block in d_counters
  // pcp_submit_frames_device (many frames, no per-frame outputs): a slot's stream forks from ctx->stream once per call, not once per
  // frame, and (SUM) its done-event is recorded once at the end of the call: 6 -> 3-4 driver calls per frame (C1 is bound by them)
  bool in_batch = false;
  uint32_t batch_forked = 0;    // slots whose stream already waited for ctx->stream in this call
  uint32_t batch_pending = 0;   // slots whose ev_done is still to be recorded
  void end_batch();
  // the last single frame, for pcp_frame_mask_cloud: its device image / mask, its float transform, and whether d_pix holds its pixels
  const uint8_t* last_bgr = nullptr; size_t last_pitch = 0;
  const uint8_t* last_mask = nullptr; size_t last_mask_pitch = 0;
  pcp::Xf34 last_w2c = {};
  bool pix_valid = false;
  bool mask_idx_valid = false;    // d_idx / mask_idx_total hold the compacted indices of that frame (a count-only call came first)
  uint32_t mask_idx_total = 0;
  cudaEvent_t last_done = nullptr;  // ev_done of the last submitted frame (null before the first)
  unsigned long long* d_zgrid = nullptr;       // kMaxFlight x Wc*Hc packed (fl32 depth, point index) keys: grid 0 of every slot
  unsigned long long* d_zprev = nullptr;       // kMaxFlight x Wc*Hc: grid 1 of every slot (frames of a slot alternate; the other grid
                                               // holds the previous frame's final state = the seed points)
  pcp::FrameCounters* d_counters = nullptr;    // 2 x kMaxFlight entries (slot, parity)
  pcp::FrameCounters* h_counters = nullptr;  // pinned
  uint32_t* d_total = nullptr;

  int scan_grid = 148 * 4, resolve_grid = 148 * 4;
  int frame_index = 0;
  int frame_first = 0;            // value frame_index restarts from at pcp_set_cloud / pcp_reset_accumulators
  int frame_stride = 1;           // pcp_set_frame_numbering: global keyframe number of the next frame and its increment
  bool use_seed = true;   // last frame's nearest points go through the exact path first (pcp_params.debug_flags & PCP_DEBUG_NO_SEED disables it)
  bool prefilter = true;  // FP32 pre-filter of the scan kernel (PCP_DEBUG_NO_PREFILTER disables it: exact path for every point)
  uint64_t launches = 0;
  std::string err;

  // optional per-kernel timing
  bool profiling = false;
  std::vector<cudaEvent_t> prof_events;  // 4 per frame
  size_t prof_used = 0;

  // NCCL (resolved with dlopen at first use; see pcp_nccl.cpp)
  void* nccl_comm = nullptr;
  int nccl_rank = 0, nccl_nranks = 1;

  void fail(const char* fmt, ...);
  void free_cloud();
  cudaError_t join();   // ctx->stream waits (asynchronously) for the last submitted frame
  cudaError_t drain();  // host waits for every frame slot and ctx->stream
};

int scan_ctas_per_sm();
void nccl_teardown(pcp_ctx* ctx);
// in-place all-gather of `slice` float4 points per rank on ctx->stream (rank r's points sit at d_cloud + r * slice); PCP_* code
int nccl_all_gather_cloud(pcp_ctx* ctx, float4* d_cloud, size_t slice);
