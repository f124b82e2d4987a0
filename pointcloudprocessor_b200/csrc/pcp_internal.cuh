// pcp_internal.cuh — device-side data layout and kernel launch prototypes (not part of the ABI).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "pcp_math.cuh"

namespace pcp {

// Initial z-grid key: fl32 FLT_MAX in the high word (view_culling.cpp:64), index 0xffffffff (= -1, :65).
static constexpr unsigned long long kZInit = (0x7f7fffffull << 32) | 0xffffffffull;
static constexpr unsigned int kZBoundInit = 0x7f7fffffu;  // FLT_MAX
#ifndef PCP_CHUNK
#define PCP_CHUNK 64
#endif
static constexpr int kChunk = PCP_CHUNK;  // points per pre-binned chunk = one warp iteration of the scan kernel (kChunk/32 x 32)

// One candidate that may still be visible after pass 1 (32 bytes, written/read as two 16-byte words).
struct __align__(16) Record {
  float x, y, z;      // camera-frame float point (cloudInCameraPose, cpp:521)
  uint32_t idx;       // original point index
  double dist;        // ||p|| in double (view_culling.cpp:102)
  int32_t cell;       // cy*Wc+cx, or -1 when outside the grid (only emitted when depth culling is off)
  int32_t pix;        // generateColorMap pixel v*W+u or -1 (cpp:752-754)
};
static_assert(sizeof(Record) == 32, "Record must be 32 bytes");

// Per-frame device counters (one 64-byte block per frame slot).
struct FrameCounters {
  unsigned int n_records;     // records emitted by the scan kernel
  unsigned int n_candidates;  // gate-passers (view_culling.cpp:96)
  unsigned int n_visible;     // kept by pass 2 (view_culling.cpp:167)
  unsigned int n_colored;     // generateColorMap pushes (cpp:763)
  unsigned int n_active_blocks;
  unsigned int stats[3];      // PCP_STATS builds: chunks skipped whole, points kept by stage A0, points kept by stage A1
  unsigned int pool_next;     // scan_ring_kernel: next entry of the common pool of the active list (relative to its start)
  unsigned int pad[7];
};
static_assert(sizeof(FrameCounters) == 64, "FrameCounters layout");

// FP32 pre-filter constants (pcp_kernels.cu: prefilter()): the camera in float, the same coefficients as magnitudes, the
// decision boundaries in pixels and the error-bound factors.
struct PreF {
  float fx, fy, cx, cy, k1, k2, k3, p1, p2;
  float ak1, ak2, ak3, apt;      // |k1|,|k2|,|k3| and 3(|p1|+|p2|), rounded up
  float afmax, acmax;            // max(|fx|,|fy|), max(|cx|,|cy|), rounded up
  float inv_ds, wc, hc;          // 1/ds, (float)Wc, (float)Hc: cells are trunc(u/ds), (-1,1) -> 0
  int Wc;
  float lo;                      // -ds
  float u_hi_cand, v_hi_cand;    // ds * cullW, ds * cullH (pass-1 gate, candidate counting only)
  float err, slack;              // 64 * 2^-24 and an absolute slack in pixels
  float dscale, tol_up;          // (1 - 8 eps): lower bound of the float norm; tolerance rounded up to float
};

struct FrameParams {
  Xf34 w2c;          // float pre-transform (identity when absent)
  Xd34 T;            // double T_camera_lidar (view_culling.cpp:37)
  CamD cam;
  double tol;        // 0.05
  double cam_pos[3]; // keyframe->pose x,y,z for the orientation score (hpp:208)
  int use_w2c, use_T;
  int cullW, cullH;  // image_size given to ViewCulling
  int Wc, Hc;        // cullW/ds, cullH/ds (integer division, view_culling.cpp:64)
  float dsf;         // (float)ds
  int imgW, imgH;    // actual image size
  int depth_culling;
  int frame_index;
  int use_prefilter;  // 0 when a double T is applied before the gate (pcp_cull with T16) or when disabled for testing
  PreF pre;
};

struct ScanArgs {
  const float4* __restrict__ cloud;   // N x {x,y,z,intensity}, possibly block-sorted
  const uint32_t* __restrict__ perm;  // sorted position -> original index (null when not pre-binned)
  const float4* __restrict__ aabb;    // pre-binned: 2 float4 per VISIT SLOT {lo.xyz, nonfinite}, {hi.xyz, count} (boxes stored in visiting order); null otherwise
  const uint32_t* __restrict__ order; // pre-binned: visit slot -> chunk id (a fixed pseudo-random permutation of the chunks)
  int seed_blocks;                    // classify_kernel: its first seed_blocks CTAs run the seed points, the others classify chunks
  uint4* active;                      // pre-binned: {chunk id, ChunkInfo rect, org, d2lo bits} of the chunks worth processing this frame
  uint32_t n_chunks;                  // ceil(n / kChunk)
  unsigned long long* zgrid;          // Wc*Hc packed keys
  unsigned long long* zprev;          // pre-binned: the slot's other z-grid = the previous frame's final grid (its index words name the
                                      // seed points); classify_kernel clears it for the slot's next frame
  FrameCounters* counters_next;       // pre-binned: the counter block of the slot's next frame (cleared by classify_kernel)
  int seed;                           // 1: the previous frame's winners go through the exact path first
  int cells;                          // Wc*Hc
  Record* records;
  FrameCounters* counters;
  uint32_t n;
};

struct ResolveArgs {
  const Record* __restrict__ records;
  const unsigned long long* __restrict__ zgrid;
  FrameCounters* counters;
  const uint32_t* __restrict__ perm;  // pre-binned: record.idx is a sorted position; outputs in caller order go through perm
  const uint8_t* __restrict__ bgr; size_t pitch;
  const uint8_t* __restrict__ mask; size_t mask_pitch;
  uint4* acc;                 // N x {sum_r,sum_g,sum_b,count}
  uint2* mask_acc;            // N x {hits(==255), max}
  float* top_score;           // top_m x N   (TOP5 mode)
  uint32_t* top_rgbf;         // top_m x N : r | g<<8 | b<<16  (frame index kept in top_frame)
  uint32_t* top_frame;        // top_m x N
  uint32_t* vis_bits;         // optional
  int32_t* pix_out;           // optional
  uint32_t n;
  int top_m;
  int accum_mode;
  int sample_mode;
};

// Arguments of up to kGroupMax frames for one launch (classify / scan / resolve _multi_kernel: blockIdx.y = frame of the group)
template <typename A, int K>
struct FramePack {
  A a[K];
  FrameParams fp[K];
};
constexpr int kGroupMax = 4;  // 4 x (112 + 400) bytes = half of the 4 KB a kernel may take as parameters

void launch_zgrid_init(unsigned long long* zgrid, int cells, FrameCounters* counters, cudaStream_t s);
// k = 2 or 4 frames of the same cloud on different slots in one launch each (pre-binned cloud only)
int launch_scan_group(const ScanArgs* a, const FrameParams* fp, int k, int sm_count, bool count_candidates, int device, cudaStream_t s);
int launch_resolve_group(const ResolveArgs* a, const FrameParams* fp, int k, int grid, cudaStream_t s);
int launch_scan(const ScanArgs& a, const FrameParams& fp, int grid, int sm_count, bool count_candidates, int device, cudaStream_t s);
void launch_resolve(const ResolveArgs& a, const FrameParams& fp, int grid, cudaStream_t s);
void launch_aos_to_soa(const uint8_t* d_aos, size_t n, size_t stride, size_t ioff, float4* out, cudaStream_t s);
void launch_finalize_sum(const uint4* acc, const uint32_t* inv, uint32_t n, uint8_t* rgb, uint32_t* count, int fixed88, cudaStream_t s);
void launch_pack_sum(const uint4* acc, uint32_t j0, uint32_t j1, uint2* packed, int fixed88, cudaStream_t s);
void launch_unpack_colours(const uint2* packed, const uint32_t* inv, uint32_t n, uint8_t* rgb, uint32_t* count, cudaStream_t s);
void launch_pixel_bits(const int32_t* pix, uint32_t n, uint32_t* bits, cudaStream_t s);
void launch_gather_mask_cloud(const int32_t* idx, uint32_t m, const float4* cloud, const uint32_t* inv, const int32_t* pix, const uint8_t* bgr,
                              size_t pitch, const uint8_t* mask, size_t mask_pitch, int W, const Xf34& w2c, uint32_t* out, cudaStream_t s);
void launch_finalize_top(const float* score, const uint32_t* rgbf, const uint32_t* frame, const uint32_t* perm, uint32_t n, int top_m,
                         uint8_t* rgb, uint32_t* count, cudaStream_t s);
// generateColorMap's 8-bit HSV round trip (pcp_image.cu); src and dst may alias
void launch_hsv_roundtrip(const uint8_t* src, size_t spitch, uint8_t* dst, size_t dpitch, int W, int H, float sat_scale, float val_scale, int block,
                          int sm_count, cudaStream_t s);
// pre-binning (pcp_prebin.cu) and the un-permuting copy of per-point arrays (elem_words u32 words per point)
cudaError_t prebin_cloud(float4* d_cloud, uint32_t n, uint32_t* d_perm, float4* d_aabb, uint32_t* d_order, void** scratch, size_t* scratch_bytes,
                         int sm_count, cudaStream_t s);
void launch_invert_perm(const uint32_t* perm, uint32_t n, uint32_t* inv, cudaStream_t s);
void launch_inv_locality(const uint32_t* inv, uint32_t n, uint32_t* out2, cudaStream_t s);
void launch_unpermute(const uint32_t* src, const uint32_t* perm, uint32_t n, int elem_words, uint32_t* dst, cudaStream_t s);
// ascending index list from a visibility bitmask; returns through d_total the number of set bits
void launch_compact_bits(const uint32_t* bits, uint32_t n, uint32_t* word_offsets, uint32_t* block_sums, int32_t* idx_out,
                         uint32_t* d_total, cudaStream_t s);
size_t compact_scratch_words(uint32_t n);
// removePointsWithNoColor on the device: coloured-point bitmask, then the gather of {x, y, z, packed bgra} records (pcp_kernels.cu)
void launch_coloured_bits(const uint8_t* rgb, uint32_t n, uint32_t* bits, cudaStream_t s);
void launch_gather_coloured(const int32_t* idx, uint32_t m, const float4* cloud, const uint32_t* inv, const uint8_t* rgb, uint4* out, cudaStream_t s);
// distance / 3-D point masks of a frame from the resolve kernel's pixel indices (pcp_splat.cu); `winner` = imgW*imgH zeroed u64 words
void launch_splat(const int32_t* pix, const uint32_t* perm, uint32_t n, const float4* cloud, const FrameParams& fp, unsigned long long* winner,
                  float* dist, float* xyz, int32_t* index, unsigned int* n_filled, int sm_count, cudaStream_t s);

}  // namespace pcp
