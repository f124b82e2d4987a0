/* pcp.h — C ABI of libpcp_b200.so: the B200-native (sm_100a) per-frame colorization hot path.
 *
 * The reference (ChunLI-666/PointCloudProcessor) has no FFI/plugin layer: its boundary is a C++ class API.
 * Each entry point below names the reference interface it replaces (paths relative to
 * /root/reference/PointCloudProcessor/).  The C++ facade in include/pcp/pcp.hpp re-creates those classes on
 * top of this ABI; INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions: plain pointers and sizes only; every function returns a pcp_status (never throws);
 * pcp_last_error() gives the message.  All matrices are 4x4 column-major (Eigen's default storage).
 * One context per GPU; a context is not thread-safe.  There is NO CPU fallback: without a CUDA device
 * pcp_create fails with PCP_ECUDA.
 */
#ifndef PCP_PCP_H_
#define PCP_PCP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCP_ABI_VERSION 1

typedef enum { PCP_OK = 0, PCP_EINVAL = 1, PCP_EIO = 2, PCP_ECUDA = 3, PCP_ENCCL = 4, PCP_ENOMEM = 5 } pcp_status;

typedef struct pcp_ctx pcp_ctx;
typedef struct pcp_nid pcp_nid;

/* camera::create_camera("pinhole", intrinsics{fx,fy,cx,cy}, distortion{k1,k2,p1,p2,k3})
 * include/camera/create_camera.hpp:14, include/camera/pinhole.hpp:12-51 */
typedef struct {
  double fx, fy, cx, cy;
  double dist[5]; /* k1, k2, p1, p2, k3 (plumb_bob order) */
} pcp_camera;

enum { PCP_ACCUM_SUM = 0, PCP_ACCUM_TOP5 = 1 };
enum { PCP_SAMPLE_NEAREST = 0, PCP_SAMPLE_BILINEAR = 1 };

/* Every constant the reference hard-codes on this path, lifted into one POD.  pcp_default_params() fills the
 * reference values. */
typedef struct {
  int image_width, image_height;   /* actual image size: bounds of the colour gather (src/PointCloudProcessor.cpp:754) */
  int cull_width, cull_height;     /* image_size handed to ViewCulling (reference hard-codes 4096x3000, cpp:525) */
  int downsample;                  /* z-buffer downsample factor, 14 (src/vlcal/calib/view_culling.cpp:63) */
  double depth_tolerance;          /* 0.05 (view_culling.cpp:157) */
  int enable_depth_buffer_culling; /* ViewCullingParams (include/vlcal/calib/view_culling.hpp:13) */
  int accum_mode;                  /* PCP_ACCUM_SUM (u32 sums+count, all-reducible) | PCP_ACCUM_TOP5 (cpp:604-631) */
  int sample_mode;                 /* PCP_SAMPLE_NEAREST (reference, cpp:752-762) | PCP_SAMPLE_BILINEAR (extension) */
  int top_m;                       /* 5 (cpp:615) */
  int prebin;                      /* 1 (default): Morton-sort the device copy of the cloud into 64-point chunks with bounding
                                    * boxes at set_cloud; chunks certainly behind the camera / outside the z-grid are skipped per
                                    * frame.  Every output stays in the caller's point order.  0: keep the caller's order. */
  int frames_in_flight;            /* 1..24; 0 (default) = chosen per cloud: 12 for clouds of <= 4e6 points, else 3.  Frames run on that many internal slots, each with its own z-grid / record
                                    * scratch, so that the scan of frame f+1 overlaps the tail of frame f and its resolve pass.
                                    * Accumulator updates stay in frame order; results do not depend on this value. */
  int hsv_adjust;                  /* 1: pcp_submit_frame applies generateColorMap's 8-bit BGR->HSV->BGR round trip (cpp:722-741) to the
                                    * uploaded image on the device before sampling; 0 (default): the caller passes the adjusted image */
  float saturation_scale;          /* 1.0 (cpp:726) */
  float brightness_scale;          /* 1.0 (cpp:727) */
  int hsv_vector_block;            /* pixels per vector iteration of the OpenCV build to reproduce (its scalar row tail rounds
                                    * differently from its vector body): 32 = cv2 4.13 AVX2 build; 0 = whole row is vector body */
  int debug_flags;                 /* PCP_DEBUG_*: 0 in production.  Test hooks that select a slower path with identical results. */
  int frames_per_launch;           /* pcp_submit_frames_device, SUM mode, pre-binned cloud: 1, 2 or 4 consecutive frames share one launch of
                                    * each of the frame's three kernels (they run on neighbouring slots; results do not depend on it).
                                    * 0 (default) = chosen per cloud: for <= 4e6 points 4 when frames_in_flight >= 8, 2 when >= 4; else 1.
                                    * For small clouds the launches, not the GPU, bound the frame rate. */
} pcp_params;
/* PCP_DEBUG_NO_PREFILTER: every point takes the exact FP64 path of pass 1 (no FP32 pre-filter, no chunk skipping by distance).
 * PCP_DEBUG_NO_SEED: the previous frame's nearest points are not sent through pass 1 first.
 * PCP_DEBUG_FINALIZE_ONEPASS / _TWOPASS: pcp_finalize always gathers the 16-byte sums in one kernel / always packs them to 8 bytes
 * first (by default the cloud size decides). */
enum { PCP_DEBUG_NO_PREFILTER = 1, PCP_DEBUG_NO_SEED = 2, PCP_DEBUG_FINALIZE_ONEPASS = 4, PCP_DEBUG_FINALIZE_TWOPASS = 8 };

/* Optional per-frame outputs (host buffers, caller-owned; any pointer may be NULL). */
typedef struct {
  uint32_t* visible_bits; /* ceil(N/32) words; bit i = point i survived ViewCulling::view_culling (view_culling.cpp:52-174) */
  int32_t* pixel_index;   /* N entries: v*W+u of the sampled pixel (cpp:752-754) or -1 */
  int64_t n_candidates;   /* points recorded by pass 1 (view_culling.cpp:96); -1 with PCP_OUT_SKIP_CANDIDATES */
  int64_t n_visible;      /* points kept by pass 2 (view_culling.cpp:167) */
  int64_t n_colored;      /* points pushed by generateColorMap (cpp:763) */
  uint32_t flags;         /* in: PCP_OUT_* */
  uint32_t reserved;
} pcp_frame_out;
/* n_candidates is an internal of the reference (indices.size() before its second pass) that nothing downstream reads;
 * counting it exactly costs a wider FP32 classification per point.  Frames submitted without a pcp_frame_out, or with
 * this flag, use the kernel variant that skips the count.  Every other output is identical. */
enum { PCP_OUT_SKIP_CANDIDATES = 1, PCP_OUT_PIXELS_ON_DEVICE = 2 /* keep the frame's pixel indices on the device (for pcp_frame_mask_cloud) without copying them out */ };

void pcp_default_params(pcp_params* p);
int pcp_abi_version(void);

/* PointCloudProcessor::PointCloudProcessor (include/PointCloudProcessor.hpp:17-25): owns the device state. */
int pcp_create(pcp_ctx** out, int device, const pcp_camera* cam, const pcp_params* params);
void pcp_destroy(pcp_ctx* ctx);
const char* pcp_last_error(const pcp_ctx* ctx); /* ctx may be NULL: message of the last failed pcp_create */

/* Use an existing CUDA stream (cudaStream_t) for all work of this context; NULL restores the internal one. */
int pcp_set_stream(pcp_ctx* ctx, void* cuda_stream);
void* pcp_get_stream(pcp_ctx* ctx);
int pcp_synchronize(pcp_ctx* ctx);
/* The context's stream waits (asynchronously, on the device) for every frame submitted so far; nothing blocks on the host. */
int pcp_join(pcp_ctx* ctx);

/* Pinned host memory for images/clouds, so that submit_frame's copies are asynchronous. */
int pcp_host_alloc(void** out, size_t bytes);
int pcp_host_free(void* p);

/* loadPointCloud (cpp:92-154) hands over the map: `n` points, x,y,z floats at byte offsets 0,4,8 of each
 * `stride_bytes` record and intensity at `intensity_offset` (pcl::PointXYZI: stride 32, intensity at 16).
 * Copies (asynchronously when `points` is pinned memory: keep it valid until the next synchronising call, e.g.
 * pcp_synchronize or pcp_finalize); converts to float4 SoA on the device; resets the accumulators. */
int pcp_set_cloud(pcp_ctx* ctx, const void* points, size_t n, size_t stride_bytes, size_t intensity_offset);
/* Multi-GPU hand-over of the same map (after pcp_nccl_init; no counterpart in the reference, which is one process): the map crosses the
 * host link ONCE in total instead of once per GPU.  Rank r of G passes only ITS slice of the records — points [first, first + count) as
 * pcp_cloud_slice computes them (G equal slices of ceil(n_total / G) points, the last one shorter) — and the slices are all-gathered
 * over NVLink (ncclAllGather, in place) into every rank's full copy; everything else is pcp_set_cloud.  Collective: every rank calls
 * it, with the same n_total and layout.  Measured host-link bandwidth per rank halves from 2 to 4 ranks on these boxes. */
int pcp_set_cloud_sharded(pcp_ctx* ctx, const void* slice_points, size_t n_total, size_t stride_bytes, size_t intensity_offset);
int pcp_cloud_slice(size_t n_total, int rank, int nranks, size_t* first, size_t* count);
/* Same with a device-resident float4 {x,y,z,intensity} array (no host copy). */
int pcp_set_cloud_device(pcp_ctx* ctx, const void* d_xyzi, size_t n);
size_t pcp_num_points(const pcp_ctx* ctx);

/* One iteration of the keyframe loop of pcdColorizationAndSmooth (cpp:488-595):
 * pose7 = {x,y,z,qw,qx,qy,qz} (include/FrameData.hpp:9-12); T_opt16 = T_camera_lidar_optimized or
 * initTLidarCamera (cpp:504-519), NULL when neither NID nor manual guess is enabled; bgr = the (already
 * HSV-adjusted, cpp:716-741) 8-bit image, `pitch` bytes per row; mask = 8-bit segmentation image or NULL
 * (generateSegmentMap, cpp:769-816).  Host buffers; the copy is asynchronous when they are pinned.
 * Returns after enqueueing unless `out` is given. */
int pcp_submit_frame(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, const uint8_t* bgr, size_t pitch,
                     const uint8_t* mask, size_t mask_pitch, pcp_frame_out* out);
/* Same with device-resident image/mask (used to time the kernels with inputs already in HBM). */
int pcp_submit_frame_device(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, const uint8_t* d_bgr, size_t pitch,
                            const uint8_t* d_mask, size_t mask_pitch, pcp_frame_out* out);
/* The same loop body for n_frames device-resident images in one call: frame i takes poses7[7 i ..] and the image d_bgr[i] (a host
 * array of device pointers; masks likewise when d_mask is not NULL).  Identical to n_frames calls of pcp_submit_frame_device
 * without per-frame outputs; the per-call host overhead is paid once (tiny frames are bound by it).
 * Replaces: the keyframe loop header, src/PointCloudProcessor.cpp:488-493. */
int pcp_submit_frames_device(pcp_ctx* ctx, size_t n_frames, const double* poses7, const double* T_opt16, const uint8_t* const* d_bgr, size_t pitch,
                             const uint8_t* const* d_mask, size_t mask_pitch);

/* vlcal::ViewCulling::cull(points, T_camera_lidar) (include/vlcal/calib/view_culling.hpp:34, z-buffer variant
 * view_culling.cpp:52-174) on the context's cloud.  w2c16: optional float pre-transform
 * (pcl::transformPointCloud, cpp:196/521); T16: optional double T_camera_lidar (view_culling.cpp:37).
 * idx_out: room for N int32; receives the kept indices in ascending order; *m_out their number. */
int pcp_cull(pcp_ctx* ctx, const float* w2c16, const double* T16, int32_t* idx_out, size_t* m_out);
/* Loop body of viewCullingAndSaveFilteredPcds (cpp:178-224): pose -> w2c -> cull. */
int pcp_cull_pose(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, int32_t* idx_out, size_t* m_out);
/* Distance / 3-D point masks of one frame: generate_distance_masks of scripts/genNormAndDistanceMask.py:233-265 applied to the
 * frame's culled cloud (the filtered_pcd/<ts>_beforeNID.pcd it reads, cpp:213-217), without the file round trip.  Every culled
 * point whose truncated projection lies inside the image writes its range ||p|| (double norm of the float camera-frame point,
 * stored as float32) and its camera-frame coordinates at that pixel; points later in the file overwrite earlier ones, i.e. the
 * largest point index wins.  distance_out: H*W floats, 0 where no point landed; xyz_out: H*W*3 floats or NULL; index_out
 * (extension): H*W int32 original index of the winning point, -1 where empty, or NULL; n_filled_out: pixels written, or NULL.
 * Host buffers; synchronous; does not touch the accumulators or the frame numbering. */
int pcp_splat_masks_pose(pcp_ctx* ctx, const double pose7[7], const double* T_opt16, float* distance_out, float* xyz_out,
                         int32_t* index_out, size_t* n_filled_out);
/* Host helper: the float matrices the reference derives from a pose (cpp:495-519). */
int pcp_pose_matrices(const double pose7[7], const double* T_opt16, float w2c16[16], float c2w16[16]);

/* cv::cvtColor(BGR2HSV) -> S *= saturation_scale, V *= brightness_scale (saturate_cast) -> cv::cvtColor(HSV2BGR), 8-bit, of
 * generateColorMap (src/PointCloudProcessor.cpp:722-741) on the device; host buffers in and out (may alias).  Uses the
 * context's hsv_vector_block. */
int pcp_hsv_roundtrip(pcp_ctx* ctx, const uint8_t* bgr_in, size_t pitch_in, uint8_t* bgr_out, size_t pitch_out, int width, int height,
                      float saturation_scale, float brightness_scale);

/* Raw accumulators: acc4 = N x {sum_r, sum_g, sum_b, count} u32; mask_hits/mask_max = N u32 (may be NULL). */
int pcp_get_accumulators(pcp_ctx* ctx, uint32_t* acc4, uint32_t* mask_hits, uint32_t* mask_max);
/* Device pointer of the N x 4 u32 accumulator array, e.g. for an external all-reduce.  With prebin = 1 the array is in the
 * library's sorted order (identical on every rank that holds the same cloud); pcp_get_accumulators un-permutes. */
void* pcp_accumulators_device(pcp_ctx* ctx);
int pcp_reset_accumulators(pcp_ctx* ctx);

/* Multi-GPU: frames are sharded across ranks (every rank holds the whole cloud); one exchange of the accumulators at the end.
 * pcp_nccl_unique_id fills 128 bytes on rank 0; every rank then calls pcp_nccl_init with the same bytes.
 * pcp_allreduce, called by every rank after its last frame, leaves the accumulators of the WHOLE job on every rank:
 *   PCP_ACCUM_SUM   ncclAllReduce(u32, sum) — bit-identical for any rank count;
 *   PCP_ACCUM_TOP5  ncclAllGather of the slot lists + a merge kernel (top_m best of the union, same order as the insertion);
 *                   identical to the single-rank run when frames carry global numbers (pcp_set_frame_numbering);
 *   mask accumulators {hits, max}: all-reduced with sum and max (allocated on ranks that saw no mask). */
int pcp_nccl_unique_id(uint8_t id128[128]);
int pcp_nccl_init(pcp_ctx* ctx, const uint8_t id128[128], int rank, int nranks);
int pcp_allreduce(pcp_ctx* ctx);
/* The exchange fused with finalize (SUM accumulators): ncclReduceScatter of the u32 sums (every rank then owns the whole job's sums
 * of 1/G of the points), finalize of that slice only, ncclAllGather of the packed colour + count (8 B per point instead of 16 B of
 * sums), outputs in the caller's point order on EVERY rank.  21 instead of 28 bytes per point cross NVLink and the finalize
 * arithmetic is done once per point instead of G times; the results equal pcp_allreduce + pcp_finalize bit for bit.  TOP5 mode
 * falls back to exactly that pair.  Afterwards the accumulators count as consumed: pcp_get_accumulators / pcp_finalize /
 * pcp_submit_frame return PCP_EINVAL until pcp_reset_accumulators or pcp_set_cloud.  One exchange per pass: a second
 * pcp_allreduce / pcp_exchange_finalize without a reset in between returns PCP_EINVAL on every rank (it would double-count).
 * Replaces: smoothColors over the union of all keyframes' observations (src/PointCloudProcessor.cpp:604-631). */
int pcp_exchange_finalize(pcp_ctx* ctx, uint8_t* rgb_out, uint32_t* count_out);                 /* host outputs; synchronises */
int pcp_exchange_finalize_device(pcp_ctx* ctx, uint8_t* d_rgb_out, uint32_t* d_count_out);      /* device outputs; asynchronous */
/* Global keyframe numbers of the frames this context submits (the tie-break of equal TOP5 scores, cpp:604-631 keeps the earlier
 * observation): the next frame is number `next`, then next + stride, ...  Rank r of G with interleaved sharding: (r, G).
 * Default (0, 1).  The numbering restarts from `next` at pcp_set_cloud / pcp_reset_accumulators. */
int pcp_set_frame_numbering(pcp_ctx* ctx, int next, int stride);

/* smoothColors + removePointsWithNoColor inputs (cpp:604-631, hpp:238-252): rgb_out = N x 3 (r,g,b) u8,
 * count_out = N u32 observations (either may be NULL).  SUM mode: integer mean; TOP5: the reference's top-M
 * weighted mean. */
int pcp_finalize(pcp_ctx* ctx, uint8_t* rgb_out, uint32_t* count_out);

/* smoothColors + removePointsWithNoColor (hpp:238-252) with the compaction on the device: only the points whose finalized colour is
 * not (0,0,0), in the caller's point order, as pcl::PointXYZRGB payloads {x, y, z, b | g << 8 | r << 16 | 255 << 24} (16 bytes each).
 * *n_out = their number; at most `cap` records are written to points_out (NULL: count only).  Synchronises. */
int pcp_finalize_compact(pcp_ctx* ctx, void* points_out, size_t cap, size_t* n_out);

/* generateSegmentMap (cpp:769-816) and the per-keyframe cloud of cpp:533-546 for the frame JUST submitted, built on the device: the
 * points that received a pixel, in the map's order, as 20-byte PointXYZRGBMask records (include/FrameData.hpp:68-87): {float x, y, z in
 * the camera frame (pcl::transformPointCloud, cpp:521); uint8 b, g, r of the (adjusted) image; uint8 255; uint16 mask value; 2 bytes
 * of padding}; mask == 255 paints the point red (cpp:806-811).  The frame must have been submitted alone (pcp_submit_frame or
 * pcp_submit_frame_device) with a mask and with pcp_frame_out.pixel_index or flags & PCP_OUT_PIXELS_ON_DEVICE; anything else returns
 * PCP_EINVAL.  *n_out = the number of such points; at most `cap` records are written (records_out NULL: count only).  Synchronises. */
int pcp_frame_mask_cloud(pcp_ctx* ctx, void* records_out, size_t cap, size_t* n_out);

/* pcp_compact_bits_noalloc — NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Bitmask -> ascending index list (the order of ViewCulling::sample's output), i.e. the compaction inside pcp_cull, with its
 * scratch in a file-scope static __device__ array instead of a context-owned buffer: the call allocates NOTHING (no cudaMalloc,
 * cudaFree or allocating library routine).  All pointers are DEVICE pointers owned by the caller:
 *   d_bits       ceil(n_bits / 32) words, bit i of word i / 32 set = element i is kept (the layout of pcp_frame_out.visible_bits)
 *   d_idx_out    room for as many int32 as bits are set (n_bits is always enough): the set positions, ascending
 *   d_count_out  one u32: the number of set bits (always written)
 * Asynchronous on the context's stream.  Limit: n_bits <= PCP_NOALLOC_MAX_BITS, larger sizes return PCP_EINVAL with a message
 * (the static scratch is 16.6 KB per GPU, held until the process exits).  Every call on a device shares that scratch: calls
 * that overlap on the GPU (other streams or contexts of the same device) race — serialise them.  The first call on a device
 * resolves the scratch's device address (cudaGetSymbolAddress); make it once outside a graph capture. */
#define PCP_NOALLOC_MAX_BITS (1ull << 27)
int pcp_compact_bits_noalloc(pcp_ctx* ctx, const uint32_t* d_bits, size_t n_bits, int32_t* d_idx_out, uint32_t* d_count_out);

/* Same as pcp_finalize with device-resident outputs (no host copy; asynchronous on the context's stream). */
int pcp_finalize_device(pcp_ctx* ctx, uint8_t* d_rgb_out, uint32_t* d_count_out);

/* Per-kernel device timing (CUDA events on the context's stream around the scan and resolve kernels of every frame
 * submitted while enabled).  pcp_profile_read synchronises and returns the accumulated milliseconds
 * {zgrid init, scan, resolve} and the number of frames, then resets the accumulation. */
int pcp_profile_enable(pcp_ctx* ctx, int on);
int pcp_profile_read(pcp_ctx* ctx, double ms3[3], uint64_t* frames);
/* Debug: milliseconds of the phases of the last pcp_exchange_finalize[_device] made while profiling was enabled:
 * {reduce-scatter, finalize of the rank's slice, all-gather, unpack into the caller's order}.  Synchronises. */
int pcp_debug_exchange_times(pcp_ctx* ctx, double ms4[4]);

/* vlcal::NIDCost (include/vlcal/costs/nid_cost.hpp:24-126).  points: m x {x,y,z,intensity} floats with
 * `stride_bytes` per record; image: W x H doubles in [0,1] (the CV_64F plane of
 * visual_camera_calibration.cpp:173), row-major. */
int pcp_nid_create(pcp_ctx* ctx, const void* points, size_t m, size_t stride_bytes, size_t intensity_offset, const double* image,
                   int width, int height, int bins, pcp_nid** out);
/* The same cost from the 8-bit image itself (visual_camera_calibration.cpp:171-173: imread, then convertTo(CV_64FC1, 1.0 / 255.0)):
 * `image` is a host plane of `height` rows of `width` bytes, `pitch` bytes apart.  The conversion to [0,1] doubles and the binning
 * happen on the device (the cost uses a pixel only through its bin, a function of the byte: results equal pcp_nid_create on
 * double(byte) * (1.0 / 255.0) bit for bit) and 8x fewer bytes cross the host link.  Grey plane: pitch = width.  The reference's
 * actual read (a 3-channel CV_64FC3 image indexed with at<double>(y, x), x < cols: the first `cols` interleaved B,G,R values of every
 * row, nid_cost.hpp:61, :88) is image = the BGR image, pitch = 3 * width. */
int pcp_nid_create_u8(pcp_ctx* ctx, const void* points, size_t m, size_t stride_bytes, size_t intensity_offset, const uint8_t* image,
                      size_t pitch, int width, int height, int bins, pcp_nid** out);
void pcp_nid_destroy(pcp_nid* nid);
/* NIDCost::operator()<double> for B candidate poses at once.  se3: B x 7 doubles in Sophus order
 * {qx,qy,qz,qw,tx,ty,tz}; nid_out: B doubles; ok_out: B bytes, 0 where the reference returns false
 * (non-finite NID, nid_cost.hpp:107-111). */
int pcp_nid_eval(pcp_nid* nid, const double* se3, int B, double* nid_out, uint8_t* ok_out);
/* MultiNIDCost::operator()<double> (src/vlcal/calib/visual_camera_calibration.cpp:93-124): pose gate
 * (0.2 m / 2 deg from init_se3) then the sum over `n_costs` frames. */
int pcp_nid_eval_multi(pcp_nid* const* costs, int n_costs, const double init_se3[7], const double* se3, int B, double* sum_out,
                       uint8_t* ok_out);

/* PointCloudProcessor(pointCloudPath, odometryPath, imagesFolder, maskImageFolder, outputPath, enableMLS, enableNIDOptimize,
 * enableInitialGuessManual) + process()  (include/PointCloudProcessor.hpp:17-27, src/PointCloudProcessor.cpp:1007-1032).
 * Folders are used as prefixes exactly as the reference does (imagesFolder + std::to_string(ts) + ".jpg", cpp:981; outputPath +
 * "filtered_pcd/", cpp:1038), so they normally end with '/'.  Files read: PCD (ascii / binary / binary_compressed), the
 * odometry text "ts x y z qw qx qy qz", <ts>.jpg images (nvJPEG; <ts>.ppm / .png accepted as a lossless extension) and <ts>.png
 * masks (<ts>.pgm accepted).  Files written (ASCII PCD, PCL 1.10 layout): scans-crop.pcd, filtered_pcd/<ts>_beforeNID.pcd,
 * filtered_pcd/<ts>_rgb-mask.pcd, cloudInWorldWithRGB.pcd, cloudInWorldWithRGBandMask.pcd.
 * The constants the reference hard-codes (cpp:51-62, :525) are the defaults of the optional fields. */
typedef struct {
  const char* point_cloud_path;
  const char* odometry_path;
  const char* images_folder;
  const char* mask_image_folder;       /* NULL or "" disables the mask path (cpp:89) */
  const char* output_path;
  int enable_mls;                      /* must be 0: pcl::MovingLeastSquares is out of scope (-> PCP_EINVAL) */
  int enable_nid_optimize;             /* applyNIDBasedPoseOptimization (cpp:156-164): pcp_calibrate over the keyframes' NID costs, from Identity */
  int enable_initial_guess_manual;     /* must be 0: the GUI needs a graphics stack (unavailable on this box) */
  int device;
  const pcp_camera* camera;            /* NULL: the reference camera (cpp:57-60) */
  int cull_width, cull_height;         /* 0: 4096 x 3000 (cpp:525) */
  const double* T_camera_lidar_optimized; /* optional 4x4 column-major: the result NID / the manual guess would have produced (cpp:504-519);
                                          * with enable_nid_optimize it is the initial guess instead of Identity */
  int skip_outputs;                    /* 1: compute everything but write no PCD (benchmarks) */
  int sample_mode;                     /* PCP_SAMPLE_NEAREST (reference) | PCP_SAMPLE_BILINEAR */
  int nid_grey_plane;                  /* 0 (reference): the NID cost reads the image as NIDCost does, the first `cols` interleaved B,G,R values of
                                        * every row (nid_cost.hpp:61, :88); 1: the grey image a single-channel read would have seen */
  int binary_output;                   /* 0 (reference, cpp:920 / :939 / :954): cloudInWorldWithRGB.pcd is written as ASCII; 1: DATA binary (PCL's
                                        * writeBinary layout of pcl::PointXYZRGB: x y z rgb, 16 bytes per point) — at 1e8 points the text file is
                                        * the wall time of process() */
  int reserved[7];
} pcp_process_args;

typedef struct {
  int64_t n_points, n_cropped, n_frames, n_keyframes, n_colored_points, n_mask_points;
  int64_t n_visible_total;             /* sum over keyframes of the culled cloud sizes (loop #1) */
  double seconds_load, seconds_cull_pass, seconds_color_pass, seconds_write, seconds_total;
  double seconds_nid;                  /* enable_nid_optimize: image planes + NID costs + the solver */
  int64_t nid_outer_iterations, nid_inner_iterations, nid_cost_evaluations, nid_converged;
  double nid_initial_cost, nid_final_cost;
  double T_camera_lidar[16];           /* the extrinsic the colour pass used (column-major; identity when none) */
} pcp_process_stats;

/* Returns a pcp_status; on failure `err` (if given) receives the message the reference would have thrown. */
int pcp_process(const pcp_process_args* args, pcp_process_stats* stats, char* err, size_t err_len);

/* File formats at the boundary (host side; replace pcl::io::loadPCDFile<PointXYZI>, pcl::io::savePCDFileASCII and cv::imread at
 * src/PointCloudProcessor.cpp:112, :135, :716, :774).  Buffers returned through *_out are released with pcp_io_free.
 * pcp_pcd_read_xyzi: n x {x,y,z,intensity} floats from an ascii / binary / binary_compressed PCD.
 * pcp_image_read: 8-bit image, grayscale != 0 -> 1 channel, else 3 channels B,G,R; JPEG is decoded with nvJPEG (needs a GPU). */
int pcp_pcd_read_xyzi(const char* path, float** xyzi_out, size_t* n_out);
int pcp_pcd_write_xyzi(const char* path, const float* xyzi, size_t n);
/* saveColorizedPointCloud (cpp:912-923) from 16-byte pcl::PointXYZRGB payloads {x, y, z, b | g << 8 | r << 16 | a << 24} as
 * pcp_finalize_compact delivers them; binary = 0: pcl::io::savePCDFileASCII's layout, 1: pcl::io::savePCDFileBinary's. */
int pcp_pcd_write_xyzrgb(const char* path, const void* points16, size_t n, int binary);
int pcp_image_read(const char* path, int grayscale, uint8_t** data_out, int* width, int* height, int* channels);
void pcp_io_free(void* p);

/* Debug/bench: the 16 u32 device counters of the last submitted frame: {records emitted by the scan, candidates, visible,
 * coloured, active chunks, then (library built with -DPCP_STATS only) chunks skipped whole, points kept by the chunk-level
 * depth test, points kept by the FP32 projection pre-filter}.  Words 14 and 15 (pre-binned cloud): of [15] sampled neighbours in the
 * caller's point order, [14] are also neighbours in the library's sorted order (pcp_finalize picks its gather by that ratio).
 * No reference counterpart. */
int pcp_debug_counters(pcp_ctx* ctx, uint32_t out16[16]);

/* NIDCost::operator()<ceres::Jet<double, 7>> — the instantiation ceres::AutoDiffFirstOrderFunction<MultiNIDCost, 7> evaluates
 * (visual_camera_calibration.cpp:204-233): value and the 7 partial derivatives over the ambient Sophus parameters
 * {qx,qy,qz,qw,tx,ty,tz} (the manifold projection stays with the caller, as with ceres).  grad_out: B x 7. */
int pcp_nid_eval_grad(pcp_nid* nid, const double* se3, int B, double* nid_out, double* grad_out, uint8_t* ok_out);
/* MultiNIDCost::operator()<Jet>: pose gate, then sums of values and gradients over the frames. */
int pcp_nid_eval_multi_grad(pcp_nid* const* costs, int n_costs, const double init_se3[7], const double* se3, int B, double* sum_out,
                            double* grad_sum_out, uint8_t* ok_out);

/* ---- pose refinement: the caller of the NID cost ------------------------------------------------------
 * ceres::GradientProblemSolver with line_search_direction_type = BFGS on Sophus::Manifold<SE3>
 * (src/vlcal/calib/visual_camera_calibration.cpp:204-233).  Ceres / Sophus are not vendored by the reference; the
 * solver is a restatement of their published algorithms with ceres' default constants (csrc/pcp_calib.cu). */
typedef struct {
  int max_num_iterations;                        /* 50 */
  double function_tolerance;                     /* 1e-6 */
  double gradient_tolerance;                     /* 1e-10 */
  double parameter_tolerance;                    /* 1e-8 */
  double sufficient_function_decrease;           /* 1e-4 */
  double sufficient_curvature_decrease;          /* 0.9 */
  double max_line_search_step_expansion;         /* 10 */
  double min_line_search_step_contraction;       /* 0.6 */
  double max_line_search_step_contraction;       /* 1e-3 */
  double min_line_search_step_size;              /* 1e-9 */
  int max_num_line_search_step_size_iterations;  /* 20 */
  int max_num_line_search_direction_restarts;    /* 5 */
  int verbose;                                   /* 1: one line per iteration on stderr (minimizer_progress_to_stdout) */
  int reserved[5];
} pcp_bfgs_options;

enum { PCP_BFGS_CONVERGENCE = 0, PCP_BFGS_NO_CONVERGENCE = 1, PCP_BFGS_FAILURE = 2 };

typedef struct {
  int iterations, termination;                   /* termination: PCP_BFGS_* */
  int num_cost_evaluations, num_line_search_steps;
  double initial_cost, final_cost;
  char message[96];
} pcp_bfgs_summary;

/* A first-order function over SE3 in Sophus' ambient order {qx,qy,qz,qw,tx,ty,tz}: writes the value and the 7 ambient partial
 * derivatives (what AutoDiffFirstOrderFunction<F, 7> produces); returns 0 where the functor returns false. */
typedef int (*pcp_se3_cost_fn)(void* user, const double se3[7], double* value, double grad7[7]);

void pcp_bfgs_default_options(pcp_bfgs_options* options);
/* ceres::Solve(options, GradientProblem(cost, Sophus::Manifold<SE3>), x) for any first-order function (host callback). */
int pcp_se3_bfgs(pcp_se3_cost_fn fn, void* user, const double init_se3[7], const pcp_bfgs_options* options, double out_se3[7],
                 pcp_bfgs_summary* summary);

/* vlcal::VisualCameraCalibrationParams (include/vlcal/calib/visual_camera_calibration.hpp:14-48). */
typedef struct {
  int max_outer_iterations;                      /* 10 */
  double delta_trans_thresh;                     /* 0.01 m */
  double delta_rot_thresh;                       /* 1 degree in radians */
  pcp_bfgs_options inner;
} pcp_calib_params;

typedef struct {
  int outer_iterations, converged;               /* converged: the outer loop met both thresholds */
  int inner_iterations, cost_evaluations, last_inner_termination;
  double initial_cost, final_cost;               /* sum of the frames' NID at the first / last pose */
  double delta_t, delta_r;                       /* last outer step */
} pcp_calib_summary;

void pcp_calib_default_params(pcp_calib_params* params);
/* VisualCameraCalibration::calibrate(init_T_camera_lidar) (visual_camera_calibration.cpp:49-84) over the per-keyframe NID
 * costs (estimate_pose_bfgs, :142-252): outer loop of BFGS solves, each with the pose gate re-centred on its start, until
 * the step between two solves is below both thresholds.  Every cost / gradient evaluation runs on the device
 * (pcp_nid_eval_multi_grad).  Matrices 4x4 column-major. */
int pcp_calibrate(pcp_nid* const* costs, int n_costs, const double init_T_camera_lidar[16], const pcp_calib_params* params,
                  double out_T_camera_lidar[16], pcp_calib_summary* summary);

/* Counters for benches: kernels launched by this context since creation. */
uint64_t pcp_launch_count(const pcp_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* PCP_PCP_H_ */
