// pcp_process.cu — PointCloudProcessor::process() (src/PointCloudProcessor.cpp:1007-1032) on top of the C ABI: load odometry +
// image list, load / crop the PCD, select keyframes, cull every keyframe and write filtered_pcd/<ts>_beforeNID.pcd, colorize
// (+ masks), smooth, drop uncoloured points, write cloudInWorldWithRGB.pcd (+ cloudInWorldWithRGBandMask.pcd).
//
// Host control flow only; every per-point computation of the hot path runs in the CUDA kernels behind pcp_cull_pose /
// pcp_submit_frame / pcp_finalize.  The host re-derives camera-frame coordinates of the (few) points it writes to disk with
// the same float arithmetic (PCL order, no FMA: this file is compiled with -ffp-contract=off).
// enableNIDOptimize runs the extrinsic refinement of csrc/pcp_calib.cu over the device NID costs.  Not available here (SURVEY.md §2
// rows 6, 12): MLS smoothing and the manual-guess GUI; an already optimised T_camera_lidar can be passed in instead
// (pcp_process_args.T_camera_lidar_optimized).
#include <chrono>
#include <cfloat>
#include <cmath>
#include <cstdarg>
#include <filesystem>
#include <memory>

#include "../../include/pcp/pcp.h"
#include "pcp_host_math.hpp"
#include "pcp_imgio.hpp"
#include "pcp_io.hpp"

namespace pcp {
namespace io {
std::string decode_jpeg(const std::vector<uint8_t>& file, HostImage& im, cudaStream_t stream);
}
}  // namespace pcp

namespace {

using namespace pcp;
using pcp::io::HostImage;

struct Frame {  // FrameData (include/FrameData.hpp:89-126), the fields this path uses
  double ts;
  double pose[7];
  std::string image_path, mask_path;
};

struct ProcError : std::runtime_error {
  int code;
  ProcError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

// pcl::transformPointCloud per point, PCL 1.10 order x*c0 + (y*c1 + (z*c2 + c3)) (oracle decision D1; csrc/pcp_math.cuh)
inline void xform(const host::M4f& M, float x, float y, float z, float& ox, float& oy, float& oz) {
  ox = x * host::at(M, 0, 0) + (y * host::at(M, 0, 1) + (z * host::at(M, 0, 2) + host::at(M, 0, 3)));
  oy = x * host::at(M, 1, 0) + (y * host::at(M, 1, 1) + (z * host::at(M, 1, 2) + host::at(M, 1, 3)));
  oz = x * host::at(M, 2, 0) + (y * host::at(M, 2, 1) + (z * host::at(M, 2, 2) + host::at(M, 2, 3)));
}

// cv::imread(path) / cv::imread(path, IMREAD_GRAYSCALE): .jpg/.jpeg through nvJPEG, .png / .ppm / .pgm on the host.
bool load_image(const std::string& path, bool grey, HostImage& im, std::string& why) {
  std::vector<uint8_t> file;
  if (!io::read_file(path, file) || file.empty()) { why = "cannot open"; return false; }
  bool ok = false;
  if (file.size() > 2 && file[0] == 0xff && file[1] == 0xd8) {
    why = io::decode_jpeg(file, im, nullptr);
    ok = why.empty();
  } else if (file[0] == 'P') {
    ok = io::decode_pnm(file, im);
  } else {
    ok = io::decode_png(file, im, grey);
  }
  if (!ok) { if (why.empty()) why = "unsupported or corrupt image"; return false; }
  if (grey) io::to_grey(im); else io::to_bgr(im);
  return true;
}

std::string first_existing(const std::string& stem, std::initializer_list<const char*> exts) {
  for (const char* e : exts)
    if (io::file_exists(stem + e)) return stem + e;
  return std::string();
}

void check(pcp_ctx* ctx, int rc, const char* what) {
  if (rc != PCP_OK) throw ProcError(rc, std::string(what) + ": " + pcp_last_error(ctx));
}

struct CtxGuard {
  pcp_ctx* ctx = nullptr;
  ~CtxGuard() { if (ctx) pcp_destroy(ctx); }
};

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

static void run_process(const pcp_process_args& A, pcp_process_stats& S) {
  namespace fs = std::filesystem;
  const double t_start = now_s();
  if (!A.point_cloud_path || !A.odometry_path || !A.images_folder || !A.output_path) throw ProcError(PCP_EINVAL, "pcp_process: missing path");
  if (A.enable_mls) throw ProcError(PCP_EINVAL, "pcp_process: enableMLS is not available in this build (pcl::MovingLeastSquares is out of scope)");
  if (A.enable_initial_guess_manual) throw ProcError(PCP_EINVAL, "pcp_process: enableInitialGuessManual needs a graphics stack (unavailable)");
  const std::string images_folder = A.images_folder, out_path = A.output_path;
  const std::string mask_folder = A.mask_image_folder ? A.mask_image_folder : "";
  const bool masks_on = !mask_folder.empty();  // cpp:89

  // ---- loadImagesAndOdometry (cpp:965-1005) ----------------------------------------------------------------------------
  std::vector<Frame> frames;
  for (const io::OdomLine& o : io::read_odometry(A.odometry_path)) {
    const std::string stem = images_folder + io::to_string_f(o.ts);
    const std::string img = first_existing(stem, {".jpg", ".ppm", ".png"});  // .ppm / .png: lossless extension of cpp:981
    if (img.empty()) continue;                                                 // cpp:984-987
    Frame f;
    f.ts = o.ts;
    std::memcpy(f.pose, o.pose, sizeof(f.pose));
    f.image_path = img;
    if (masks_on) {
      const std::string mstem = mask_folder + io::to_string_f(o.ts);
      f.mask_path = first_existing(mstem, {".png", ".pgm"});
      if (f.mask_path.empty()) f.mask_path = mstem + ".png";  // cpp:991: the path is recorded even when the file is missing
    }
    frames.push_back(f);
  }
  S.n_frames = (int64_t)frames.size();

  // ---- loadPointCloud (cpp:92-154) -------------------------------------------------------------------------------------
  double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {DBL_MIN, DBL_MIN, DBL_MIN};  // DBL_MIN as in cpp:96 (smallest positive double)
  for (const Frame& f : frames)
    for (int k = 0; k < 3; k++) { mn[k] = std::fmin(mn[k], f.pose[k]); mx[k] = std::fmax(mx[k], f.pose[k]); }
  for (int k = 0; k < 3; k++) { mn[k] -= 2.0; mx[k] += 2.0; }  // padding, cpp:106-108
  std::vector<io::XYZI> cloud;
  io::read_pcd_xyzi(A.point_cloud_path, cloud);
  S.n_points = (int64_t)cloud.size();
  {
    // pcl::CropBox with Vector4f(min), Vector4f(max): a point is dropped when it is not finite or outside [min, max] (cpp:120-125)
    const float fmn[3] = {(float)mn[0], (float)mn[1], (float)mn[2]}, fmx[3] = {(float)mx[0], (float)mx[1], (float)mx[2]};
    std::vector<io::XYZI> cropped;
    for (const io::XYZI& p : cloud) {
      if (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z)) continue;
      if (p.x < fmn[0] || p.y < fmn[1] || p.z < fmn[2] || p.x > fmx[0] || p.y > fmx[1] || p.z > fmx[2]) continue;
      cropped.push_back(p);
    }
    S.n_cropped = (int64_t)cropped.size();
    if (!A.skip_outputs) {
      std::error_code ec;
      fs::create_directories(fs::path(out_path + "x").parent_path(), ec);
      io::write_pcd_xyzi(out_path + "scans-crop.pcd", cropped.data(), cropped.size());  // cpp:131-135 (the FULL cloud is used below, cpp:146)
    }
  }

  // ---- generateResultStorageFolder (cpp:1034-1048) ---------------------------------------------------------------------
  if (!A.skip_outputs) {
    const fs::path dir(out_path + "filtered_pcd/");
    std::error_code ec;
    if (fs::exists(dir, ec)) fs::remove_all(dir, ec);
    fs::create_directories(dir, ec);
  }

  // ---- selectKeyframes (cpp:1050-1075, hpp:151-191): first frame, then every frame >= 0.1 m from the last keyframe ---------
  std::vector<const Frame*> keyframes;
  const Frame* prev = nullptr;
  for (const Frame& f : frames) {
    bool is_key = prev == nullptr;
    if (prev) {
      const double dx = f.pose[0] - prev->pose[0], dy = f.pose[1] - prev->pose[1], dz = f.pose[2] - prev->pose[2];
      is_key = std::sqrt(dx * dx + dy * dy + dz * dz) >= 0.1;  // hpp:166, :181 (the angle is computed but unused)
    }
    if (is_key) { keyframes.push_back(&f); prev = &f; }
  }
  S.n_keyframes = (int64_t)keyframes.size();
  S.seconds_load = now_s() - t_start;
  if (keyframes.empty() || cloud.empty()) { S.seconds_total = now_s() - t_start; return; }

  // ---- device context ---------------------------------------------------------------------------------------------------
  HostImage first;
  std::string why;
  if (!load_image(keyframes[0]->image_path, false, first, why)) throw ProcError(PCP_EIO, "Failed to read image from: " + keyframes[0]->image_path);  // cpp:719
  pcp_camera cam;
  if (A.camera) cam = *A.camera;
  else cam = pcp_camera{4818.200388954926, 4819.10345841615, 2032.4178620390019, 1535.1895959282901,
                        {0.003043514741045163, 0.06634739187544138, -0.000217681797407554, -0.0006654964142658197, 0.0}};  // cpp:57-60
  pcp_params prm;
  pcp_default_params(&prm);
  prm.image_width = first.width;
  prm.image_height = first.height;
  prm.cull_width = A.cull_width > 0 ? A.cull_width : 4096;    // cpp:525
  prm.cull_height = A.cull_height > 0 ? A.cull_height : 3000;
  prm.accum_mode = PCP_ACCUM_TOP5;                            // smoothColors, cpp:604-631
  prm.sample_mode = A.sample_mode;
  prm.hsv_adjust = 1;  // generateColorMap's HSV round trip runs on the uploaded image; the per-frame mask clouds are built on the device from it
  CtxGuard G;
  {
    const int rc = pcp_create(&G.ctx, A.device, &cam, &prm);
    if (rc != PCP_OK) throw ProcError(rc, std::string("pcp_create: ") + pcp_last_error(nullptr));
  }
  pcp_ctx* ctx = G.ctx;
  check(ctx, pcp_set_cloud(ctx, cloud.data(), cloud.size(), sizeof(io::XYZI), 12), "pcp_set_cloud");
  const size_t N = cloud.size();

  // ---- viewCullingAndSaveFilteredPcds (cpp:178-224) ---------------------------------------------------------------------
  const double t_cull = now_s();
  std::vector<int32_t> idx(N);
  std::vector<io::XYZI> culled;
  std::vector<std::vector<io::XYZI>> culled_for_nid;  // what estimate_pose_bfgs reads back from <ts>_beforeNID.pcd
  for (const Frame* kf : keyframes) {
    size_t m = 0;
    check(ctx, pcp_cull_pose(ctx, kf->pose, nullptr, idx.data(), &m), "pcp_cull_pose");  // the un-optimised pose, cpp:188-196
    S.n_visible_total += (int64_t)m;
    if (A.skip_outputs && !A.enable_nid_optimize) continue;
    host::M4f w2c, c2w;
    host::pose_matrices(kf->pose, nullptr, w2c, c2w);
    culled.resize(m);
    for (size_t j = 0; j < m; j++) {
      const io::XYZI& p = cloud[(size_t)idx[j]];
      xform(w2c, p.x, p.y, p.z, culled[j].x, culled[j].y, culled[j].z);
      culled[j].intensity = p.intensity;
    }
    if (!A.skip_outputs) {
      const std::string path = out_path + "filtered_pcd/" + io::to_string_f(kf->ts) + "_beforeNID.pcd";  // cpp:213
      if (!io::write_pcd_xyzi(path, culled.data(), m)) throw ProcError(PCP_EIO, "Couldn't save filtered point cloud to PCD file.");  // cpp:220
    }
    if (A.enable_nid_optimize) {
      // the optimiser re-reads the ASCII file (visual_camera_calibration.cpp:163): 8 significant digits do not round-trip every
      // float, so the in-memory copy goes through the same text form
      for (size_t j = 0; j < m; j++) {
        culled[j].x = io::through_text(culled[j].x); culled[j].y = io::through_text(culled[j].y);
        culled[j].z = io::through_text(culled[j].z); culled[j].intensity = io::through_text(culled[j].intensity);
      }
      culled_for_nid.push_back(culled);
    }
  }
  S.seconds_cull_pass = now_s() - t_cull;

  // ---- applyNIDBasedPoseOptimization (cpp:156-164, src/calibrate.cpp:42-79, visual_camera_calibration.cpp:49-252) ------------
  double T_nid[16];
  const double* T_opt = A.T_camera_lidar_optimized;
  if (A.enable_nid_optimize) {
    const double t_nid = now_s();
    struct NidGuard { std::vector<pcp_nid*> v; ~NidGuard() { for (pcp_nid* n : v) pcp_nid_destroy(n); } } nids;
    for (size_t k = 0; k < keyframes.size(); k++) {
      HostImage img;
      if (k == 0) img = first;
      else if (!load_image(keyframes[k]->image_path, false, img, why)) throw ProcError(PCP_EIO, "Failed to read image from: " + keyframes[k]->image_path);
      // cv::imread gives a 3-channel image and convertTo(CV_64FC1, 1/255) keeps the channels (only the depth changes), yet NIDCost reads
      // it with at<double>(y, x) and cols x rows bounds (nid_cost.hpp:61, :88): element x of row y of the INTERLEAVED doubles, i.e. the
      // first cols values B0 G0 R0 B1 ... of each row.  Restated literally: the 8-bit BGR image goes to the device as it is, with the
      // row pitch of the interleaved image and `cols` bytes per row (nid_grey_plane = 1 takes the grey image a single-channel read
      // would have seen instead).  The conversion to doubles / bins happens on the device (pcp_nid_create_u8).
      const int w = img.width, h = img.height;
      if (A.nid_grey_plane) io::to_grey(img);
      pcp_nid* nid = nullptr;
      const std::vector<io::XYZI>& pts = culled_for_nid[k];
      if (pts.empty()) continue;  // a frame that sees nothing adds no cost
      check(ctx, pcp_nid_create_u8(ctx, pts.data(), pts.size(), sizeof(io::XYZI), 12, img.data.data(), (size_t)w * (size_t)img.channels, w, h, 16, &nid),
            "pcp_nid_create_u8");  // nid_bins = 16, calibrate.cpp:52
      nids.v.push_back(nid);
    }
    if (nids.v.empty()) throw ProcError(PCP_EINVAL, "pcp_process: NID optimisation has no visible points in any keyframe");
    pcp_calib_params cp;
    pcp_calib_default_params(&cp);
    cp.inner.verbose = std::getenv("PCP_VERBOSE") != nullptr;
    double T0[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};  // init_T_camera_lidar = Identity, calibrate.cpp:46
    if (A.T_camera_lidar_optimized) std::memcpy(T0, A.T_camera_lidar_optimized, sizeof(T0));  // extension: a caller-supplied initial guess
    pcp_calib_summary cs;
    check(ctx, pcp_calibrate(nids.v.data(), (int)nids.v.size(), T0, &cp, T_nid, &cs), "pcp_calibrate");
    T_opt = T_nid;
    S.nid_outer_iterations = cs.outer_iterations;
    S.nid_inner_iterations = cs.inner_iterations;
    S.nid_cost_evaluations = cs.cost_evaluations;
    S.nid_converged = cs.converged;
    S.nid_initial_cost = cs.initial_cost;
    S.nid_final_cost = cs.final_cost;
    S.seconds_nid = now_s() - t_nid;
  }
  for (int i = 0; i < 16; i++) S.T_camera_lidar[i] = T_opt ? T_opt[i] : (i % 5 == 0 ? 1.0 : 0.0);

  // ---- pcdColorizationAndSmooth (cpp:474-602) ---------------------------------------------------------------------------
  const double t_col = now_s();
  const int W = prm.image_width, H = prm.image_height;
  std::vector<io::XYZRGBMask> world_mask_cloud, frame_mask_cloud;
  static_assert(sizeof(io::XYZRGBMask) == 20, "pcp_frame_mask_cloud writes 20-byte records");
  // pinned staging for the images (two buffers: the upload of one frame overlaps the decode of the next)
  uint8_t* h_img[2] = {nullptr, nullptr};
  uint8_t* h_msk[2] = {nullptr, nullptr};
  struct PinGuard { uint8_t** a; uint8_t** b; ~PinGuard() { for (int i = 0; i < 2; i++) { if (a[i]) pcp_host_free(a[i]); if (b[i]) pcp_host_free(b[i]); } } } pg{h_img, h_msk};
  for (int i = 0; i < 2; i++) {
    if (pcp_host_alloc(reinterpret_cast<void**>(&h_img[i]), (size_t)W * H * 3) != PCP_OK) throw ProcError(PCP_ENOMEM, "pcp_process: pinned allocation failed");
    if (masks_on && pcp_host_alloc(reinterpret_cast<void**>(&h_msk[i]), (size_t)W * H) != PCP_OK) throw ProcError(PCP_ENOMEM, "pcp_process: pinned allocation failed");
  }
  double write_s = 0.0;
  for (size_t k = 0; k < keyframes.size(); k++) {
    const Frame* kf = keyframes[k];
    const int sb = (int)(k & 1);
    HostImage img;
    if (k == 0) img = std::move(first);
    else if (!load_image(kf->image_path, false, img, why)) throw ProcError(PCP_EIO, "Failed to read image from: " + kf->image_path);  // cpp:719
    if (img.width != W || img.height != H) throw ProcError(PCP_EINVAL, "pcp_process: image size differs from the first keyframe's: " + kf->image_path);
    if (k >= 2) check(ctx, pcp_synchronize(ctx), "pcp_synchronize");  // the staging buffer of frame k-2 is free again
    std::memcpy(h_img[sb], img.data.data(), (size_t)W * H * 3);
    bool have_mask = false;
    if (masks_on) {
      HostImage m;
      why.clear();
      if (load_image(kf->mask_path, true, m, why)) {
        if (m.width != W || m.height != H)
          throw ProcError(PCP_EIO, "mask image " + kf->mask_path + " is " + std::to_string(m.width) + "x" + std::to_string(m.height) + ", the camera image " +
                                       std::to_string(W) + "x" + std::to_string(H));
        std::memcpy(h_msk[sb], m.data.data(), (size_t)W * H);
        have_mask = true;
      } else if (why == "cannot open") {
        std::fprintf(stdout, "Failed to read image from: %s\n", kf->mask_path.c_str());  // cpp:776-781: the file is missing: log and carry on without the mask
      } else {
        // the file is there but this build cannot decode it: cv::imread would have read it, so going on without the mask would silently
        // drop <ts>_rgb-mask.pcd and the mask cloud.  Fail with the path instead.
        throw ProcError(PCP_EIO, "Failed to decode mask image: " + kf->mask_path + " (" + why + ")");
      }
    }
    pcp_frame_out fo;
    std::memset(&fo, 0, sizeof(fo));
    fo.flags = PCP_OUT_SKIP_CANDIDATES | PCP_OUT_PIXELS_ON_DEVICE;  // the pixel indices stay on the device for pcp_frame_mask_cloud
    check(ctx, pcp_submit_frame(ctx, kf->pose, T_opt, h_img[sb], (size_t)W * 3, have_mask ? h_msk[sb] : nullptr, (size_t)W,
                                have_mask ? &fo : nullptr),
          "pcp_submit_frame");
    if (!have_mask) continue;
    // generateSegmentMap (cpp:769-816) for this keyframe, gathered on the device: only the points that received a pixel come back,
    // as PointXYZRGBMask records in the camera frame (cloudInCameraPose, cpp:521), in the map's order
    host::M4f w2c, c2w;
    host::pose_matrices(kf->pose, T_opt, w2c, c2w);
    {
      size_t m = 0;
      check(ctx, pcp_frame_mask_cloud(ctx, nullptr, 0, &m), "pcp_frame_mask_cloud");
      frame_mask_cloud.resize(m);
      if (m) check(ctx, pcp_frame_mask_cloud(ctx, frame_mask_cloud.data(), m, &m), "pcp_frame_mask_cloud");
    }
    S.n_mask_points += (int64_t)frame_mask_cloud.size();
    const double tw = now_s();
    if (!A.skip_outputs) {
      const std::string path = out_path + "filtered_pcd/" + io::to_string_f(kf->ts) + "_rgb-mask.pcd";  // cpp:539
      if (!io::write_pcd_xyzrgbmask(path, frame_mask_cloud.data(), frame_mask_cloud.size()))
        throw ProcError(PCP_EIO, "Couldn't save filtered point cloud to PCD file.");  // cpp:544
    }
    for (io::XYZRGBMask q : frame_mask_cloud) {  // pcl::transformPointCloud(scanInBody, scanInWorld, c2w) (cpp:549)
      float x, y, z;
      xform(c2w, q.x, q.y, q.z, x, y, z);
      q.x = x; q.y = y; q.z = z;
      world_mask_cloud.push_back(q);
    }
    write_s += now_s() - tw;
  }
  // smoothColors + removePointsWithNoColor (cpp:604-631, hpp:238-252)
  // the compaction runs on the device (pcp_finalize_compact): only the coloured points cross the host link, already in
  // pcl::PointXYZRGB's 16-byte layout and in the map's point order
  std::vector<io::XYZRGB> colored;
  {
    size_t m = 0;
    check(ctx, pcp_finalize_compact(ctx, nullptr, 0, &m), "pcp_finalize_compact");
    colored.resize(m);
    if (m) check(ctx, pcp_finalize_compact(ctx, colored.data(), m, &m), "pcp_finalize_compact");
  }
  S.n_colored_points = (int64_t)colored.size();
  S.seconds_color_pass = now_s() - t_col - write_s;

  // ---- saveColorizedPointCloud (cpp:912-963) ----------------------------------------------------------------------------
  const double tw = now_s();
  if (!A.skip_outputs) {
    if (masks_on && !world_mask_cloud.empty() &&
        !io::write_pcd_xyzrgbmask(out_path + "cloudInWorldWithRGBandMask.pcd", world_mask_cloud.data(), world_mask_cloud.size()))
      throw ProcError(PCP_EIO, "Couldn't save colorized and segment colored point cloud.");  // cpp:941
    if (!colored.empty() && !(A.binary_output ? io::write_pcd_xyzrgb_binary(out_path + "cloudInWorldWithRGB.pcd", colored.data(), colored.size())
                                              : io::write_pcd_xyzrgb(out_path + "cloudInWorldWithRGB.pcd", colored.data(), colored.size())))
      throw ProcError(PCP_EIO, "Couldn't save colorized point cloud.");  // cpp:922, :956
  }
  S.seconds_write = write_s + (now_s() - tw);
  S.seconds_total = now_s() - t_start;
}

extern "C" int pcp_process(const pcp_process_args* args, pcp_process_stats* stats, char* err, size_t err_len) {
  pcp_process_stats local;
  std::memset(&local, 0, sizeof(local));
  auto report = [&](const char* m) { if (err && err_len) { std::strncpy(err, m, err_len - 1); err[err_len - 1] = 0; } };
  if (err && err_len) err[0] = 0;
  int rc = PCP_OK;
  if (!args) { report("pcp_process: null arguments"); return PCP_EINVAL; }
  try {
    run_process(*args, local);
  } catch (const ProcError& e) {
    report(e.what());
    rc = e.code;
  } catch (const std::exception& e) {  // std::runtime_error from the readers: the reference's messages
    report(e.what());
    rc = PCP_EIO;
  }
  if (stats) *stats = local;
  return rc;
}

extern "C" int pcp_pcd_read_xyzi(const char* path, float** xyzi_out, size_t* n_out) {
  if (!path || !xyzi_out || !n_out) return PCP_EINVAL;
  *xyzi_out = nullptr;
  *n_out = 0;
  try {
    std::vector<pcp::io::XYZI> pts;
    pcp::io::read_pcd_xyzi(path, pts);
    float* out = static_cast<float*>(std::malloc(sizeof(float) * 4 * (pts.size() ? pts.size() : 1)));
    if (!out) return PCP_ENOMEM;
    std::memcpy(out, pts.data(), sizeof(float) * 4 * pts.size());
    *xyzi_out = out;
    *n_out = pts.size();
  } catch (const std::exception&) {
    return PCP_EIO;
  }
  return PCP_OK;
}

extern "C" int pcp_pcd_write_xyzi(const char* path, const float* xyzi, size_t n) {
  if (!path || (!xyzi && n)) return PCP_EINVAL;
  return pcp::io::write_pcd_xyzi(path, reinterpret_cast<const pcp::io::XYZI*>(xyzi), n) ? PCP_OK : PCP_EIO;
}

extern "C" int pcp_image_read(const char* path, int grayscale, uint8_t** data_out, int* width, int* height, int* channels) {
  if (!path || !data_out || !width || !height || !channels) return PCP_EINVAL;
  *data_out = nullptr;
  HostImage im;
  std::string why;
  if (!load_image(path, grayscale != 0, im, why)) return PCP_EIO;
  uint8_t* out = static_cast<uint8_t*>(std::malloc(im.data.size() ? im.data.size() : 1));
  if (!out) return PCP_ENOMEM;
  std::memcpy(out, im.data.data(), im.data.size());
  *data_out = out;
  *width = im.width; *height = im.height; *channels = im.channels;
  return PCP_OK;
}

extern "C" int pcp_pcd_write_xyzrgb(const char* path, const void* points16, size_t n, int binary) {
  if (!path || (!points16 && n)) return PCP_EINVAL;
  const pcp::io::XYZRGB* p = static_cast<const pcp::io::XYZRGB*>(points16);
  return (binary ? pcp::io::write_pcd_xyzrgb_binary(path, p, n) : pcp::io::write_pcd_xyzrgb(path, p, n)) ? PCP_OK : PCP_EIO;
}

extern "C" void pcp_io_free(void* p) { std::free(p); }
