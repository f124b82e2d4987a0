// pcp.hpp — header-only C++ facade over the C ABI of pcp.h with the reference's class and method names.
//
// PCL / OpenCV / Eigen are not available on this box, so their types are replaced by layout-compatible PODs:
//   pcp::PointXYZI        == pcl::PointXYZI        (32 bytes: x y z pad | intensity pad pad pad)
//   pcp::PointXYZRGB      == pcl::PointXYZRGB      (32 bytes: x y z pad | b g r a pad pad pad)
//   pcp::PointXYZRGBMask  == PointXYZRGBMask       (include/FrameData.hpp:68-87)
//   pcp::Isometry3d       == Eigen::Isometry3d     (double[16], column-major)
//   pcp::Image            ~  cv::Mat               (8-bit, data/pitch/width/height/channels)
// Errors: every failure of the C ABI is rethrown as std::runtime_error(pcp_last_error()), the convention main.cpp:64-68
// already handles.  Citations are relative to /root/reference/PointCloudProcessor/.
#ifndef PCP_PCP_HPP_
#define PCP_PCP_HPP_

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "pcp.h"

namespace pcp {

struct alignas(16) PointXYZI { float x, y, z, pad0; float intensity, pad1, pad2, pad3; };
struct alignas(16) PointXYZRGB { float x, y, z, pad0; uint8_t b, g, r, a; float pad1, pad2, pad3; };
struct alignas(16) PointXYZRGBMask { float x, y, z, pad0; uint8_t b, g, r, a; uint16_t segmentMask; uint16_t pad1; float pad2, pad3; };
static_assert(sizeof(PointXYZI) == 32 && sizeof(PointXYZRGB) == 32 && sizeof(PointXYZRGBMask) == 32, "PCL point layouts");

struct Isometry3d {
  double m[16];  // column-major, Eigen's default
  static Isometry3d Identity() { Isometry3d T{}; T.m[0] = T.m[5] = T.m[10] = T.m[15] = 1.0; return T; }
  const double* data() const { return m; }
};

struct Image {
  const uint8_t* data = nullptr;
  int width = 0, height = 0, channels = 3;
  size_t pitch = 0;  // bytes per row
};

// include/FrameData.hpp:9-12
struct Pose { double x, y, z, qw, qx, qy, qz; };

inline void check(int rc, pcp_ctx* ctx, const char* what) {
  if (rc != PCP_OK) {
    const char* msg = pcp_last_error(ctx);
    throw std::runtime_error(std::string(what) + ": " + (msg ? msg : "error"));
  }
}

}  // namespace pcp

namespace camera {

// camera::GenericCameraBase / create_camera (include/camera/create_camera.hpp:14, src/camera/create_camera.cpp:33-49)
struct GenericCameraBase {
  using ConstPtr = std::shared_ptr<const GenericCameraBase>;
  pcp_camera cam{};
};

inline GenericCameraBase::ConstPtr create_camera(const std::string& camera_model, const std::vector<double>& intrinsics,
                                                 const std::vector<double>& distortion_coeffs) {
  if (camera_model != "pinhole") throw std::runtime_error("error: unknown camera model " + camera_model);
  if (intrinsics.size() != 4) throw std::runtime_error("error: num of intrinsic parameters mismatch!!");
  auto c = std::make_shared<GenericCameraBase>();
  c->cam.fx = intrinsics[0]; c->cam.fy = intrinsics[1]; c->cam.cx = intrinsics[2]; c->cam.cy = intrinsics[3];
  for (size_t i = 0; i < 5; i++) c->cam.dist[i] = i < distortion_coeffs.size() ? distortion_coeffs[i] : 0.0;  // create_camera.cpp:23-26
  return c;
}

}  // namespace camera

namespace pcp {

// Device-resident state of pcdColorizationAndSmooth (src/PointCloudProcessor.cpp:474-602) on one GPU.
class Colorizer {
 public:
  Colorizer(const camera::GenericCameraBase::ConstPtr& proj, const pcp_params& params, int device = 0) : params_(params) {
    if (pcp_create(&ctx_, device, &proj->cam, &params) != PCP_OK) throw std::runtime_error(pcp_last_error(nullptr));
  }
  ~Colorizer() { pcp_destroy(ctx_); }
  Colorizer(const Colorizer&) = delete;
  Colorizer& operator=(const Colorizer&) = delete;

  static pcp_params default_params(int image_width, int image_height) {
    pcp_params p;
    pcp_default_params(&p);
    p.image_width = image_width; p.image_height = image_height;
    return p;  // cull size stays the reference's hard-coded 4096x3000 (cpp:525) unless the caller changes it
  }

  // generate_distance_masks (scripts/genNormAndDistanceMask.py:233-265) of the frame's culled cloud: H x W ranges (0 = empty),
  // optional H x W x 3 camera-frame points and H x W winning point indices (-1 = empty).  Returns the number of pixels written.
  size_t distance_masks(const Pose& pose, std::vector<float>& distance_mask, std::vector<float>* points_3d_mask = nullptr,
                        std::vector<int32_t>* index_mask = nullptr, const Isometry3d* T_opt = nullptr) {
    const double p7[7] = {pose.x, pose.y, pose.z, pose.qw, pose.qx, pose.qy, pose.qz};
    const size_t npix = (size_t)params_.image_width * params_.image_height;
    distance_mask.resize(npix);
    if (points_3d_mask) points_3d_mask->resize(3 * npix);
    if (index_mask) index_mask->resize(npix);
    size_t filled = 0;
    check(pcp_splat_masks_pose(ctx_, p7, T_opt ? T_opt->data() : nullptr, distance_mask.data(), points_3d_mask ? points_3d_mask->data() : nullptr,
                               index_mask ? index_mask->data() : nullptr, &filled), ctx_, "pcp_splat_masks_pose");
    return filled;
  }
  pcp_ctx* handle() const { return ctx_; }
  size_t size() const { return pcp_num_points(ctx_); }

  // loadPointCloud (cpp:92-154)
  void setCloud(const std::vector<PointXYZI>& cloud) {
    check(pcp_set_cloud(ctx_, cloud.data(), cloud.size(), sizeof(PointXYZI), offsetof(PointXYZI, intensity)), ctx_, "pcp_set_cloud");
  }
  // one keyframe of the loop at cpp:488-595
  void addFrame(const Pose& pose, const Image& bgr, const Image* mask = nullptr, const Isometry3d* T_opt = nullptr,
                pcp_frame_out* out = nullptr) {
    const double p7[7] = {pose.x, pose.y, pose.z, pose.qw, pose.qx, pose.qy, pose.qz};
    check(pcp_submit_frame(ctx_, p7, T_opt ? T_opt->data() : nullptr, bgr.data, bgr.pitch, mask ? mask->data : nullptr,
                           mask ? mask->pitch : 0, out), ctx_, "pcp_submit_frame");
  }
  // loop body of viewCullingAndSaveFilteredPcds (cpp:186-207): indices kept by the z-buffer culler, ascending
  std::vector<int32_t> cullPose(const Pose& pose, const Isometry3d* T_opt = nullptr) {
    const double p7[7] = {pose.x, pose.y, pose.z, pose.qw, pose.qx, pose.qy, pose.qz};
    std::vector<int32_t> idx(size() ? size() : 1);
    size_t m = 0;
    check(pcp_cull_pose(ctx_, p7, T_opt ? T_opt->data() : nullptr, idx.data(), &m), ctx_, "pcp_cull_pose");
    idx.resize(m);
    return idx;
  }
  // smoothColors (cpp:604-631) + the information removePointsWithNoColor needs (hpp:238-252)
  void smoothColors(std::vector<uint8_t>& rgb, std::vector<uint32_t>& count) {
    rgb.resize(3 * size()); count.resize(size());
    check(pcp_finalize(ctx_, rgb.data(), count.data()), ctx_, "pcp_finalize");
  }
  void allreduce() { check(pcp_allreduce(ctx_), ctx_, "pcp_allreduce"); }

  // removePointsWithNoColor after smoothColors (hpp:238-252, cpp:604-631) with the compaction on the device: the coloured points
  // only, as pcl::PointXYZRGB, in the map's order
  std::vector<PointXYZRGB> coloredCloud() {
    struct Rec { float x, y, z; uint8_t b, g, r, a; };
    static_assert(sizeof(Rec) == 16, "pcp_finalize_compact writes 16-byte records");
    size_t m = 0;
    check(pcp_finalize_compact(ctx_, nullptr, 0, &m), ctx_, "pcp_finalize_compact");
    std::vector<Rec> rec(m);
    if (m) check(pcp_finalize_compact(ctx_, rec.data(), m, &m), ctx_, "pcp_finalize_compact");
    std::vector<PointXYZRGB> out(m);
    for (size_t i = 0; i < m; i++) out[i] = PointXYZRGB{rec[i].x, rec[i].y, rec[i].z, 1.0f, rec[i].b, rec[i].g, rec[i].r, rec[i].a, 0.0f, 0.0f, 0.0f};
    return out;
  }
  // generateSegmentMap (cpp:769-816) for the keyframe just added: the points that received a pixel, in the camera frame, with the
  // image's colour (red where the mask is 255) and the mask value.  Add that keyframe with a mask and `keepPixels(out)`.
  static void keepPixels(pcp_frame_out& out) { out.flags |= PCP_OUT_PIXELS_ON_DEVICE | PCP_OUT_SKIP_CANDIDATES; }
  std::vector<PointXYZRGBMask> segmentCloud() {
    struct Rec { float x, y, z; uint8_t b, g, r, a; uint16_t mask, pad; };
    static_assert(sizeof(Rec) == 20, "pcp_frame_mask_cloud writes 20-byte records");
    size_t m = 0;
    check(pcp_frame_mask_cloud(ctx_, nullptr, 0, &m), ctx_, "pcp_frame_mask_cloud");
    std::vector<Rec> rec(m);
    if (m) check(pcp_frame_mask_cloud(ctx_, rec.data(), m, &m), ctx_, "pcp_frame_mask_cloud");
    std::vector<PointXYZRGBMask> out(m);
    for (size_t i = 0; i < m; i++)
      out[i] = PointXYZRGBMask{rec[i].x, rec[i].y, rec[i].z, 1.0f, rec[i].b, rec[i].g, rec[i].r, rec[i].a, rec[i].mask, 0, 0.0f, 0.0f};
    return out;
  }

  // ---- several GPUs, one Colorizer per GPU (no counterpart in the reference, which is one process) ------------------------------
  static void ncclUniqueId(uint8_t id128[128]) {
    if (pcp_nccl_unique_id(id128) != PCP_OK) throw std::runtime_error("pcp_nccl_unique_id failed");
  }
  void ncclInit(const uint8_t id128[128], int rank, int nranks) { check(pcp_nccl_init(ctx_, id128, rank, nranks), ctx_, "pcp_nccl_init"); }
  // keyframe numbers of this rank's share (global numbers break ties of equal scores as one process would)
  void setFrameNumbering(int next, int stride) { check(pcp_set_frame_numbering(ctx_, next, stride), ctx_, "pcp_set_frame_numbering"); }
  // every rank passes the WHOLE map and uploads only its slice of it; the slices are all-gathered over NVLink
  void setCloudSharded(const std::vector<PointXYZI>& cloud, int rank, int nranks) {
    size_t first = 0, count = 0;
    if (pcp_cloud_slice(cloud.size(), rank, nranks, &first, &count) != PCP_OK) throw std::runtime_error("pcp_cloud_slice: bad rank");
    check(pcp_set_cloud_sharded(ctx_, cloud.data() + first, cloud.size(), sizeof(PointXYZI), offsetof(PointXYZI, intensity)), ctx_, "pcp_set_cloud_sharded");
  }
  // smoothColors of the whole job on every rank: reduce-scatter of the sums, finalize of the rank's slice, all-gather
  void exchangeAndSmoothColors(std::vector<uint8_t>& rgb, std::vector<uint32_t>& count) {
    rgb.resize(3 * size()); count.resize(size());
    check(pcp_exchange_finalize(ctx_, rgb.data(), count.data()), ctx_, "pcp_exchange_finalize");
  }

 private:
  pcp_ctx* ctx_ = nullptr;
  pcp_params params_;
};

}  // namespace pcp

namespace vlcal {

// include/vlcal/calib/view_culling.hpp:10-19
struct ViewCullingParams {
  bool enable_depth_buffer_culling = true;
  double hidden_points_removal_max_z = 200.0;  // Katz HPR parameters: unused, the z-buffer culler replaces HPR (north_star b)
};

// vlcal::ViewCulling(proj, image_size, params).cull(points, T_camera_lidar) — include/vlcal/calib/view_culling.hpp:23-34,
// z-buffer variant src/vlcal/calib/view_culling.cpp:52-174.
class ViewCulling {
 public:
  ViewCulling(const camera::GenericCameraBase::ConstPtr& proj, int image_width, int image_height, const ViewCullingParams& params,
              int device = 0) {
    pcp_params p;
    pcp_default_params(&p);
    p.image_width = p.cull_width = image_width;
    p.image_height = p.cull_height = image_height;
    p.enable_depth_buffer_culling = params.enable_depth_buffer_culling ? 1 : 0;
    c_.reset(new pcp::Colorizer(proj, p, device));
  }
  std::vector<pcp::PointXYZI> cull(const std::vector<pcp::PointXYZI>& points, const pcp::Isometry3d& T_camera_lidar) const {
    c_->setCloud(points);
    std::vector<int32_t> idx(points.size() ? points.size() : 1);
    size_t m = 0;
    pcp::check(pcp_cull(c_->handle(), nullptr, T_camera_lidar.data(), idx.data(), &m), c_->handle(), "pcp_cull");
    std::vector<pcp::PointXYZI> out;
    out.reserve(m);
    for (size_t k = 0; k < m; k++) out.push_back(points[idx[k]]);  // ViewCulling::sample (view_culling.cpp:176-185)
    return out;
  }

 private:
  std::unique_ptr<pcp::Colorizer> c_;
};

// vlcal::NIDCost(proj, normalized_image, points, bins) — include/vlcal/costs/nid_cost.hpp:26-43 (value path, T = double).
class NIDCost {
 public:
  NIDCost(pcp::Colorizer& ctx, const double* normalized_image, int width, int height, const std::vector<pcp::PointXYZI>& points,
          int bins = 16) : ctx_(ctx.handle()) {
    pcp::check(pcp_nid_create(ctx_, points.data(), points.size(), sizeof(pcp::PointXYZI), offsetof(pcp::PointXYZI, intensity),
                              normalized_image, width, height, bins, &nid_), ctx_, "pcp_nid_create");
  }
  ~NIDCost() { pcp_nid_destroy(nid_); }
  NIDCost(const NIDCost&) = delete;
  NIDCost& operator=(const NIDCost&) = delete;
  // bool operator()(const T* T_camera_lidar_params, T* residual) with Sophus parameter order {qx,qy,qz,qw,tx,ty,tz}
  bool operator()(const double* T_camera_lidar_params, double* residual) const {
    uint8_t ok = 0;
    pcp::check(pcp_nid_eval(nid_, T_camera_lidar_params, 1, residual, &ok), ctx_, "pcp_nid_eval");
    return ok != 0;
  }
  // B candidate poses per launch (what a batched line search or derivative-free search should call)
  void evaluate(const double* params_Bx7, int B, double* nid_out, uint8_t* ok_out) const {
    pcp::check(pcp_nid_eval(nid_, params_Bx7, B, nid_out, ok_out), ctx_, "pcp_nid_eval");
  }
  // operator()<ceres::Jet<double, 7>>: value and the 7 partials over the ambient parameters (what AutoDiffFirstOrderFunction
  // hands to the solver before the manifold projection), visual_camera_calibration.cpp:204-233.
  bool operator()(const double* T_camera_lidar_params, double* residual, double* gradient7) const {
    uint8_t ok = 0;
    pcp::check(pcp_nid_eval_grad(nid_, T_camera_lidar_params, 1, residual, gradient7, &ok), ctx_, "pcp_nid_eval_grad");
    return ok != 0;
  }
  pcp_nid* handle() const { return nid_; }

 private:
  pcp_ctx* ctx_;
  pcp_nid* nid_ = nullptr;
};

// MultiNIDCost (src/vlcal/calib/visual_camera_calibration.cpp:86-129): pose gate (0.2 m / 2 deg from init) + sum over the frames.
// Evaluate() has the signature of ceres::FirstOrderFunction, so the reference's GradientProblem can wrap it directly.
class MultiNIDCost {
 public:
  explicit MultiNIDCost(const double init_T_camera_lidar_params[7]) { for (int i = 0; i < 7; i++) init_[i] = init_T_camera_lidar_params[i]; }
  void add(const std::shared_ptr<NIDCost>& cost) { costs_.push_back(cost); handles_.push_back(cost->handle()); }
  bool operator()(const double* params, double* residual) const {
    uint8_t ok = 0;
    if (pcp_nid_eval_multi(handles_.data(), (int)handles_.size(), init_, params, 1, residual, &ok) != PCP_OK) throw std::runtime_error("pcp_nid_eval_multi failed");
    return ok != 0;
  }
  bool Evaluate(const double* parameters, double* cost, double* gradient) const {
    if (!gradient) return (*this)(parameters, cost);
    uint8_t ok = 0;
    if (pcp_nid_eval_multi_grad(handles_.data(), (int)handles_.size(), init_, parameters, 1, cost, gradient, &ok) != PCP_OK)
      throw std::runtime_error("pcp_nid_eval_multi_grad failed");
    return ok != 0;
  }
  int NumParameters() const { return 7; }

 private:
  double init_[7];
  std::vector<std::shared_ptr<NIDCost>> costs_;
  std::vector<pcp_nid*> handles_;
};

// VisualCameraCalibrationParams / VisualCameraCalibration (include/vlcal/calib/visual_camera_calibration.hpp:14-70).  The keyframes'
// culled clouds and normalised images are handed over as NIDCost objects (estimate_pose_bfgs builds exactly these, :150-181).
struct VisualCameraCalibrationParams {
  int max_outer_iterations = 10;
  double delta_trans_thresh = 0.01;
  double delta_rot_thresh = 1.0 * 3.14159265358979323846 / 180.0;
  bool verbose = false;
};

class VisualCameraCalibration {
 public:
  VisualCameraCalibration(const std::vector<std::shared_ptr<NIDCost>>& keyframe_costs, const VisualCameraCalibrationParams& params = VisualCameraCalibrationParams())
      : costs_(keyframe_costs), params_(params) {}
  // Eigen::Isometry3d calibrate(const Eigen::Isometry3d& init_T_camera_lidar)  (visual_camera_calibration.cpp:49-84)
  pcp::Isometry3d calibrate(const pcp::Isometry3d& init_T_camera_lidar) {
    std::vector<pcp_nid*> h;
    for (const auto& c : costs_) h.push_back(c->handle());
    pcp_calib_params p;
    pcp_calib_default_params(&p);
    p.max_outer_iterations = params_.max_outer_iterations;
    p.delta_trans_thresh = params_.delta_trans_thresh;
    p.delta_rot_thresh = params_.delta_rot_thresh;
    p.inner.verbose = params_.verbose;
    pcp::Isometry3d out;
    if (pcp_calibrate(h.data(), (int)h.size(), init_T_camera_lidar.data(), &p, out.m, &summary_) != PCP_OK) throw std::runtime_error("pcp_calibrate failed");
    return out;
  }
  const pcp_calib_summary& summary() const { return summary_; }

 private:
  std::vector<std::shared_ptr<NIDCost>> costs_;
  VisualCameraCalibrationParams params_;
  pcp_calib_summary summary_{};
};

}  // namespace vlcal

// PointCloudProcessor(pointCloudPath, odometryPath, imagesFolder, maskImageFolder, outputPath, enableMLS, enableNIDOptimize,
// enableInitialGuessManual) + process() — include/PointCloudProcessor.hpp:17-27, src/PointCloudProcessor.cpp:24-90, :1007-1032.
// Same argument order and the same exception type / messages (std::runtime_error) as the reference class.
class PointCloudProcessor {
 public:
  PointCloudProcessor(const std::string& pointCloudPath, const std::string& odometryPath, const std::string& imagesFolder,
                      const std::string& maskImageFolder, const std::string& outputPath, const bool& enableMLS, const bool& enableNIDOptimize,
                      const bool& enableInitialGuessManual)
      : pointCloudPath_(pointCloudPath), odometryPath_(odometryPath), imagesFolder_(imagesFolder), maskImageFolder_(maskImageFolder),
        outputPath_(outputPath), enableMLS_(enableMLS), enableNIDOptimize_(enableNIDOptimize), enableInitialGuessManual_(enableInitialGuessManual) {}

  // Optional overrides of what the reference hard-codes (cpp:51-62, :525); defaults reproduce it.
  void setCamera(const pcp_camera& cam) { cam_ = cam; have_cam_ = true; }
  void setCullSize(int width, int height) { cull_w_ = width; cull_h_ = height; }
  void setOptimizedExtrinsic(const pcp::Isometry3d& T_camera_lidar_optimized) { T_ = T_camera_lidar_optimized; have_T_ = true; }
  void setDevice(int device) { device_ = device; }

  void process() {
    pcp_process_args a{};
    a.point_cloud_path = pointCloudPath_.c_str();
    a.odometry_path = odometryPath_.c_str();
    a.images_folder = imagesFolder_.c_str();
    a.mask_image_folder = maskImageFolder_.c_str();
    a.output_path = outputPath_.c_str();
    a.enable_mls = enableMLS_;
    a.enable_nid_optimize = enableNIDOptimize_;
    a.enable_initial_guess_manual = enableInitialGuessManual_;
    a.device = device_;
    a.camera = have_cam_ ? &cam_ : nullptr;
    a.cull_width = cull_w_;
    a.cull_height = cull_h_;
    a.T_camera_lidar_optimized = have_T_ ? T_.m : nullptr;
    char err[1024];
    const int rc = pcp_process(&a, &stats_, err, sizeof(err));
    if (rc != PCP_OK) throw std::runtime_error(err[0] ? err : "pcp_process failed");
  }
  const pcp_process_stats& stats() const { return stats_; }

 private:
  std::string pointCloudPath_, odometryPath_, imagesFolder_, maskImageFolder_, outputPath_;
  bool enableMLS_, enableNIDOptimize_, enableInitialGuessManual_;
  pcp_camera cam_{};
  bool have_cam_ = false, have_T_ = false;
  int cull_w_ = 0, cull_h_ = 0, device_ = 0;
  pcp::Isometry3d T_{};
  pcp_process_stats stats_{};
};

#endif  // PCP_PCP_HPP_
