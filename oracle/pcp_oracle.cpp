// pcp_oracle.cpp — CPU ORACLE. TEST INFRASTRUCTURE ONLY.
//
// A dependency-free C++17 restatement of the per-frame colorization hot path of
// ChunLI-666/PointCloudProcessor (transform -> project -> z-buffer cull -> sample -> accumulate,
// plus the NID joint-histogram cost).  Nothing in the product (pointcloudprocessor_b200/, include/)
// may link, import or call this file: only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs do, and there only as the checker / the timed CPU baseline.
//
// PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path and cannot be
// compiled here (needs PCL, OpenCV, Eigen, Ceres, Sophus, qhull, Boost, ROS tf, Iridescence — none
// installed, no network).  Every function below cites the reference file:line it follows (paths relative
// to /root/reference/PointCloudProcessor/).  Where the arithmetic lives in a third-party library that is
// not vendored, the choice made here is listed in the DECISION LOG and in DESIGN.md.
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fPIC -shared [-fopenmp]  (no -ffast-math, no -march=native:
// strict IEEE binary32/binary64, SSE2 scalar code, no FMA contraction — the same arithmetic the
// reference's Release build performs on x86-64).
//
// DECISION LOG (third-party behaviour restated, not verifiable on this box)
//  D1 pcl::transformPointCloud (PCL 1.10, SSE path of pcl::detail::Transformer<float>::se3):
//     out = x*c0 + (y*c1 + (z*c2 + c3)), every product and sum rounded to binary32, no FMA.
//  D2 Eigen::Quaterniond::toRotationMatrix(): the tx/ty/tz .. tzz formulation, quaternion NOT normalised.
//  D3 Eigen Isometry3d::inverse(): linear = R^T, translation = -(R^T) * t, 3-term sums left to right.
//  D4 Eigen Affine3f::inverse() (general 3x3 float inverse): cofactors, det = sum(cof_col0 * m_col0)
//     left to right, invdet = 1/det, result(r,c) = cofactor(c,r) * invdet; translation = -(Linv * t).
//  D5 Eigen Vector3d::norm(): sqrt((x*x + y*y) + z*z).
//  D6 Eigen Isometry3d * Vector4d / Isometry3d * Isometry3d: row sums left to right.
//  D7 x86 cvttss2si/cvttsd2si give INT_MIN for NaN/out-of-range; CUDA saturates and maps NaN->0.  The
//     oracle (and the kernels) reject any point whose projected coordinates are not finite; all other
//     out-of-range values are rejected identically by both conventions.
//  D8 std::sort (unstable) on finalScore in smoothColors: ties broken by ascending keyframe index.
//  D9 static_cast<uint8_t>(NaN) for points never observed: defined as 0 (what x86 produces in practice).
//  D10 Sophus SE3d * point: p + w*uv + qv x uv with uv = 2*(qv x p), then + translation.
//  D12 cv::cvtColor BGR<->HSV 8-bit: restated from OpenCV 4.x and pinned against cv2 4.13.0 over all 2^24 inputs; the
//      backward conversion differs between OpenCV's vector body and its scalar row tail (see hsv_roundtrip).
//  D11 kd-tree re-indexing (PointCloudProcessor.cpp:557-593) is replaced by carrying the point index
//      ("reindex=carry", the default and what the CUDA path does); scores are computed from the camera-frame float point
//      directly.  "reindex=radius" (orc_colorizer_set_reindex, small clouds only: brute force) restates the reference
//      literally: float c2w' of the coloured camera-frame point, every map point within 1e-5 m of it receives the
//      observation (none if the float round trip drifted further; coincident points all do), scores from the point taken
//      back through Affine3f::inverse() [D4].  tests/test_oracle_cpu.py and tools/reindex_divergence.py measure the difference.
//  D13 view_culling.cpp:88 `pt_2d_f / downsample_factor` is Eigen::Vector2f / int: Eigen promotes the int to float and
//      evaluates scalar_quotient_op, a TRUE binary32 division per component (Eigen 3.2 and 3.3 alike; ROS noetic ships
//      3.3.7) — not a multiplication by the reciprocal, which would move cell borders by one ulp for some u.
//  D14 nid_cost.hpp:95-103 Eigen .sum(): `hist_points` holds integers, so `sum` is exact whatever the order.  The three
//      entropy sums reduce `x * log(x + 1e-6)`: Eigen 3.3.7 (ROS noetic) has no packet log for double (packet_traits<double>::
//      HasLog == 0 on SSE), and ceres::Jet is never vectorised, so for both T = double and T = Jet the reduction takes
//      DefaultTraversal: std::log per coefficient, summed one coefficient at a time in storage order — i = 0..bins-1 for the two
//      marginals, and COLUMN-major for `hist` (rows = image bin, columns = point bin): the oracle walks its row-major copy in
//      that order.  (Eigen >= 3.4 vectorises log for doubles and would differ in the last bits; tolerance 1e-5 covers it.)
//  D15 pcl::KdTreeFLANN::radiusSearch (cpp:571): squared distances accumulated in binary32 in x,y,z order (FLANN L2_Simple),
//      compared against float(radius * radius) with radius = double(1e-5f); a neighbour is kept when dist2 < radius2... the
//      boundary case cannot be pinned without FLANN; the oracle uses <=, which differs only for a point EXACTLY at 1e-5 m.

#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct Camera {
  double fx, fy, cx, cy;
  double D[5];  // k1, k2, p1, p2, k3 — plumb_bob order (include/camera/pinhole.hpp:14-19)
};

// ---------------------------------------------------------------------------------------------------
// Pose -> matrices.  src/PointCloudProcessor.cpp:495-519 (and :186-194).  All matrices column-major 4x4.
// ---------------------------------------------------------------------------------------------------
struct Mat4d { double m[16]; double& operator()(int r, int c) { return m[c * 4 + r]; } double operator()(int r, int c) const { return m[c * 4 + r]; } };
struct Mat4f { float m[16]; float& operator()(int r, int c) { return m[c * 4 + r]; } float operator()(int r, int c) const { return m[c * 4 + r]; } };

// Eigen::Quaterniond(w,x,y,z).toRotationMatrix()  [D2]
void quat_to_rot(double w, double x, double y, double z, double R[9] /*row-major*/) {
  const double tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1.0 - (tyy + tzz); R[1] = txy - twz;         R[2] = txz + twy;
  R[3] = txy + twz;         R[4] = 1.0 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;         R[7] = tyz + twx;         R[8] = 1.0 - (txx + tyy);
}

// t_c2w = Identity; translate(x,y,z); rotate(q)   (cpp:496-498) -> [R | t] exactly.
Mat4d pose_to_c2w(const double p[7]) {
  double R[9];
  quat_to_rot(p[3], p[4], p[5], p[6], R);  // Pose = {x,y,z,qw,qx,qy,qz} (include/FrameData.hpp:9-12)
  Mat4d T{};
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) T(r, c) = R[r * 3 + c];
  T(0, 3) = p[0]; T(1, 3) = p[1]; T(2, 3) = p[2];
  T(3, 0) = T(3, 1) = T(3, 2) = 0.0; T(3, 3) = 1.0;
  return T;
}

// Isometry3d::inverse()  [D3]
Mat4d isometry_inverse(const Mat4d& T) {
  Mat4d I{};
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) I(r, c) = T(c, r);
  for (int r = 0; r < 3; r++) {
    const double a0 = -I(r, 0), a1 = -I(r, 1), a2 = -I(r, 2);
    I(r, 3) = (a0 * T(0, 3) + a1 * T(1, 3)) + a2 * T(2, 3);
  }
  I(3, 0) = I(3, 1) = I(3, 2) = 0.0; I(3, 3) = 1.0;
  return I;
}

// Isometry3d * Isometry3d  [D6]   (cpp:507, :516)
Mat4d isometry_mul(const Mat4d& A, const Mat4d& B) {
  Mat4d C{};
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) C(r, c) = (A(r, 0) * B(0, c) + A(r, 1) * B(1, c)) + A(r, 2) * B(2, c);
    C(r, 3) = ((A(r, 0) * B(0, 3) + A(r, 1) * B(1, 3)) + A(r, 2) * B(2, 3)) + A(r, 3);
  }
  C(3, 0) = C(3, 1) = C(3, 2) = 0.0; C(3, 3) = 1.0;
  return C;
}

Mat4f cast_f(const Mat4d& T) { Mat4f F{}; for (int i = 0; i < 16; i++) F.m[i] = (float)T.m[i]; return F; }

// Affine3f::inverse()  [D4]   (cpp:509, :518)
Mat4f affine_inverse_f(const Mat4f& A) {
  auto cof = [&](int i, int j) -> float {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return A(i1, j1) * A(i2, j2) - A(i1, j2) * A(i2, j1);
  };
  const float c00 = cof(0, 0), c10 = cof(1, 0), c20 = cof(2, 0);
  const float det = (c00 * A(0, 0) + c10 * A(1, 0)) + c20 * A(2, 0);
  const float invdet = 1.0f / det;
  Mat4f I{};
  I(0, 0) = c00 * invdet; I(0, 1) = c10 * invdet; I(0, 2) = c20 * invdet;
  I(1, 0) = cof(0, 1) * invdet; I(1, 1) = cof(1, 1) * invdet; I(1, 2) = cof(2, 1) * invdet;
  I(2, 0) = cof(0, 2) * invdet; I(2, 1) = cof(1, 2) * invdet; I(2, 2) = cof(2, 2) * invdet;
  for (int r = 0; r < 3; r++) {
    const float a0 = -I(r, 0), a1 = -I(r, 1), a2 = -I(r, 2);
    I(r, 3) = (a0 * A(0, 3) + a1 * A(1, 3)) + a2 * A(2, 3);
  }
  I(3, 0) = I(3, 1) = I(3, 2) = 0.0f; I(3, 3) = 1.0f;
  return I;
}

void pose_matrices(const double pose[7], const double* T_opt /*col-major 4x4 or null*/, Mat4f& w2c, Mat4f& c2w) {
  const Mat4d t_c2w = pose_to_c2w(pose);
  if (!T_opt) {
    w2c = cast_f(isometry_inverse(t_c2w));  // cpp:499-500
    c2w = cast_f(t_c2w);                    // cpp:501-502
  } else {
    Mat4d Topt; std::memcpy(Topt.m, T_opt, sizeof(Topt.m));
    c2w = cast_f(isometry_mul(t_c2w, Topt));  // cpp:507-508
    w2c = affine_inverse_f(c2w);              // cpp:509
  }
}

// pcl::transformPointCloud per point  [D1]   (cpp:196, :521)
inline void transform_point(const Mat4f& M, float x, float y, float z, float& ox, float& oy, float& oz) {
  ox = x * M(0, 0) + (y * M(0, 1) + (z * M(0, 2) + M(0, 3)));
  oy = x * M(1, 0) + (y * M(1, 1) + (z * M(1, 2) + M(1, 3)));
  oz = x * M(2, 0) + (y * M(2, 1) + (z * M(2, 2) + M(2, 3)));
}

// ---------------------------------------------------------------------------------------------------
// Pinhole + plumb_bob.  include/camera/pinhole.hpp:12-51 == include/PointCloudProcessor.hpp:100-123 +
// src/PointCloudProcessor.cpp:750-753.  Operation order kept literally.
// ---------------------------------------------------------------------------------------------------
inline void project(const Camera& cam, double X, double Y, double Z, double& u, double& v) {
  const double x = X / Z, y = Y / Z;
  const double k1 = cam.D[0], k2 = cam.D[1], p1 = cam.D[2], p2 = cam.D[3], k3 = cam.D[4];
  const double x2 = x * x, y2 = y * y;
  const double r2 = x2 + y2, r4 = r2 * r2, r6 = r2 * r4;
  const double r_coeff = 1.0 + k1 * r2 + k2 * r4 + k3 * r6;
  const double t1 = 2.0 * x * y, t2 = r2 + 2.0 * x2, t3 = r2 + 2.0 * y2;
  const double xd = r_coeff * x + p1 * t1 + p2 * t2;
  const double yd = r_coeff * y + p1 * t3 + p2 * t1;
  u = cam.fx * xd + cam.cx;
  v = cam.fy * yd + cam.cy;
}

// ---------------------------------------------------------------------------------------------------
// z-buffer culling.  src/vlcal/calib/view_culling.cpp:52-174 (+ cull() :23-50, sample() :176-185).
// Literal two sequential passes.  Returns the kept indices in ascending order.
// ---------------------------------------------------------------------------------------------------
struct CullGate { bool pass; int cx, cy; double dist; };

inline CullGate cull_gate(const Camera& cam, int W, int H, int ds, double X, double Y, double Z) {
  CullGate g{false, 0, 0, 0.0};
  // view_culling.cpp:76  pt.normalized().z() < 0  (pt = (x,y,z,1); the sign test equals z < 0)
  if (Z < 0.0) return g;
  double u, v;
  project(cam, X, Y, Z, u, v);                 // :86
  const float uf = (float)u, vf = (float)v;     // .cast<float>()
  if (!std::isfinite(uf) || !std::isfinite(vf)) return g;  // [D7]
  const float qx = uf / (float)ds, qy = vf / (float)ds;     // :88  Vector2f / int
  if (!(qx > -2147483648.0f && qx < 2147483648.0f && qy > -2147483648.0f && qy < 2147483648.0f)) return g;  // [D7]
  const int px = (int)qx, py = (int)qy;         // .cast<int>() truncation
  if (px < 0 || py < 0 || px >= W || py >= H) return g;     // :90 (full-resolution image_size)
  g.pass = true; g.cx = px; g.cy = py;
  g.dist = std::sqrt((X * X + Y * Y) + Z * Z);  // :102  [D5]
  return g;
}

int view_culling(const Camera& cam, int W, int H, int ds, double tol, bool depth_culling, const float* pts_cam /*n x 3*/,
                 size_t n, const double* T /*col-major 4x4 or null*/, int* out_idx, float* dist_map, int* index_map,
                 int* n_candidates) {
  const int Wc = W / ds, Hc = H / ds;  // :64-65 integer division
  std::vector<float> dm_local; std::vector<int> im_local;
  if (!dist_map) { dm_local.resize((size_t)Wc * Hc); dist_map = dm_local.data(); }
  if (!index_map) { im_local.resize((size_t)Wc * Hc); index_map = im_local.data(); }
  for (size_t i = 0; i < (size_t)Wc * Hc; i++) { dist_map[i] = FLT_MAX; index_map[i] = -1; }

  std::vector<int> indices; std::vector<int> cells; std::vector<double> dists;
  indices.reserve(n / 4); cells.reserve(n / 4); dists.reserve(n / 4);
  for (size_t i = 0; i < n; i++) {
    double X = pts_cam[3 * i], Y = pts_cam[3 * i + 1], Z = pts_cam[3 * i + 2];
    if (T) {  // cull(): T_camera_lidar * (x,y,z,1)   view_culling.cpp:33-37  [D6]
      const double x = X, y = Y, z = Z;
      X = ((T[0] * x + T[4] * y) + T[8] * z) + T[12];
      Y = ((T[1] * x + T[5] * y) + T[9] * z) + T[13];
      Z = ((T[2] * x + T[6] * y) + T[10] * z) + T[14];
    }
    const CullGate g = cull_gate(cam, W, H, ds, X, Y, Z);
    if (!g.pass) continue;
    indices.push_back((int)i);                       // :96  recorded BEFORE the depth test
    const bool in_grid = g.cx < Wc && g.cy < Hc;
    cells.push_back(in_grid ? g.cy * Wc + g.cx : -1);  // :97
    dists.push_back(g.dist);
    if (depth_culling && in_grid) {                  // :100-131
      float& d = dist_map[g.cy * Wc + g.cx];
      if (g.dist > (double)d) continue;              // :118
      d = (float)g.dist;                             // :123
      index_map[g.cy * Wc + g.cx] = (int)i;          // :124
    }
  }
  if (n_candidates) *n_candidates = (int)indices.size();
  int m = 0;
  if (depth_culling) {                               // :135-171
    for (size_t k = 0; k < indices.size(); k++) {
      if (cells[k] < 0) continue;                    // :162-165
      if (dists[k] > (double)dist_map[cells[k]] + tol) continue;  // :157
      out_idx[m++] = indices[k];
    }
  } else {
    for (size_t k = 0; k < indices.size(); k++) out_idx[m++] = indices[k];
  }
  return m;
}

// ---------------------------------------------------------------------------------------------------
// Colour / mask gather for one camera-frame point.  src/PointCloudProcessor.cpp:743-766, :783-815.
// Returns pixel index v*W+u or -1.
// ---------------------------------------------------------------------------------------------------
inline int color_pixel(const Camera& cam, int W, int H, float xf, float yf, float zf) {
  const double X = xf, Y = yf, Z = zf;
  if (!(Z > 0)) return -1;                         // cpp:748
  double ud, vd;
  project(cam, X, Y, Z, ud, vd);                   // cpp:750-753
  if (!std::isfinite(ud) || !std::isfinite(vd)) return -1;            // [D7]
  if (!(ud > -2147483648.0 && ud < 2147483648.0 && vd > -2147483648.0 && vd < 2147483648.0)) return -1;  // [D7]
  const int u = (int)ud, v = (int)vd;              // static_cast<int> truncation
  if (u >= 0 && u < W && v >= 0 && v < H) return v * W + u;           // cpp:754
  return -1;
}

// BILINEAR sampling (north_star (c); an EXTENSION: the reference samples the nearest pixel, cpp:752-762).  Restated here as
// the definition the CUDA path must reproduce bit for bit.  The point is coloured iff the nearest-pixel gate passes (same
// coloured set as the reference); (ud, vd) is the double projection; x0 = int(ud), y0 = int(vd) (truncation == floor since
// both are > -1 inside the gate; a fraction below 0 is clamped to 0); x1 = min(x0 + 1, W - 1), y1 = min(y0 + 1, H - 1); fx = ud - x0, fy = vd - y0;
//   val = (p00 * (1 - fx) + p10 * fx) * (1 - fy) + (p01 * (1 - fx) + p11 * fx) * fy      (double, every operation rounded)
// accumulated as 8.8 fixed point q = (uint32)(val * 256 + 0.5) in the same u32 sums as the nearest mode; the finalize step
// rounds the mean: min(255, (sum / count + 128) >> 8).  Error against real-valued bilinear interpolation: < 1/256 of a level
// per observation plus the final rounding, i.e. well inside the 1/255 per channel of north_star.
inline uint32_t bilinear_q(const uint8_t* img, size_t pitch, int W, int H, int ch, double ud, double vd) {
  const int x0 = (int)ud, y0 = (int)vd;
  const int x1 = x0 + 1 < W ? x0 + 1 : W - 1, y1 = y0 + 1 < H ? y0 + 1 : H - 1;
  double fx = ud - (double)x0, fy = vd - (double)y0;
  if (fx < 0.0) fx = 0.0;   // ud in (-1, 0) truncates to column 0 (the reference's gate accepts it): no extrapolation
  if (fy < 0.0) fy = 0.0;
  const double gx = 1.0 - fx, gy = 1.0 - fy;
  const double p00 = img[(size_t)y0 * pitch + (size_t)x0 * 3 + ch], p10 = img[(size_t)y0 * pitch + (size_t)x1 * 3 + ch];
  const double p01 = img[(size_t)y1 * pitch + (size_t)x0 * 3 + ch], p11 = img[(size_t)y1 * pitch + (size_t)x1 * 3 + ch];
  const double top = p00 * gx + p10 * fx;
  const double bot = p01 * gx + p11 * fx;
  const double val = top * gy + bot * fy;
  return (uint32_t)(val * 256.0 + 0.5);
}

// Scores.  include/PointCloudProcessor.hpp:205-236, src/PointCloudProcessor.cpp:586-588.  [D11]
inline float orientation_score(float px, float py, float pz, const double pose[7]) {
  const double vx = (double)px - pose[0], vy = (double)py - pose[1], vz = (double)pz - pose[2];
  const double n2 = (vx * vx + vy * vy) + vz * vz;
  double cz = vz;
  if (n2 > 0.0) cz = vz / std::sqrt(n2);           // Eigen normalized()
  float s = (float)((cz + 1.0) / 2.0);
  s = 0.2f + 0.8f * s;
  return s;
}
inline float distance_score(float px, float py, float pz) {
  const float distance = std::sqrt(px * px + py * py + pz * pz);
  const float optimal = 2.0f;
  const float diff = std::fabs(distance - optimal);
  const float nd = std::min(diff / optimal, 1.0f);
  float s = 1.0f - nd;
  s = 0.2f + 0.8f * s;
  return s;
}

struct Obs { float r, g, b, score; int frame; };

// ---------------------------------------------------------------------------------------------------
// Colorizer state: the loop body of pcdColorizationAndSmooth (cpp:488-595) with two accumulator modes.
// ---------------------------------------------------------------------------------------------------
struct Colorizer {
  Camera cam; int W, H, cullW, cullH, ds; double tol;
  size_t n = 0;
  std::vector<float> xyzi;             // n x 4
  std::vector<uint32_t> acc;           // n x 4 : sum r, sum g, sum b, count      ("sum" mode)
  std::vector<uint32_t> mask_hits;     // n : #frames whose mask pixel == 255 (cpp:806)
  std::vector<uint32_t> mask_max;      // n : max mask value seen (cpp:804)
  bool keep_obs = false;
  int sample_mode = 0;                 // 0 nearest (reference), 1 bilinear (extension, see bilinear_q)
  std::vector<std::vector<Obs>> obs;   // per-point observation lists (include/RGBCloud.hpp:29-41)
  int reindex = 0;                     // 0 carry the point index [D11]; 1 radius: the reference's kd-tree re-indexing, brute force
  int frame_counter = 0;
};

struct FrameResult { int n_candidates, n_visible, n_colored; };

// One keyframe.  vis: optional n bytes (1 = kept by culling); pix: optional n int32 (pixel index or -1).
FrameResult colorize_frame(Colorizer& S, const double pose[7], const double* T_opt, const uint8_t* bgr, size_t pitch,
                           const uint8_t* mask, size_t mask_pitch, uint8_t* vis, int32_t* pix, int threads) {
  Mat4f w2c, c2w;
  pose_matrices(pose, T_opt, w2c, c2w);
  const int Wc = S.cullW / S.ds, Hc = S.cullH / S.ds;
  const size_t n = S.n;
  FrameResult R{0, 0, 0};
  if (vis) std::memset(vis, 0, n);
  if (pix) for (size_t i = 0; i < n; i++) pix[i] = -1;

  std::vector<float> zmin((size_t)Wc * Hc, FLT_MAX);
  // candidate records (kept in ascending point order inside each thread chunk)
  struct Cand { int idx; int cell; double dist; float x, y, z; };
  int T = 1;
#ifdef _OPENMP
  T = threads > 0 ? threads : 1;
#endif
  std::vector<std::vector<Cand>> cands(T);
  std::vector<std::vector<float>> zl(T);
#pragma omp parallel num_threads(T)
  {
    int t0 = 0, tstep = 1;
#ifdef _OPENMP
    t0 = omp_get_thread_num(); tstep = omp_get_num_threads();
#endif
    for (int t = t0; t < T; t += tstep) {
    std::vector<float>& z = zl[t];
    z.assign((size_t)Wc * Hc, FLT_MAX);
    std::vector<Cand>& C = cands[t];
    C.reserve(n / (4 * T) + 16);
    const size_t lo = n * t / T, hi = n * (t + 1) / T;
    for (size_t i = lo; i < hi; i++) {
      float x, y, zc;
      transform_point(w2c, S.xyzi[4 * i], S.xyzi[4 * i + 1], S.xyzi[4 * i + 2], x, y, zc);  // cpp:521
      const CullGate g = cull_gate(S.cam, S.cullW, S.cullH, S.ds, x, y, zc);                 // cpp:527 (T = Identity)
      if (!g.pass) continue;
      const bool in_grid = g.cx < Wc && g.cy < Hc;
      C.push_back({(int)i, in_grid ? g.cy * Wc + g.cx : -1, g.dist, x, y, zc});
      if (in_grid) {
        float& d = z[g.cy * Wc + g.cx];
        if (!(g.dist > (double)d)) d = (float)g.dist;  // view_culling.cpp:118-123
      }
    }
    }
  }
  // merge per-thread grids: the final map is min over fl32(dist) whatever the order
  for (int t = 0; t < T; t++) for (size_t c = 0; c < zmin.size(); c++) zmin[c] = std::min(zmin[c], zl[t][c]);

  struct Colored { float x, y, z; uint8_t r, g, b; };
  std::vector<std::vector<Colored>> colored(T);
  int n_cand = 0, n_vis = 0, n_col = 0;
#pragma omp parallel for num_threads(T) reduction(+ : n_cand, n_vis, n_col) schedule(static, 1)
  for (int t = 0; t < T; t++) {
    for (const Cand& c : cands[t]) {
      n_cand++;
      if (c.cell < 0) continue;
      if (c.dist > (double)zmin[c.cell] + S.tol) continue;   // view_culling.cpp:157
      n_vis++;
      if (vis) vis[c.idx] = 1;
      const int p = color_pixel(S.cam, S.W, S.H, c.x, c.y, c.z);  // generateColorMap cpp:743-766
      if (p < 0) continue;
      n_col++;
      if (pix) pix[c.idx] = p;
      const int v = p / S.W, u = p % S.W;
      const uint8_t* px = bgr + (size_t)v * pitch + (size_t)u * 3;
      uint8_t b = px[0], g = px[1], r = px[2];                // cpp:760-762
      uint32_t* a = &S.acc[4 * (size_t)c.idx];
      if (S.sample_mode == 1) {
        double ud, vd;
        project(S.cam, (double)c.x, (double)c.y, (double)c.z, ud, vd);   // the same doubles color_pixel truncated
        const uint32_t qr = bilinear_q(bgr, pitch, S.W, S.H, 2, ud, vd), qg = bilinear_q(bgr, pitch, S.W, S.H, 1, ud, vd),
                       qb = bilinear_q(bgr, pitch, S.W, S.H, 0, ud, vd);
        a[0] += qr; a[1] += qg; a[2] += qb; a[3] += 1;
        r = (uint8_t)((qr + 128) >> 8 > 255 ? 255 : (qr + 128) >> 8);   // 8-bit colour of the observation (TOP5 lists)
        g = (uint8_t)((qg + 128) >> 8 > 255 ? 255 : (qg + 128) >> 8);
        b = (uint8_t)((qb + 128) >> 8 > 255 ? 255 : (qb + 128) >> 8);
      } else {
        a[0] += r; a[1] += g; a[2] += b; a[3] += 1;
      }
      if (mask) {                                             // generateSegmentMap cpp:783-815
        const uint8_t mv = mask[(size_t)v * mask_pitch + u];
        if (mv == 255) S.mask_hits[c.idx] += 1;
        S.mask_max[c.idx] = std::max<uint32_t>(S.mask_max[c.idx], mv);
      }
      if (S.keep_obs && S.reindex == 1) {
        colored[t].push_back({c.x, c.y, c.z, r, g, b});      // re-indexed below, serially (observations land on OTHER points' lists)
      } else if (S.keep_obs) {                                // cpp:586-591
        const float o = orientation_score(c.x, c.y, c.z, pose);
        const float d = distance_score(c.x, c.y, c.z);
        const float f = (float)((double)(o + d) / 2.0);
        S.obs[c.idx].push_back({(float)r, (float)g, (float)b, f, S.frame_counter});
      }
    }
  }
  if (S.keep_obs && S.reindex == 1) {
    // src/PointCloudProcessor.cpp:554-594, literally: coloredCloud -> world with c2w' in float (:555), radiusSearch(eps = 1e-5f)
    // over the map cloud (:571) [D15], and for EVERY index found the scores of the point taken back through
    // transformation_c2w_optimized.inverse() (:579, Affine3f::inverse() [D4], Matrix4f * Vector4f as column sums left to right).
    const Mat4f inv = affine_inverse_f(c2w);
    const float r2 = (float)((double)1e-5f * (double)1e-5f);
    for (int t = 0; t < T; t++)
      for (const Colored& c : colored[t]) {
        float wx, wy, wz;
        transform_point(c2w, c.x, c.y, c.z, wx, wy, wz);                                  // :555 [D1]
        const float cx = ((inv(0, 0) * wx + inv(0, 1) * wy) + inv(0, 2) * wz) + inv(0, 3) * 1.0f;   // :579
        const float cy = ((inv(1, 0) * wx + inv(1, 1) * wy) + inv(1, 2) * wz) + inv(1, 3) * 1.0f;
        const float cz = ((inv(2, 0) * wx + inv(2, 1) * wy) + inv(2, 2) * wz) + inv(2, 3) * 1.0f;
        const float o = orientation_score(cx, cy, cz, pose);                              // :586
        const float d = distance_score(cx, cy, cz);                                       // :587
        const float f = (float)((double)(o + d) / 2.0);                                   // :588
        for (size_t j = 0; j < n; j++) {                                                  // :571, brute force
          const float dx = S.xyzi[4 * j] - wx, dy = S.xyzi[4 * j + 1] - wy, dz = S.xyzi[4 * j + 2] - wz;
          float d2 = dx * dx; d2 += dy * dy; d2 += dz * dz;
          if (d2 <= r2) S.obs[j].push_back({(float)c.r, (float)c.g, (float)c.b, f, S.frame_counter});   // :590-591
        }
      }
  }
  R.n_candidates = n_cand; R.n_visible = n_vis; R.n_colored = n_col;
  S.frame_counter++;
  return R;
}

// ---------------------------------------------------------------------------------------------------
// NID.  include/vlcal/costs/nid_cost.hpp:42-116.   se3 = {qx,qy,qz,qw,tx,ty,tz} (Sophus order).
// image: single-channel double plane (pixel/255), W x H, row-major.
// ---------------------------------------------------------------------------------------------------
bool nid_eval(const Camera& cam, const double* image, int W, int H, const float* xyzi, size_t n, int bins,
              const double se3[7], double* nid_out, double* hist_out /*bins*bins or null*/, double* counts_out) {
  const double qx = se3[0], qy = se3[1], qz = se3[2], qw = se3[3], tx = se3[4], ty = se3[5], tz = se3[6];
  std::vector<double> hist((size_t)bins * bins, 0.0), hist_image(bins, 0.0), hist_points(bins, 0.0);
  static const double C[4][4] = {{1.0, -3.0, 3.0, -1.0}, {4.0, 0.0, -6.0, 3.0}, {1.0, 3.0, 3.0, -3.0}, {0.0, 0.0, 0.0, 1.0}};
  double Cn[4][4];
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) Cn[i][j] = C[i][j] / 6.0;   // :35-39
  for (size_t i = 0; i < n; i++) {
    const double px = xyzi[4 * i], py = xyzi[4 * i + 1], pz = xyzi[4 * i + 2];
    // Sophus SE3 * p  [D10]   (:55)
    double uvx = qy * pz - qz * py, uvy = qz * px - qx * pz, uvz = qx * py - qy * px;
    uvx += uvx; uvy += uvy; uvz += uvz;
    const double X = (px + qw * uvx + (qy * uvz - qz * uvy)) + tx;
    const double Y = (py + qw * uvy + (qz * uvx - qx * uvz)) + ty;
    const double Z = (pz + qw * uvz + (qx * uvy - qy * uvx)) + tz;
    const double intensity = xyzi[4 * i + 3];                                          // :57
    const double ib = intensity * bins;
    int bin_p;                                                                         // :58
    if (!(ib == ib)) bin_p = 0; else if (ib >= (double)(bins - 1)) bin_p = bins - 1; else if (ib <= 0.0) bin_p = 0; else bin_p = (int)ib;
    double u, v;
    project(cam, X, Y, Z, u, v);                                                       // :60  (no z>0 gate)
    if (!std::isfinite(u) || !std::isfinite(v)) continue;                              // [D7] counted as outlier
    const double fu = std::floor(u), fv = std::floor(v);                               // :61
    if (fu < 0 || fv < 0 || fu >= (double)W || fv >= (double)H) continue;              // :64-67
    const int kx = (int)fu, ky = (int)fv;
    const double sx = u - (double)kx, sy = v - (double)ky;                             // :62
    hist_points[bin_p] += 1.0;                                                         // :69
    const double sex[4] = {1.0, sx, sx * sx, (sx * sx) * sx};                          // :71-75
    const double sey[4] = {1.0, sy, sy * sy, (sy * sy) * sy};
    double bx[4], by[4];
    for (int r = 0; r < 4; r++) {                                                      // :77
      bx[r] = ((Cn[r][0] * sex[0] + Cn[r][1] * sex[1]) + Cn[r][2] * sex[2]) + Cn[r][3] * sex[3];
      by[r] = ((Cn[r][0] * sey[0] + Cn[r][1] * sey[1]) + Cn[r][2] * sey[2]) + Cn[r][3] * sey[3];
    }
    for (int a = 0; a < 4; a++) {
      const int x = std::min(std::max(kx - 1 + a, 0), W - 1);                          // :79-82
      for (int b = 0; b < 4; b++) {
        const int y = std::min(std::max(ky - 1 + b, 0), H - 1);
        const double w = bx[a] * by[b];                                                // :86
        const double pixv = image[(size_t)y * W + x];                                  // :87
        const int bin_i = std::min<int>((int)(pixv * bins), bins - 1);                 // :88
        hist[(size_t)bin_i * bins + bin_p] += w;                                       // :89
        hist_image[bin_i] += w;                                                        // :90
      }
    }
  }
  double sum = 0.0;
  for (int i = 0; i < bins; i++) sum += hist_points[i];                                // :95
  if (counts_out) for (int i = 0; i < bins; i++) counts_out[i] = hist_points[i];
  if (hist_out) std::memcpy(hist_out, hist.data(), sizeof(double) * bins * bins);
  double Hi = 0.0, Hp = 0.0, Hj = 0.0;
  for (int i = 0; i < bins; i++) { const double p = hist_image[i] / sum; Hi += p * std::log(p + 1e-6); }
  for (int i = 0; i < bins; i++) { const double p = hist_points[i] / sum; Hp += p * std::log(p + 1e-6); }
  for (int c = 0; c < bins; c++)          // Eigen's column-major walk of hist(row = image bin, column = point bin)  [D14]
    for (int r = 0; r < bins; r++) { const double p = hist[(size_t)r * bins + c] / sum; Hj += p * std::log(p + 1e-6); }
  Hi = -Hi; Hp = -Hp; Hj = -Hj;                                                        // :101-103
  const double MI = Hi + Hp - Hj;                                                      // :104
  const double NID = (Hj - MI) / Hj;                                                   // :105
  *nid_out = NID;
  return std::isfinite(NID);                                                           // :107-111
}

// ---------------------------------------------------------------------------------------------------
// NID value + gradient: NIDCost::operator()<ceres::Jet<double, 7>> (nid_cost.hpp:42-116) as instantiated by
// ceres::AutoDiffFirstOrderFunction<MultiNIDCost, 7> (visual_camera_calibration.cpp:204-233): forward-mode dual numbers over
// the 7 ambient Sophus parameters {qx,qy,qz,qw,tx,ty,tz}.  Jet7 restates ceres::Jet's arithmetic (ceres/jet.h: +, -, *, /, log;
// third party, not vendored): f(a + v eps) = f(a) + f'(a) v eps.  floor()/knots/bins are taken on the real part (get_real, :61),
// hist_points and H_points are plain doubles (:50, :102).  The statement order is that of nid_eval above.
// ---------------------------------------------------------------------------------------------------
struct Jet7 {
  double a;
  double v[7];
  Jet7() : a(0.0) { for (double& x : v) x = 0.0; }
  Jet7(double s) : a(s) { for (double& x : v) x = 0.0; }  // NOLINT: implicit like ceres::Jet(const T&)
  static Jet7 var(double s, int k) { Jet7 j(s); j.v[k] = 1.0; return j; }
};
inline Jet7 operator+(const Jet7& f, const Jet7& g) { Jet7 h; h.a = f.a + g.a; for (int k = 0; k < 7; k++) h.v[k] = f.v[k] + g.v[k]; return h; }
inline Jet7 operator-(const Jet7& f, const Jet7& g) { Jet7 h; h.a = f.a - g.a; for (int k = 0; k < 7; k++) h.v[k] = f.v[k] - g.v[k]; return h; }
inline Jet7 operator-(const Jet7& f) { Jet7 h; h.a = -f.a; for (int k = 0; k < 7; k++) h.v[k] = -f.v[k]; return h; }
inline Jet7 operator*(const Jet7& f, const Jet7& g) { Jet7 h; h.a = f.a * g.a; for (int k = 0; k < 7; k++) h.v[k] = f.a * g.v[k] + f.v[k] * g.a; return h; }
inline Jet7 operator/(const Jet7& f, const Jet7& g) {  // ceres: g_a_inverse = 1 / g.a; f_a_by_g_a = f.a * g_a_inverse; (f.v - f_a_by_g_a * g.v) * g_a_inverse
  Jet7 h; const double gi = 1.0 / g.a; const double fg = f.a * gi; h.a = fg; for (int k = 0; k < 7; k++) h.v[k] = (f.v[k] - fg * g.v[k]) * gi; return h;
}
inline Jet7 jlog(const Jet7& f) { Jet7 h; h.a = std::log(f.a); const double i = 1.0 / f.a; for (int k = 0; k < 7; k++) h.v[k] = f.v[k] * i; return h; }
inline Jet7& operator+=(Jet7& f, const Jet7& g) { f = f + g; return f; }

inline void project_jet(const Camera& cam, const Jet7& X, const Jet7& Y, const Jet7& Z, Jet7& u, Jet7& v) {  // pinhole.hpp:12-51 with T = Jet
  const Jet7 x = X / Z, y = Y / Z;
  const Jet7 x2 = x * x, y2 = y * y;
  const Jet7 r2 = x2 + y2, r4 = r2 * r2, r6 = r2 * r4;
  const Jet7 rc = ((Jet7(1.0) + Jet7(cam.D[0]) * r2) + Jet7(cam.D[1]) * r4) + Jet7(cam.D[4]) * r6;
  const Jet7 t1 = (Jet7(2.0) * x) * y;
  const Jet7 t2 = r2 + Jet7(2.0) * x2, t3 = r2 + Jet7(2.0) * y2;
  const Jet7 xd = (rc * x + Jet7(cam.D[2]) * t1) + Jet7(cam.D[3]) * t2;
  const Jet7 yd = (rc * y + Jet7(cam.D[2]) * t3) + Jet7(cam.D[3]) * t1;
  u = Jet7(cam.fx) * xd + Jet7(cam.cx);
  v = Jet7(cam.fy) * yd + Jet7(cam.cy);
}

bool nid_eval_grad(const Camera& cam, const double* image, int W, int H, const float* xyzi, size_t n, int bins, const double se3[7], double* nid_out,
                   double* grad_out /*7*/) {
  const Jet7 qx = Jet7::var(se3[0], 0), qy = Jet7::var(se3[1], 1), qz = Jet7::var(se3[2], 2), qw = Jet7::var(se3[3], 3), tx = Jet7::var(se3[4], 4),
             ty = Jet7::var(se3[5], 5), tz = Jet7::var(se3[6], 6);
  std::vector<Jet7> hist((size_t)bins * bins), hist_image(bins);
  std::vector<double> hist_points(bins, 0.0);
  static const double C[4][4] = {{1.0, -3.0, 3.0, -1.0}, {4.0, 0.0, -6.0, 3.0}, {1.0, 3.0, 3.0, -3.0}, {0.0, 0.0, 0.0, 1.0}};
  double Cn[4][4];
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) Cn[i][j] = C[i][j] / 6.0;
  for (size_t i = 0; i < n; i++) {
    const Jet7 px(xyzi[4 * i]), py(xyzi[4 * i + 1]), pz(xyzi[4 * i + 2]);
    Jet7 uvx = qy * pz - qz * py, uvy = qz * px - qx * pz, uvz = qx * py - qy * px;  // [D10]
    uvx += uvx; uvy += uvy; uvz += uvz;
    const Jet7 X = (px + qw * uvx + (qy * uvz - qz * uvy)) + tx;
    const Jet7 Y = (py + qw * uvy + (qz * uvx - qx * uvz)) + ty;
    const Jet7 Z = (pz + qw * uvz + (qx * uvy - qy * uvx)) + tz;
    const double ib = (double)xyzi[4 * i + 3] * bins;
    int bin_p;
    if (!(ib == ib)) bin_p = 0; else if (ib >= (double)(bins - 1)) bin_p = bins - 1; else if (ib <= 0.0) bin_p = 0; else bin_p = (int)ib;
    Jet7 u, v;
    project_jet(cam, X, Y, Z, u, v);
    if (!std::isfinite(u.a) || !std::isfinite(v.a)) continue;
    const double fu = std::floor(u.a), fv = std::floor(v.a);
    if (fu < 0 || fv < 0 || fu >= (double)W || fv >= (double)H) continue;
    const int kx = (int)fu, ky = (int)fv;
    const Jet7 sx = u - Jet7((double)kx), sy = v - Jet7((double)ky);
    hist_points[bin_p] += 1.0;
    const Jet7 sex[4] = {Jet7(1.0), sx, sx * sx, (sx * sx) * sx}, sey[4] = {Jet7(1.0), sy, sy * sy, (sy * sy) * sy};
    Jet7 bx[4], by[4];
    for (int r = 0; r < 4; r++) {
      bx[r] = ((Jet7(Cn[r][0]) * sex[0] + Jet7(Cn[r][1]) * sex[1]) + Jet7(Cn[r][2]) * sex[2]) + Jet7(Cn[r][3]) * sex[3];
      by[r] = ((Jet7(Cn[r][0]) * sey[0] + Jet7(Cn[r][1]) * sey[1]) + Jet7(Cn[r][2]) * sey[2]) + Jet7(Cn[r][3]) * sey[3];
    }
    for (int a = 0; a < 4; a++) {
      const int x = std::min(std::max(kx - 1 + a, 0), W - 1);
      for (int b = 0; b < 4; b++) {
        const int y = std::min(std::max(ky - 1 + b, 0), H - 1);
        const Jet7 w = bx[a] * by[b];
        const int bin_i = std::min<int>((int)(image[(size_t)y * W + x] * bins), bins - 1);
        hist[(size_t)bin_i * bins + bin_p] += w;
        hist_image[bin_i] += w;
      }
    }
  }
  double sum = 0.0;
  for (int i = 0; i < bins; i++) sum += hist_points[i];
  Jet7 Hi, Hj;
  double Hp = 0.0;
  for (int i = 0; i < bins; i++) { const Jet7 p = hist_image[i] / Jet7(sum); Hi += p * jlog(p + Jet7(1e-6)); }
  for (int i = 0; i < bins; i++) { const double p = hist_points[i] / sum; Hp += p * std::log(p + 1e-6); }
  for (int c = 0; c < bins; c++)          // column-major, as in nid_eval  [D14]
    for (int r = 0; r < bins; r++) { const Jet7 p = hist[(size_t)r * bins + c] / Jet7(sum); Hj += p * jlog(p + Jet7(1e-6)); }
  Hi = -Hi; Hp = -Hp; Hj = -Hj;
  const Jet7 MI = (Hi + Jet7(Hp)) - Hj;
  const Jet7 NID = (Hj - MI) / Hj;
  *nid_out = NID.a;
  for (int k = 0; k < 7; k++) grad_out[k] = NID.v[k];
  return std::isfinite(NID.a);
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// generateColorMap's image preparation: BGR -> HSV -> (S * saturation_scale, V * brightness_scale, saturate_cast) -> BGR,
// all 8-bit (src/PointCloudProcessor.cpp:722-741).  cv::cvtColor is third-party (OpenCV, unpinned by the reference's
// CMakeLists.txt:19); restated from OpenCV 4.x imgproc color_hsv and PINNED against cv2 4.13.0 in this container over all
// 2^24 inputs of each direction (tests/test_oracle_cpu.py).  [D12]
//   forward  (RGB2HSV_b, hrange 180): integer, hsv_shift = 12, sdiv_table[i] = cvRound((255 << 12) / i),
//            hdiv_table180[i] = cvRound((180 << 12) / (6 i)).
//   backward (HSV2RGB_b): float.  OpenCV's vector body and its scalar row tail are DIFFERENT functions of the pixel:
//            body  (x < floor(W / block) * block; block = 32 pixels in the cv2 4.13 AVX2 build measured here):
//                  h = H * (6/180); sector = trunc(h) mod 6; f = h - trunc(h); tab = {v, v(1-s), v*fma(-s,f,1), v*fma(-s,1-f,1)};
//                  channel = trunc(tab * 255)            (no s == 0 shortcut, no fmod)
//            tail: h = fmod(H * (6/180), 6); sector = floor(h); s == 0 -> grey; same tab; channel = cvRound(tab * 255).
// ---------------------------------------------------------------------------------------------------
inline uint8_t sat_u8_round(float x) { long v = lrintf(x); return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }  // cvRound + saturate
inline uint8_t sat_u8_trunc(float x) { int v = (int)x; return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }

void hsv_roundtrip(const uint8_t* src, size_t spitch, uint8_t* dst, size_t dpitch, int W, int H, float sat_scale, float val_scale, int block) {
  static int sdiv[256], hdiv[256];
  static bool init = false;
  if (!init) {
    sdiv[0] = hdiv[0] = 0;
    for (int i = 1; i < 256; i++) {
      sdiv[i] = (int)lrint((255 << 12) / (1. * i));
      hdiv[i] = (int)lrint((180 << 12) / (6. * i));
    }
    init = true;
  }
  static const int sector_data[6][3] = {{1, 3, 0}, {1, 0, 2}, {3, 0, 1}, {0, 2, 1}, {0, 1, 3}, {2, 1, 0}};
  const int body = block > 0 ? (W / block) * block : W;
  for (int y = 0; y < H; y++) {
    const uint8_t* sp = src + (size_t)y * spitch;
    uint8_t* dp = dst + (size_t)y * dpitch;
    for (int x = 0; x < W; x++) {
      const int b = sp[3 * x], g = sp[3 * x + 1], r = sp[3 * x + 2];
      int v = b, vmin = b;
      if (g > v) v = g;
      if (r > v) v = r;
      if (g < vmin) vmin = g;
      if (r < vmin) vmin = r;
      const int diff = v - vmin;
      int sI = (diff * sdiv[v] + (1 << 11)) >> 12;
      int hI = v == r ? (g - b) : (v == g ? (b - r + 2 * diff) : (r - g + 4 * diff));
      hI = (hI * hdiv[diff] + (1 << 11)) >> 12;
      if (hI < 0) hI += 180;
      const uint8_t h8 = (uint8_t)(hI < 0 ? 0 : hI > 255 ? 255 : hI);
      const uint8_t s8 = sat_u8_round((float)(uint8_t)sI * sat_scale);   // cpp:733
      const uint8_t v8 = sat_u8_round((float)(uint8_t)v * val_scale);    // cpp:734
      const float hs = 6.0f / 180.0f;
      const float sf = s8 * (1.0f / 255.0f), vf = v8 * (1.0f / 255.0f);
      float h = h8 * hs, f, tab[4];
      int sector;
      const bool in_body = x < body;
      if (in_body) {
        const float pre = truncf(h);
        f = h - pre;
        sector = (int)pre % 6;
      } else {
        h = fmodf(h, 6.0f);
        sector = (int)floorf(h);
        f = h - (float)sector;
        if ((unsigned)sector >= 6u) { sector = 0; f = 0.0f; }
      }
      tab[0] = vf;
      tab[1] = vf * (1.0f - sf);
      tab[2] = vf * fmaf(-sf, f, 1.0f);
      tab[3] = vf * fmaf(-sf, 1.0f - f, 1.0f);
      for (int k = 0; k < 3; k++) {
        float c = tab[sector_data[sector][k]];
        if (!in_body && sf == 0.0f) c = vf;
        dp[3 * x + k] = in_body ? sat_u8_trunc(c * 255.0f) : sat_u8_round(c * 255.0f);
      }
    }
  }
}

// ===================================================================================================
// C interface for ctypes (tests / bench cpu_baseline only)
// ===================================================================================================
extern "C" {

struct orc_camera { double fx, fy, cx, cy, k1, k2, p1, p2, k3; };

static Camera to_cam(const orc_camera* c) { return Camera{c->fx, c->fy, c->cx, c->cy, {c->k1, c->k2, c->p1, c->p2, c->k3}}; }

int orc_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void orc_pose_matrices(const double* pose7, const double* T_opt16, float* w2c16, float* c2w16) {
  Mat4f a, b;
  pose_matrices(pose7, T_opt16, a, b);
  std::memcpy(w2c16, a.m, sizeof(a.m)); std::memcpy(c2w16, b.m, sizeof(b.m));
}

void orc_transform(const float* xyz, size_t n, size_t stride_floats, const float* M16, float* out_xyz /*n x 3*/) {
  Mat4f M; std::memcpy(M.m, M16, sizeof(M.m));
  for (size_t i = 0; i < n; i++)
    transform_point(M, xyz[i * stride_floats], xyz[i * stride_floats + 1], xyz[i * stride_floats + 2], out_xyz[3 * i], out_xyz[3 * i + 1], out_xyz[3 * i + 2]);
}

void orc_project(const orc_camera* c, const double* xyz, size_t n, double* uv) {
  const Camera cam = to_cam(c);
  for (size_t i = 0; i < n; i++) project(cam, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], uv[2 * i], uv[2 * i + 1]);
}

// ViewCulling::cull (z-buffer variant).  Returns number of kept points; out_idx needs room for n ints.
int orc_view_culling(const orc_camera* c, int W, int H, int ds, double tol, int depth_culling, const float* pts_cam, size_t n,
                     const double* T16, int* out_idx, float* dist_map, int* index_map, int* n_candidates) {
  return view_culling(to_cam(c), W, H, ds, tol, depth_culling != 0, pts_cam, n, T16, out_idx, dist_map, index_map, n_candidates);
}

void orc_color_pixels(const orc_camera* c, int W, int H, const float* pts_cam, size_t n, int32_t* pix) {
  const Camera cam = to_cam(c);
  for (size_t i = 0; i < n; i++) pix[i] = color_pixel(cam, W, H, pts_cam[3 * i], pts_cam[3 * i + 1], pts_cam[3 * i + 2]);
}

// generate_distance_masks of scripts/genNormAndDistanceMask.py:233-265 on one frame's culled cloud (the file
// filtered_pcd/<ts>_beforeNID.pcd the script reads, written at src/PointCloudProcessor.cpp:213-217): the points are visited
// in file order (:252), projected with cv2.projectPoints(rvec = tvec = 0) (:245-246; the same plumb_bob model as
// pinhole.hpp:12-51, pinned against cv2 in tests/test_oracle_cpu.py), truncated by astype(int) (:247), and every point
// inside the mask (:258) overwrites distance_mask[y, x] = np.linalg.norm(xyz) stored as float32 (:251,:254,:259) and
// points_3d_mask[y, x] = xyz (:260): the LAST point of the file that lands on a pixel wins.
// Decisions: [M1] open3d holds the file's float32 coordinates as float64, so the norm is sqrt((x*x + y*y) + z*z) in double,
// rounded once to float32 (numpy's dot may fuse the sum on some builds: third-party detail, invisible after the float32
// rounding except on exact ties); [M2] astype(int) of a non-finite / out-of-range value is INT64_MIN in numpy: rejected.
// index_out (extension): file position of the winning point, -1 where the pixel is empty.
void orc_distance_masks(const orc_camera* c, int W, int H, const float* pts_cam /*n x 3*/, size_t n, float* dist /*H x W*/,
                        float* xyz /*H x W x 3 or null*/, int32_t* index_out /*H x W or null*/) {
  const Camera cam = to_cam(c);
  const size_t np_ = (size_t)W * H;
  for (size_t p = 0; p < np_; p++) dist[p] = 0.0f;                                     // np.zeros_like  :251
  if (xyz) for (size_t p = 0; p < 3 * np_; p++) xyz[p] = 0.0f;                         // :252
  if (index_out) for (size_t p = 0; p < np_; p++) index_out[p] = -1;
  for (size_t i = 0; i < n; i++) {                                                     // zip(points_2d, points)  :255
    const double X = pts_cam[3 * i], Y = pts_cam[3 * i + 1], Z = pts_cam[3 * i + 2];
    double ud, vd;
    project(cam, X, Y, Z, ud, vd);
    if (!std::isfinite(ud) || !std::isfinite(vd)) continue;                            // [M2]
    if (!(ud > -2147483648.0 && ud < 2147483648.0 && vd > -2147483648.0 && vd < 2147483648.0)) continue;
    const long x = (long)ud, y = (long)vd;                                             // astype(int)  :247
    if (!(0 <= x && x < W && 0 <= y && y < H)) continue;                               // :258
    const size_t p = (size_t)y * W + (size_t)x;
    dist[p] = (float)std::sqrt((X * X + Y * Y) + Z * Z);                               // :256,:259  [M1]
    if (xyz) { xyz[3 * p] = pts_cam[3 * i]; xyz[3 * p + 1] = pts_cam[3 * i + 1]; xyz[3 * p + 2] = pts_cam[3 * i + 2]; }  // :260
    if (index_out) index_out[p] = (int32_t)i;
  }
}

void orc_scores(const float* pts_cam, size_t n, const double* pose7, float* out3 /*n x 3: orient, dist, final*/) {
  for (size_t i = 0; i < n; i++) {
    const float o = orientation_score(pts_cam[3 * i], pts_cam[3 * i + 1], pts_cam[3 * i + 2], pose7);
    const float d = distance_score(pts_cam[3 * i], pts_cam[3 * i + 1], pts_cam[3 * i + 2]);
    out3[3 * i] = o; out3[3 * i + 1] = d; out3[3 * i + 2] = (float)((double)(o + d) / 2.0);
  }
}

void* orc_colorizer_create(const orc_camera* c, int W, int H, int cullW, int cullH, int ds, double tol, const float* xyzi, size_t n,
                           int keep_obs) {
  Colorizer* S = new Colorizer();
  S->cam = to_cam(c); S->W = W; S->H = H; S->cullW = cullW; S->cullH = cullH; S->ds = ds; S->tol = tol; S->n = n;
  S->xyzi.assign(xyzi, xyzi + 4 * n);
  S->acc.assign(4 * n, 0u); S->mask_hits.assign(n, 0u); S->mask_max.assign(n, 0u);
  S->keep_obs = keep_obs != 0;
  if (S->keep_obs) S->obs.resize(n);
  return S;
}
void orc_hsv_roundtrip(const uint8_t* src, size_t spitch, uint8_t* dst, size_t dpitch, int W, int H, float sat_scale, float val_scale, int block) {
  hsv_roundtrip(src, spitch, dst, dpitch, W, H, sat_scale, val_scale, block);
}
void orc_colorizer_destroy(void* h) { delete (Colorizer*)h; }
void orc_colorizer_set_sample_mode(void* h, int mode) { ((Colorizer*)h)->sample_mode = mode; }

void orc_colorizer_frame(void* h, const double* pose7, const double* T_opt16, const uint8_t* bgr, size_t pitch, const uint8_t* mask,
                         size_t mask_pitch, uint8_t* vis, int32_t* pix, int threads, int* counts3) {
  const FrameResult r = colorize_frame(*(Colorizer*)h, pose7, T_opt16, bgr, pitch, mask, mask_pitch, vis, pix, threads);
  if (counts3) { counts3[0] = r.n_candidates; counts3[1] = r.n_visible; counts3[2] = r.n_colored; }
}

// reindex = 0: carry the index (default, D11); 1: the reference's kd-tree re-indexing by brute force (needs keep_obs; small clouds)
void orc_colorizer_set_reindex(void* h, int mode) { ((Colorizer*)h)->reindex = mode; }
// number of observations every point's list holds (RGBCloud::getPointData(i).rgbScores.size())
void orc_colorizer_obs_counts(void* h, uint32_t* counts) {
  Colorizer* S = (Colorizer*)h;
  for (size_t i = 0; i < S->n; i++) counts[i] = S->keep_obs ? (uint32_t)S->obs[i].size() : 0u;
}

void orc_colorizer_get_acc(void* h, uint32_t* acc4, uint32_t* mask_hits, uint32_t* mask_max) {
  Colorizer* S = (Colorizer*)h;
  if (acc4) std::memcpy(acc4, S->acc.data(), sizeof(uint32_t) * 4 * S->n);
  if (mask_hits) std::memcpy(mask_hits, S->mask_hits.data(), sizeof(uint32_t) * S->n);
  if (mask_max) std::memcpy(mask_max, S->mask_max.data(), sizeof(uint32_t) * S->n);
}

// "sum" finalize: integer mean (truncating), uncoloured points -> 0.
void orc_colorizer_finalize_sum(void* h, uint8_t* rgb /*n x 3*/) {
  Colorizer* S = (Colorizer*)h;
  for (size_t i = 0; i < S->n; i++) {
    const uint32_t* a = &S->acc[4 * i];
    for (int k = 0; k < 3; k++) {
      uint32_t m = a[3] ? a[k] / a[3] : 0;
      if (S->sample_mode == 1) { m = (m + 128) >> 8; if (m > 255) m = 255; }   // 8.8 fixed point mean, rounded
      rgb[3 * i + k] = (uint8_t)m;
    }
  }
}

// smoothColors: top-5 by score, weighted mean, truncation.  src/PointCloudProcessor.cpp:604-631  [D8, D9]
void orc_colorizer_finalize_top5(void* h, uint8_t* rgb /*n x 3*/, int top_m) {
  Colorizer* S = (Colorizer*)h;
  for (size_t i = 0; i < S->n; i++) {
    std::vector<Obs> v = S->obs[i];
    std::stable_sort(v.begin(), v.end(), [](const Obs& a, const Obs& b) { return a.score > b.score; });  // frames already ascending
    const size_t M = std::min(v.size(), (size_t)top_m);
    float total = 0, r = 0, g = 0, b = 0;
    for (size_t j = 0; j < M; j++) {
      const float s = v[j].score;
      r += v[j].r * s; g += v[j].g * s; b += v[j].b * s; total += s;
    }
    if (M == 0) { rgb[3 * i] = rgb[3 * i + 1] = rgb[3 * i + 2] = 0; continue; }
    rgb[3 * i] = (uint8_t)(r / total); rgb[3 * i + 1] = (uint8_t)(g / total); rgb[3 * i + 2] = (uint8_t)(b / total);
  }
}

int orc_nid(const orc_camera* c, const double* image, int W, int H, const float* xyzi, size_t n, int bins, const double* se3,
            double* nid_out, double* hist_out, double* counts_out) {
  return nid_eval(to_cam(c), image, W, H, xyzi, n, bins, se3, nid_out, hist_out, counts_out) ? 1 : 0;
}

int orc_nid_grad(const orc_camera* c, const double* image, int W, int H, const float* xyzi, size_t n, int bins, const double* se3,
                 double* nid_out, double* grad7_out) {
  return nid_eval_grad(to_cam(c), image, W, H, xyzi, n, bins, se3, nid_out, grad7_out) ? 1 : 0;
}

// MultiNIDCost pose gate.  src/vlcal/calib/visual_camera_calibration.cpp:99-105.
// delta = init^-1 * T;  reject if |t| > 0.2 or angle > 2 deg.  Returns 1 if the candidate passes.
int orc_nid_gate(const double* init_se3, const double* se3) {
  auto qmul = [](const double a[4], const double b[4], double o[4]) {  // (x,y,z,w)
    o[3] = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
    o[0] = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
    o[1] = a[3] * b[1] + a[1] * b[3] + a[2] * b[0] - a[0] * b[2];
    o[2] = a[3] * b[2] + a[2] * b[3] + a[0] * b[1] - a[1] * b[0];
  };
  auto rot = [](const double q[4], const double p[3], double o[3]) {
    double uv[3] = {q[1] * p[2] - q[2] * p[1], q[2] * p[0] - q[0] * p[2], q[0] * p[1] - q[1] * p[0]};
    for (double& u : uv) u += u;
    o[0] = p[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
    o[1] = p[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
    o[2] = p[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
  };
  const double qi[4] = {-init_se3[0], -init_se3[1], -init_se3[2], init_se3[3]};  // conjugate
  double dq[4]; qmul(qi, se3, dq);
  const double dt0[3] = {se3[4] - init_se3[4], se3[5] - init_se3[5], se3[6] - init_se3[6]};
  double dt[3]; rot(qi, dt0, dt);
  const double tn = std::sqrt((dt[0] * dt[0] + dt[1] * dt[1]) + dt[2] * dt[2]);
  const double vn = std::sqrt((dq[0] * dq[0] + dq[1] * dq[1]) + dq[2] * dq[2]);
  const double angle = 2.0 * std::atan2(vn, std::fabs(dq[3]));
  return !(tn > 0.2 || angle > 2.0 * M_PI / 180.0);
}

}  // extern "C"
