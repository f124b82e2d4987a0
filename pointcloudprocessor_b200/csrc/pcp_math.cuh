// pcp_math.cuh — bit-exact device arithmetic for the colorization hot path (sm_100a).
//
// Every operation that decides a visibility bit or a pixel index is written with explicit round-to-nearest
// intrinsics (__fmul_rn/__fadd_rn/__dmul_rn/__dadd_rn/__ddiv_rn/__dsqrt_rn), which nvcc never contracts into
// FMA, so the results equal the reference's x86-64 SSE2 Release build operation by operation.
// Reference citations are relative to /root/reference/PointCloudProcessor/.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace pcp {

struct CamD {
  double fx, fy, cx, cy;
  double k1, k2, p1, p2, k3;  // include/camera/pinhole.hpp:14-19
};

// Row-major 3x4 float transform [R|t]: pcl::transformPointCloud (src/PointCloudProcessor.cpp:521).
struct Xf34 {
  float r[12];
};

// PCL 1.10 pcl::detail::Transformer<float>::se3 summation order: x*c0 + (y*c1 + (z*c2 + c3)).
__device__ __forceinline__ void transform_point(const Xf34& M, float x, float y, float z, float& ox, float& oy, float& oz) {
  ox = __fadd_rn(__fmul_rn(x, M.r[0]), __fadd_rn(__fmul_rn(y, M.r[1]), __fadd_rn(__fmul_rn(z, M.r[2]), M.r[3])));
  oy = __fadd_rn(__fmul_rn(x, M.r[4]), __fadd_rn(__fmul_rn(y, M.r[5]), __fadd_rn(__fmul_rn(z, M.r[6]), M.r[7])));
  oz = __fadd_rn(__fmul_rn(x, M.r[8]), __fadd_rn(__fmul_rn(y, M.r[9]), __fadd_rn(__fmul_rn(z, M.r[10]), M.r[11])));
}

// Optional double-precision T_camera_lidar * (x,y,z,1) of ViewCulling::cull (view_culling.cpp:33-37); row-major 3x4.
struct Xd34 {
  double r[12];
};
__device__ __forceinline__ void transform_point_d(const Xd34& T, double x, double y, double z, double& ox, double& oy, double& oz) {
  ox = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T.r[0], x), __dmul_rn(T.r[1], y)), __dmul_rn(T.r[2], z)), T.r[3]);
  oy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T.r[4], x), __dmul_rn(T.r[5], y)), __dmul_rn(T.r[6], z)), T.r[7]);
  oz = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T.r[8], x), __dmul_rn(T.r[9], y)), __dmul_rn(T.r[10], z)), T.r[11]);
}

// Pinhole + plumb_bob in double, operation order of include/camera/pinhole.hpp:12-51
// (== include/PointCloudProcessor.hpp:100-123 + src/PointCloudProcessor.cpp:750-753).
__device__ __forceinline__ void project(const CamD& c, double X, double Y, double Z, double& u, double& v) {
  const double x = __ddiv_rn(X, Z), y = __ddiv_rn(Y, Z);
  const double x2 = __dmul_rn(x, x), y2 = __dmul_rn(y, y);
  const double r2 = __dadd_rn(x2, y2);
  const double r4 = __dmul_rn(r2, r2);
  const double r6 = __dmul_rn(r2, r4);
  const double rc = __dadd_rn(__dadd_rn(__dadd_rn(1.0, __dmul_rn(c.k1, r2)), __dmul_rn(c.k2, r4)), __dmul_rn(c.k3, r6));
  const double t1 = __dmul_rn(__dmul_rn(2.0, x), y);
  const double t2 = __dadd_rn(r2, __dmul_rn(2.0, x2));
  const double t3 = __dadd_rn(r2, __dmul_rn(2.0, y2));
  const double xd = __dadd_rn(__dadd_rn(__dmul_rn(rc, x), __dmul_rn(c.p1, t1)), __dmul_rn(c.p2, t2));
  const double yd = __dadd_rn(__dadd_rn(__dmul_rn(rc, y), __dmul_rn(c.p1, t3)), __dmul_rn(c.p2, t1));
  u = __dadd_rn(__dmul_rn(c.fx, xd), c.cx);
  v = __dadd_rn(__dmul_rn(c.fy, yd), c.cy);
}

__device__ __forceinline__ bool finite_f(float a) { return fabsf(a) <= 3.402823466e+38f; }  // false for NaN
__device__ __forceinline__ bool finite_d(double a) { return fabs(a) <= 1.7976931348623157e+308; }

// Pass-1 gate of ViewCulling::view_culling (view_culling.cpp:76-97).  Returns true when the reference records
// the point; (px,py) is its down-sampled cell (may lie outside the Wc x Hc grid: such points are recorded by
// the reference and dropped by its second pass, :162-165).
__device__ __forceinline__ bool cull_gate(double u, double v, float dsf, int W, int H, int& px, int& py) {
  const float uf = __double2float_rn(u), vf = __double2float_rn(v);  // .cast<float>()  :86
  if (!finite_f(uf) || !finite_f(vf)) return false;                  // oracle decision D7
  const float qx = __fdiv_rn(uf, dsf), qy = __fdiv_rn(vf, dsf);      // Vector2f / int  :88
  if (!(qx > -2147483648.0f && qx < 2147483648.0f && qy > -2147483648.0f && qy < 2147483648.0f)) return false;
  px = __float2int_rz(qx);
  py = __float2int_rz(qy);
  return !(px < 0 || py < 0 || px >= W || py >= H);                  // :90 (full-resolution image_size)
}

// Eigen Vector3d::norm() (view_culling.cpp:102): sqrt((x*x + y*y) + z*z).
__device__ __forceinline__ double norm3(double X, double Y, double Z) {
  return __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(X, X), __dmul_rn(Y, Y)), __dmul_rn(Z, Z)));
}

// generateColorMap's pixel (src/PointCloudProcessor.cpp:748-754): z > 0, truncation of the double coordinates,
// bounds against the actual image.  Returns v*W+u or -1.
__device__ __forceinline__ int color_pixel(double Z, double u, double v, int W, int H) {
  if (!(Z > 0.0)) return -1;
  if (!finite_d(u) || !finite_d(v)) return -1;
  if (!(u > -2147483648.0 && u < 2147483648.0 && v > -2147483648.0 && v < 2147483648.0)) return -1;
  const int iu = __double2int_rz(u), iv = __double2int_rz(v);
  if (iu >= 0 && iu < W && iv >= 0 && iv < H) return iv * W + iu;
  return -1;
}

// Scores of include/PointCloudProcessor.hpp:205-236 and src/PointCloudProcessor.cpp:588.
__device__ __forceinline__ float orientation_score(float px, float py, float pz, double cx, double cy, double cz) {
  const double vx = __dsub_rn((double)px, cx), vy = __dsub_rn((double)py, cy), vz = __dsub_rn((double)pz, cz);
  const double n2 = __dadd_rn(__dadd_rn(__dmul_rn(vx, vx), __dmul_rn(vy, vy)), __dmul_rn(vz, vz));
  double c = vz;
  if (n2 > 0.0) c = __ddiv_rn(vz, __dsqrt_rn(n2));
  float s = __double2float_rn(__ddiv_rn(__dadd_rn(c, 1.0), 2.0));
  return __fadd_rn(0.2f, __fmul_rn(0.8f, s));
}
__device__ __forceinline__ float distance_score(float px, float py, float pz) {
  const float d = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)), __fmul_rn(pz, pz)));
  const float diff = fabsf(__fsub_rn(d, 2.0f));
  const float nd = fminf(__fdiv_rn(diff, 2.0f), 1.0f);
  const float s = __fsub_rn(1.0f, nd);
  return __fadd_rn(0.2f, __fmul_rn(0.8f, s));
}
__device__ __forceinline__ float final_score(float o, float d) { return __fmul_rn(__fadd_rn(o, d), 0.5f); }

}  // namespace pcp
