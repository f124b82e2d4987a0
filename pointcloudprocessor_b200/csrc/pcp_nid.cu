// pcp_nid.cu — (d) the NID joint-histogram cost of vlcal::NIDCost (include/vlcal/costs/nid_cost.hpp:42-116) for a batch of
// candidate poses: per-CTA shared-memory bins (one private copy per warp), per-thread aggregation of the 4x4 B-spline
// taps that fall in the same bin, one flush of each CTA's bins to the per-pose global histogram, then a tiny entropy
// kernel.  Everything is FP64; no tensor cores (nothing here is a contraction).
#include <cmath>
#include <cstring>
#include <vector>

#include "../../include/pcp/pcp.h"
#include "pcp_ctx.hpp"

using namespace pcp;

struct pcp_nid {
  pcp_ctx* ctx = nullptr;
  size_t m = 0;
  int W = 0, H = 0, bins = 16;
  float4* d_pts = nullptr;
  uint8_t* d_binimg = nullptr;  // per-pixel image bin: min(int(pix*bins), bins-1)  (nid_cost.hpp:88)
  double* d_se3 = nullptr;      // B x 7
  double* d_hist = nullptr;     // B x (bins*bins + bins): joint histogram then point-bin counts
  double* d_out = nullptr;      // B x 2: NID, ok flag
  int cap_B = 0;
  double* d_hist_g = nullptr;   // gradient path: B x (8 * bins*bins + bins)
  double* d_out_g = nullptr;    // B x 9: NID, ok flag, 7 partials
  int cap_Bg = 0;
  // batched MultiNIDCost evaluation (pcp_nid_eval_multi_grad), owned by the FIRST cost of the batch: one descriptor per cost, one
  // contiguous histogram / output buffer for all costs x poses, so that an evaluation is two launches and one read-back
  void* d_multi_args = nullptr;  // K x NidArgs
  double* d_multi_hist = nullptr;
  double* d_multi_out = nullptr;
  double* d_multi_se3 = nullptr;
  size_t multi_cap_costs = 0, multi_cap_rows = 0;
  int multi_cap_B = 0;
  std::vector<const void*> multi_key;  // the costs the descriptors were built for
  int multi_key_B = 0;
};

namespace {

constexpr int kNidThreads = 256;
constexpr int kNidSmemDoubles = 2048;  // 16 KB: 8 private copies of a 16x16 histogram

__global__ void nid_bin_image_kernel(const double* __restrict__ img, size_t npix, int bins, uint8_t* __restrict__ out) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
    const double pb = __dmul_rn(img[i], (double)bins);
    int b = __double2int_rz(pb);
    if (!(pb == pb)) b = 0;
    b = b < bins - 1 ? b : bins - 1;
    out[i] = (uint8_t)(b < 0 ? 0 : b);
  }
}

// The same bins from the 8-bit image itself: the cost uses a pixel only through min(int(pix * bins), bins - 1) with pix = the CV_64F
// value convertTo(CV_64FC1, 1.0 / 255.0) makes of the byte (visual_camera_calibration.cpp:173), a function of the byte alone: `lut`.
__global__ void nid_bin_image_u8_kernel(const uint8_t* __restrict__ img, size_t pitch, int W, int H, const uint8_t* __restrict__ lut,
                                        uint8_t* __restrict__ out) {
  __shared__ uint8_t s_lut[256];
  s_lut[threadIdx.x & 255] = lut[threadIdx.x & 255];
  __syncthreads();
  const size_t npix = (size_t)W * H;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
    const size_t y = i / (size_t)W, x = i - y * (size_t)W;
    out[i] = s_lut[img[y * pitch + x]];
  }
}

struct NidArgs {
  const float4* __restrict__ pts;
  const uint8_t* __restrict__ binimg;
  const double* __restrict__ se3;
  double* hist;
  uint32_t m;
  int W, H, bins, copies;
  CamD cam;
};

__device__ __forceinline__ void bspline(double s, double b[4]) {
  // beta = (1/6) * [[1,-3,3,-1],[4,0,-6,3],[1,3,3,-3],[0,0,0,1]] * [1,s,s^2,s^3]^T   (nid_cost.hpp:35-39, :71-77)
  const double s2 = s * s, s3 = s2 * s;
  const double k = 1.0 / 6.0;
  b[0] = ((k + (-3.0 * k) * s) + (3.0 * k) * s2) + (-k) * s3;
  b[1] = ((4.0 * k) + (-6.0 * k) * s2) + (3.0 * k) * s3;
  b[2] = ((k + (3.0 * k) * s) + (3.0 * k) * s2) + (-3.0 * k) * s3;
  b[3] = k * s3;
}

__global__ void __launch_bounds__(kNidThreads) nid_hist_kernel(const NidArgs a) {
#ifndef PCP_NID_DOUBLE_ATOMICS
  // Joint-histogram weights accumulate in shared memory as 2^-31 fixed point in (hi, lo) 32-bit pairs: a shared-memory
  // atomicAdd(double) is an ATOMS.CAST.SPIN loop on this architecture (so is a 64-bit integer add), and the 32 lanes of a
  // warp collide on the 256 bins all the time; ATOMS.ADD on 32 bits is native.  `lo` takes the weight, the lane whose add
  // wraps carries one into `hi`: exact, and independent of the order of the adds.  A flush holds <= 16 taps whose weights sum
  // to <= 1, so it fits in 32 bits; the quantisation is 2^-32 per flush against sums normalised by the point count
  // (relative error of the NID ~1e-9; the tolerance is 1e-5).
  __shared__ unsigned int sh_lo[kNidSmemDoubles], sh_hi[kNidSmemDoubles];
#else
  __shared__ double sh[kNidSmemDoubles];
#endif
  __shared__ unsigned int sh_pts[8 * 32];
  const int bins = a.bins, nb2 = bins * bins;
  const int pose = blockIdx.y;
#ifndef PCP_NID_DOUBLE_ATOMICS
  for (int i = threadIdx.x; i < a.copies * nb2; i += blockDim.x) { sh_lo[i] = 0u; sh_hi[i] = 0u; }
#else
  for (int i = threadIdx.x; i < a.copies * nb2; i += blockDim.x) sh[i] = 0.0;
#endif
  for (int i = threadIdx.x; i < 8 * 32; i += blockDim.x) sh_pts[i] = 0u;
  __syncthreads();
  const int warp = threadIdx.x >> 5;
#ifndef PCP_NID_DOUBLE_ATOMICS
  unsigned int* mylo = sh_lo + (warp % a.copies) * nb2;
  unsigned int* myhi = sh_hi + (warp % a.copies) * nb2;
  auto flush = [&](int key, double acc) {
    const unsigned int q = __double2uint_rn(acc * 2147483648.0);
    const unsigned int old = atomicAdd(&mylo[key], q);
    if (old + q < old) atomicAdd(&myhi[key], 1u);
  };
#else
  double* myh = sh + (warp % a.copies) * nb2;
  auto flush = [&](int key, double acc) { atomicAdd(&myh[key], acc); };
#endif
  unsigned int* myp = sh_pts + warp * 32;

  const double* q = a.se3 + (size_t)pose * 7;
  const double qx = q[0], qy = q[1], qz = q[2], qw = q[3], tx = q[4], ty = q[5], tz = q[6];

  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.m; i += gridDim.x * blockDim.x) {
    const float4 p4 = __ldg(a.pts + i);
    const double px = p4.x, py = p4.y, pz = p4.z;
    // Sophus SE3 * p : p + w*uv + qv x uv, uv = 2*(qv x p)   (nid_cost.hpp:55)
    double ux = qy * pz - qz * py, uy = qz * px - qx * pz, uz = qx * py - qy * px;
    ux += ux; uy += uy; uz += uz;
    const double X = (px + qw * ux + (qy * uz - qz * uy)) + tx;
    const double Y = (py + qw * uy + (qz * ux - qx * uz)) + ty;
    const double Z = (pz + qw * uz + (qx * uy - qy * ux)) + tz;
    const double ib = (double)p4.w * (double)bins;                 // :57-58
    int bin_p;
    if (!(ib == ib)) bin_p = 0; else if (ib >= (double)(bins - 1)) bin_p = bins - 1; else if (ib <= 0.0) bin_p = 0; else bin_p = (int)ib;
    double u, v;
    project(a.cam, X, Y, Z, u, v);                                 // :60 (no z > 0 gate)
    if (!finite_d(u) || !finite_d(v)) continue;
    const double fu = floor(u), fv = floor(v);                     // :61
    if (fu < 0.0 || fv < 0.0 || fu >= (double)a.W || fv >= (double)a.H) continue;  // :64-67
    const int kx = (int)fu, ky = (int)fv;
    atomicAdd(&myp[bin_p], 1u);                                    // :69
    double bx[4], by[4];
    bspline(u - fu, bx);
    bspline(v - fv, by);
    int cur = -1;
    double acc = 0.0;
    #pragma unroll
    for (int ia = 0; ia < 4; ia++) {
      const int x = min(max(kx - 1 + ia, 0), a.W - 1);             // :79-82
      #pragma unroll
      for (int jb = 0; jb < 4; jb++) {
        const int y = min(max(ky - 1 + jb, 0), a.H - 1);
        const int bi = __ldg(a.binimg + (size_t)y * a.W + x);      // :87-88
        const double w = bx[ia] * by[jb];                          // :86
        if (bi != cur) {
          if (cur >= 0) flush(cur * bins + bin_p, acc);  // :89
          cur = bi;
          acc = 0.0;
        }
        acc += w;
      }
    }
    if (cur >= 0) flush(cur * bins + bin_p, acc);
  }
  __syncthreads();
  double* gh = a.hist + (size_t)pose * (nb2 + bins);
  for (int i = threadIdx.x; i < nb2; i += blockDim.x) {
#ifndef PCP_NID_DOUBLE_ATOMICS
    unsigned long long t = 0ull;
    for (int c = 0; c < a.copies; c++) t += ((unsigned long long)sh_hi[c * nb2 + i] << 32) | (unsigned long long)sh_lo[c * nb2 + i];
    const double s = (double)t * (1.0 / 2147483648.0);
#else
    double s = 0.0;
    for (int c = 0; c < a.copies; c++) s += sh[c * nb2 + i];
#endif
    if (s != 0.0) atomicAdd(&gh[i], s);
  }
  for (int i = threadIdx.x; i < bins; i += blockDim.x) {
    unsigned int s = 0;
    for (int w = 0; w < 8; w++) s += sh_pts[w * 32 + i];
    if (s) atomicAdd(&gh[nb2 + i], (double)s);
  }
}

// Entropies and NID (nid_cost.hpp:95-111); one warp per pose.
__global__ void nid_entropy_kernel(const double* __restrict__ hist, int bins, int B, double* __restrict__ out) {
  const int pose = blockIdx.x;
  if (pose >= B) return;
  const int nb2 = bins * bins;
  const double* h = hist + (size_t)pose * (nb2 + bins);
  const int lane = threadIdx.x;
  double sum = 0.0;
  for (int i = lane; i < bins; i += 32) sum += h[nb2 + i];
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  double Hi = 0.0, Hp = 0.0, Hj = 0.0;
  for (int i = lane; i < bins; i += 32) {
    double hi = 0.0;
    for (int j = 0; j < bins; j++) hi += h[i * bins + j];        // hist_image[bin_i] = row sum (:90)
    const double p = hi / sum;
    Hi += p * log(p + 1e-6);
    const double pp = h[nb2 + i] / sum;
    Hp += pp * log(pp + 1e-6);
  }
  for (int i = lane; i < nb2; i += 32) {
    const double p = h[i] / sum;
    Hj += p * log(p + 1e-6);
  }
  for (int o = 16; o > 0; o >>= 1) {
    Hi += __shfl_xor_sync(0xffffffffu, Hi, o);
    Hp += __shfl_xor_sync(0xffffffffu, Hp, o);
    Hj += __shfl_xor_sync(0xffffffffu, Hj, o);
  }
  if (lane == 0) {
    Hi = -Hi; Hp = -Hp; Hj = -Hj;
    const double MI = Hi + Hp - Hj;
    const double nid = (Hj - MI) / Hj;
    out[2 * pose] = nid;
    out[2 * pose + 1] = finite_d(nid) ? 1.0 : 0.0;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Value + gradient: NIDCost::operator()<ceres::Jet<double, 7>> (what ceres::AutoDiffFirstOrderFunction<MultiNIDCost, 7>
// evaluates, visual_camera_calibration.cpp:204-233).  The dual parts are propagated analytically: per point the 3x7 Jacobian
// of Sophus' SE3 * p over the ambient parameters {qx,qy,qz,qw,tx,ty,tz}, the 2x3 Jacobian of the plumb_bob projection, the
// derivative of the cubic B-spline weights; floor / knots / bins carry no derivative (get_real, nid_cost.hpp:61).  Per bin
// the kernel accumulates {w, dw/dtheta_0..6}: 8 planes of bins x bins doubles in shared memory per CTA.
// Same mathematics as the Jets, different rounding order: tolerance 1e-5 relative (north_star) against the oracle's Jet7.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void bspline_d(double s, double b[4], double d[4]) {
  const double s2 = s * s, s3 = s2 * s;
  const double k = 1.0 / 6.0;
  b[0] = ((k + (-3.0 * k) * s) + (3.0 * k) * s2) + (-k) * s3;
  b[1] = ((4.0 * k) + (-6.0 * k) * s2) + (3.0 * k) * s3;
  b[2] = ((k + (3.0 * k) * s) + (3.0 * k) * s2) + (-3.0 * k) * s3;
  b[3] = k * s3;
  d[0] = ((-3.0 * k) + (6.0 * k) * s) + (-3.0 * k) * s2;
  d[1] = (-12.0 * k) * s + (9.0 * k) * s2;
  d[2] = ((3.0 * k) + (6.0 * k) * s) + (-9.0 * k) * s2;
  d[3] = (3.0 * k) * s2;
}

__device__ __forceinline__ void nid_hist_grad_body(const NidArgs& a) {
  extern __shared__ double shg[];  // 8 planes x bins x bins
  __shared__ unsigned int sh_pts[8 * 32];
  const int bins = a.bins, nb2 = bins * bins;
  const int pose = blockIdx.y;
  for (int i = threadIdx.x; i < 8 * nb2; i += blockDim.x) shg[i] = 0.0;
  for (int i = threadIdx.x; i < 8 * 32; i += blockDim.x) sh_pts[i] = 0u;
  __syncthreads();
  unsigned int* myp = sh_pts + (threadIdx.x >> 5) * 32;
  const double* q = a.se3 + (size_t)pose * 7;
  const double qx = q[0], qy = q[1], qz = q[2], qw = q[3], tx = q[4], ty = q[5], tz = q[6];
  const CamD& c = a.cam;

  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.m; i += gridDim.x * blockDim.x) {
    const float4 p4 = __ldg(a.pts + i);
    const double px = p4.x, py = p4.y, pz = p4.z;
    double ux = qy * pz - qz * py, uy = qz * px - qx * pz, uz = qx * py - qy * px;
    ux += ux; uy += uy; uz += uz;
    const double X = (px + qw * ux + (qy * uz - qz * uy)) + tx;
    const double Y = (py + qw * uy + (qz * ux - qx * uz)) + ty;
    const double Z = (pz + qw * uz + (qx * uy - qy * ux)) + tz;
    const double ib = (double)p4.w * (double)bins;
    int bin_p;
    if (!(ib == ib)) bin_p = 0; else if (ib >= (double)(bins - 1)) bin_p = bins - 1; else if (ib <= 0.0) bin_p = 0; else bin_p = (int)ib;
    double u, v;
    project(c, X, Y, Z, u, v);
    if (!finite_d(u) || !finite_d(v)) continue;
    const double fu = floor(u), fv = floor(v);
    if (fu < 0.0 || fv < 0.0 || fu >= (double)a.W || fv >= (double)a.H) continue;
    const int kx = (int)fu, ky = (int)fv;
    atomicAdd(&myp[bin_p], 1u);

    // d(pc)/d(theta): columns qx, qy, qz (d uv = 2 e_k x p; d pc = w d uv + e_k x uv + qv x d uv), qw (uv), t (identity)
    double JX[7], JY[7], JZ[7];
    {
      // k = x: e = (1,0,0): e x p = (0, -pz, py)
      double dx = 0.0, dy = -2.0 * pz, dz = 2.0 * py;
      JX[0] = qw * dx + (0.0) + (qy * dz - qz * dy);
      JY[0] = qw * dy + (-uz) + (qz * dx - qx * dz);
      JZ[0] = qw * dz + (uy) + (qx * dy - qy * dx);
      // k = y: e = (0,1,0): e x p = (pz, 0, -px); e x uv = (uz, 0, -ux)
      dx = 2.0 * pz; dy = 0.0; dz = -2.0 * px;
      JX[1] = qw * dx + (uz) + (qy * dz - qz * dy);
      JY[1] = qw * dy + (0.0) + (qz * dx - qx * dz);
      JZ[1] = qw * dz + (-ux) + (qx * dy - qy * dx);
      // k = z: e = (0,0,1): e x p = (-py, px, 0); e x uv = (-uy, ux, 0)
      dx = -2.0 * py; dy = 2.0 * px; dz = 0.0;
      JX[2] = qw * dx + (-uy) + (qy * dz - qz * dy);
      JY[2] = qw * dy + (ux) + (qz * dx - qx * dz);
      JZ[2] = qw * dz + (0.0) + (qx * dy - qy * dx);
      JX[3] = ux; JY[3] = uy; JZ[3] = uz;
      JX[4] = 1.0; JY[4] = 0.0; JZ[4] = 0.0;
      JX[5] = 0.0; JY[5] = 1.0; JZ[5] = 0.0;
      JX[6] = 0.0; JY[6] = 0.0; JZ[6] = 1.0;
    }
    // projection Jacobian: (u, v) over the normalised (x, y), then (x, y) over (X, Y, Z)
    const double iz = 1.0 / Z, x = X * iz, y = Y * iz;
    const double x2 = x * x, y2 = y * y, r2 = x2 + y2, r4 = r2 * r2;
    const double rc = 1.0 + c.k1 * r2 + c.k2 * r4 + c.k3 * (r2 * r4);
    const double drc = c.k1 + 2.0 * c.k2 * r2 + 3.0 * c.k3 * r4;  // d rc / d r2
    // d xd / dx, d xd / dy, d yd / dx, d yd / dy   (r2_x = 2x, r2_y = 2y)
    const double xd_x = drc * 2.0 * x * x + rc + 2.0 * c.p1 * y + c.p2 * (2.0 * x + 4.0 * x);
    const double xd_y = drc * 2.0 * y * x + 2.0 * c.p1 * x + c.p2 * (2.0 * y);
    const double yd_x = drc * 2.0 * x * y + c.p1 * (2.0 * x) + 2.0 * c.p2 * y;
    const double yd_y = drc * 2.0 * y * y + rc + c.p1 * (2.0 * y + 4.0 * y) + 2.0 * c.p2 * x;
    double gu[7], gv[7];
#pragma unroll
    for (int k = 0; k < 7; k++) {
      const double dxn = (JX[k] - x * JZ[k]) * iz, dyn = (JY[k] - y * JZ[k]) * iz;
      gu[k] = c.fx * (xd_x * dxn + xd_y * dyn);
      gv[k] = c.fy * (yd_x * dxn + yd_y * dyn);
    }
    double bx[4], by[4], dbx[4], dby[4];
    bspline_d(u - fu, bx, dbx);
    bspline_d(v - fv, by, dby);
    int cur = -1;
    double acc = 0.0, accA = 0.0, accB = 0.0;
    auto flush = [&]() {
      double* h = shg + cur * bins + bin_p;
      atomicAdd(h, acc);
#pragma unroll
      for (int k = 0; k < 7; k++) atomicAdd(h + (k + 1) * nb2, accA * gu[k] + accB * gv[k]);
    };
#pragma unroll
    for (int ia = 0; ia < 4; ia++) {
      const int xx = min(max(kx - 1 + ia, 0), a.W - 1);
#pragma unroll
      for (int jb = 0; jb < 4; jb++) {
        const int yy = min(max(ky - 1 + jb, 0), a.H - 1);
        const int bi = __ldg(a.binimg + (size_t)yy * a.W + xx);
        if (bi != cur) {
          if (cur >= 0) flush();
          cur = bi;
          acc = accA = accB = 0.0;
        }
        acc += bx[ia] * by[jb];
        accA += dbx[ia] * by[jb];   // d w / d u
        accB += bx[ia] * dby[jb];   // d w / d v
      }
    }
    if (cur >= 0) flush();
  }
  __syncthreads();
  double* gh = a.hist + (size_t)pose * (8 * nb2 + bins);
  for (int i = threadIdx.x; i < 8 * nb2; i += blockDim.x)
    if (shg[i] != 0.0) atomicAdd(&gh[i], shg[i]);
  for (int i = threadIdx.x; i < bins; i += blockDim.x) {
    unsigned int s = 0;
    for (int w = 0; w < 8; w++) s += sh_pts[w * 32 + i];
    if (s) atomicAdd(&gh[8 * nb2 + i], (double)s);
  }
}

__global__ void __launch_bounds__(kNidThreads) nid_hist_grad_kernel(const NidArgs a) { nid_hist_grad_body(a); }

// All the keyframes of a MultiNIDCost evaluation in one launch: blockIdx.z = cost (keyframe), blockIdx.y = pose, blockIdx.x = point
// tile; the grid's x extent is sized for the largest cloud, tiles beyond a cost's own cloud leave at once.
__global__ void __launch_bounds__(kNidThreads) nid_hist_grad_multi_kernel(const NidArgs* __restrict__ items) {
  const NidArgs a = items[blockIdx.z];
  if (blockIdx.x * blockDim.x >= a.m) return;
  nid_hist_grad_body(a);
}

// Entropies, NID and d NID / d theta; one warp per pose.  out: {NID, ok, g0..g6} per pose.
//   H = -sum p log(p + eps), dH = -sum dp (log(p + eps) + p / (p + eps));  NID = (2 Hj - Hi - Hp) / Hj
__global__ void nid_entropy_grad_kernel(const double* __restrict__ hist, int bins, int B, double* __restrict__ out) {
  const int pose = blockIdx.x;
  if (pose >= B) return;
  const int nb2 = bins * bins;
  const double* h = hist + (size_t)pose * (8 * nb2 + bins);
  const int lane = threadIdx.x;
  double sum = 0.0;
  for (int i = lane; i < bins; i += 32) sum += h[8 * nb2 + i];
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  double Hi = 0.0, Hp = 0.0, Hj = 0.0, dHi[7], dHj[7];
  for (int k = 0; k < 7; k++) dHi[k] = dHj[k] = 0.0;
  for (int i = lane; i < bins; i += 32) {
    double hi = 0.0, dhi[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int j = 0; j < bins; j++) {
      hi += h[i * bins + j];
      for (int k = 0; k < 7; k++) dhi[k] += h[(k + 1) * nb2 + i * bins + j];
    }
    const double p = hi / sum, lg = log(p + 1e-6);
    Hi += p * lg;
    const double f = lg + p / (p + 1e-6);
    for (int k = 0; k < 7; k++) dHi[k] += (dhi[k] / sum) * f;
    const double pp = h[8 * nb2 + i] / sum;
    Hp += pp * log(pp + 1e-6);
  }
  for (int i = lane; i < nb2; i += 32) {
    const double p = h[i] / sum, lg = log(p + 1e-6);
    Hj += p * lg;
    const double f = lg + p / (p + 1e-6);
    for (int k = 0; k < 7; k++) dHj[k] += (h[(k + 1) * nb2 + i] / sum) * f;
  }
  for (int o = 16; o > 0; o >>= 1) {
    Hi += __shfl_xor_sync(0xffffffffu, Hi, o);
    Hp += __shfl_xor_sync(0xffffffffu, Hp, o);
    Hj += __shfl_xor_sync(0xffffffffu, Hj, o);
    for (int k = 0; k < 7; k++) {
      dHi[k] += __shfl_xor_sync(0xffffffffu, dHi[k], o);
      dHj[k] += __shfl_xor_sync(0xffffffffu, dHj[k], o);
    }
  }
  if (lane == 0) {
    Hi = -Hi; Hp = -Hp; Hj = -Hj;
    const double MI = Hi + Hp - Hj;
    const double num = Hj - MI;
    const double nid = num / Hj;
    out[9 * pose] = nid;
    out[9 * pose + 1] = finite_d(nid) ? 1.0 : 0.0;
    for (int k = 0; k < 7; k++) {
      const double dHi_k = -dHi[k], dHj_k = -dHj[k];
      const double dnum = 2.0 * dHj_k - dHi_k;          // Hp is a constant
      out[9 * pose + 2 + k] = (dnum * Hj - num * dHj_k) / (Hj * Hj);
    }
  }
}

#define CUN(call)                                                                                 \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess) {                                                                      \
      nid->ctx->fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return PCP_ECUDA;                                                                           \
    }                                                                                             \
  } while (0)

int ensure_batch(pcp_nid* nid, int B) {
  if (B <= nid->cap_B) return PCP_OK;
  if (nid->d_se3) cudaFree(nid->d_se3);
  if (nid->d_hist) cudaFree(nid->d_hist);
  if (nid->d_out) cudaFree(nid->d_out);
  nid->d_se3 = nullptr; nid->d_hist = nullptr; nid->d_out = nullptr; nid->cap_B = 0;
  const int nb = nid->bins * nid->bins + nid->bins;
  CUN(cudaMalloc(&nid->d_se3, sizeof(double) * 7 * B));
  CUN(cudaMalloc(&nid->d_hist, sizeof(double) * nb * B));
  CUN(cudaMalloc(&nid->d_out, sizeof(double) * 2 * B));
  nid->cap_B = B;
  return PCP_OK;
}

}  // namespace

extern "C" {

static int nid_create_impl(pcp_ctx* ctx, const void* points, size_t m, size_t stride_bytes, size_t intensity_offset, const double* image,
                           const uint8_t* image_u8, size_t pitch_u8, int width, int height, int bins, pcp_nid** out) {
  if (!ctx || !out) return PCP_EINVAL;
  *out = nullptr;
  if ((!points && m) || (!image && !image_u8) || (image_u8 && pitch_u8 < (size_t)(width > 0 ? width : 0)) || width <= 0 || height <= 0 || bins < 2 || bins > 32 || stride_bytes < 12 || (stride_bytes & 3) ||
      (intensity_offset & 3) || intensity_offset + 4 > stride_bytes || m >= 0xfffffff0ull) {
    ctx->fail("pcp_nid_create: invalid argument");
    return PCP_EINVAL;
  }
  if (cudaSetDevice(ctx->device) != cudaSuccess) { ctx->fail("cudaSetDevice failed"); return PCP_ECUDA; }
  pcp_nid* nid = new pcp_nid();
  nid->ctx = ctx; nid->m = m; nid->W = width; nid->H = height; nid->bins = bins;
  cudaStream_t s = ctx->stream;
  const size_t npix = (size_t)width * height;
  // One persistent allocation per cost (points, then the binned image) and NO temporary ones: the staging buffers (8-bit or double
  // plane, bin table, non-float4 point records) live in a context-owned scratch that is kept until pcp_destroy.  cudaFree synchronises
  // the device: three of them per cost were most of the 9.6 ms that setting up 20 keyframes took.
  const size_t pts_bytes = (sizeof(float4) * (m ? m : 1) + 255) & ~(size_t)255;
  const size_t img_bytes = image_u8 ? ((npix + 256 + 255) & ~(size_t)255) : ((npix * sizeof(double) + 255) & ~(size_t)255);
  const bool aos = m && !(stride_bytes == 16 && intensity_offset == 12);
  const size_t need = img_bytes + (aos ? m * stride_bytes : 0);
  uint8_t lut[256];
  cudaError_t e;
  if (ctx->tmp_bytes < need) {
    if (ctx->d_tmp) { if ((e = cudaStreamSynchronize(s)) != cudaSuccess) goto fail; cudaFree(ctx->d_tmp); ctx->d_tmp = nullptr; ctx->tmp_bytes = 0; }
    const size_t want = need + need / 4;
    if ((e = cudaMalloc(&ctx->d_tmp, want)) != cudaSuccess) goto fail;
    ctx->tmp_bytes = want;
  }
  {
    uint8_t* block = nullptr;
    if ((e = cudaMalloc(&block, pts_bytes + npix)) != cudaSuccess) goto fail;
    nid->d_pts = reinterpret_cast<float4*>(block);
    nid->d_binimg = block + pts_bytes;
  }
  if (image_u8) {
    // byte -> bin, evaluated on the host in the operations of the double path: (double)v * (1.0 / 255.0), then * bins, truncation, clamp
    for (int v = 0; v < 256; v++) {
      const double pix = (double)v * (1.0 / 255.0);
      int b = (int)(pix * (double)bins);
      b = b < bins - 1 ? b : bins - 1;
      lut[v] = (uint8_t)(b < 0 ? 0 : b);
    }
    uint8_t* d_u8 = ctx->d_tmp;  // the 8-bit plane (W x H, packed) followed by the 256-entry bin table
    if ((e = cudaMemcpy2DAsync(d_u8, (size_t)width, image_u8, pitch_u8, (size_t)width, (size_t)height, cudaMemcpyHostToDevice, s)) != cudaSuccess) goto fail;
    if ((e = cudaMemcpyAsync(d_u8 + npix, lut, 256, cudaMemcpyHostToDevice, s)) != cudaSuccess) goto fail;
    nid_bin_image_u8_kernel<<<ctx->sm_count * 8, 256, 0, s>>>(d_u8, (size_t)width, width, height, d_u8 + npix, nid->d_binimg);
  } else {
    double* d_img = reinterpret_cast<double*>(ctx->d_tmp);
    if ((e = cudaMemcpyAsync(d_img, image, npix * sizeof(double), cudaMemcpyHostToDevice, s)) != cudaSuccess) goto fail;
    nid_bin_image_kernel<<<ctx->sm_count * 8, 256, 0, s>>>(d_img, npix, bins, nid->d_binimg);
  }
  ctx->launches++;
  if (m) {
    if (!aos) {
      if ((e = cudaMemcpyAsync(nid->d_pts, points, sizeof(float4) * m, cudaMemcpyHostToDevice, s)) != cudaSuccess) goto fail;
    } else {
      uint8_t* d_aos = ctx->d_tmp + img_bytes;
      if ((e = cudaMemcpyAsync(d_aos, points, m * stride_bytes, cudaMemcpyHostToDevice, s)) != cudaSuccess) goto fail;
      launch_aos_to_soa(d_aos, m, stride_bytes, intensity_offset, nid->d_pts, s);
      ctx->launches++;
    }
  }
  if ((e = cudaStreamSynchronize(s)) != cudaSuccess) goto fail;  // the caller's host buffers (and `lut`) may go away; the scratch is free again
  *out = nid;
  return PCP_OK;
fail:
  ctx->fail("pcp_nid_create: %s", cudaGetErrorString(e));
  pcp_nid_destroy(nid);
  return PCP_ECUDA;
}

int pcp_nid_create(pcp_ctx* ctx, const void* points, size_t m, size_t stride_bytes, size_t intensity_offset, const double* image,
                   int width, int height, int bins, pcp_nid** out) {
  if (ctx && !image) { ctx->fail("pcp_nid_create: invalid argument"); return PCP_EINVAL; }
  return nid_create_impl(ctx, points, m, stride_bytes, intensity_offset, image, nullptr, 0, width, height, bins, out);
}

int pcp_nid_create_u8(pcp_ctx* ctx, const void* points, size_t m, size_t stride_bytes, size_t intensity_offset, const uint8_t* image,
                      size_t pitch, int width, int height, int bins, pcp_nid** out) {
  if (ctx && !image) { ctx->fail("pcp_nid_create_u8: invalid argument"); return PCP_EINVAL; }
  return nid_create_impl(ctx, points, m, stride_bytes, intensity_offset, nullptr, image, pitch, width, height, bins, out);
}

void pcp_nid_destroy(pcp_nid* nid) {
  if (!nid) return;
  if (nid->ctx) cudaSetDevice(nid->ctx->device);
  if (nid->d_pts) cudaFree(nid->d_pts);  // one block: the points, then the binned image
  if (nid->d_se3) cudaFree(nid->d_se3);
  if (nid->d_hist) cudaFree(nid->d_hist);
  if (nid->d_out) cudaFree(nid->d_out);
  if (nid->d_hist_g) cudaFree(nid->d_hist_g);
  if (nid->d_out_g) cudaFree(nid->d_out_g);
  if (nid->d_multi_args) cudaFree(nid->d_multi_args);
  if (nid->d_multi_hist) cudaFree(nid->d_multi_hist);
  if (nid->d_multi_out) cudaFree(nid->d_multi_out);
  if (nid->d_multi_se3) cudaFree(nid->d_multi_se3);
  delete nid;
}

int pcp_nid_eval(pcp_nid* nid, const double* se3, int B, double* nid_out, uint8_t* ok_out) {
  if (!nid || !se3 || B <= 0 || !nid_out) return PCP_EINVAL;
  pcp_ctx* ctx = nid->ctx;
  CUN(cudaSetDevice(ctx->device));
  int rc = ensure_batch(nid, B);
  if (rc != PCP_OK) return rc;
  cudaStream_t s = ctx->stream;
  const int nb = nid->bins * nid->bins + nid->bins;
  CUN(cudaMemcpyAsync(nid->d_se3, se3, sizeof(double) * 7 * B, cudaMemcpyHostToDevice, s));
  CUN(cudaMemsetAsync(nid->d_hist, 0, sizeof(double) * nb * B, s));
  NidArgs a;
  a.pts = nid->d_pts; a.binimg = nid->d_binimg; a.se3 = nid->d_se3; a.hist = nid->d_hist;
  a.m = (uint32_t)nid->m; a.W = nid->W; a.H = nid->H; a.bins = nid->bins;
  const int nb2 = nid->bins * nid->bins;
  a.copies = kNidSmemDoubles / nb2 >= 8 ? 8 : (kNidSmemDoubles / nb2 >= 2 ? 2 : 1);
  const pcp_camera& c = ctx->cam;
  a.cam = CamD{c.fx, c.fy, c.cx, c.cy, c.dist[0], c.dist[1], c.dist[2], c.dist[3], c.dist[4]};
  // enough CTAs per pose to fill the GPU: down to one point per thread when few poses are evaluated
  int per_pose = (int)((nid->m + kNidThreads - 1) / kNidThreads);
  const int want = (ctx->sm_count * 4 + B - 1) / B;
  if (per_pose > want) per_pose = want;
  if (per_pose < 1) per_pose = 1;
  for (int b0 = 0; b0 < B; b0 += 65535) {
    const int bn = B - b0 < 65535 ? B - b0 : 65535;
    NidArgs ab = a;
    ab.se3 = a.se3 + (size_t)b0 * 7;
    ab.hist = a.hist + (size_t)b0 * nb;
    nid_hist_kernel<<<dim3(per_pose, bn), kNidThreads, 0, s>>>(ab);
    ctx->launches++;
  }
  nid_entropy_kernel<<<B, 32, 0, s>>>(nid->d_hist, nid->bins, B, nid->d_out);
  ctx->launches++;
  std::vector<double> h(2 * (size_t)B);
  CUN(cudaMemcpyAsync(h.data(), nid->d_out, sizeof(double) * 2 * B, cudaMemcpyDeviceToHost, s));
  CUN(cudaStreamSynchronize(s));
  CUN(cudaGetLastError());
  for (int b = 0; b < B; b++) {
    nid_out[b] = h[2 * b];
    if (ok_out) ok_out[b] = h[2 * b + 1] != 0.0 ? 1 : 0;
  }
  return PCP_OK;
}

// Enqueues the upload of the poses, the histogram / entropy kernels of one cost on the context's stream; no host synchronisation.
static int nid_grad_enqueue(pcp_nid* nid, const double* se3, int B) {
  pcp_ctx* ctx = nid->ctx;
  CUN(cudaSetDevice(ctx->device));
  const int nb2 = nid->bins * nid->bins, nb = 8 * nb2 + nid->bins;
  if (B > nid->cap_Bg) {
    if (nid->d_hist_g) cudaFree(nid->d_hist_g);
    if (nid->d_out_g) cudaFree(nid->d_out_g);
    nid->d_hist_g = nullptr; nid->d_out_g = nullptr; nid->cap_Bg = 0;
    CUN(cudaMalloc(&nid->d_hist_g, sizeof(double) * nb * B));
    CUN(cudaMalloc(&nid->d_out_g, sizeof(double) * 9 * B));
    nid->cap_Bg = B;
  }
  int rc = ensure_batch(nid, B);  // d_se3
  if (rc != PCP_OK) return rc;
  cudaStream_t s = ctx->stream;
  CUN(cudaMemcpyAsync(nid->d_se3, se3, sizeof(double) * 7 * B, cudaMemcpyHostToDevice, s));
  CUN(cudaMemsetAsync(nid->d_hist_g, 0, sizeof(double) * nb * B, s));
  NidArgs a;
  a.pts = nid->d_pts; a.binimg = nid->d_binimg; a.se3 = nid->d_se3; a.hist = nid->d_hist_g;
  a.m = (uint32_t)nid->m; a.W = nid->W; a.H = nid->H; a.bins = nid->bins; a.copies = 1;
  const pcp_camera& c = ctx->cam;
  a.cam = CamD{c.fx, c.fy, c.cx, c.cy, c.dist[0], c.dist[1], c.dist[2], c.dist[3], c.dist[4]};
  const size_t smem = sizeof(double) * 8 * nb2;
  CUN(cudaFuncSetAttribute(nid_hist_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // a line-search evaluation is ONE pose over one keyframe's culled cloud (~4e4 points): down to one point per thread so that it
  // still spreads over the GPU (2048 points per CTA left 22 CTAs on 148 SMs)
  int per_pose = (int)((nid->m + kNidThreads - 1) / kNidThreads);
  const int want = (ctx->sm_count * 4 + B - 1) / B;
  if (per_pose > want) per_pose = want;
  if (per_pose < 1) per_pose = 1;
  for (int b0 = 0; b0 < B; b0 += 65535) {
    const int bn = B - b0 < 65535 ? B - b0 : 65535;
    NidArgs ab = a;
    ab.se3 = a.se3 + (size_t)b0 * 7;
    ab.hist = a.hist + (size_t)b0 * nb;
    nid_hist_grad_kernel<<<dim3(per_pose, bn), kNidThreads, smem, s>>>(ab);
    ctx->launches++;
  }
  nid_entropy_grad_kernel<<<B, 32, 0, s>>>(nid->d_hist_g, nid->bins, B, nid->d_out_g);
  ctx->launches++;
  CUN(cudaGetLastError());
  return PCP_OK;
}

// Reads the B x {NID, ok, 7 partials} of an enqueued evaluation back (synchronises the stream).
static int nid_grad_collect(pcp_nid* nid, int B, double* nid_out, double* grad_out, uint8_t* ok_out) {
  pcp_ctx* ctx = nid->ctx;
  cudaStream_t s = ctx->stream;
  std::vector<double> h(9 * (size_t)B);
  CUN(cudaMemcpyAsync(h.data(), nid->d_out_g, sizeof(double) * 9 * B, cudaMemcpyDeviceToHost, s));
  CUN(cudaStreamSynchronize(s));
  CUN(cudaGetLastError());
  for (int b = 0; b < B; b++) {
    nid_out[b] = h[9 * b];
    if (ok_out) ok_out[b] = h[9 * b + 1] != 0.0 ? 1 : 0;
    for (int k = 0; k < 7; k++) grad_out[7 * (size_t)b + k] = h[9 * b + 2 + k];
  }
  return PCP_OK;
}

int pcp_nid_eval_grad(pcp_nid* nid, const double* se3, int B, double* nid_out, double* grad_out, uint8_t* ok_out) {
  if (!nid || !se3 || B <= 0 || !nid_out || !grad_out) return PCP_EINVAL;
  const int rc = nid_grad_enqueue(nid, se3, B);
  if (rc != PCP_OK) return rc;
  return nid_grad_collect(nid, B, nid_out, grad_out, ok_out);
}

// Pose gate of MultiNIDCost (visual_camera_calibration.cpp:99-105): delta = init^-1 * T.
static bool nid_pose_gate(const double* i7, const double* t7) {
  const double ax = -i7[0], ay = -i7[1], az = -i7[2], aw = i7[3];  // conjugate of the (unit) init quaternion
  const double bx = t7[0], by = t7[1], bz = t7[2], bw = t7[3];
  const double dw = aw * bw - ax * bx - ay * by - az * bz;
  const double dx = aw * bx + ax * bw + ay * bz - az * by;
  const double dy = aw * by + ay * bw + az * bx - ax * bz;
  const double dz = aw * bz + az * bw + ax * by - ay * bx;
  const double p[3] = {t7[4] - i7[4], t7[5] - i7[5], t7[6] - i7[6]};
  double uv[3] = {ay * p[2] - az * p[1], az * p[0] - ax * p[2], ax * p[1] - ay * p[0]};
  for (double& u : uv) u += u;
  const double rx = p[0] + aw * uv[0] + (ay * uv[2] - az * uv[1]);
  const double ry = p[1] + aw * uv[1] + (az * uv[0] - ax * uv[2]);
  const double rz = p[2] + aw * uv[2] + (ax * uv[1] - ay * uv[0]);
  const double tn = std::sqrt((rx * rx + ry * ry) + rz * rz);
  const double vn = std::sqrt((dx * dx + dy * dy) + dz * dz);
  const double angle = 2.0 * std::atan2(vn, std::fabs(dw));
  return !(tn > 0.2 || angle > 2.0 * M_PI / 180.0);
}

int pcp_nid_eval_multi(pcp_nid* const* costs, int n_costs, const double init_se3[7], const double* se3, int B, double* sum_out,
                       uint8_t* ok_out) {
  if (!costs || n_costs <= 0 || !init_se3 || !se3 || B <= 0 || !sum_out || !ok_out) return PCP_EINVAL;
  std::vector<double> vals(B);
  std::vector<uint8_t> oks(B);
  std::vector<uint8_t> gate(B);
  for (int b = 0; b < B; b++) { gate[b] = nid_pose_gate(init_se3, se3 + 7 * (size_t)b) ? 1 : 0; sum_out[b] = 0.0; ok_out[b] = gate[b]; }
  for (int k = 0; k < n_costs; k++) {
    const int rc = pcp_nid_eval(costs[k], se3, B, vals.data(), oks.data());
    if (rc != PCP_OK) return rc;
    for (int b = 0; b < B; b++) {
      sum_out[b] += vals[b];                       // visual_camera_calibration.cpp:116-119
      if (!oks[b]) ok_out[b] = 0;                  // :123
    }
  }
  return PCP_OK;
}

int pcp_nid_eval_multi_grad(pcp_nid* const* costs, int n_costs, const double init_se3[7], const double* se3, int B, double* sum_out,
                            double* grad_sum_out, uint8_t* ok_out) {
  if (!costs || n_costs <= 0 || !init_se3 || !se3 || B <= 0 || !sum_out || !grad_sum_out || !ok_out) return PCP_EINVAL;
  for (int b = 0; b < B; b++) {
    ok_out[b] = nid_pose_gate(init_se3, se3 + 7 * (size_t)b) ? 1 : 0;
    sum_out[b] = 0.0;
    for (int k = 0; k < 7; k++) grad_sum_out[7 * (size_t)b + k] = 0.0;
  }
  for (int c = 0; c < n_costs; c++)
    if (!costs[c]) return PCP_EINVAL;
  pcp_nid* head = costs[0];
  pcp_nid* nid = head;  // CUN reports through nid->ctx
  pcp_ctx* ctx = head->ctx;
  bool same = n_costs <= 65535 && B <= 65535;
  for (int c = 1; c < n_costs; c++) same = same && costs[c]->ctx == ctx && costs[c]->bins == head->bins;
  if (!same) {
    // costs of different contexts / bin counts: one cost at a time (every cost's kernels are enqueued before the first read-back)
    std::vector<double> vals(B), grads(7 * (size_t)B);
    std::vector<uint8_t> oks(B);
    for (int c = 0; c < n_costs; c++) {
      const int rc = nid_grad_enqueue(costs[c], se3, B);
      if (rc != PCP_OK) return rc;
    }
    for (int c = 0; c < n_costs; c++) {
      const int rc = nid_grad_collect(costs[c], B, vals.data(), grads.data(), oks.data());
      if (rc != PCP_OK) return rc;
      for (int b = 0; b < B; b++) {
        sum_out[b] += vals[b];
        for (int k = 0; k < 7; k++) grad_sum_out[7 * (size_t)b + k] += grads[7 * (size_t)b + k];
        if (!oks[b]) ok_out[b] = 0;
      }
    }
    return PCP_OK;
  }
  // ---- batched: every keyframe in ONE histogram launch and ONE entropy launch, one pose upload, one read-back ----------------
  CUN(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->stream;
  const int bins = head->bins, nb2 = bins * bins, nb = 8 * nb2 + bins;
  const size_t rows = (size_t)n_costs * B;
  if (head->multi_cap_costs < (size_t)n_costs || head->multi_cap_rows < rows || head->multi_cap_B < B) {
    if (head->d_multi_args) cudaFree(head->d_multi_args);
    if (head->d_multi_hist) cudaFree(head->d_multi_hist);
    if (head->d_multi_out) cudaFree(head->d_multi_out);
    if (head->d_multi_se3) cudaFree(head->d_multi_se3);
    head->d_multi_args = nullptr; head->d_multi_hist = head->d_multi_out = head->d_multi_se3 = nullptr;
    head->multi_cap_costs = head->multi_cap_rows = 0; head->multi_cap_B = 0; head->multi_key.clear();
    CUN(cudaMalloc(&head->d_multi_args, sizeof(NidArgs) * n_costs));
    CUN(cudaMalloc(&head->d_multi_hist, sizeof(double) * nb * rows));
    CUN(cudaMalloc(&head->d_multi_out, sizeof(double) * 9 * rows));
    CUN(cudaMalloc(&head->d_multi_se3, sizeof(double) * 7 * B));
    head->multi_cap_costs = n_costs; head->multi_cap_rows = rows; head->multi_cap_B = B;
  }
  uint32_t m_max = 0;
  for (int c = 0; c < n_costs; c++) m_max = costs[c]->m > m_max ? (uint32_t)costs[c]->m : m_max;
  bool fresh = head->multi_key.size() != (size_t)n_costs || head->multi_key_B != B;
  for (int c = 0; c < n_costs && !fresh; c++) fresh = head->multi_key[c] != (const void*)costs[c];
  if (fresh) {  // the descriptors change only when the set of costs (or B) does: a BFGS solve uploads them once
    std::vector<NidArgs> items(n_costs);
    const pcp_camera& cm = ctx->cam;
    for (int c = 0; c < n_costs; c++) {
      NidArgs& a = items[c];
      a.pts = costs[c]->d_pts; a.binimg = costs[c]->d_binimg; a.se3 = head->d_multi_se3;
      a.hist = head->d_multi_hist + (size_t)c * B * nb;
      a.m = (uint32_t)costs[c]->m; a.W = costs[c]->W; a.H = costs[c]->H; a.bins = bins; a.copies = 1;
      a.cam = CamD{cm.fx, cm.fy, cm.cx, cm.cy, cm.dist[0], cm.dist[1], cm.dist[2], cm.dist[3], cm.dist[4]};
    }
    CUN(cudaMemcpyAsync(head->d_multi_args, items.data(), sizeof(NidArgs) * n_costs, cudaMemcpyHostToDevice, s));
    CUN(cudaStreamSynchronize(s));  // `items` is pageable stack data
    head->multi_key.assign(n_costs, nullptr);
    for (int c = 0; c < n_costs; c++) head->multi_key[c] = costs[c];
    head->multi_key_B = B;
  }
  CUN(cudaMemcpyAsync(head->d_multi_se3, se3, sizeof(double) * 7 * B, cudaMemcpyHostToDevice, s));
  CUN(cudaMemsetAsync(head->d_multi_hist, 0, sizeof(double) * nb * rows, s));
  const size_t smem = sizeof(double) * 8 * nb2;
  CUN(cudaFuncSetAttribute(nid_hist_grad_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_pose = (int)((m_max + kNidThreads - 1) / kNidThreads);
  const int want = (int)((ctx->sm_count * 4 + rows - 1) / rows);
  if (per_pose > want) per_pose = want;
  if (per_pose < 1) per_pose = 1;
  if (m_max) {
    nid_hist_grad_multi_kernel<<<dim3(per_pose, B, n_costs), kNidThreads, smem, s>>>(static_cast<const NidArgs*>(head->d_multi_args));
    ctx->launches++;
  }
  nid_entropy_grad_kernel<<<(unsigned)rows, 32, 0, s>>>(head->d_multi_hist, bins, (int)rows, head->d_multi_out);
  ctx->launches++;
  CUN(cudaGetLastError());
  std::vector<double> h(9 * rows);
  CUN(cudaMemcpyAsync(h.data(), head->d_multi_out, sizeof(double) * 9 * rows, cudaMemcpyDeviceToHost, s));
  CUN(cudaStreamSynchronize(s));
  for (int c = 0; c < n_costs; c++)  // the sums keep the reference's order of the keyframes (visual_camera_calibration.cpp:116-119)
    for (int b = 0; b < B; b++) {
      const double* r = &h[9 * ((size_t)c * B + b)];
      sum_out[b] += r[0];  // residuals[0] += residuals[i] on Jets: values and dual parts add
      for (int k = 0; k < 7; k++) grad_sum_out[7 * (size_t)b + k] += r[2 + k];
      if (r[1] == 0.0) ok_out[b] = 0;
    }
  return PCP_OK;
}

}  // extern "C"
