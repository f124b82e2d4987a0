// pcp_host_math.hpp — host-side pose algebra of the keyframe loop (src/PointCloudProcessor.cpp:495-519), written out
// scalar by scalar so that the float matrices handed to the kernels equal what the reference's Eigen expressions
// produce on x86-64 (no FMA: this translation unit is compiled with -ffp-contract=off).
// All 4x4 matrices are column-major (element (r,c) at [c*4+r]).
#pragma once
#include <cmath>
#include <cstring>

namespace pcp {
namespace host {

struct M4d { double a[16]; };
struct M4f { float a[16]; };

inline double& at(M4d& m, int r, int c) { return m.a[c * 4 + r]; }
inline double at(const M4d& m, int r, int c) { return m.a[c * 4 + r]; }
inline float& at(M4f& m, int r, int c) { return m.a[c * 4 + r]; }
inline float at(const M4f& m, int r, int c) { return m.a[c * 4 + r]; }

// Isometry3d::Identity().translate(t).rotate(Quaterniond(qw,qx,qy,qz))  (cpp:496-498).
// The quaternion is used as given (Eigen's toRotationMatrix does not normalise).
inline M4d camera_to_world(const double pose[7]) {
  const double px = pose[0], py = pose[1], pz = pose[2];
  const double w = pose[3], x = pose[4], y = pose[5], z = pose[6];
  const double x2 = 2.0 * x, y2 = 2.0 * y, z2 = 2.0 * z;
  const double wx = x2 * w, wy = y2 * w, wz = z2 * w;
  const double xx = x2 * x, xy = y2 * x, xz = z2 * x;
  const double yy = y2 * y, yz = z2 * y, zz = z2 * z;
  M4d T;
  std::memset(&T, 0, sizeof(T));
  at(T, 0, 0) = 1.0 - (yy + zz); at(T, 0, 1) = xy - wz;         at(T, 0, 2) = xz + wy;
  at(T, 1, 0) = xy + wz;         at(T, 1, 1) = 1.0 - (xx + zz); at(T, 1, 2) = yz - wx;
  at(T, 2, 0) = xz - wy;         at(T, 2, 1) = yz + wx;         at(T, 2, 2) = 1.0 - (xx + yy);
  at(T, 0, 3) = px; at(T, 1, 3) = py; at(T, 2, 3) = pz; at(T, 3, 3) = 1.0;
  return T;
}

// Isometry3d::inverse(): R^T and -(R^T) t, three-term sums left to right.
inline M4d rigid_inverse(const M4d& T) {
  M4d I;
  std::memset(&I, 0, sizeof(I));
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) at(I, r, c) = at(T, c, r);
  for (int r = 0; r < 3; r++)
    at(I, r, 3) = ((-at(I, r, 0)) * at(T, 0, 3) + (-at(I, r, 1)) * at(T, 1, 3)) + (-at(I, r, 2)) * at(T, 2, 3);
  at(I, 3, 3) = 1.0;
  return I;
}

// Isometry3d * Isometry3d (cpp:507, :516).
inline M4d rigid_compose(const M4d& A, const M4d& B) {
  M4d C;
  std::memset(&C, 0, sizeof(C));
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) at(C, r, c) = (at(A, r, 0) * at(B, 0, c) + at(A, r, 1) * at(B, 1, c)) + at(A, r, 2) * at(B, 2, c);
    at(C, r, 3) = ((at(A, r, 0) * at(B, 0, 3) + at(A, r, 1) * at(B, 1, 3)) + at(A, r, 2) * at(B, 2, 3)) + at(A, r, 3);
  }
  at(C, 3, 3) = 1.0;
  return C;
}

inline M4f narrow(const M4d& T) {
  M4f F;
  for (int i = 0; i < 16; i++) F.a[i] = (float)T.a[i];
  return F;
}

// Affine3f::inverse() (cpp:509, :518): cofactor inverse of the 3x3 block in float, then -(Linv) t.
inline M4f affine_inverse(const M4f& A) {
  float cof[3][3];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      cof[i][j] = at(A, i1, j1) * at(A, i2, j2) - at(A, i1, j2) * at(A, i2, j1);
    }
  const float det = (cof[0][0] * at(A, 0, 0) + cof[1][0] * at(A, 1, 0)) + cof[2][0] * at(A, 2, 0);
  const float inv = 1.0f / det;
  M4f I;
  std::memset(&I, 0, sizeof(I));
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) at(I, r, c) = cof[c][r] * inv;
  for (int r = 0; r < 3; r++)
    at(I, r, 3) = ((-at(I, r, 0)) * at(A, 0, 3) + (-at(I, r, 1)) * at(A, 1, 3)) + (-at(I, r, 2)) * at(A, 2, 3);
  at(I, 3, 3) = 1.0f;
  return I;
}

// The pair (transformation_w2c_optimized, transformation_c2w_optimized) of cpp:499-519.
inline void pose_matrices(const double pose[7], const double* T_opt16, M4f& w2c, M4f& c2w) {
  const M4d c2w_d = camera_to_world(pose);
  if (!T_opt16) {
    w2c = narrow(rigid_inverse(c2w_d));
    c2w = narrow(c2w_d);
  } else {
    M4d Topt;
    std::memcpy(Topt.a, T_opt16, sizeof(Topt.a));
    c2w = narrow(rigid_compose(c2w_d, Topt));
    w2c = affine_inverse(c2w);
  }
}

}  // namespace host
}  // namespace pcp
