// pcp_calib.cu — the caller of the NID cost: vlcal::VisualCameraCalibration::calibrate (outer loop,
// src/vlcal/calib/visual_camera_calibration.cpp:49-84) and estimate_pose_bfgs (:142-252), which hands
// AutoDiffFirstOrderFunction<MultiNIDCost, 7> on Sophus::Manifold<SE3> to ceres::GradientProblemSolver with
// line_search_direction_type = BFGS and every other option at its default.
//
// Ceres and Sophus are third-party dependencies that are absent from the reference tree (find_package(Ceres), CMakeLists.txt:22,
// version not pinned; thirdparty/Sophus is an empty submodule).  This file restates their published algorithms:
//   * Sophus::Manifold<SE3>: Plus(x, d) = x * SE3::exp(d), tangent order {upsilon, omega}; PlusJacobian = Dx_this_mul_exp_x_at_0;
//   * ceres::internal::LineSearchMinimizer with the BFGS direction (dense inverse-Hessian update, secant tolerance 1e-14) and a
//     strong-Wolfe line search (bracketing then zoom, cubic interpolation, halving on invalid evaluations) with ceres' default
//     constants (function_tolerance 1e-6, gradient_tolerance 1e-10, parameter_tolerance 1e-8, max_num_iterations 50,
//     sufficient decrease 1e-4, curvature 0.9, expansion <= 10, contraction in [1e-3, 0.6], 20 line-search iterations).
// The iterates are not claimed bit-equal to a particular ceres release (parity unpinned); the cost and gradient
// every iteration evaluates are the parity-tested device kernels of pcp_nid.cu, one batched launch pair per frame.
// The solver itself is host code (6 unknowns); all the arithmetic over points runs on the GPU.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/pcp/pcp.h"

namespace pcp {
namespace calib {

// ---- SE3 in Sophus' ambient parameterisation {qx,qy,qz,qw,tx,ty,tz} ---------------------------------------------------------
static void quat_rotate(const double q[4], const double v[3], double out[3]) {
  // Eigen's Quaternion::_transformVector: v + 2w (u x v) + 2 u x (u x v), with uv = 2 (u x v)
  const double ux = q[0], uy = q[1], uz = q[2], w = q[3];
  double uv[3] = {uy * v[2] - uz * v[1], uz * v[0] - ux * v[2], ux * v[1] - uy * v[0]};
  for (double& c : uv) c += c;
  out[0] = v[0] + w * uv[0] + (uy * uv[2] - uz * uv[1]);
  out[1] = v[1] + w * uv[1] + (uz * uv[0] - ux * uv[2]);
  out[2] = v[2] + w * uv[2] + (ux * uv[1] - uy * uv[0]);
}

static void quat_mul(const double a[4], const double b[4], double out[4]) {
  const double ax = a[0], ay = a[1], az = a[2], aw = a[3], bx = b[0], by = b[1], bz = b[2], bw = b[3];
  out[3] = aw * bw - ax * bx - ay * by - az * bz;
  out[0] = aw * bx + ax * bw + ay * bz - az * by;
  out[1] = aw * by + ay * bw + az * bx - ax * bz;
  out[2] = aw * bz + az * bw + ax * by - ay * bx;
}

// SO3::operator*: the product is pulled back to the unit sphere with one first-order step when it has drifted.
static void so3_mul(const double a[4], const double b[4], double out[4]) {
  quat_mul(a, b, out);
  const double n2 = out[0] * out[0] + out[1] * out[1] + out[2] * out[2] + out[3] * out[3];
  if (n2 != 1.0) {
    const double scale = 2.0 / (1.0 + n2);
    for (int i = 0; i < 4; i++) out[i] *= scale;
  }
}

// SE3::exp: tangent {upsilon (3), omega (3)}.
static void se3_exp(const double a[6], double out[7]) {
  const double* up = a;
  const double* om = a + 3;
  const double theta_sq = om[0] * om[0] + om[1] * om[1] + om[2] * om[2];
  double imag, real, theta = 0.0;
  const double eps = 1e-10;
  if (theta_sq < eps * eps) {
    const double theta_po4 = theta_sq * theta_sq;
    imag = 0.5 - theta_sq / 48.0 + theta_po4 / 3840.0;
    real = 1.0 - theta_sq / 8.0 + theta_po4 / 384.0;
  } else {
    theta = std::sqrt(theta_sq);
    const double half = 0.5 * theta;
    imag = std::sin(half) / theta;
    real = std::cos(half);
  }
  out[0] = imag * om[0]; out[1] = imag * om[1]; out[2] = imag * om[2]; out[3] = real;
  // t = V upsilon, V = I + (1 - cos)/theta^2 Omega + (theta - sin)/theta^3 Omega^2  (V = R for tiny theta)
  if (theta < eps) {
    quat_rotate(out, up, out + 4);
  } else {
    const double A = (1.0 - std::cos(theta)) / theta_sq, B = (theta - std::sin(theta)) / (theta_sq * theta);
    const double w1[3] = {om[1] * up[2] - om[2] * up[1], om[2] * up[0] - om[0] * up[2], om[0] * up[1] - om[1] * up[0]};      // omega x u
    const double w2[3] = {om[1] * w1[2] - om[2] * w1[1], om[2] * w1[0] - om[0] * w1[2], om[0] * w1[1] - om[1] * w1[0]};      // omega x (omega x u)
    for (int i = 0; i < 3; i++) out[4 + i] = up[i] + A * w1[i] + B * w2[i];
  }
}

static void se3_mul(const double a[7], const double b[7], double out[7]) {
  double q[4], rt[3];
  so3_mul(a, b, q);
  quat_rotate(a, b + 4, rt);
  for (int i = 0; i < 4; i++) out[i] = q[i];
  for (int i = 0; i < 3; i++) out[4 + i] = a[4 + i] + rt[i];
}

static void se3_inverse(const double a[7], double out[7]) {
  out[0] = -a[0]; out[1] = -a[1]; out[2] = -a[2]; out[3] = a[3];
  double t[3];
  quat_rotate(out, a + 4, t);
  for (int i = 0; i < 3; i++) out[4 + i] = -t[i];
}

// Manifold::Plus.
static void se3_plus(const double x[7], const double d[6], double out[7]) {
  double e[7];
  se3_exp(d, e);
  se3_mul(x, e, out);
}

// g6 = PlusJacobian(x)^T g7, PlusJacobian = Dx_this_mul_exp_x_at_0 (7 x 6).
static void tangent_gradient(const double x[7], const double g7[7], double g6[6]) {
  // rotation matrix of the (unit) quaternion, Eigen's toRotationMatrix
  const double qx = x[0], qy = x[1], qz = x[2], qw = x[3];
  const double tx = 2 * qx, ty = 2 * qy, tz = 2 * qz;
  const double twx = tx * qw, twy = ty * qw, twz = tz * qw, txx = tx * qx, txy = ty * qx, txz = tz * qx, tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
  const double R[3][3] = {{1 - (tyy + tzz), txy - twz, txz + twy}, {txy + twz, 1 - (txx + tzz), tyz - twx}, {txz - twy, tyz + twx, 1 - (txx + tyy)}};
  for (int c = 0; c < 3; c++) g6[c] = R[0][c] * g7[4] + R[1][c] * g7[5] + R[2][c] * g7[6];  // d t / d upsilon = R
  g6[3] = 0.5 * (qw * g7[0] + qz * g7[1] - qy * g7[2] - qx * g7[3]);
  g6[4] = 0.5 * (-qz * g7[0] + qw * g7[1] + qx * g7[2] - qy * g7[3]);
  g6[5] = 0.5 * (qy * g7[0] - qx * g7[1] + qw * g7[2] - qz * g7[3]);
}

static void se3_from_matrix(const double T[16], double out[7]) {
  // Eigen::Quaterniond(Matrix3d) (Shepperd's branches as Eigen orders them), column-major input
  const double m00 = T[0], m10 = T[1], m20 = T[2], m01 = T[4], m11 = T[5], m21 = T[6], m02 = T[8], m12 = T[9], m22 = T[10];
  double t = m00 + m11 + m22, q[4];
  if (t > 0.0) {
    t = std::sqrt(t + 1.0);
    q[3] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (m21 - m12) * t; q[1] = (m02 - m20) * t; q[2] = (m10 - m01) * t;
  } else {
    const double m[3][3] = {{m00, m01, m02}, {m10, m11, m12}, {m20, m21, m22}};
    int i = 0;
    if (m11 > m00) i = 1;
    if (m22 > m[i][i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(m[i][i] - m[j][j] - m[k][k] + 1.0);
    q[i] = 0.5 * t;
    t = 0.5 / t;
    q[3] = (m[k][j] - m[j][k]) * t;
    q[j] = (m[j][i] + m[i][j]) * t;
    q[k] = (m[k][i] + m[i][k]) * t;
  }
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);  // Sophus::SE3d(Matrix4d) normalises
  for (int i = 0; i < 4; i++) out[i] = q[i] / n;
  out[4] = T[12]; out[5] = T[13]; out[6] = T[14];
}

static void se3_to_matrix(const double x[7], double T[16]) {
  const double qx = x[0], qy = x[1], qz = x[2], qw = x[3];
  const double tx = 2 * qx, ty = 2 * qy, tz = 2 * qz;
  const double twx = tx * qw, twy = ty * qw, twz = tz * qw, txx = tx * qx, txy = ty * qx, txz = tz * qx, tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
  T[0] = 1 - (tyy + tzz); T[4] = txy - twz; T[8] = txz + twy; T[12] = x[4];
  T[1] = txy + twz; T[5] = 1 - (txx + tzz); T[9] = tyz - twx; T[13] = x[5];
  T[2] = txz - twy; T[6] = tyz + twx; T[10] = 1 - (txx + tyy); T[14] = x[6];
  T[3] = 0; T[7] = 0; T[11] = 0; T[15] = 1;
}

// Rotation angle of a unit quaternion as Eigen::AngleAxisd(Matrix3d).angle() reports it (in [0, pi]).
static double quat_angle(const double q[4]) {
  const double vn = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  return 2.0 * std::atan2(vn, std::fabs(q[3]));
}

// ---- the first-order problem over the manifold ----------------------------------------------------------------------------
struct Problem {
  pcp_se3_cost_fn fn;
  void* user;
  int evals = 0;
  // value and TANGENT gradient at x; false where the functor returns false or produces non-finite numbers (ceres treats both
  // as an invalid evaluation)
  bool eval(const double x[7], double* value, double g6[6]) {
    double g7[7] = {0, 0, 0, 0, 0, 0, 0};
    evals++;
    const int ok = fn(user, x, value, g7);
    if (!ok || !std::isfinite(*value)) return false;
    tangent_gradient(x, g7, g6);
    for (int i = 0; i < 6; i++)
      if (!std::isfinite(g6[i])) return false;
    return true;
  }
};

struct Sample {
  double x = 0.0, value = 0.0, gradient = 0.0;  // step along the direction, cost, directional derivative
  bool valid = false;
  double state[7] = {0, 0, 0, 1, 0, 0, 0};      // Plus(x0, x * direction)
  double g6[6] = {0, 0, 0, 0, 0, 0};            // tangent gradient there
};

struct LineFn {
  Problem* p;
  const double* x0;
  const double* dir;
  void eval(double step, Sample* s) const {
    double d[6];
    for (int i = 0; i < 6; i++) d[i] = step * dir[i];
    s->x = step;
    se3_plus(x0, d, s->state);
    s->valid = p->eval(s->state, &s->value, s->g6);
    s->gradient = 0.0;
    if (s->valid)
      for (int i = 0; i < 6; i++) s->gradient += s->g6[i] * dir[i];
  }
};

// Minimiser over [lo, hi] of the cubic through two samples with values and slopes (falls back to the quadratic through
// the valid data, then to bisection).
static double interpolate_min(const Sample& a, const Sample& b, double lo, double hi) {
  if (!a.valid || !b.valid || a.x == b.x) return std::fmin(std::fmax(0.5 * (lo + hi), lo), hi);
  const double h = b.x - a.x;
  // Hermite form in s = (x - a.x) / h: p(s) = c0 + c1 s + c2 s^2 + c3 s^3
  const double c0 = a.value, c1 = a.gradient * h;
  const double df = b.value - a.value;
  const double c2 = 3.0 * df - (2.0 * a.gradient + b.gradient) * h;
  const double c3 = -2.0 * df + (a.gradient + b.gradient) * h;
  auto poly = [&](double x) { const double s = (x - a.x) / h; return c0 + s * (c1 + s * (c2 + s * c3)); };
  double best_x = lo, best_v = poly(lo);
  auto consider = [&](double x) {
    if (!(x >= lo && x <= hi)) return;
    const double v = poly(x);
    if (v < best_v) { best_v = v; best_x = x; }
  };
  consider(hi);
  // p'(s) = c1 + 2 c2 s + 3 c3 s^2
  if (std::fabs(c3) > 1e-300) {
    const double disc = c2 * c2 - 3.0 * c3 * c1;
    if (disc >= 0.0) {
      const double r = std::sqrt(disc);
      consider(a.x + h * ((-c2 + r) / (3.0 * c3)));
      consider(a.x + h * ((-c2 - r) / (3.0 * c3)));
    }
  } else if (std::fabs(c2) > 1e-300) {
    consider(a.x + h * (-c1 / (2.0 * c2)));
  }
  return best_x;
}

struct LineSearchResult {
  bool success = false;
  Sample point;
  int iterations = 0;
};

// Strong-Wolfe search: bracketing then zoom.  f0 / g0: cost and directional derivative (< 0) at step 0.
static LineSearchResult wolfe_search(const LineFn& f, double step0, double f0, double g0, double dir_max_norm, const pcp_bfgs_options& o) {
  LineSearchResult res;
  Sample initial;
  initial.x = 0.0; initial.value = f0; initial.gradient = g0; initial.valid = true;
  Sample previous = initial, current;
  Sample best;  // lowest valid sample that satisfies the sufficient-decrease condition
  auto armijo = [&](const Sample& s) { return s.valid && s.value <= f0 + o.sufficient_function_decrease * g0 * s.x; };
  auto wolfe = [&](const Sample& s) { return s.valid && std::fabs(s.gradient) <= -o.sufficient_curvature_decrease * g0; };
  auto note = [&](const Sample& s) { if (armijo(s) && s.x > 0.0 && (!best.valid || s.value < best.value)) best = s; };
  auto finish = [&](bool ok, const Sample& s) { res.success = ok; res.point = s; return res; };

  f.eval(step0, &current);
  note(current);
  Sample low, high;
  bool do_zoom = false;
  for (;;) {
    res.iterations++;
    if (current.valid && !armijo(current)) { low = previous; high = current; do_zoom = true; break; }
    if (current.valid && previous.x > 0.0 && current.value >= previous.value) { low = previous; high = current; do_zoom = true; break; }
    if (wolfe(current)) return finish(true, current);
    if (current.valid && current.gradient >= 0.0) { low = current; high = previous; do_zoom = true; break; }
    if (current.valid && std::fabs(current.x - previous.x) * dir_max_norm < o.min_line_search_step_size) break;
    if (res.iterations >= o.max_num_line_search_step_size_iterations) break;
    double step;
    if (current.valid) {
      step = interpolate_min(previous, current, current.x, current.x * o.max_line_search_step_expansion);
      step = std::fmax(step, 1.1 * current.x);  // always move forward while no bracket exists
      previous = current;
    } else {
      step = 0.5 * current.x;                   // invalid evaluation (pose gate, non-finite NID): halve
      if (step * dir_max_norm < o.min_line_search_step_size) break;
    }
    f.eval(step, &current);
    note(current);
  }
  if (!do_zoom) return best.valid ? finish(true, best) : finish(false, current);

  // zoom: `low` satisfies sufficient decrease and has the lower cost of the two ends
  for (;;) {
    if (res.iterations >= o.max_num_line_search_step_size_iterations) break;
    if (std::fabs(high.x - low.x) * dir_max_norm < o.min_line_search_step_size) break;
    res.iterations++;
    const Sample& lower = low.x < high.x ? low : high;
    const Sample& upper = low.x < high.x ? high : low;
    double step = interpolate_min(lower, upper, lower.x, upper.x);
    // keep the trial inside the bracket by the contraction limits (a cubic minimum on an end makes no progress)
    const double width = upper.x - lower.x;
    const double margin = o.max_line_search_step_contraction * width;
    if (step < lower.x + margin || step > upper.x - margin) step = lower.x + 0.5 * width;
    Sample s;
    f.eval(step, &s);
    if (!s.valid) { high = s; continue; }  // an invalid trial bounds the bracket; the next trial bisects
    note(s);
    if (!armijo(s) || s.value >= low.value) { high = s; continue; }
    if (wolfe(s)) return finish(true, s);
    if (s.gradient * (high.x - low.x) >= 0.0) high = low;
    low = s;
  }
  return best.valid ? finish(true, best) : finish(false, low);
}

// ---- LineSearchMinimizer with the dense BFGS direction ----------------------------------------------------------------------
static double max_norm_of_gradient_step(const double x[7], const double g6[6]) {
  // ceres: gradient_max_norm = || x - Plus(x, -gradient) ||_inf
  double neg[6], xp[7], m = 0.0;
  for (int i = 0; i < 6; i++) neg[i] = -g6[i];
  se3_plus(x, neg, xp);
  for (int i = 0; i < 7; i++) m = std::fmax(m, std::fabs(x[i] - xp[i]));
  return m;
}

static void set_message(pcp_bfgs_summary* s, int termination, const char* msg) {
  s->termination = termination;
  std::snprintf(s->message, sizeof(s->message), "%s", msg);
}

static int minimize(Problem& prob, const double init[7], const pcp_bfgs_options& o, double out[7], pcp_bfgs_summary* sum) {
  std::memset(sum, 0, sizeof(*sum));
  double x[7];
  for (int i = 0; i < 7; i++) out[i] = x[i] = init[i];
  double cost, g[6];
  if (!prob.eval(x, &cost, g)) {
    set_message(sum, PCP_BFGS_FAILURE, "Initial cost and gradient evaluation failed");
    sum->num_cost_evaluations = prob.evals;
    return PCP_OK;
  }
  sum->initial_cost = sum->final_cost = cost;
  double gmax = max_norm_of_gradient_step(x, g);
  if (gmax <= o.gradient_tolerance) {
    set_message(sum, PCP_BFGS_CONVERGENCE, "Initial gradient max norm below the gradient tolerance");
    sum->num_cost_evaluations = prob.evals;
    return PCP_OK;
  }
  double H[6][6];
  auto reset_H = [&] { for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) H[i][j] = i == j ? 1.0 : 0.0; };
  reset_H();
  double dir[6], prev_g[6], prev_dir[6], prev_cost = cost, prev_step = 0.0;
  bool have_prev = false, direction_ok = false;
  int restarts = 0;
  for (int it = 1;; it++) {
    if (it > o.max_num_iterations) { set_message(sum, PCP_BFGS_NO_CONVERGENCE, "Maximum number of iterations reached"); break; }
    sum->iterations = it;
    direction_ok = false;
    if (have_prev) {
      // BFGS::NextDirection: s = step * previous direction, y = g - previous g
      double sv[6], yv[6], sy = 0.0;
      for (int i = 0; i < 6; i++) { sv[i] = prev_step * prev_dir[i]; yv[i] = g[i] - prev_g[i]; sy += sv[i] * yv[i]; }
      if (sy > 1e-14) {
        double Hy[6], yHy = 0.0;
        for (int i = 0; i < 6; i++) { Hy[i] = 0.0; for (int j = 0; j < 6; j++) Hy[i] += H[i][j] * yv[j]; }
        for (int i = 0; i < 6; i++) yHy += yv[i] * Hy[i];
        const double rho = 1.0 / sy;
        for (int i = 0; i < 6; i++)
          for (int j = 0; j < 6; j++) H[i][j] += rho * ((1.0 + rho * yHy) * sv[i] * sv[j] - (sv[i] * Hy[j] + Hy[i] * sv[j]));
      }
      double gd = 0.0;
      for (int i = 0; i < 6; i++) { dir[i] = 0.0; for (int j = 0; j < 6; j++) dir[i] -= H[i][j] * g[j]; gd += dir[i] * g[i]; }
      direction_ok = gd < 0.0;
      if (!direction_ok) {
        if (++restarts > o.max_num_line_search_direction_restarts) { set_message(sum, PCP_BFGS_FAILURE, "Line search direction failure: too many restarts"); break; }
        reset_H();
      }
    }
    if (!direction_ok) for (int i = 0; i < 6; i++) dir[i] = -g[i];
    double dd = 0.0, dmax = 0.0;
    for (int i = 0; i < 6; i++) { dd += g[i] * dir[i]; dmax = std::fmax(dmax, std::fabs(dir[i])); }
    const double step0 = (it == 1 || !direction_ok) ? std::fmin(1.0, 1.0 / gmax) : std::fmin(1.0, 2.0 * (cost - prev_cost) / dd);
    if (!(step0 > 0.0)) { set_message(sum, PCP_BFGS_FAILURE, "Initial line-search step size is not positive"); break; }
    LineFn lf{&prob, x, dir};
    const LineSearchResult ls = wolfe_search(lf, step0, cost, dd, dmax, o);
    sum->num_line_search_steps += ls.iterations;
    if (!ls.success) {
      // ceres: a failed search with the BFGS direction restarts from steepest descent; with steepest descent it ends the solve
      if (direction_ok && ++restarts <= o.max_num_line_search_direction_restarts) { reset_H(); have_prev = false; it--; continue; }
      set_message(sum, PCP_BFGS_FAILURE, "Line search failed to find a step with sufficient decrease");
      break;
    }
    for (int i = 0; i < 6; i++) { prev_g[i] = g[i]; prev_dir[i] = dir[i]; }
    prev_cost = cost; prev_step = ls.point.x; have_prev = true;
    double xn = 0.0, step_norm = 0.0;
    for (int i = 0; i < 7; i++) { x[i] = ls.point.state[i]; xn += x[i] * x[i]; }
    for (int i = 0; i < 6; i++) { g[i] = ls.point.g6[i]; step_norm += prev_step * dir[i] * prev_step * dir[i]; }
    cost = ls.point.value;
    xn = std::sqrt(xn); step_norm = std::sqrt(step_norm);
    sum->final_cost = cost;
    if (o.verbose) std::fprintf(stderr, "[pcp bfgs] it %d cost %.9g step %.3g |g|inf %.3g ls %d\n", it, cost, prev_step, gmax, ls.iterations);
    gmax = max_norm_of_gradient_step(x, g);
    if (gmax <= o.gradient_tolerance) { set_message(sum, PCP_BFGS_CONVERGENCE, "Gradient tolerance reached"); break; }
    if (std::fabs(prev_cost - cost) <= o.function_tolerance * std::fabs(prev_cost)) { set_message(sum, PCP_BFGS_CONVERGENCE, "Function tolerance reached"); break; }
    if (step_norm <= o.parameter_tolerance * (xn + o.parameter_tolerance)) { set_message(sum, PCP_BFGS_CONVERGENCE, "Parameter tolerance reached"); break; }
  }
  for (int i = 0; i < 7; i++) out[i] = x[i];
  sum->num_cost_evaluations = prob.evals;
  return PCP_OK;
}

}  // namespace calib
}  // namespace pcp

// ---- C ABI -------------------------------------------------------------------------------------------------------------
extern "C" {

void pcp_bfgs_default_options(pcp_bfgs_options* o) {
  if (!o) return;
  std::memset(o, 0, sizeof(*o));
  o->max_num_iterations = 50;
  o->function_tolerance = 1e-6;
  o->gradient_tolerance = 1e-10;
  o->parameter_tolerance = 1e-8;
  o->sufficient_function_decrease = 1e-4;
  o->sufficient_curvature_decrease = 0.9;
  o->max_line_search_step_expansion = 10.0;
  o->min_line_search_step_contraction = 0.6;
  o->max_line_search_step_contraction = 1e-3;
  o->min_line_search_step_size = 1e-9;
  o->max_num_line_search_step_size_iterations = 20;
  o->max_num_line_search_direction_restarts = 5;
  o->verbose = 0;
}

int pcp_se3_bfgs(pcp_se3_cost_fn fn, void* user, const double init_se3[7], const pcp_bfgs_options* options, double out_se3[7],
                 pcp_bfgs_summary* summary) {
  if (!fn || !init_se3 || !out_se3) return PCP_EINVAL;
  pcp_bfgs_options o;
  if (options) o = *options; else pcp_bfgs_default_options(&o);
  pcp_bfgs_summary local;
  pcp::calib::Problem prob{fn, user};
  return pcp::calib::minimize(prob, init_se3, o, out_se3, summary ? summary : &local);
}

void pcp_calib_default_params(pcp_calib_params* p) {
  if (!p) return;
  std::memset(p, 0, sizeof(*p));
  p->max_outer_iterations = 10;             // include/vlcal/calib/visual_camera_calibration.hpp:17
  p->delta_trans_thresh = 0.01;             // :23
  p->delta_rot_thresh = 1.0 * M_PI / 180.0; // :24
  pcp_bfgs_default_options(&p->inner);
}

namespace {
struct MultiNidUser {
  pcp_nid* const* costs;
  int n_costs;
  const double* init;  // the pose gate's centre (MultiNIDCost::init_T_camera_lidar)
  int rc = PCP_OK;
};
int multi_nid_cb(void* user, const double se3[7], double* value, double grad7[7]) {
  MultiNidUser* u = static_cast<MultiNidUser*>(user);
  uint8_t ok = 0;
  const int rc = pcp_nid_eval_multi_grad(u->costs, u->n_costs, u->init, se3, 1, value, grad7, &ok);
  if (rc != PCP_OK) { u->rc = rc; return 0; }
  return ok ? 1 : 0;
}
}  // namespace

int pcp_calibrate(pcp_nid* const* costs, int n_costs, const double init_T_camera_lidar[16], const pcp_calib_params* params,
                  double out_T_camera_lidar[16], pcp_calib_summary* summary) {
  using namespace pcp::calib;
  if (!costs || n_costs <= 0 || !init_T_camera_lidar || !out_T_camera_lidar) return PCP_EINVAL;
  pcp_calib_params p;
  if (params) p = *params; else pcp_calib_default_params(&p);
  pcp_calib_summary local;
  pcp_calib_summary* s = summary ? summary : &local;
  std::memset(s, 0, sizeof(*s));
  double T[7];
  se3_from_matrix(init_T_camera_lidar, T);
  for (int i = 0; i < p.max_outer_iterations; i++) {  // visual_camera_calibration.cpp:53-81
    double Tn[7];
    MultiNidUser user{costs, n_costs, T};            // estimate_pose_bfgs: the gate is centred on this iteration's start (:201)
    Problem prob{multi_nid_cb, &user};
    pcp_bfgs_summary inner;
    minimize(prob, T, p.inner, Tn, &inner);
    if (user.rc != PCP_OK) return user.rc;
    if (i == 0) s->initial_cost = inner.initial_cost;
    s->final_cost = inner.final_cost;
    s->inner_iterations += inner.iterations;
    s->cost_evaluations += inner.num_cost_evaluations;
    s->last_inner_termination = inner.termination;
    s->outer_iterations = i + 1;
    double inv[7], delta[7];
    se3_inverse(Tn, inv);
    se3_mul(inv, T, delta);                           // new_T.inverse() * T  (:67)
    for (int k = 0; k < 7; k++) T[k] = Tn[k];
    s->delta_t = std::sqrt(delta[4] * delta[4] + delta[5] * delta[5] + delta[6] * delta[6]);
    s->delta_r = quat_angle(delta);
    const bool converged = s->delta_t < p.delta_trans_thresh && s->delta_r < p.delta_rot_thresh;  // :72
    if (p.inner.verbose)
      std::fprintf(stderr, "[pcp calibrate] outer %d: cost %.9g -> %.9g, delta_t %.4f m delta_r %.6f rad, %s (%s)\n", i, inner.initial_cost,
                   inner.final_cost, s->delta_t, s->delta_r, converged ? "converged" : "not converged", inner.message);
    if (converged) { s->converged = 1; break; }
  }
  se3_to_matrix(T, out_T_camera_lidar);
  return PCP_OK;
}

}  // extern "C"
