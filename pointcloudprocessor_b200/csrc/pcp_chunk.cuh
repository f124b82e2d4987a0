// pcp_chunk.cuh — conservative per-chunk classification for the pre-binned scan (one lane classifies one chunk).
//
// A chunk is kChunk Morton-consecutive points with an axis-aligned bounding box.  The box is pushed through the float
// world-to-camera transform (centre + per-axis radius, with a rounding margin that covers both this evaluation and the
// per-point evaluation in PCL's order) and, when it lies entirely in front of the camera, through an INTERVAL evaluation
// of the plumb_bob projection.  The result is only ever used to skip work whose outcome is certain for every point of
// the chunk; anything else is processed point by point by the exact machinery.
#pragma once
#include "pcp_internal.cuh"

namespace pcp {

enum { kChunkSkip = 0, kChunkProcess = 1, kChunkAllCandidates = 2 };

struct Ival { float lo, hi; };

__device__ __forceinline__ Ival iv_add(Ival a, Ival b) { return Ival{a.lo + b.lo, a.hi + b.hi}; }
__device__ __forceinline__ Ival iv_scale(float k, Ival a) { return k >= 0.0f ? Ival{k * a.lo, k * a.hi} : Ival{k * a.hi, k * a.lo}; }
__device__ __forceinline__ Ival iv_mul(Ival a, Ival b) {
  const float p0 = a.lo * b.lo, p1 = a.lo * b.hi, p2 = a.hi * b.lo, p3 = a.hi * b.hi;
  return Ival{fminf(fminf(p0, p1), fminf(p2, p3)), fmaxf(fmaxf(p0, p1), fmaxf(p2, p3))};
}
__device__ __forceinline__ Ival iv_sq(Ival a) {
  const float l2 = a.lo * a.lo, h2 = a.hi * a.hi;
  if (a.lo >= 0.0f) return Ival{l2, h2};
  if (a.hi <= 0.0f) return Ival{h2, l2};
  return Ival{0.0f, fmaxf(l2, h2)};
}
__device__ __forceinline__ float iv_absmax(Ival a) { return fmaxf(fabsf(a.lo), fabsf(a.hi)); }

// What the scan kernel needs to test a whole chunk (and then its points one by one) against the z-grid BEFORE projecting
// anything: the rectangle of z-grid cells the chunk can fall into and a lower bound of the squared distance of its points.
// A point whose squared distance exceeds (min + tol)^2 of EVERY cell of the rectangle is rejected by the reference's second
// pass whatever its own cell is (view_culling.cpp:157; cell minima only decrease), and it cannot lower a minimum itself.
struct ChunkInfo {
  uint32_t rect;   // 0: no test possible; else w | h << 6 | point count << 12   (w * h <= 32)
  uint32_t org;    // cx0 | cy0 << 16
  float d2lo;      // lower bound of ||p||^2 over the chunk
};

// a0 = {lo.xyz, nonfinite flag}, a1 = {hi.xyz, count}.  `count_candidates`: the caller needs n_candidates exact.
__device__ __forceinline__ int classify_chunk(const FrameParams& fp, float4 a0, float4 a1, bool count_candidates, ChunkInfo& info) {
  info.rect = 0u; info.org = 0u; info.d2lo = 0.0f;
  if (a0.w != 0.0f) return kChunkProcess;  // holds a non-finite coordinate
  const float eps = 5.9604644775390625e-08f;
  const float cx = 0.5f * (a0.x + a1.x), cy = 0.5f * (a0.y + a1.y), cz = 0.5f * (a0.z + a1.z);
  // half extents, inflated so that [c - h, c + h] contains the box despite the rounding of c and h
  const float hx = 0.5f * (a1.x - a0.x) + 4.0f * eps * (fabsf(a0.x) + fabsf(a1.x)) + 1e-30f;
  const float hy = 0.5f * (a1.y - a0.y) + 4.0f * eps * (fabsf(a0.y) + fabsf(a1.y)) + 1e-30f;
  const float hz = 0.5f * (a1.z - a0.z) + 4.0f * eps * (fabsf(a0.z) + fabsf(a1.z)) + 1e-30f;
  float cc[3], rr[3];
  if (fp.use_w2c) {
#pragma unroll
    for (int r = 0; r < 3; r++) {
      const float m0 = fp.w2c.r[4 * r], m1 = fp.w2c.r[4 * r + 1], m2 = fp.w2c.r[4 * r + 2], m3 = fp.w2c.r[4 * r + 3];
      cc[r] = fmaf(m0, cx, fmaf(m1, cy, fmaf(m2, cz, m3)));
      const float rad = fabsf(m0) * hx + fabsf(m1) * hy + fabsf(m2) * hz;
      const float mag = fabsf(m0) * fabsf(cx) + fabsf(m1) * fabsf(cy) + fabsf(m2) * fabsf(cz) + fabsf(m3) + rad;
      rr[r] = rad + 16.0f * eps * mag;  // this evaluation and the per-point one (PCL order) each err by < 4 eps * mag
    }
  } else {
    cc[0] = cx; cc[1] = cy; cc[2] = cz; rr[0] = hx; rr[1] = hy; rr[2] = hz;
  }
  const float zlo = cc[2] - rr[2], zhi = cc[2] + rr[2];
  if (zhi < 0.0f) return kChunkSkip;         // every point has z < 0: rejected at view_culling.cpp:76, not even a candidate
  if (!(zlo > 1e-30f)) return kChunkProcess;  // touches the camera plane (or NaN)
  // normalised coordinates x/z, y/z over the box (z > 0): interval division
  const float izl = (1.0f / zhi) * (1.0f - 4.0f * eps), izh = (1.0f / zlo) * (1.0f + 4.0f * eps);
  const float xl = cc[0] - rr[0], xh = cc[0] + rr[0], yl = cc[1] - rr[1], yh = cc[1] + rr[1];
  const Ival X{xl < 0.0f ? xl * izh : xl * izl, xh < 0.0f ? xh * izl : xh * izh};
  const Ival Y{yl < 0.0f ? yl * izh : yl * izl, yh < 0.0f ? yh * izl : yh * izh};
  const PreF& c = fp.pre;
  const Ival x2 = iv_sq(X), y2 = iv_sq(Y);
  const Ival r2 = iv_add(x2, y2);                                   // >= 0
  const Ival r4{r2.lo * r2.lo, r2.hi * r2.hi}, r6{r4.lo * r2.lo, r4.hi * r2.hi};
  Ival rc = iv_add(iv_add(iv_scale(c.k1, r2), iv_scale(c.k2, r4)), iv_scale(c.k3, r6));
  rc.lo += 1.0f; rc.hi += 1.0f;
  const Ival t1 = iv_scale(2.0f, iv_mul(X, Y));
  const Ival t2 = iv_add(r2, iv_scale(2.0f, x2)), t3 = iv_add(r2, iv_scale(2.0f, y2));
  const Ival xd = iv_add(iv_add(iv_mul(rc, X), iv_scale(c.p1, t1)), iv_scale(c.p2, t2));
  const Ival yd = iv_add(iv_add(iv_mul(rc, Y), iv_scale(c.p1, t3)), iv_scale(c.p2, t1));
  Ival u = iv_scale(c.fx, xd), v = iv_scale(c.fy, yd);
  u.lo += c.cx; u.hi += c.cx; v.lo += c.cy; v.hi += c.cy;
  // rounding of this interval evaluation (round-to-nearest, ~25 operations): relative to the magnitude polynomial
  const float arc = fmaf(r2.hi, fmaf(r2.hi, fmaf(r2.hi, c.ak3, c.ak2), c.ak1), 1.0f);
  const float m = fmaf(arc, fmaxf(iv_absmax(X), iv_absmax(Y)), c.apt * r2.hi);
  const float e = fmaf(fmaf(c.afmax, m, c.acmax), 4.0f * c.err, 0.5f);  // 256 eps * magnitude + half a pixel
  const float ulo = u.lo - e, uhi = u.hi + e, vlo = v.lo - e, vhi = v.hi + e;
  const float u_hi_grid = -c.lo * c.wc, v_hi_grid = -c.lo * c.hc;      // ds * Wc, ds * Hc
  const bool out_grid = (uhi <= c.lo) || (vhi <= c.lo) || (ulo >= u_hi_grid) || (vlo >= v_hi_grid);
  const bool none_cand = (uhi <= c.lo) || (vhi <= c.lo) || (ulo >= c.u_hi_cand) || (vlo >= c.v_hi_cand);
  const bool all_cand = (ulo > c.lo) && (vlo > c.lo) && (uhi < c.u_hi_cand) && (vhi < c.v_hi_cand);
  if (none_cand) return kChunkSkip;  // no point passes the gate of :90, whatever the mode
  if (!fp.depth_culling) return kChunkProcess;  // every candidate is kept (:135)
  if (!out_grid) {
    // Cell rectangle (clamped to the grid: points outside the grid are never kept, :162-165) and distance bound.  The 0.5 px
    // of `e` dwarfs the rounding of the reference's float division by ds.  When n_candidates must stay exact the test is
    // only offered for chunks whose points are all certain candidates (the scan kernel then counts the points it drops).
    if (c.tol_up < 3.0e38f && (!count_candidates || all_cand)) {
      const int x0 = max((int)floorf(ulo * c.inv_ds), 0), x1 = min((int)floorf(uhi * c.inv_ds), fp.Wc - 1);
      const int y0 = max((int)floorf(vlo * c.inv_ds), 0), y1 = min((int)floorf(vhi * c.inv_ds), fp.Hc - 1);
      const int w = x1 - x0 + 1, h = y1 - y0 + 1;
      if (w >= 1 && h >= 1 && w * h <= 32) {
        const float dx = fmaxf(fabsf(cc[0]) - rr[0], 0.0f), dy = fmaxf(fabsf(cc[1]) - rr[1], 0.0f), dz = fmaxf(fabsf(cc[2]) - rr[2], 0.0f);
        info.rect = (uint32_t)w | ((uint32_t)h << 6) | ((uint32_t)a1.w << 12);
        info.org = (uint32_t)x0 | ((uint32_t)y0 << 16);
        info.d2lo = fmaf(dz, dz, fmaf(dy, dy, dx * dx)) * c.dscale;
      }
    }
    return kChunkProcess;
  }
  // certainly outside the z-grid: never kept by pass 2 (:162-165)
  if (!count_candidates) return kChunkSkip;
  return all_cand ? kChunkAllCandidates : kChunkProcess;
}

}  // namespace pcp
