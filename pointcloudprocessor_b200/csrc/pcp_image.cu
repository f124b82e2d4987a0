// pcp_image.cu — generateColorMap's image preparation on the device: the 8-bit BGR -> HSV -> (S, V scaled) -> BGR round trip
// of src/PointCloudProcessor.cpp:722-741, bit-exact against OpenCV 4.x (cv2 4.13.0: all 2^24 inputs, tests/).
//
// The forward conversion is OpenCV's integer RGB2HSV_b (hsv_shift = 12, hrange = 180).  The backward conversion is float and
// OpenCV evaluates it with two different formulas: the vector body of a row (pixels x < floor(W / block) * block) truncates and
// has no s == 0 shortcut; the scalar tail uses fmod / floor and rounds to nearest even.  Both are reproduced; `block` is the
// number of pixels per vector iteration of the OpenCV build to match (32 for the AVX2 build measured, 0 = whole row is body).
// One thread per pixel, in place; HBM-bound (3 B read + 3 B written per pixel).
#include "pcp_internal.cuh"

namespace pcp {

__global__ void __launch_bounds__(256) hsv_roundtrip_kernel(const uint8_t* __restrict__ src, size_t spitch, uint8_t* __restrict__ dst, size_t dpitch,
                                                            int W, int H, float sat_scale, float val_scale, int body) {
  __shared__ int sdiv[256], hdiv[256];
  {
    const int i = threadIdx.x;  // blockDim.x == 256
    // cv::saturate_cast<int>(double) = cvRound: round to nearest even
    sdiv[i] = i ? __double2int_rn(__ddiv_rn((double)(255 << 12), (double)i)) : 0;
    hdiv[i] = i ? __double2int_rn(__ddiv_rn((double)(180 << 12), __dmul_rn(6.0, (double)i))) : 0;
  }
  __syncthreads();
  const size_t total = (size_t)W * H;
  for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < total; p += (size_t)gridDim.x * blockDim.x) {
    const int y = (int)(p / W), x = (int)(p - (size_t)y * W);
    const uint8_t* sp = src + (size_t)y * spitch + (size_t)x * 3;
    const int b = sp[0], g = sp[1], r = sp[2];
    const int v = max(b, max(g, r)), vmin = min(b, min(g, r));
    const int diff = v - vmin;
    const int sI = (diff * sdiv[v] + (1 << 11)) >> 12;
    int hI = v == r ? (g - b) : (v == g ? (b - r + 2 * diff) : (r - g + 4 * diff));
    hI = (hI * hdiv[diff] + (1 << 11)) >> 12;
    if (hI < 0) hI += 180;
    const int h8 = min(max(hI, 0), 255);
    // hsv[1] = saturate_cast<uchar>(hsv[1] * saturation_scale), hsv[2] likewise (cpp:733-734): float product, cvRound, clamp
    const int s8 = min(max(__float2int_rn(__fmul_rn((float)(sI & 255), sat_scale)), 0), 255);
    const int v8 = min(max(__float2int_rn(__fmul_rn((float)v, val_scale)), 0), 255);
    const float sf = __fmul_rn((float)s8, 1.0f / 255.0f), vf = __fmul_rn((float)v8, 1.0f / 255.0f);
    float h = __fmul_rn((float)h8, 6.0f / 180.0f), f;
    int sector;
    const bool in_body = x < body;
    if (in_body) {
      const float pre = truncf(h);
      f = __fsub_rn(h, pre);
      sector = (int)pre % 6;
    } else {
      h = fmodf(h, 6.0f);
      sector = (int)floorf(h);
      f = __fsub_rn(h, (float)sector);
      if ((unsigned)sector >= 6u) { sector = 0; f = 0.0f; }
    }
    const float t0 = vf;
    const float t1 = __fmul_rn(vf, __fsub_rn(1.0f, sf));
    const float t2 = __fmul_rn(vf, __fmaf_rn(-sf, f, 1.0f));
    const float t3 = __fmul_rn(vf, __fmaf_rn(-sf, __fsub_rn(1.0f, f), 1.0f));
    // sector_data = {{1,3,0},{1,0,2},{3,0,1},{0,2,1},{0,1,3},{2,1,0}} (b, g, r pick tab[.])
    float cb, cg, cr;
    switch (sector) {
      case 0: cb = t1; cg = t3; cr = t0; break;
      case 1: cb = t1; cg = t0; cr = t2; break;
      case 2: cb = t3; cg = t0; cr = t1; break;
      case 3: cb = t0; cg = t2; cr = t1; break;
      case 4: cb = t0; cg = t1; cr = t3; break;
      default: cb = t2; cg = t1; cr = t0; break;
    }
    int ob, og, orr;
    if (in_body) {
      ob = __float2int_rz(__fmul_rn(cb, 255.0f)); og = __float2int_rz(__fmul_rn(cg, 255.0f)); orr = __float2int_rz(__fmul_rn(cr, 255.0f));
    } else {
      if (sf == 0.0f) cb = cg = cr = vf;
      ob = __float2int_rn(__fmul_rn(cb, 255.0f)); og = __float2int_rn(__fmul_rn(cg, 255.0f)); orr = __float2int_rn(__fmul_rn(cr, 255.0f));
    }
    uint8_t* dp = dst + (size_t)y * dpitch + (size_t)x * 3;
    dp[0] = (uint8_t)min(max(ob, 0), 255);
    dp[1] = (uint8_t)min(max(og, 0), 255);
    dp[2] = (uint8_t)min(max(orr, 0), 255);
  }
}

void launch_hsv_roundtrip(const uint8_t* src, size_t spitch, uint8_t* dst, size_t dpitch, int W, int H, float sat_scale, float val_scale, int block,
                          int sm_count, cudaStream_t s) {
  if (W <= 0 || H <= 0) return;
  const int body = block > 0 ? (W / block) * block : W;
  const size_t total = (size_t)W * H;
  const int want = (int)((total + 255) / 256);
  const int grid = want < sm_count * 8 ? want : sm_count * 8;
  hsv_roundtrip_kernel<<<grid, 256, 0, s>>>(src, spitch, dst, dpitch, W, H, sat_scale, val_scale, body);
}

}  // namespace pcp
