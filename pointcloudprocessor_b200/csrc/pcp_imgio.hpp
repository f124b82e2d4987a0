// pcp_imgio.hpp — host-side decoding of the lossless image formats at the boundary: binary PGM / PPM and PNG (8-bit, non-interlaced;
// zlib's inflate is the only library call).  JPEG goes through nvJPEG on the device (pcp_jpeg.cu).  The reference reads everything
// with cv::imread (src/PointCloudProcessor.cpp:716, :774); OpenCV is not installed here.
#pragma once
#include <zlib.h>

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace pcp {
namespace io {

struct HostImage {
  int width = 0, height = 0, channels = 0;  // channels: 1 (grey) or 3 (B, G, R interleaved, OpenCV's order)
  bool from_png = false;                    // selects the colour -> grey rule of cv::imread(IMREAD_GRAYSCALE), see to_grey
  std::vector<uint8_t> data;
};

// cv::imread(path, IMREAD_GRAYSCALE) on a colour file, measured against cv2 4.13 (tests/test_io_cpu.py):
//   PPM (and the fallback for JPEG): OpenCV's own 8-bit BGR2GRAY, (R*4899 + G*9617 + B*1868 + 8192) >> 14;
//   PNG: the conversion happens inside libpng (png_set_rgb_to_gray), (R*9797 + G*19234 + B*3737) >> 15, truncating.
inline uint8_t grey_of(uint8_t r, uint8_t g, uint8_t b) { return (uint8_t)((r * 4899 + g * 9617 + b * 1868 + 8192) >> 14); }
inline uint8_t grey_of_png(uint8_t r, uint8_t g, uint8_t b) { return (uint8_t)((r * 9797 + g * 19234 + b * 3737) >> 15); }

inline void to_grey(HostImage& im) {
  if (im.channels == 1) return;
  std::vector<uint8_t> g((size_t)im.width * im.height);
  for (size_t i = 0; i < g.size(); i++)
    g[i] = im.from_png ? grey_of_png(im.data[3 * i + 2], im.data[3 * i + 1], im.data[3 * i]) : grey_of(im.data[3 * i + 2], im.data[3 * i + 1], im.data[3 * i]);
  im.data.swap(g);
  im.channels = 1;
}
inline void to_bgr(HostImage& im) {
  if (im.channels == 3) return;
  std::vector<uint8_t> c((size_t)im.width * im.height * 3);
  for (size_t i = 0; i < (size_t)im.width * im.height; i++) c[3 * i] = c[3 * i + 1] = c[3 * i + 2] = im.data[i];
  im.data.swap(c);
  im.channels = 3;
}

// ---- binary PGM (P5) / PPM (P6), maxval 255 -------------------------------------------------------------------------
inline bool decode_pnm(const std::vector<uint8_t>& f, HostImage& im) {
  if (f.size() < 7 || f[0] != 'P' || (f[1] != '5' && f[1] != '6')) return false;
  size_t p = 2;
  auto next_int = [&](int& v) -> bool {
    for (;;) {
      while (p < f.size() && (f[p] == ' ' || f[p] == '\n' || f[p] == '\r' || f[p] == '\t')) p++;
      if (p < f.size() && f[p] == '#') { while (p < f.size() && f[p] != '\n') p++; continue; }
      break;
    }
    if (p >= f.size() || f[p] < '0' || f[p] > '9') return false;
    long x = 0;
    while (p < f.size() && f[p] >= '0' && f[p] <= '9') { x = x * 10 + (f[p] - '0'); p++; if (x > (1 << 30)) return false; }
    v = (int)x;
    return true;
  };
  int w, h, maxv;
  if (!next_int(w) || !next_int(h) || !next_int(maxv) || maxv != 255 || w <= 0 || h <= 0) return false;
  p++;  // the single whitespace after maxval
  const int ch = f[1] == '6' ? 3 : 1;
  const size_t need = (size_t)w * h * ch;
  if (f.size() < p + need) return false;
  im.width = w; im.height = h; im.channels = ch;
  im.data.assign(f.begin() + (long)p, f.begin() + (long)(p + need));
  if (ch == 3)  // PPM is R, G, B: store B, G, R
    for (size_t i = 0; i < (size_t)w * h; i++) { const uint8_t t = im.data[3 * i]; im.data[3 * i] = im.data[3 * i + 2]; im.data[3 * i + 2] = t; }
  return true;
}

// ---- PNG ------------------------------------------------------------------------------------------------------------
inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline bool decode_png(const std::vector<uint8_t>& f, HostImage& im) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (f.size() < 33 || std::memcmp(f.data(), sig, 8) != 0) return false;
  size_t p = 8;
  uint32_t w = 0, h = 0;
  int depth = 0, ctype = -1, interlace = 0;
  std::vector<uint8_t> idat, plte;
  while (p + 12 <= f.size()) {
    const uint32_t len = be32(&f[p]);
    const char* type = reinterpret_cast<const char*>(&f[p + 4]);
    if (p + 12 + len > f.size()) return false;
    const uint8_t* d = &f[p + 8];
    if (!std::memcmp(type, "IHDR", 4)) {
      if (len < 13) return false;
      w = be32(d); h = be32(d + 4); depth = d[8]; ctype = d[9]; interlace = d[12];
    } else if (!std::memcmp(type, "PLTE", 4)) plte.assign(d, d + len);
    else if (!std::memcmp(type, "IDAT", 4)) idat.insert(idat.end(), d, d + len);
    else if (!std::memcmp(type, "IEND", 4)) break;
    p += 12 + len;
  }
  if (w == 0 || h == 0 || w > (1u << 20) || h > (1u << 20) || depth != 8 || interlace != 0 || idat.empty()) return false;
  int spp;  // samples per pixel
  switch (ctype) { case 0: spp = 1; break; case 2: spp = 3; break; case 3: spp = 1; break; case 4: spp = 2; break; case 6: spp = 4; break; default: return false; }
  if (ctype == 3 && plte.size() < 3) return false;
  const size_t stride = (size_t)w * spp;
  std::vector<uint8_t> raw((stride + 1) * h);
  uLongf out_len = (uLongf)raw.size();
  if (uncompress(raw.data(), &out_len, idat.data(), (uLong)idat.size()) != Z_OK || out_len != raw.size()) return false;
  std::vector<uint8_t> pix(stride * h);
  for (uint32_t y = 0; y < h; y++) {
    const uint8_t ft = raw[(stride + 1) * y];
    const uint8_t* in = &raw[(stride + 1) * y + 1];
    uint8_t* cur = &pix[stride * y];
    const uint8_t* up = y ? &pix[stride * (y - 1)] : nullptr;
    for (size_t i = 0; i < stride; i++) {
      const int a = i >= (size_t)spp ? cur[i - spp] : 0, b = up ? up[i] : 0, c = (up && i >= (size_t)spp) ? up[i - spp] : 0;
      int pred = 0;
      switch (ft) {
        case 0: pred = 0; break;
        case 1: pred = a; break;
        case 2: pred = b; break;
        case 3: pred = (a + b) >> 1; break;
        case 4: { const int pp = a + b - c, pa = std::abs(pp - a), pb = std::abs(pp - b), pc = std::abs(pp - c);
                  pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
        default: return false;
      }
      cur[i] = (uint8_t)(in[i] + pred);
    }
  }
  im.width = (int)w; im.height = (int)h;
  im.from_png = true;
  const size_t n = (size_t)w * h;
  if (ctype == 0 || ctype == 4) {
    im.channels = 1;
    im.data.resize(n);
    for (size_t i = 0; i < n; i++) im.data[i] = pix[i * spp];
  } else {
    im.channels = 3;
    im.data.resize(n * 3);
    for (size_t i = 0; i < n; i++) {
      uint8_t r, g, b;
      if (ctype == 3) { const size_t k = (size_t)pix[i] * 3; if (k + 2 >= plte.size()) return false; r = plte[k]; g = plte[k + 1]; b = plte[k + 2]; }
      else { r = pix[i * spp]; g = pix[i * spp + 1]; b = pix[i * spp + 2]; }
      im.data[3 * i] = b; im.data[3 * i + 1] = g; im.data[3 * i + 2] = r;
    }
  }
  return true;
}

}  // namespace io
}  // namespace pcp
