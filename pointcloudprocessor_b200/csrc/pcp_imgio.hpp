// pcp_imgio.hpp — host-side decoding of the lossless image formats at the boundary: binary PGM / PPM and PNG (every colour type and bit
// depth, Adam7; zlib's inflate is the only library call).  JPEG goes through nvJPEG on the device (pcp_jpeg.cu).  The reference reads everything
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

// Every PNG cv::imread(path) / cv::imread(path, IMREAD_GRAYSCALE) reads: colour types 0 / 2 / 3 / 4 / 6, bit depths 1 / 2 / 4 / 8 / 16,
// non-interlaced and Adam7.  The conversions to 8 bits are libpng's as OpenCV 4.x configures it (pinned against cv2 4.13,
// tests/test_io_cpu.py): grey samples of 1 / 2 / 4 bits are expanded by bit replication (x255, x85, x17), palette entries become
// R,G,B, alpha (and tRNS) is dropped, 16-bit samples are chopped to their high byte (png_set_strip_16) — after the colour -> grey
// conversion when IMREAD_GRAYSCALE reads a 16-bit colour file (libpng converts first, with rounding).
// `want_grey`: the caller will ask for one channel; only matters for 16-bit colour files (see above).
inline bool decode_png(const std::vector<uint8_t>& f, HostImage& im, bool want_grey = false) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (f.size() < 33 || std::memcmp(f.data(), sig, 8) != 0) return false;
  size_t p = 8;
  uint32_t w = 0, h = 0;
  int depth = 0, ctype = -1, interlace = 0;
  std::vector<uint8_t> idat, plte;
  while (p + 12 <= f.size()) {
    const uint32_t len = be32(&f[p]);
    const char* type = reinterpret_cast<const char*>(&f[p + 4]);
    if (p + 12 + (size_t)len > f.size()) return false;
    const uint8_t* d = &f[p + 8];
    if (!std::memcmp(type, "IHDR", 4)) {
      if (len < 13) return false;
      w = be32(d); h = be32(d + 4); depth = d[8]; ctype = d[9]; interlace = d[12];
    } else if (!std::memcmp(type, "PLTE", 4)) plte.assign(d, d + len);
    else if (!std::memcmp(type, "IDAT", 4)) idat.insert(idat.end(), d, d + len);
    else if (!std::memcmp(type, "IEND", 4)) break;
    p += 12 + len;
  }
  if (w == 0 || h == 0 || w > (1u << 20) || h > (1u << 20) || interlace > 1 || idat.empty()) return false;
  int spp;  // samples per pixel
  switch (ctype) { case 0: spp = 1; break; case 2: spp = 3; break; case 3: spp = 1; break; case 4: spp = 2; break; case 6: spp = 4; break; default: return false; }
  const bool depth_ok = (ctype == 0) ? (depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)
                      : (ctype == 3) ? (depth == 1 || depth == 2 || depth == 4 || depth == 8)
                                     : (depth == 8 || depth == 16);
  if (!depth_ok) return false;
  if (ctype == 3 && plte.size() < 3) return false;
  const int bits_pp = spp * depth;               // bits per pixel
  const size_t fbpp = bits_pp >= 8 ? (size_t)bits_pp / 8 : 1;  // the filter's "bytes per pixel"
  // passes: {x0, y0, dx, dy}; one pass for a non-interlaced file, Adam7 otherwise
  static const int adam7[7][4] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4}, {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
  static const int whole[1][4] = {{0, 0, 1, 1}};
  const int (*pass)[4] = interlace ? adam7 : whole;
  const int n_pass = interlace ? 7 : 1;
  size_t raw_size = 0;
  for (int k = 0; k < n_pass; k++) {
    const size_t pw = ((size_t)w > (size_t)pass[k][0]) ? (w - pass[k][0] + pass[k][2] - 1) / pass[k][2] : 0;
    const size_t ph = ((size_t)h > (size_t)pass[k][1]) ? (h - pass[k][1] + pass[k][3] - 1) / pass[k][3] : 0;
    if (pw && ph) raw_size += (1 + (pw * bits_pp + 7) / 8) * ph;
  }
  std::vector<uint8_t> raw(raw_size);
  uLongf out_len = (uLongf)raw.size();
  if (uncompress(raw.data(), &out_len, idat.data(), (uLong)idat.size()) != Z_OK || out_len != raw.size()) return false;
  // samples of the whole image as 16-bit values, spp per pixel
  std::vector<uint16_t> smp((size_t)w * h * spp);
  size_t off = 0;
  std::vector<uint8_t> prev, cur;
  for (int k = 0; k < n_pass; k++) {
    const size_t pw = ((size_t)w > (size_t)pass[k][0]) ? (w - pass[k][0] + pass[k][2] - 1) / pass[k][2] : 0;
    const size_t ph = ((size_t)h > (size_t)pass[k][1]) ? (h - pass[k][1] + pass[k][3] - 1) / pass[k][3] : 0;
    if (!pw || !ph) continue;
    const size_t rb = (pw * bits_pp + 7) / 8;
    prev.assign(rb, 0);
    cur.resize(rb);
    for (size_t y = 0; y < ph; y++) {
      const uint8_t ft = raw[off];
      const uint8_t* in = &raw[off + 1];
      off += rb + 1;
      for (size_t i = 0; i < rb; i++) {
        const int a = i >= fbpp ? cur[i - fbpp] : 0, b = prev[i], c = i >= fbpp ? prev[i - fbpp] : 0;
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
      const size_t Y = (size_t)pass[k][1] + y * pass[k][3];
      for (size_t x = 0; x < pw; x++) {
        const size_t X = (size_t)pass[k][0] + x * pass[k][2];
        uint16_t* o = &smp[(Y * w + X) * spp];
        for (int c = 0; c < spp; c++) {
          const size_t si = x * spp + c;  // sample index in the row
          if (depth == 8) o[c] = cur[si];
          else if (depth == 16) o[c] = (uint16_t)((cur[2 * si] << 8) | cur[2 * si + 1]);
          else { const size_t bit = si * depth; o[c] = (uint16_t)((cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1)); }
        }
      }
      prev.swap(cur);
      cur.resize(rb);
    }
  }
  im.width = (int)w; im.height = (int)h;
  im.from_png = true;
  const size_t n = (size_t)w * h;
  auto to8 = [&](uint16_t v) -> uint8_t {  // a grey / colour sample to 8 bits
    switch (depth) { case 1: return v ? 255 : 0; case 2: return (uint8_t)(v * 85); case 4: return (uint8_t)(v * 17); case 16: return (uint8_t)(v >> 8); default: return (uint8_t)v; }
  };
  if (ctype == 0 || ctype == 4) {
    im.channels = 1;
    im.data.resize(n);
    for (size_t i = 0; i < n; i++) im.data[i] = to8(smp[i * spp]);
  } else if (ctype == 3) {
    im.channels = 3;
    im.data.resize(n * 3);
    for (size_t i = 0; i < n; i++) {
      const size_t k = (size_t)smp[i] * 3;
      if (k + 2 >= plte.size()) return false;
      im.data[3 * i] = plte[k + 2]; im.data[3 * i + 1] = plte[k + 1]; im.data[3 * i + 2] = plte[k];
    }
  } else if (depth == 16 && want_grey) {
    // libpng converts to grey in 16 bits before the samples are chopped: (R*9797 + G*19234 + B*3737 + 16384) >> 15 (its 16-bit path
    // rounds, its 8-bit path truncates), then the high byte
    im.channels = 1;
    im.data.resize(n);
    for (size_t i = 0; i < n; i++) {
      const uint64_t g16 = ((uint64_t)smp[i * spp] * 9797 + (uint64_t)smp[i * spp + 1] * 19234 + (uint64_t)smp[i * spp + 2] * 3737 + 16384) >> 15;
      im.data[i] = (uint8_t)(g16 >> 8);
    }
  } else {
    im.channels = 3;
    im.data.resize(n * 3);
    for (size_t i = 0; i < n; i++) {
      im.data[3 * i] = to8(smp[i * spp + 2]); im.data[3 * i + 1] = to8(smp[i * spp + 1]); im.data[3 * i + 2] = to8(smp[i * spp]);
    }
  }
  return true;
}

}  // namespace io
}  // namespace pcp
