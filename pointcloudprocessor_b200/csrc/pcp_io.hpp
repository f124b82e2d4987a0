// pcp_io.hpp — host-side file formats at the boundary of PointCloudProcessor::process() (SURVEY.md §8b "File formats"):
// PCD read (ascii / binary / binary_compressed) and ASCII write, the odometry text file, PGM/PPM/PNG images.
//
// The reference does all of this through PCL and OpenCV, neither of which is vendored or installed here; the layouts below
// follow the PCD v0.7 format as written by PCL 1.10 (pcl::PCDWriter::writeASCII, precision 8) — third-party behaviour,
// parity unpinned.  Reference call sites are cited per function, relative to
// /root/reference/PointCloudProcessor/.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace pcp {
namespace io {

struct XYZI { float x, y, z, intensity; };

// std::to_string(double) of the reference's file names (src/PointCloudProcessor.cpp:981, :991, :213): "%f".
inline std::string to_string_f(double v) {
  char buf[64];
  std::snprintf(buf, sizeof(buf), "%f", v);
  return buf;
}

// ---------------------------------------------------------------------------------------------------------------------
// LZF decompression (PCD binary_compressed; the format of Marc Lehmann's liblzf that PCL embeds).
// ---------------------------------------------------------------------------------------------------------------------
inline bool lzf_decompress(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len) {
  size_t ip = 0, op = 0;
  while (ip < in_len) {
    unsigned ctrl = in[ip++];
    if (ctrl < 32) {  // literal run of ctrl + 1 bytes
      ctrl++;
      if (op + ctrl > out_len || ip + ctrl > in_len) return false;
      std::memcpy(out + op, in + ip, ctrl);
      op += ctrl; ip += ctrl;
    } else {  // back reference
      unsigned len = ctrl >> 5;
      if (ip >= in_len) return false;
      if (len == 7) { len += in[ip++]; if (ip >= in_len) return false; }
      const size_t off = ((size_t)(ctrl & 0x1f) << 8) + in[ip++] + 1;
      len += 2;
      if (off > op || op + len > out_len) return false;
      for (unsigned i = 0; i < len; i++, op++) out[op] = out[op - off];  // may overlap
    }
  }
  return op == out_len;
}

// ---------------------------------------------------------------------------------------------------------------------
// pcl::io::loadPCDFile<pcl::PointXYZI> (src/PointCloudProcessor.cpp:112, :146; visual_camera_calibration.cpp:162).
// Any field order / extra fields; x, y, z must exist; a missing intensity reads as 0 (PCL leaves it at the default).
// ---------------------------------------------------------------------------------------------------------------------
struct PcdField { std::string name; int size = 4; char type = 'F'; int count = 1; size_t offset = 0; };

inline double pcd_value(const uint8_t* p, const PcdField& f) {
  switch (f.type) {
    case 'F': if (f.size == 4) { float v; std::memcpy(&v, p, 4); return v; } else { double v; std::memcpy(&v, p, 8); return v; }
    case 'U': if (f.size == 1) return *p; if (f.size == 2) { uint16_t v; std::memcpy(&v, p, 2); return v; }
              if (f.size == 4) { uint32_t v; std::memcpy(&v, p, 4); return v; } { uint64_t v; std::memcpy(&v, p, 8); return (double)v; }
    default:  if (f.size == 1) return *(const int8_t*)p; if (f.size == 2) { int16_t v; std::memcpy(&v, p, 2); return v; }
              if (f.size == 4) { int32_t v; std::memcpy(&v, p, 4); return v; } { int64_t v; std::memcpy(&v, p, 8); return (double)v; }
  }
}

inline void read_pcd_xyzi(const std::string& path, std::vector<XYZI>& out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("Couldn't read point cloud file.");  // cpp:114
  std::vector<PcdField> fields;
  size_t points = 0, width = 0, height = 1;
  bool have_points = false;
  std::string data_mode, line;
  while (std::getline(f, line)) {
    if (!line.empty() && line.back() == '\r') line.pop_back();
    if (line.empty() || line[0] == '#') continue;
    std::istringstream ls(line);
    std::string key;
    ls >> key;
    if (key == "FIELDS" || key == "COLUMNS") { std::string n; while (ls >> n) { PcdField fd; fd.name = n; fields.push_back(fd); } }
    else if (key == "SIZE") { for (auto& fd : fields) ls >> fd.size; }
    else if (key == "TYPE") { for (auto& fd : fields) ls >> fd.type; }
    else if (key == "COUNT") { for (auto& fd : fields) ls >> fd.count; }
    else if (key == "WIDTH") ls >> width;
    else if (key == "HEIGHT") ls >> height;
    else if (key == "POINTS") { ls >> points; have_points = true; }
    else if (key == "DATA") { ls >> data_mode; break; }
  }
  if (fields.empty() || data_mode.empty()) throw std::runtime_error("Couldn't read point cloud file.");
  if (!have_points) points = width * height;
  size_t stride = 0;
  for (auto& fd : fields) { fd.offset = stride; stride += (size_t)fd.size * fd.count; }
  int ix = -1, iy = -1, iz = -1, ii = -1;
  for (size_t k = 0; k < fields.size(); k++) {
    if (fields[k].name == "x") ix = (int)k; else if (fields[k].name == "y") iy = (int)k; else if (fields[k].name == "z") iz = (int)k;
    else if (fields[k].name == "intensity") ii = (int)k;
  }
  if (ix < 0 || iy < 0 || iz < 0) throw std::runtime_error("Couldn't read point cloud file.");
  out.assign(points, XYZI{0.f, 0.f, 0.f, 0.f});
  if (data_mode == "ascii") {
    for (size_t i = 0; i < points; i++) {
      if (!std::getline(f, line)) { out.resize(i); break; }
      std::istringstream ls(line);
      XYZI p{0.f, 0.f, 0.f, 0.f};
      for (size_t k = 0; k < fields.size(); k++)
        for (int c = 0; c < fields[k].count; c++) {
          std::string tok;
          if (!(ls >> tok)) break;
          if (c) continue;
          float v = 0.f;
          if (tok == "nan" || tok == "-nan" || tok == "NaN") v = std::nanf("");
          else v = (float)std::strtod(tok.c_str(), nullptr);
          if ((int)k == ix) p.x = v; else if ((int)k == iy) p.y = v; else if ((int)k == iz) p.z = v; else if ((int)k == ii) p.intensity = v;
        }
      out[i] = p;
    }
  } else if (data_mode == "binary") {
    std::vector<uint8_t> buf(points * stride);
    f.read(reinterpret_cast<char*>(buf.data()), (std::streamsize)buf.size());
    if ((size_t)f.gcount() != buf.size()) throw std::runtime_error("Couldn't read point cloud file.");
    for (size_t i = 0; i < points; i++) {
      const uint8_t* r = buf.data() + i * stride;
      out[i].x = (float)pcd_value(r + fields[ix].offset, fields[ix]);
      out[i].y = (float)pcd_value(r + fields[iy].offset, fields[iy]);
      out[i].z = (float)pcd_value(r + fields[iz].offset, fields[iz]);
      if (ii >= 0) out[i].intensity = (float)pcd_value(r + fields[ii].offset, fields[ii]);
    }
  } else if (data_mode == "binary_compressed") {
    uint32_t csize = 0, usize = 0;
    f.read(reinterpret_cast<char*>(&csize), 4);
    f.read(reinterpret_cast<char*>(&usize), 4);
    std::vector<uint8_t> cbuf(csize), ubuf(usize);
    f.read(reinterpret_cast<char*>(cbuf.data()), csize);
    if ((size_t)f.gcount() != csize || usize != points * stride || !lzf_decompress(cbuf.data(), csize, ubuf.data(), usize))
      throw std::runtime_error("Couldn't read point cloud file.");
    // the uncompressed block is field-major: all x, then all y, ...
    std::vector<size_t> start(fields.size());
    size_t o = 0;
    for (size_t k = 0; k < fields.size(); k++) { start[k] = o; o += (size_t)fields[k].size * fields[k].count * points; }
    auto at = [&](int k, size_t i) { return ubuf.data() + start[k] + i * (size_t)fields[k].size * fields[k].count; };
    for (size_t i = 0; i < points; i++) {
      out[i].x = (float)pcd_value(at(ix, i), fields[ix]);
      out[i].y = (float)pcd_value(at(iy, i), fields[iy]);
      out[i].z = (float)pcd_value(at(iz, i), fields[iz]);
      if (ii >= 0) out[i].intensity = (float)pcd_value(at(ii, i), fields[ii]);
    }
  } else {
    throw std::runtime_error("Couldn't read point cloud file.");
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// pcl::PCDWriter::writeASCII (cpp:135, :217, :542, :920, :939, :954), PCL 1.10 layout: precision 8 ("%.8g"), the packed rgb
// field written as an unsigned integer (TYPE U in ASCII files), one point per line, VIEWPOINT 0 0 0 1 0 0 0.
// ---------------------------------------------------------------------------------------------------------------------
struct XYZRGB { float x, y, z; uint8_t b, g, r, a; };                       // pcl::PointXYZRGB payload (a = 255)
struct XYZRGBMask { float x, y, z; uint8_t b, g, r, a; uint16_t mask; };     // PointXYZRGBMask (include/FrameData.hpp:68-87)

inline void pcd_header(FILE* f, const char* fields, const char* sizes, const char* types, const char* counts, size_t n) {
  std::fprintf(f, "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS %s\nSIZE %s\nTYPE %s\nCOUNT %s\nWIDTH %zu\nHEIGHT 1\n"
                  "VIEWPOINT 0 0 0 1 0 0 0\nPOINTS %zu\nDATA ascii\n", fields, sizes, types, counts, n, n);
}
inline int fmt_f(char* b, float v) {
  if (v != v) { std::memcpy(b, "nan", 3); return 3; }
  return std::snprintf(b, 32, "%.8g", (double)v);
}
inline uint32_t pack_rgb(uint8_t r, uint8_t g, uint8_t b, uint8_t a = 255) { return ((uint32_t)a << 24) | ((uint32_t)r << 16) | ((uint32_t)g << 8) | b; }

// A float as it comes back from the ASCII PCD (written with 8 significant digits, read with strtof).
inline float through_text(float v) {
  char b[40];
  b[fmt_f(b, v)] = 0;
  return std::strtof(b, nullptr);
}

inline bool write_pcd_xyzi(const std::string& path, const XYZI* pts, size_t n) {
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) return false;
  pcd_header(f, "x y z intensity", "4 4 4 4", "F F F F", "1 1 1 1", n);
  char line[160];
  for (size_t i = 0; i < n; i++) {
    int o = fmt_f(line, pts[i].x); line[o++] = ' ';
    o += fmt_f(line + o, pts[i].y); line[o++] = ' ';
    o += fmt_f(line + o, pts[i].z); line[o++] = ' ';
    o += fmt_f(line + o, pts[i].intensity); line[o++] = '\n';
    std::fwrite(line, 1, (size_t)o, f);
  }
  return std::fclose(f) == 0;
}
inline bool write_pcd_xyzrgb(const std::string& path, const XYZRGB* pts, size_t n) {
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) return false;
  std::setvbuf(f, nullptr, _IOFBF, 1 << 22);
  pcd_header(f, "x y z rgb", "4 4 4 4", "F F F U", "1 1 1 1", n);
  char line[160];
  for (size_t i = 0; i < n; i++) {
    int o = fmt_f(line, pts[i].x); line[o++] = ' ';
    o += fmt_f(line + o, pts[i].y); line[o++] = ' ';
    o += fmt_f(line + o, pts[i].z); line[o++] = ' ';
    o += std::snprintf(line + o, 16, "%u\n", pack_rgb(pts[i].r, pts[i].g, pts[i].b, pts[i].a));
    std::fwrite(line, 1, (size_t)o, f);
  }
  return std::fclose(f) == 0;
}
// pcl::PCDWriter::writeBinary<pcl::PointXYZRGB> (opt-in, pcp_process_args.binary_output): the header generateHeader emits for the
// point type (rgb stays the float field it is registered as) and the fields of every point packed without padding.
inline bool write_pcd_xyzrgb_binary(const std::string& path, const XYZRGB* pts, size_t n) {
  static_assert(sizeof(XYZRGB) == 16, "pcl::PointXYZRGB payload");
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) return false;
  std::fprintf(f, "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z rgb\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\nWIDTH %zu\nHEIGHT 1\n"
                  "VIEWPOINT 0 0 0 1 0 0 0\nPOINTS %zu\nDATA binary\n", n, n);
  const bool ok = n == 0 || std::fwrite(pts, sizeof(XYZRGB), n, f) == n;
  return (std::fclose(f) == 0) && ok;
}
inline bool write_pcd_xyzrgbmask(const std::string& path, const XYZRGBMask* pts, size_t n) {
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) return false;
  pcd_header(f, "x y z rgb segmentMask", "4 4 4 4 2", "F F F U U", "1 1 1 1 1", n);
  char line[192];
  for (size_t i = 0; i < n; i++) {
    int o = fmt_f(line, pts[i].x); line[o++] = ' ';
    o += fmt_f(line + o, pts[i].y); line[o++] = ' ';
    o += fmt_f(line + o, pts[i].z); line[o++] = ' ';
    o += std::snprintf(line + o, 32, "%u %u\n", pack_rgb(pts[i].r, pts[i].g, pts[i].b, pts[i].a), (unsigned)pts[i].mask);
    std::fwrite(line, 1, (size_t)o, f);
  }
  return std::fclose(f) == 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// loadImagesAndOdometry (cpp:965-1005): "ts x y z qw qx qy qz" per line, stop at the first malformed line.
// ---------------------------------------------------------------------------------------------------------------------
struct OdomLine { double ts; double pose[7]; };  // pose = {x, y, z, qw, qx, qy, qz} (include/FrameData.hpp:9-12)
inline std::vector<OdomLine> read_odometry(const std::string& path) {
  std::vector<OdomLine> out;
  std::ifstream f(path);
  std::string line;
  while (std::getline(f, line)) {
    std::istringstream iss(line);
    OdomLine o;
    if (!(iss >> o.ts >> o.pose[0] >> o.pose[1] >> o.pose[2] >> o.pose[3] >> o.pose[4] >> o.pose[5] >> o.pose[6])) break;  // cpp:974-977
    out.push_back(o);
  }
  return out;
}

inline bool file_exists(const std::string& p) { std::ifstream f(p, std::ios::binary); return (bool)f; }
inline bool read_file(const std::string& p, std::vector<uint8_t>& out) {
  std::ifstream f(p, std::ios::binary | std::ios::ate);
  if (!f) return false;
  const std::streamsize n = f.tellg();
  f.seekg(0);
  out.resize((size_t)n);
  f.read(reinterpret_cast<char*>(out.data()), n);
  return (bool)f || f.gcount() == n;
}

}  // namespace io
}  // namespace pcp
