// pcp_jpeg.cu — JPEG decode of the keyframe images (cv::imread, src/PointCloudProcessor.cpp:716) with nvJPEG on the device.
//
// nvJPEG is library code of the CUDA toolkit (like CUB for the set-up sort); it is resolved with dlopen so that libpcp_b200.so
// loads on machines without it (decoding a .jpg then fails with a message; there is no CPU decoder).  The decoded pixels of
// a lossy format depend on the decoder's IDCT and chroma up-sampling: nvJPEG and OpenCV's libjpeg-turbo differ by a few levels
// on some pixels — third-party behaviour, parity unpinned; everything downstream of the decoded buffer is exact.
#include <dlfcn.h>
#include <nvjpeg.h>

#include <mutex>
#include <string>

#include "pcp_imgio.hpp"

namespace pcp {
namespace io {

namespace {
struct NvJpegApi {
  void* lib = nullptr;
  nvjpegStatus_t (*CreateSimple)(nvjpegHandle_t*) = nullptr;
  nvjpegStatus_t (*Destroy)(nvjpegHandle_t) = nullptr;
  nvjpegStatus_t (*JpegStateCreate)(nvjpegHandle_t, nvjpegJpegState_t*) = nullptr;
  nvjpegStatus_t (*JpegStateDestroy)(nvjpegJpegState_t) = nullptr;
  nvjpegStatus_t (*GetImageInfo)(nvjpegHandle_t, const unsigned char*, size_t, int*, nvjpegChromaSubsampling_t*, int*, int*) = nullptr;
  nvjpegStatus_t (*Decode)(nvjpegHandle_t, nvjpegJpegState_t, const unsigned char*, size_t, nvjpegOutputFormat_t, nvjpegImage_t*, cudaStream_t) = nullptr;
  nvjpegHandle_t handle = nullptr;
  nvjpegJpegState_t state = nullptr;
  bool ok = false;
  std::mutex mu;
};

NvJpegApi& api() {
  static NvJpegApi a;
  static std::once_flag once;
  std::call_once(once, [] {
    for (const char* name : {"libnvjpeg.so.12", "libnvjpeg.so", "/usr/local/cuda/lib64/libnvjpeg.so.12"}) {
      a.lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
      if (a.lib) break;
    }
    if (!a.lib) return;
    auto sym = [&](const char* n) { return dlsym(a.lib, n); };
    a.CreateSimple = reinterpret_cast<decltype(a.CreateSimple)>(sym("nvjpegCreateSimple"));
    a.Destroy = reinterpret_cast<decltype(a.Destroy)>(sym("nvjpegDestroy"));
    a.JpegStateCreate = reinterpret_cast<decltype(a.JpegStateCreate)>(sym("nvjpegJpegStateCreate"));
    a.JpegStateDestroy = reinterpret_cast<decltype(a.JpegStateDestroy)>(sym("nvjpegJpegStateDestroy"));
    a.GetImageInfo = reinterpret_cast<decltype(a.GetImageInfo)>(sym("nvjpegGetImageInfo"));
    a.Decode = reinterpret_cast<decltype(a.Decode)>(sym("nvjpegDecode"));
    if (!a.CreateSimple || !a.JpegStateCreate || !a.GetImageInfo || !a.Decode) return;
    if (a.CreateSimple(&a.handle) != NVJPEG_STATUS_SUCCESS) return;
    if (a.JpegStateCreate(a.handle, &a.state) != NVJPEG_STATUS_SUCCESS) return;
    a.ok = true;
  });
  return a;
}
}  // namespace

// Decodes to interleaved B,G,R (OpenCV's order) on the device, then copies to the host image.  Returns an empty string on
// success, the error otherwise.
std::string decode_jpeg(const std::vector<uint8_t>& file, HostImage& im, cudaStream_t stream) {
  NvJpegApi& a = api();
  if (!a.ok) return "nvJPEG (libnvjpeg.so.12) is not available: cannot decode JPEG";
  std::lock_guard<std::mutex> lock(a.mu);
  int ncomp = 0, widths[NVJPEG_MAX_COMPONENT] = {0}, heights[NVJPEG_MAX_COMPONENT] = {0};
  nvjpegChromaSubsampling_t sub;
  if (a.GetImageInfo(a.handle, file.data(), file.size(), &ncomp, &sub, widths, heights) != NVJPEG_STATUS_SUCCESS) return "not a JPEG stream";
  const int W = widths[0], H = heights[0];
  if (W <= 0 || H <= 0) return "empty JPEG";
  uint8_t* d = nullptr;
  if (cudaMalloc(&d, (size_t)W * H * 3) != cudaSuccess) return "cudaMalloc failed";
  nvjpegImage_t out;
  std::memset(&out, 0, sizeof(out));
  out.channel[0] = d;
  out.pitch[0] = (size_t)W * 3;
  std::string err;
  if (a.Decode(a.handle, a.state, file.data(), file.size(), NVJPEG_OUTPUT_BGRI, &out, stream) != NVJPEG_STATUS_SUCCESS) err = "nvjpegDecode failed";
  if (err.empty()) {
    im.width = W; im.height = H; im.channels = 3;
    im.data.resize((size_t)W * H * 3);
    if (cudaMemcpyAsync(im.data.data(), d, im.data.size(), cudaMemcpyDeviceToHost, stream) != cudaSuccess || cudaStreamSynchronize(stream) != cudaSuccess)
      err = "copy of the decoded image failed";
  }
  cudaFree(d);
  return err;
}

}  // namespace io
}  // namespace pcp
