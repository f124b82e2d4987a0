// pcp_prebin.cu — one-time spatial pre-binning of the map cloud (pcp_params.prebin, SURVEY.md §8d "pre-binned").
//
// At pcp_set_cloud the points are sorted by a 30-bit Morton key of their position inside the cloud's bounding box and cut
// into chunks of kChunk = 128 consecutive points (one warp iteration of the scan kernel); every chunk gets an axis-aligned
// bounding box.  Per frame the scan kernel (a) skips, after reading 32 bytes, every chunk that is certainly behind the
// camera or certainly outside the z-grid, (b) runs with coherent warps on the rest, and (c) addresses the per-point
// accumulators by SORTED position, which turns the resolve kernel's random 16-byte read-modify-writes over the whole
// array into accesses clustered on the visible surfaces (L2-resident).  Results do not depend on the order of the
// points, and every output of the C ABI is delivered in the caller's original order through `perm`.
//
// The sort itself is set-up, not the hot path: it uses CUB's device radix sort from the CUDA toolkit (library code, like
// cuBLAS for a plain GEMM); keys, gather and AABB kernels are ours.
#include <cub/device/device_radix_sort.cuh>

#include <cfloat>

#include "pcp_ctx.hpp"

namespace pcp {

namespace {

struct BBox { float lo[3], hi[3]; };

__device__ __forceinline__ bool finite3(float4 p) {
  return fabsf(p.x) <= FLT_MAX && fabsf(p.y) <= FLT_MAX && fabsf(p.z) <= FLT_MAX;
}

__global__ void bbox_kernel(const float4* __restrict__ pts, uint32_t n, float* __restrict__ out6 /* min xyz as ordered ints, max xyz */) {
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = pts[i];
    if (!finite3(p)) continue;
    lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
    hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
  }
  for (int o = 16; o > 0; o >>= 1)
    for (int k = 0; k < 3; k++) {
      lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
      hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
    }
  if ((threadIdx.x & 31) == 0) {
    // float atomics through the order-preserving int mapping
    auto enc = [](float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; };
    int* o = reinterpret_cast<int*>(out6);
    for (int k = 0; k < 3; k++) { atomicMin(o + k, enc(lo[k])); atomicMax(o + 3 + k, enc(hi[k])); }
  }
}

__device__ __forceinline__ uint32_t spread10(uint32_t v) {  // 10 bits -> every third bit
  v &= 0x3ffu;
  v = (v | (v << 16)) & 0x030000ffu;
  v = (v | (v << 8)) & 0x0300f00fu;
  v = (v | (v << 4)) & 0x030c30c3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

__global__ void morton_kernel(const float4* __restrict__ pts, uint32_t n, const int* __restrict__ box6, uint32_t* __restrict__ keys,
                              uint32_t* __restrict__ idx) {
  auto dec = [](int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); };
  const float lx = dec(box6[0]), ly = dec(box6[1]), lz = dec(box6[2]);
  const float sx = 1023.0f / fmaxf(dec(box6[3]) - lx, 1e-20f), sy = 1023.0f / fmaxf(dec(box6[4]) - ly, 1e-20f),
              sz = 1023.0f / fmaxf(dec(box6[5]) - lz, 1e-20f);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = pts[i];
    uint32_t key = 0x3fffffffu;  // non-finite points sort to the end
    if (finite3(p)) {
      const uint32_t qx = (uint32_t)fminf(fmaxf((p.x - lx) * sx, 0.0f), 1023.0f);
      const uint32_t qy = (uint32_t)fminf(fmaxf((p.y - ly) * sy, 0.0f), 1023.0f);
      const uint32_t qz = (uint32_t)fminf(fmaxf((p.z - lz) * sz, 0.0f), 1023.0f);
      key = spread10(qx) | (spread10(qy) << 1) | (spread10(qz) << 2);
    }
    keys[i] = key;
    idx[i] = i;
  }
}

__global__ void gather_kernel(const float4* __restrict__ src, const uint32_t* __restrict__ perm, uint32_t n, float4* __restrict__ dst) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = src[perm[i]];
}

// One warp per chunk of kChunk sorted points: {lo.xyz, flag}, {hi.xyz, count}; flag = 1 when the chunk holds a non-finite
// coordinate (such a chunk is never culled as a whole).
__global__ void chunk_aabb_kernel(const float4* __restrict__ pts, uint32_t n, uint32_t n_chunks, float4* __restrict__ aabb) {
  const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_chunks) return;
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  int bad = 0, cnt = 0;
  for (int k = 0; k < kChunk / 32; k++) {
    const uint32_t i = w * kChunk + k * 32 + lane;
    if (i >= n) continue;
    const float4 p = pts[i];
    cnt++;
    if (!finite3(p)) { bad = 1; continue; }
    lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
    hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
  }
  for (int o = 16; o > 0; o >>= 1) {
    for (int k = 0; k < 3; k++) {
      lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
      hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
    }
    bad |= __shfl_xor_sync(0xffffffffu, bad, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if (lane == 0) {
    aabb[2 * w] = make_float4(lo[0], lo[1], lo[2], bad ? 1.0f : 0.0f);
    aabb[2 * w + 1] = make_float4(hi[0], hi[1], hi[2], (float)cnt);
  }
}

}  // namespace

// Sorts `d_cloud` (n points, in place through a temporary) and fills d_perm (sorted position -> original index) and
// d_aabb (2 float4 per chunk).  All buffers are allocated by the caller except the temporaries.  Synchronises the stream.
cudaError_t prebin_cloud(float4* d_cloud, uint32_t n, uint32_t* d_perm, float4* d_aabb, int sm_count, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  cudaError_t e;
  int* d_box = nullptr;
  uint32_t *d_keys = nullptr, *d_keys_out = nullptr, *d_idx = nullptr;
  float4* d_tmp = nullptr;
  void* d_temp = nullptr;
  size_t temp_bytes = 0;
  auto cleanup = [&]() {
    cudaFree(d_box); cudaFree(d_keys); cudaFree(d_keys_out); cudaFree(d_idx); cudaFree(d_tmp); cudaFree(d_temp);
  };
#define PB(call) do { e = (call); if (e != cudaSuccess) { cleanup(); return e; } } while (0)
  PB(cudaMalloc(&d_box, 6 * sizeof(int)));
  PB(cudaMalloc(&d_keys, sizeof(uint32_t) * n));
  PB(cudaMalloc(&d_keys_out, sizeof(uint32_t) * n));
  PB(cudaMalloc(&d_idx, sizeof(uint32_t) * n));
  PB(cudaMalloc(&d_tmp, sizeof(float4) * n));
  const int h_box[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
  PB(cudaMemcpyAsync(d_box, h_box, sizeof(h_box), cudaMemcpyHostToDevice, s));
  const int grid = sm_count * 8;
  bbox_kernel<<<grid, 256, 0, s>>>(d_cloud, n, reinterpret_cast<float*>(d_box));
  morton_kernel<<<grid, 256, 0, s>>>(d_cloud, n, d_box, d_keys, d_idx);
  PB(cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, d_keys, d_keys_out, d_idx, d_perm, (int)n, 0, 30, s));
  PB(cudaMalloc(&d_temp, temp_bytes ? temp_bytes : 1));
  PB(cub::DeviceRadixSort::SortPairs(d_temp, temp_bytes, d_keys, d_keys_out, d_idx, d_perm, (int)n, 0, 30, s));  // stable: deterministic
  gather_kernel<<<grid, 256, 0, s>>>(d_cloud, d_perm, n, d_tmp);
  PB(cudaMemcpyAsync(d_cloud, d_tmp, sizeof(float4) * n, cudaMemcpyDeviceToDevice, s));
  const uint32_t n_chunks = (n + kChunk - 1) / kChunk;
  chunk_aabb_kernel<<<(n_chunks * 32 + 255) / 256, 256, 0, s>>>(d_cloud, n, n_chunks, d_aabb);
  PB(cudaGetLastError());
  PB(cudaStreamSynchronize(s));
#undef PB
  cleanup();
  return cudaSuccess;
}

}  // namespace pcp
