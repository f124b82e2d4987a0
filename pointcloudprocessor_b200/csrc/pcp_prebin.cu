// pcp_prebin.cu — one-time spatial pre-binning of the map cloud (pcp_params.prebin, SURVEY.md §8d "pre-binned").
//
// At pcp_set_cloud the points are sorted by a 63-bit Morton key of their position on an isotropic grid over the cloud's bounding box and cut
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
#include <cstring>

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

__device__ __forceinline__ unsigned long long spread21(unsigned long long v) {  // 21 bits -> every third bit
  v &= 0x1fffffull;
  v = (v | (v << 32)) & 0x001f00000000ffffull;
  v = (v | (v << 16)) & 0x001f0000ff0000ffull;
  v = (v | (v << 8)) & 0x100f00f00f00f00full;
  v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}

// 63-bit Morton key on an ISOTROPIC grid (one cell size for the three axes = largest extent / 2^21): the scale of an axis must
// not depend on the extent of the cloud along it, or chunks of a long thin map (a corridor) become long thin boxes whose
// bounding boxes cull poorly.  With one cell size, 128 consecutive points of a surface are a compact patch.
__global__ void morton_kernel(const float4* __restrict__ pts, uint32_t n, const int* __restrict__ box6, unsigned long long* __restrict__ keys,
                              uint32_t* __restrict__ idx) {
  auto dec = [](int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); };
  const float lx = dec(box6[0]), ly = dec(box6[1]), lz = dec(box6[2]);
  const float ext = fmaxf(fmaxf(dec(box6[3]) - lx, dec(box6[4]) - ly), fmaxf(dec(box6[5]) - lz, 1e-20f));
  const float sc = 2097151.0f / ext;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = pts[i];
    unsigned long long key = 0x7fffffffffffffffull;  // non-finite points sort to the end
    if (finite3(p)) {
      const unsigned long long qx = (unsigned long long)fminf(fmaxf((p.x - lx) * sc, 0.0f), 2097151.0f);
      const unsigned long long qy = (unsigned long long)fminf(fmaxf((p.y - ly) * sc, 0.0f), 2097151.0f);
      const unsigned long long qz = (unsigned long long)fminf(fmaxf((p.z - lz) * sc, 0.0f), 2097151.0f);
      key = spread21(qx) | (spread21(qy) << 1) | (spread21(qz) << 2);
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
// The boxes are stored in VISITING order: chunk w sits at slot (w * mul_inv) % n_chunks, the inverse of the pseudo-random visit
// sequence slot -> (slot * mul) % n_chunks, so that classify_kernel reads boxes and chunk ids coalesced and without a modulo.
__global__ void chunk_aabb_kernel(const float4* __restrict__ pts, uint32_t n, uint32_t n_chunks, uint32_t mul_inv, float4* __restrict__ aabb,
                                  uint32_t* __restrict__ order) {
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
    const uint32_t slot = (uint32_t)(((unsigned long long)w * mul_inv) % n_chunks);
    aabb[2 * slot] = make_float4(lo[0], lo[1], lo[2], bad ? 1.0f : 0.0f);
    aabb[2 * slot + 1] = make_float4(hi[0], hi[1], hi[2], (float)cnt);
    order[slot] = w;
  }
}

}  // namespace

// Sorts `d_cloud` (n points, in place through a temporary) and fills d_perm (sorted position -> original index) and
// d_aabb (2 float4 per chunk).  All buffers are allocated by the caller.
// `scratch` is a context-owned device buffer that is grown on demand and kept (cudaMalloc/cudaFree per upload cost tens of
// milliseconds of host time on a busy box); nothing here synchronises the stream.
cudaError_t prebin_cloud(float4* d_cloud, uint32_t n, uint32_t* d_perm, float4* d_aabb, uint32_t* d_order, void** scratch, size_t* scratch_bytes,
                         int sm_count, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  cudaError_t e;
  auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
  size_t temp_bytes = 0;
  e = cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, (unsigned long long*)nullptr, (unsigned long long*)nullptr, (uint32_t*)nullptr,
                                      (uint32_t*)nullptr, (int)n, 0, 63, s);
  if (e != cudaSuccess) return e;
  const size_t o_box = 0, o_keys = up(6 * sizeof(int)), o_keys_out = o_keys + up(sizeof(unsigned long long) * n),
               o_idx = o_keys_out + up(sizeof(unsigned long long) * n), o_tmp = o_idx + up(sizeof(uint32_t) * n),
               o_temp = o_tmp + up(sizeof(float4) * n), total = o_temp + up(temp_bytes ? temp_bytes : 1);
  if (*scratch_bytes < total) {
    if (*scratch) { cudaStreamSynchronize(s); cudaFree(*scratch); *scratch = nullptr; *scratch_bytes = 0; }
    if ((e = cudaMalloc(scratch, total)) != cudaSuccess) return e;
    *scratch_bytes = total;
  }
  unsigned char* base = static_cast<unsigned char*>(*scratch);
  int* d_box = reinterpret_cast<int*>(base + o_box);
  unsigned long long* d_keys = reinterpret_cast<unsigned long long*>(base + o_keys);
  unsigned long long* d_keys_out = reinterpret_cast<unsigned long long*>(base + o_keys_out);
  uint32_t* d_idx = reinterpret_cast<uint32_t*>(base + o_idx);
  float4* d_tmp = reinterpret_cast<float4*>(base + o_tmp);
  void* d_temp = base + o_temp;
  static const int h_box[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
  if ((e = cudaMemcpyAsync(d_box, h_box, sizeof(h_box), cudaMemcpyHostToDevice, s)) != cudaSuccess) return e;
  const int grid = sm_count * 8;
  bbox_kernel<<<grid, 256, 0, s>>>(d_cloud, n, reinterpret_cast<float*>(d_box));
  morton_kernel<<<grid, 256, 0, s>>>(d_cloud, n, d_box, d_keys, d_idx);
  e = cub::DeviceRadixSort::SortPairs(d_temp, temp_bytes, d_keys, d_keys_out, d_idx, d_perm, (int)n, 0, 63, s);  // stable: deterministic
  if (e != cudaSuccess) return e;
  gather_kernel<<<grid, 256, 0, s>>>(d_cloud, d_perm, n, d_tmp);
  if ((e = cudaMemcpyAsync(d_cloud, d_tmp, sizeof(float4) * n, cudaMemcpyDeviceToDevice, s)) != cudaSuccess) return e;
  const uint32_t n_chunks = (n + kChunk - 1) / kChunk;
  // visit sequence slot -> (slot * mul) % n_chunks with an odd multiplier coprime with n_chunks (a permutation that spreads the
  // chunks in flight over the whole map); the kernel needs its inverse
  auto gcd = [](uint64_t a, uint64_t b) { while (b) { const uint64_t t = a % b; a = b; b = t; } return a; };
  uint64_t mul = 1;
  for (uint32_t cand : {40503u, 7919u, 7907u, 6007u, 5003u, 4001u, 3001u, 2003u, 1009u, 509u, 251u, 127u, 61u, 31u, 13u, 7u, 3u})
    if (cand < n_chunks && gcd(cand, n_chunks) == 1) { mul = cand; break; }
  int64_t t0 = 0, t1 = 1, r0 = (int64_t)n_chunks, r1 = (int64_t)(mul % n_chunks);  // extended Euclid: mul * t1 == 1 (mod n_chunks)
  if (n_chunks == 1) { t1 = 0; r1 = 1; }
  while (r1 > 1) { const int64_t q = r0 / r1, r2 = r0 - q * r1, t2 = t0 - q * t1; r0 = r1; r1 = r2; t0 = t1; t1 = t2; }
  const uint32_t mul_inv = (uint32_t)(((t1 % (int64_t)n_chunks) + (int64_t)n_chunks) % (int64_t)n_chunks);
  chunk_aabb_kernel<<<(n_chunks * 32 + 255) / 256, 256, 0, s>>>(d_cloud, n, n_chunks, mul_inv, d_aabb, d_order);
  return cudaGetLastError();
}

// inv[perm[j]] = j: lets the finalize kernels walk the outputs in the caller's order (coalesced writes) and gather the per-point
// state, instead of scattering 3-byte colours through perm.
__global__ void invert_perm_kernel(const uint32_t* __restrict__ perm, uint32_t n, uint32_t* __restrict__ inv) {
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) inv[perm[j]] = j;
}
void launch_invert_perm(const uint32_t* perm, uint32_t n, uint32_t* inv, cudaStream_t s) {
  if (n == 0) return;
  const uint32_t blocks = (n + 255u) / 256u;
  invert_perm_kernel<<<blocks < 148u * 16u ? blocks : 148u * 16u, 256, 0, s>>>(perm, n, inv);
}

// How local is the caller's point order?  out[0] = among up to 65536 evenly spaced neighbours (i, i + 1) of the caller's order, the number
// whose sorted positions are less than 65536 apart (1 MB of accumulators); out[1] = the number of pairs looked at.  A voxel-grid
// filter's output or a scan-ordered map scores near 1, a shuffled cloud near 0: pcp_finalize picks its gather by it.
__global__ void inv_locality_kernel(const uint32_t* __restrict__ inv, uint32_t n, uint32_t samples, uint32_t stride, uint32_t* __restrict__ out) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k == 0) out[1] = samples;
  if (k >= samples) return;
  const uint32_t i = k * stride;
  const uint32_t a = inv[i], b = inv[i + 1];
  if ((a > b ? a - b : b - a) < 65536u) atomicAdd(out, 1u);
}
void launch_inv_locality(const uint32_t* inv, uint32_t n, uint32_t* out2, cudaStream_t s) {
  cudaMemsetAsync(out2, 0, 2 * sizeof(uint32_t), s);
  if (n < 2) return;
  const uint32_t samples = n - 1 < 65536u ? n - 1 : 65536u;
  inv_locality_kernel<<<(samples + 255) / 256, 256, 0, s>>>(inv, n, samples, (n - 1) / samples, out2);
}

}  // namespace pcp
