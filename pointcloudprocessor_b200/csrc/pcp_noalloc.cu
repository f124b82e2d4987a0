// pcp_noalloc.cu — ONE entry point that allocates nothing: pcp_compact_bits_noalloc.
//
// Some callers cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The
// bitmask -> ascending index list compaction (ViewCulling::sample's order, view_culling.cpp:176-185; launch_compact_bits in
// pcp_kernels.cu) needs one word of scratch per 1024-word tile of the mask.  This variant keeps that scratch in a file-scope
// `static __device__` array of a fixed, modest size instead of a context-owned cudaMalloc buffer.  It has no counterpart in the
// reference.  No existing code path uses it.
//
//   * limit: PCP_NOALLOC_MAX_BITS = 2^27 bits (134 217 728 points): 4096 tiles -> 16.6 KB of device memory, held by every
//     GPU that loads the library until the process exits;
//   * the array's NAME is a host-side stand-in in host code, not a device address.  Its device address is taken with
//     cudaGetSymbolAddress once per device, with that device current, and that pointer is what the kernels receive;
//   * every call on a device shares the array, which keeps the last call's contents.  Nothing in it has to start at zero: each
//     call overwrites the tile words it reads (the count word is the CALLER's output and is always written).  Calls that
//     overlap on the GPU (two streams, two contexts of the same device) race on it: serialise them;
//   * no cudaMalloc / cudaFree / allocating library routine is called, here or in anything this file calls.
#include <cstdint>

#include "../../include/pcp/pcp.h"
#include "pcp_ctx.hpp"

namespace {

constexpr uint32_t kTileWords = 1024;                                  // mask words per tile = threads of the scatter CTA
constexpr uint64_t kMaxBits = PCP_NOALLOC_MAX_BITS;
constexpr uint32_t kMaxTiles = (uint32_t)((kMaxBits / 32 + kTileWords - 1) / kTileWords);
constexpr int kMaxDevices = 64;

static __device__ __align__(256) uint32_t g_tile_scratch[kMaxTiles + 64];  // exclusive tile offsets of the call in progress

uint32_t* g_scratch_ptr[kMaxDevices] = {};  // device address of g_tile_scratch on device d (cudaGetSymbolAddress), or null

// set bits of every 1024-word tile -> scratch[tile]
__global__ void na_tile_sums(const uint32_t* __restrict__ bits, uint32_t words, uint32_t* __restrict__ scratch) {
  __shared__ uint32_t sh[8];
  const uint32_t tile = blockIdx.x;
  const uint32_t end = min(words, (tile + 1) * kTileWords);
  uint32_t c = 0;
  for (uint32_t w = tile * kTileWords + threadIdx.x; w < end; w += blockDim.x) c += __popc(bits[w]);
  for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (uint32_t k = 0; k < blockDim.x / 32; k++) t += sh[k];
    scratch[tile] = t;
  }
}

// in-place exclusive scan of scratch[0 .. tiles) by one CTA; the total goes to the caller's count word
__global__ void na_scan_tiles(uint32_t* __restrict__ scratch, uint32_t tiles, uint32_t* __restrict__ count_out) {
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < tiles; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < tiles ? scratch[i] : 0u;
    uint32_t incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const uint32_t w = warp_tot[lane];
      uint32_t wi = w;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += t;
      }
      warp_tot[lane] = wi - w;  // exclusive offset of every warp
    }
    __syncthreads();
    const uint32_t excl = carry + warp_tot[warp] + incl - v;
    if (i < tiles) scratch[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *count_out = carry;
}

// one thread per mask word: its set bits go to idx_out in ascending order
__global__ void na_scatter(const uint32_t* __restrict__ bits, uint32_t words, uint32_t n, const uint32_t* __restrict__ scratch,
                           int32_t* __restrict__ idx_out) {
  __shared__ uint32_t warp_tot[32];
  const uint32_t tile = blockIdx.x;
  const uint32_t w = tile * kTileWords + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t word = w < words ? bits[w] : 0u;
  const uint32_t c = __popc(word);
  uint32_t incl = c;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const uint32_t v = warp_tot[lane];
    uint32_t wi = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    warp_tot[lane] = wi - v;
  }
  __syncthreads();
  uint32_t off = scratch[tile] + warp_tot[warp] + incl - c;
  uint32_t m = word;
  while (m) {
    const int b = __ffs(m) - 1;
    m &= m - 1;
    const uint32_t i = w * 32 + (uint32_t)b;
    if (i < n) idx_out[off++] = (int32_t)i;
  }
}

__global__ void na_zero_count(uint32_t* count_out) { *count_out = 0u; }

}  // namespace

extern "C" int pcp_compact_bits_noalloc(pcp_ctx* ctx, const uint32_t* d_bits, size_t n_bits, int32_t* d_idx_out, uint32_t* d_count_out) {
  if (!ctx) return PCP_EINVAL;
  if (!d_count_out || (n_bits && (!d_bits || !d_idx_out))) { ctx->fail("pcp_compact_bits_noalloc: null argument"); return PCP_EINVAL; }
  if ((uint64_t)n_bits > kMaxBits) {
    ctx->fail("pcp_compact_bits_noalloc: %zu bits exceed the static scratch's limit of %llu (PCP_NOALLOC_MAX_BITS); use pcp_cull / a context-owned buffer",
              n_bits, (unsigned long long)kMaxBits);
    return PCP_EINVAL;
  }
  if (ctx->device < 0 || ctx->device >= kMaxDevices) { ctx->fail("pcp_compact_bits_noalloc: device index out of range"); return PCP_EINVAL; }
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess) { ctx->fail("pcp_compact_bits_noalloc: cudaSetDevice: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  // set-up, once per device, with that device current: the array's device address (its name is only a host-side stand-in here)
  uint32_t* scratch = g_scratch_ptr[ctx->device];
  if (!scratch) {
    e = cudaGetSymbolAddress((void**)&scratch, g_tile_scratch);
    if (e != cudaSuccess || !scratch) {
      ctx->fail("pcp_compact_bits_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return PCP_ECUDA;
    }
    g_scratch_ptr[ctx->device] = scratch;
  }
  cudaStream_t st = ctx->stream;
  const uint32_t n = (uint32_t)n_bits;
  const uint32_t words = (n + 31) / 32;
  const uint32_t tiles = (words + kTileWords - 1) / kTileWords;
  if (tiles == 0) {
    na_zero_count<<<1, 1, 0, st>>>(d_count_out);
  } else {
    na_tile_sums<<<tiles, 256, 0, st>>>(d_bits, words, scratch);
    na_scan_tiles<<<1, 1024, 0, st>>>(scratch, tiles, d_count_out);
    na_scatter<<<tiles, kTileWords, 0, st>>>(d_bits, words, n, scratch, d_idx_out);
  }
  ctx->launches += tiles ? 3 : 1;
  e = cudaGetLastError();
  if (e != cudaSuccess) { ctx->fail("pcp_compact_bits_noalloc: launch failed: %s", cudaGetErrorString(e)); return PCP_ECUDA; }
  return PCP_OK;
}
