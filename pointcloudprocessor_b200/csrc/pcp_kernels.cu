// pcp_kernels.cu — hand-written sm_100a kernels of the colorization hot path.
//
//   scan_kernel     (a)+(b): fused float transform + double pinhole projection over the float4 SoA cloud, pass 1 of the
//                   z-buffer (64-bit atomicMin on packed (fl32 depth, point index) keys) and emission of the candidates
//                   that can still be visible (early-z: a point farther than the current cell minimum + tol can never
//                   pass the reference's second pass because the minimum only decreases).
//   resolve_kernel  (b)+(c): pass 2 of the z-buffer on the surviving candidates against the final grid, colour / mask
//                   gather from the L2-resident image, per-point accumulation (u32 sums + count, or top-M by score).
//   finalize_*      smoothColors (src/PointCloudProcessor.cpp:604-631) as an integer mean or the top-M weighted mean.
//
// Compiled with -fmad=false; every decision-making operation additionally uses explicit _rn intrinsics (pcp_math.cuh).
#include "pcp_internal.cuh"

namespace pcp {

static constexpr int kScanThreads = 256;
static constexpr int kWarpsPerCta = kScanThreads / 32;
static constexpr int kQueueCap = 64;  // per-warp staging queue (records), flushed 32 at a time

__global__ void zgrid_init_kernel(unsigned long long* zgrid, int cells, FrameCounters* counters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < cells) zgrid[i] = kZInit;
  if (i < (int)(sizeof(FrameCounters) / sizeof(unsigned int))) reinterpret_cast<unsigned int*>(counters)[i] = 0u;
}

void launch_zgrid_init(unsigned long long* zgrid, int cells, FrameCounters* counters, cudaStream_t s) {
  const int n = cells > 32 ? cells : 32;
  zgrid_init_kernel<<<(n + 255) / 256, 256, 0, s>>>(zgrid, cells, counters);
}

// ---------------------------------------------------------------------------------------------------------------------
// Per-point body of pass 1.  Returns true when a record must be emitted.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool scan_point(const FrameParams& fp, unsigned long long* __restrict__ zgrid, float4 p, uint32_t idx,
                                           Record& rec, bool& candidate) {
  candidate = false;
  float x = p.x, y = p.y, z = p.z;
  if (fp.use_w2c) transform_point(fp.w2c, p.x, p.y, p.z, x, y, z);  // pcl::transformPointCloud (cpp:521)
  double X = (double)x, Y = (double)y, Z = (double)z;                // view_culling.cpp:33
  if (fp.use_T) transform_point_d(fp.T, (double)x, (double)y, (double)z, X, Y, Z);  // :37
  if (Z < 0.0) return false;                                          // :76
  double u, v;
  project(fp.cam, X, Y, Z, u, v);                                     // :86
  int px, py;
  if (!cull_gate(u, v, fp.dsf, fp.cullW, fp.cullH, px, py)) return false;  // :88-94
  candidate = true;                                                   // :96
  const bool in_grid = px < fp.Wc && py < fp.Hc;                      // :116
  rec.x = x; rec.y = y; rec.z = z; rec.idx = idx;
  // generateColorMap re-projects the same float point (cpp:746-753): identical doubles when T is the identity.
  rec.pix = fp.use_T ? -2 : color_pixel(Z, u, v, fp.imgW, fp.imgH);
  if (!fp.depth_culling) {  // view_culling.cpp:135: every recorded point is kept
    rec.cell = in_grid ? py * fp.Wc + px : -1;
    rec.dist = 0.0;
    return true;
  }
  if (!in_grid) return false;                                         // :126-131, :162-165
  const double dist = norm3(X, Y, Z);                                 // :102
  const int cell = py * fp.Wc + px;
  const float df = __double2float_rn(dist);                           // :123
  const unsigned long long key = ((unsigned long long)__float_as_uint(df) << 32) | (unsigned long long)idx;
  unsigned long long cur = __ldcg(zgrid + cell);                      // L2-coherent read: an upper bound of the final minimum
  if (key < cur && dist <= 3.4028234663852886e+38) {                  // :118 (dist > FLT_MAX never enters the map)
    const unsigned long long old = atomicMin(zgrid + cell, key);
    cur = old < key ? old : key;
  }
  const float minf = __uint_as_float((unsigned int)(cur >> 32));
  if (dist > __dadd_rn((double)minf, fp.tol)) return false;           // :157 can only get stricter as the minimum decreases
  rec.cell = cell;
  rec.dist = dist;
  return true;
}

// ---------------------------------------------------------------------------------------------------------------------
// scan_kernel: one point per lane per step; warp-private shared-memory queue, flushed 32 records (1 KB) at a time with a
// single global atomicAdd, so record emission costs one global atomic per 32 records.
// ---------------------------------------------------------------------------------------------------------------------
template <bool kPrebin>
__global__ void __launch_bounds__(kScanThreads) scan_kernel(const ScanArgs a, const __grid_constant__ FrameParams fp) {
  __shared__ Record queue[kWarpsPerCta][kQueueCap];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  Record* q = queue[warp];
  int qn = 0;
  unsigned int ncand = 0;

  auto step = [&](bool valid, float4 p, uint32_t idx) {
    Record rec;
    bool cand = false;
    bool emit = false;
    if (valid) emit = scan_point(fp, a.zgrid, p, idx, rec, cand);
    ncand += cand ? 1u : 0u;
    const unsigned int m = __ballot_sync(0xffffffffu, emit);
    if (m) {
      if (emit) q[qn + __popc(m & ((1u << lane) - 1u))] = rec;
      qn += __popc(m);
      __syncwarp();
      if (qn >= 32) {
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(&a.counters->n_records, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        const uint4* src = reinterpret_cast<const uint4*>(q);
        uint4* dst = reinterpret_cast<uint4*>(a.records + base);
        dst[lane] = src[lane];            // 64 x 16 B = the first 32 records, fully coalesced
        dst[lane + 32] = src[lane + 32];
        const int rem = qn - 32;
        Record tmp;
        if (lane < rem) tmp = q[32 + lane];
        __syncwarp();
        if (lane < rem) q[lane] = tmp;
        qn = rem;
        __syncwarp();
      }
    }
  };

  if (!kPrebin) {
    const uint32_t stride = gridDim.x * kScanThreads;
    for (uint32_t base = blockIdx.x * kScanThreads + warp * 32; base < a.n; base += stride) {
      const uint32_t i = base + lane;
      const bool valid = i < a.n;
      float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
      if (valid) p = __ldg(a.cloud + i);
      step(valid, p, i);
    }
  } else {
    const uint32_t nact = a.counters->n_active_blocks;
    const uint32_t warps_per_block = a.block_points / 32;
    const uint32_t total = nact * warps_per_block;
    for (uint32_t w = blockIdx.x * kWarpsPerCta + warp; w < total; w += gridDim.x * kWarpsPerCta) {
      const uint32_t blk = a.active_blocks[w / warps_per_block];
      const uint32_t i = blk * a.block_points + (w % warps_per_block) * 32 + lane;
      const bool valid = i < a.n;
      float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
      uint32_t idx = 0;
      if (valid) { p = __ldg(a.cloud + i); idx = __ldg(a.perm + i); }
      step(valid, p, idx);
    }
  }

  // flush the tail of the queue
  if (qn > 0) {
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(&a.counters->n_records, (unsigned int)qn);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (lane < qn) a.records[base + lane] = q[lane];
  }
  // candidate count: warp reduce, one atomic per warp
  for (int o = 16; o > 0; o >>= 1) ncand += __shfl_down_sync(0xffffffffu, ncand, o);
  if (lane == 0 && ncand) atomicAdd(&a.counters->n_candidates, ncand);
}

void launch_scan(const ScanArgs& a, const FrameParams& fp, int grid, cudaStream_t s) {
  if (a.active_blocks) scan_kernel<true><<<grid, kScanThreads, 0, s>>>(a, fp);
  else scan_kernel<false><<<grid, kScanThreads, 0, s>>>(a, fp);
}

// ---------------------------------------------------------------------------------------------------------------------
// resolve_kernel: pass 2 (view_culling.cpp:140-168) + generateColorMap (cpp:743-766) + generateSegmentMap (cpp:783-815)
// + scoring / accumulation (cpp:586-591).  One record per thread.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void top_insert(const ResolveArgs& a, uint32_t idx, float score, uint32_t rgb, uint32_t frame) {
  // slots sorted by (score desc, frame asc); an empty slot has frame == 0xffffffff
  float s = score; uint32_t c = rgb, f = frame;
  for (int k = 0; k < a.top_m; k++) {
    const size_t o = (size_t)k * a.n + idx;
    const float ks = a.top_score[o];
    const uint32_t kf = a.top_frame[o];
    const bool empty = kf == 0xffffffffu;
    const bool before = empty || s > ks || (s == ks && f < kf);
    if (before) {
      const uint32_t kc = a.top_rgbf[o];
      a.top_score[o] = s; a.top_rgbf[o] = c; a.top_frame[o] = f;
      if (empty) return;
      s = ks; c = kc; f = kf;
    }
  }
}

__global__ void __launch_bounds__(256) resolve_kernel(const ResolveArgs a, const __grid_constant__ FrameParams fp) {
  const unsigned int nrec = a.counters->n_records;
  unsigned int nvis = 0, ncol = 0;
  for (unsigned int r = blockIdx.x * blockDim.x + threadIdx.x; r < nrec; r += gridDim.x * blockDim.x) {
    const uint4* rp = reinterpret_cast<const uint4*>(a.records + r);
    const uint4 w0 = __ldcs(rp), w1 = __ldcs(rp + 1);
    Record rec;
    *reinterpret_cast<uint4*>(&rec) = w0;
    *(reinterpret_cast<uint4*>(&rec) + 1) = w1;
    bool visible = true;
    if (fp.depth_culling) {
      const float minf = __uint_as_float((unsigned int)(__ldg(a.zgrid + rec.cell) >> 32));
      visible = !(rec.dist > __dadd_rn((double)minf, fp.tol));  // view_culling.cpp:157
    }
    if (!visible) continue;
    nvis++;
    if (a.vis_bits) atomicOr(a.vis_bits + (rec.idx >> 5), 1u << (rec.idx & 31));
    int pix = rec.pix;
    if (pix == -2) {  // a double T was applied for culling: generateColorMap projects the float point itself
      double u, v;
      project(fp.cam, (double)rec.x, (double)rec.y, (double)rec.z, u, v);
      pix = color_pixel((double)rec.z, u, v, fp.imgW, fp.imgH);
    }
    if (pix < 0) continue;
    ncol++;
    if (a.pix_out) a.pix_out[rec.idx] = pix;
    if (!a.bgr) continue;
    const int pv = pix / fp.imgW, pu = pix - pv * fp.imgW;
    const uint8_t* px = a.bgr + (size_t)pv * a.pitch + (size_t)pu * 3;
    const uint32_t b = __ldg(px), g = __ldg(px + 1), rr = __ldg(px + 2);  // cpp:760-762
    if (a.accum_mode == 0) {
      uint4 s = a.acc[rec.idx];
      s.x += rr; s.y += g; s.z += b; s.w += 1u;
      a.acc[rec.idx] = s;
    } else {
      const float o = orientation_score(rec.x, rec.y, rec.z, fp.cam_pos[0], fp.cam_pos[1], fp.cam_pos[2]);
      const float d = distance_score(rec.x, rec.y, rec.z);
      top_insert(a, rec.idx, final_score(o, d), rr | (g << 8) | (b << 16), (uint32_t)fp.frame_index);
      uint4 s = a.acc[rec.idx];
      s.x += rr; s.y += g; s.z += b; s.w += 1u;
      a.acc[rec.idx] = s;
    }
    if (a.mask) {
      const uint32_t mv = __ldg(a.mask + (size_t)pv * a.mask_pitch + pu);  // cpp:804
      uint2 m = a.mask_acc[rec.idx];
      m.x += (mv == 255u) ? 1u : 0u;                                       // cpp:806
      m.y = mv > m.y ? mv : m.y;
      a.mask_acc[rec.idx] = m;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    nvis += __shfl_down_sync(0xffffffffu, nvis, o);
    ncol += __shfl_down_sync(0xffffffffu, ncol, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (nvis) atomicAdd(&a.counters->n_visible, nvis);
    if (ncol) atomicAdd(&a.counters->n_colored, ncol);
  }
}

void launch_resolve(const ResolveArgs& a, const FrameParams& fp, int grid, cudaStream_t s) {
  resolve_kernel<<<grid, 256, 0, s>>>(a, fp);
}

// ---------------------------------------------------------------------------------------------------------------------
// Cloud ingest: AoS records (pcl::PointXYZI: 32 B, intensity at byte 16) -> float4 SoA.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void aos_to_soa_kernel(const uint8_t* __restrict__ aos, size_t n, size_t stride, size_t ioff, float4* __restrict__ out) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float* p = reinterpret_cast<const float*>(aos + i * stride);
    const float inten = *reinterpret_cast<const float*>(aos + i * stride + ioff);
    out[i] = make_float4(p[0], p[1], p[2], inten);
  }
}
void launch_aos_to_soa(const uint8_t* d_aos, size_t n, size_t stride, size_t ioff, float4* out, cudaStream_t s) {
  if (n == 0) return;
  const int grid = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
  aos_to_soa_kernel<<<grid, 256, 0, s>>>(d_aos, n, stride, ioff, out);
}

// ---------------------------------------------------------------------------------------------------------------------
// Finalize.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void finalize_sum_kernel(const uint4* __restrict__ acc, uint32_t n, uint8_t* __restrict__ rgb, uint32_t* __restrict__ count) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint4 s = acc[i];
    uint32_t r = 0, g = 0, b = 0;
    if (s.w) { r = s.x / s.w; g = s.y / s.w; b = s.z / s.w; }
    if (rgb) { rgb[3 * (size_t)i] = (uint8_t)r; rgb[3 * (size_t)i + 1] = (uint8_t)g; rgb[3 * (size_t)i + 2] = (uint8_t)b; }
    if (count) count[i] = s.w;
  }
}
void launch_finalize_sum(const uint4* acc, uint32_t n, uint8_t* rgb, uint32_t* count, cudaStream_t s) {
  if (n == 0) return;
  const int grid = (n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16;
  finalize_sum_kernel<<<grid, 256, 0, s>>>(acc, n, rgb, count);
}

// smoothColors (cpp:612-629): slots are already sorted by (score desc, frame asc); weighted mean in float, truncation.
__global__ void finalize_top_kernel(const float* __restrict__ score, const uint32_t* __restrict__ rgbf, const uint32_t* __restrict__ frame,
                                    uint32_t n, int top_m, uint8_t* __restrict__ rgb, uint32_t* __restrict__ count) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float total = 0.f, r = 0.f, g = 0.f, b = 0.f;
    uint32_t m = 0;
    for (int k = 0; k < top_m; k++) {
      const size_t o = (size_t)k * n + i;
      if (frame[o] == 0xffffffffu) break;
      const float s = score[o];
      const uint32_t c = rgbf[o];
      r = __fadd_rn(r, __fmul_rn((float)(c & 255u), s));
      g = __fadd_rn(g, __fmul_rn((float)((c >> 8) & 255u), s));
      b = __fadd_rn(b, __fmul_rn((float)((c >> 16) & 255u), s));
      total = __fadd_rn(total, s);
      m++;
    }
    uint8_t R = 0, G = 0, B = 0;
    if (m) {
      R = (uint8_t)__float2int_rz(__fdiv_rn(r, total));
      G = (uint8_t)__float2int_rz(__fdiv_rn(g, total));
      B = (uint8_t)__float2int_rz(__fdiv_rn(b, total));
    }
    if (rgb) { rgb[3 * (size_t)i] = R; rgb[3 * (size_t)i + 1] = G; rgb[3 * (size_t)i + 2] = B; }
    if (count) count[i] = m;
  }
}
void launch_finalize_top(const float* score, const uint32_t* rgbf, const uint32_t* frame, uint32_t n, int top_m, uint8_t* rgb,
                         uint32_t* count, cudaStream_t s) {
  if (n == 0) return;
  const int grid = (n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16;
  finalize_top_kernel<<<grid, 256, 0, s>>>(score, rgbf, frame, n, top_m, rgb, count);
}

// ---------------------------------------------------------------------------------------------------------------------
// Bitmask -> ascending index list (ViewCulling::sample order, view_culling.cpp:176-185).  Three small kernels:
// per-word popcounts reduced per 1024-word tile, scan of the tile sums by one CTA, scatter.
// ---------------------------------------------------------------------------------------------------------------------
static constexpr int kTileWords = 1024;

size_t compact_scratch_words(uint32_t n) {
  const size_t words = ((size_t)n + 31) / 32;
  const size_t tiles = (words + kTileWords - 1) / kTileWords;
  return tiles + 1;
}

__global__ void compact_tile_sums(const uint32_t* __restrict__ bits, uint32_t words, uint32_t* __restrict__ tile_sums) {
  __shared__ uint32_t sh[32];
  const uint32_t tile = blockIdx.x;
  uint32_t c = 0;
  for (uint32_t w = tile * kTileWords + threadIdx.x; w < min(words, (tile + 1) * kTileWords); w += blockDim.x) c += __popc(bits[w]);
  for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x < 32) {
    c = threadIdx.x < blockDim.x / 32 ? sh[threadIdx.x] : 0;
    for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if (threadIdx.x == 0) tile_sums[tile] = c;
  }
}

__global__ void compact_scan_tiles(uint32_t* tile_sums, uint32_t tiles, uint32_t* d_total) {
  // single CTA, sequential chunks of 1024 with a running carry (tiles <= a few thousand)
  __shared__ uint32_t sh[1024];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < tiles; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < tiles ? tile_sums[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
      uint32_t t = threadIdx.x >= (unsigned)o ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    const uint32_t incl = sh[threadIdx.x];
    if (i < tiles) tile_sums[i] = carry + incl - v;  // exclusive
    __syncthreads();
    if (threadIdx.x == 1023) carry += incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *d_total = carry;
}

__global__ void compact_scatter(const uint32_t* __restrict__ bits, uint32_t words, uint32_t n, const uint32_t* __restrict__ tile_offsets,
                                int32_t* __restrict__ idx_out) {
  // one warp per 32 words; warp-level exclusive scan of popcounts; CTA = 1024 words = 32 warps
  __shared__ uint32_t warp_sums[32];
  const uint32_t tile = blockIdx.x;
  const uint32_t w = tile * kTileWords + threadIdx.x;
  const uint32_t word = w < words ? bits[w] : 0u;
  const uint32_t c = __popc(word);
  uint32_t incl = c;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t v = warp_sums[lane];
    uint32_t iv = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, iv, o);
      if (lane >= o) iv += t;
    }
    warp_sums[lane] = iv - v;
  }
  __syncthreads();
  uint32_t off = tile_offsets[tile] + warp_sums[warp] + incl - c;
  uint32_t m = word;
  while (m) {
    const int b = __ffs(m) - 1;
    m &= m - 1;
    const uint32_t i = w * 32 + b;
    if (i < n) idx_out[off++] = (int32_t)i;
  }
}

void launch_compact_bits(const uint32_t* bits, uint32_t n, uint32_t* word_offsets, uint32_t* tile_sums, int32_t* idx_out,
                         uint32_t* d_total, cudaStream_t s) {
  (void)word_offsets;
  const uint32_t words = (n + 31) / 32;
  const uint32_t tiles = (words + kTileWords - 1) / kTileWords;
  if (tiles == 0) { cudaMemsetAsync(d_total, 0, sizeof(uint32_t), s); return; }
  compact_tile_sums<<<tiles, 256, 0, s>>>(bits, words, tile_sums);
  compact_scan_tiles<<<1, 1024, 0, s>>>(tile_sums, tiles, d_total);
  compact_scatter<<<tiles, kTileWords, 0, s>>>(bits, words, n, tile_sums, idx_out);
}

}  // namespace pcp

int scan_ctas_per_sm() {
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pcp::scan_kernel<false>, pcp::kScanThreads, 0) != cudaSuccess || occ < 1) occ = 2;
  return occ;
}
