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
#include <type_traits>
#include <utility>

#include "pcp_internal.cuh"
#include "pcp_chunk.cuh"

namespace pcp {

#ifndef PCP_UNROLL
#define PCP_UNROLL_CAP 4
#else
#define PCP_UNROLL_CAP PCP_UNROLL
#endif
static constexpr int kScanThreads = 256;
static constexpr int kWarpsPerCta = kScanThreads / 32;
static constexpr int kQueueCap = 64;  // per-warp record queue, flushed 32 at a time
static constexpr int kCandCap = 32 * (PCP_UNROLL_CAP + 1); // per-warp candidate queue: < 32 left over + kUnroll x 32 new entries

// ---- PTX wrappers: mbarrier, 1-D bulk asynchronous copy (TMA engine), programmatic dependent launch ------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
// 1-D bulk asynchronous copy global -> shared (TMA engine, SASS UBLKCP); `bytes` a multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
               "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Clears a z-grid and a counter block.  Frames of a pre-binned cloud do not need it per frame: every frame slot owns TWO grids and
// counter blocks, and classify_kernel of frame k clears the pair frame k + 1 of the slot will use while it reads the seed points out
// of it (see classify_kernel).  Used at set-up, and per frame by the unbinned path.
__global__ void zgrid_init_kernel(unsigned long long* zgrid, int cells, FrameCounters* counters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  pdl_launch_dependents();
  if (i < cells) zgrid[i] = kZInit;
  if (i < (int)(sizeof(FrameCounters) / sizeof(unsigned int))) reinterpret_cast<unsigned int*>(counters)[i] = 0u;
}

void launch_zgrid_init(unsigned long long* zgrid, int cells, FrameCounters* counters, cudaStream_t s) {
  const int n = cells > 32 ? cells : 32;
  zgrid_init_kernel<<<(n + 255) / 256, 256, 0, s>>>(zgrid, cells, counters);
}

// ---------------------------------------------------------------------------------------------------------------------
// Pass 1, stage A: conservative FP32 pre-filter (branch-free, so that the kUnroll points of a lane interleave).
//
// The reference's decisions (view_culling.cpp:76-168) are taken on the DOUBLE projection / norm of the float camera-frame
// point.  Evaluating them in FP32 (with FMAs) costs a fraction of the FP64 path, and a running error bound tells when the
// FP32 value is far enough from every decision boundary for the FP64 evaluation to be unnecessary:
//   |u_f32 - u_exact| <= E = 64 eps * (max|f| * M + max|c|) + slack,
//   M = (1 + |k1| r2 + |k2| r4 + |k3| r6) * max(|x|,|y|) + 3 (|p1| + |p2|) r2  >=  the polynomial on absolute values
// (each term carries a relative error of at most ~31 eps: 3 eps from the reciprocal times the degree <= 7 of the term,
// plus one eps per operation; the FP64 evaluation, its cast to float and the multiplication by 1/ds add < 4 eps; 64 eps
// leaves a factor ~2; `slack` absorbs the 1-ulp fuzz of the reference's float division by ds).
// A point is dropped here only when one of these is CERTAIN:
//   * its cell lies outside the Wc x Hc z-grid (never kept by pass 2, :162-165), or
//   * its cell is known and ||p|| (lower bound) exceeds the cell's current minimum + tol: the minimum only decreases, so
//     pass 2 (:157) rejects it and it can never lower the minimum itself.
// Anything non-finite makes the comparisons false, which sends the point to the exact path.
// Returns: bit 0 = needs the exact FP64 path; bit 1 = certainly a pass-1 candidate (:96) that is dropped here (kCount only).
// ---------------------------------------------------------------------------------------------------------------------
struct PreA {
  int cell;            // known cell, or 0
  float dlo;           // lower bound of ||p||^2
  unsigned int flags;  // 1 cell known, 2 certainly outside the z-grid, 4 certainly not a candidate, 8 certainly a candidate
};

template <bool kCount>
__device__ __forceinline__ PreA prefilter_a(const PreF& c, float X, float Y, float Z) {
  // rcp.approx.ftz: 1 ulp; Z == 0 / denormal -> inf -> NaN or inf below -> exact path; |Z| >= 2^126 would flush the
  // reciprocal to 0 and give a finite but wrong x, so such Z are turned into NaN explicitly
  float rz;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rz) : "f"(Z));
  rz = (Z < 1.0e30f) ? rz : __int_as_float(0x7fc00000);
  const float x = X * rz, y = Y * rz;
  const float ax = fabsf(x), ay = fabsf(y);
  const float x2 = x * x, y2 = y * y;
  const float r2 = x2 + y2;
  const float rc = fmaf(r2, fmaf(r2, fmaf(r2, c.k3, c.k2), c.k1), 1.0f);
  const float arc = fmaf(r2, fmaf(r2, fmaf(r2, c.ak3, c.ak2), c.ak1), 1.0f);
  const float t1 = (x + x) * y;
  const float t2 = fmaf(2.0f, x2, r2), t3 = fmaf(2.0f, y2, r2);
  const float xd = fmaf(rc, x, fmaf(c.p1, t1, c.p2 * t2));
  const float yd = fmaf(rc, y, fmaf(c.p1, t3, c.p2 * t1));
  const float u = fmaf(c.fx, xd, c.cx), v = fmaf(c.fy, yd, c.cy);
  const float m = fmaf(arc, fmaxf(ax, ay), c.apt * r2);
  const float e = fmaf(fmaf(c.afmax, m, c.acmax), c.err, c.slack);
  const float ulo = u - e, uhi = u + e, vlo = v - e, vhi = v + e;
  const float qlo = ulo * c.inv_ds, qhi = uhi * c.inv_ds, rlo = vlo * c.inv_ds, rhi = vhi * c.inv_ds;
  // certainly inside / certainly outside the z-grid (cells are trunc(u/ds): (-1,1) -> 0)
  const bool in_grid = (qlo > -1.0f) && (qhi < c.wc) && (rlo > -1.0f) && (rhi < c.hc);
  const bool out_grid = (qhi <= -1.0f) || (qlo >= c.wc) || (rhi <= -1.0f) || (rlo >= c.hc);
  const int ix0 = __float2int_rz(qlo), ix1 = __float2int_rz(qhi), iy0 = __float2int_rz(rlo), iy1 = __float2int_rz(rhi);
  const bool cell_known = in_grid && ix0 == ix1 && iy0 == iy1;
  PreA o;
  o.cell = cell_known ? iy0 * c.Wc + ix0 : 0;
  o.dlo = fmaf(Z, Z, fmaf(Y, Y, X * X)) * c.dscale;  // <= the exact squared norm
  o.flags = (cell_known ? 1u : 0u) | (out_grid ? 2u : 0u);
  if (kCount) {
    const bool not_cand = (uhi <= c.lo) || (vhi <= c.lo) || (ulo >= c.u_hi_cand) || (vlo >= c.v_hi_cand);
    const bool is_cand = (ulo > c.lo) && (vlo > c.lo) && (uhi < c.u_hi_cand) && (vhi < c.v_hi_cand);
    o.flags |= (not_cand ? 4u : 0u) | (is_cand ? 8u : 0u);
  }
  return o;
}

// Second half: `curbits` = high word of the cell's packed key (any value the cell held during this frame is an upper bound of
// its final minimum).  Returns bit 0 = needs the exact path, bit 1 = dropped here but certainly a pass-1 candidate.
template <bool kCount>
__device__ __forceinline__ unsigned int prefilter_b(const PreF& c, const PreA& a, unsigned int curbits, bool depth_culling) {
  const float lim = __fadd_ru(__uint_as_float(curbits), c.tol_up);  // >= current minimum + tol >= 0  (FLT_MAX -> inf)
  const bool occluded = (a.flags & 1u) && (a.dlo > __fmul_ru(lim, lim));
  const bool dropped = (a.flags & 2u) || occluded;
  bool need;
  if (!kCount) {
    need = depth_culling ? !dropped : true;
    return need ? 1u : 0u;
  }
  const bool cand_known = (a.flags & 12u) != 0u;
  need = depth_culling ? !(cand_known && dropped) : !(a.flags & 4u);
  return (need ? 1u : 0u) | ((!need && (a.flags & 8u)) ? 2u : 0u);
}

// ---------------------------------------------------------------------------------------------------------------------
// Pass 1, stage B: the exact FP64 body for one camera-frame float point.  Returns true when a record must be emitted.
// ---------------------------------------------------------------------------------------------------------------------
// `idx` goes into the record (the accumulator slot: original index, or sorted position when pre-binned); `key_idx` goes
// into the packed z-grid key (tie-break among equal fl32 distances; the position of the point in the device cloud).
__device__ __forceinline__ bool exact_point(const FrameParams& fp, unsigned long long* __restrict__ zgrid, float x, float y, float z,
                                            uint32_t idx, uint32_t key_idx, Record& rec, bool& candidate) {
  candidate = false;
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
  const unsigned long long key = ((unsigned long long)__float_as_uint(df) << 32) | (unsigned long long)key_idx;
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
// scan_kernel: per warp, kUnroll x 32 points are loaded (coalesced float4), transformed in float (exact, PCL order) and
// pre-filtered; survivors are compacted into a warp-private shared-memory queue so that stage B runs with all 32 lanes
// busy.  Records leave through a second warp-private queue, 32 records (1 KB) per global atomicAdd.
// ---------------------------------------------------------------------------------------------------------------------
#ifndef PCP_UNROLL
#define PCP_UNROLL 4
#endif
#ifndef PCP_SCAN_CTAS
#define PCP_SCAN_CTAS 4
#endif
// The early-z read of stage A goes through L1: a stale value is still an upper bound of the cell's final minimum (it only
// decreases), and per-SM copies remove the L2 hot spot of the cells around the vanishing point (measured: 109 -> 77 us).
#ifdef PCP_ZLOAD_L2
#define PCP_ZLOAD(p) __ldcg(p)
#else
#define PCP_ZLOAD(p) __ldca(p)
#endif
// The per-chunk read of the pre-binned scan (one load per chunk and cell) goes to L2: it is rare enough not to create a
// hot spot, and a fresh minimum is what lets whole chunks be skipped.
#ifdef PCP_ZLOAD0_L1
#define PCP_ZLOAD0(p) __ldca(p)
#else
#define PCP_ZLOAD0(p) __ldcg(p)
#endif
static constexpr int kUnroll = PCP_UNROLL;
static constexpr int kGroups = kChunk / 32;  // pre-binned scan: one chunk = one warp iteration of kGroups x 32 points
static_assert(kGroups >= 1 && kGroups * 32 == kChunk, "chunk size must be a multiple of 32");

// ---------------------------------------------------------------------------------------------------------------------
// classify_kernel (pre-binned cloud): one lane per chunk runs classify_chunk; chunks that must be processed are appended to
// `active` (one atomicAdd per warp), chunks that are certainly "all candidates, none in the grid" only add their point
// count to n_candidates.  Slot s of the sequence is chunk (s * chunk_mul) % n_chunks.
// ---------------------------------------------------------------------------------------------------------------------
template <bool kCount>
__device__ __forceinline__ void classify_body(const ScanArgs& a, const FrameParams& fp) {
  const int lane = threadIdx.x & 31;
  unsigned int ncand = 0;
  pdl_launch_dependents();  // the scan kernel's CTAs may take their places (they wait for this grid before reading its list)
  pdl_wait();               // the previous kernel of the stream (the slot's previous frame, or the set-up)
  // The first seed_blocks CTAs own one cell each of the slot's OTHER z-grid: the grid the slot's previous frame ended with, and
  // the one its next frame will start from.  Each thread reads its cell's winner, puts the cell back to "empty", and — seed
  // points — sends that winner through the exact path of THIS frame first.  The point that was nearest in a cell of the previous
  // frame is very likely near the front of the new view too (consecutive keyframes are centimetres apart), and processing it first
  // is nothing but a different processing order (the final grid is min over all in-grid candidates whatever the order,
  // SURVEY.md 3.4), but the scan kernel then starts against minima that are already close to final: more chunks fail the
  // distance test without being loaded and fewer points reach the FP64 stage or leave a record.  Seeds emit no record and are not
  // counted; the scan kernel meets the same points again in their chunks.  A stale or foreign index (first frame of a slot, new
  // cloud) names some other point of the cloud, which is just as valid.  The same CTAs clear the next frame's counter block, so a
  // frame is three launches (classify, scan, resolve) with no separate clearing kernel.
  if ((int)blockIdx.x < a.seed_blocks) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < (int)(sizeof(FrameCounters) / sizeof(unsigned int))) reinterpret_cast<unsigned int*>(a.counters_next)[c] = 0u;
    if (c < a.cells) {
      const uint32_t pos = (uint32_t)__ldcg(a.zprev + c);  // index word = position in the (sorted) cloud
      a.zprev[c] = kZInit;
      if (a.seed && pos < a.n) {
        const float4 p = __ldg(a.cloud + pos);
        float x = p.x, y = p.y, z = p.z;
        if (fp.use_w2c) transform_point(fp.w2c, p.x, p.y, p.z, x, y, z);
        Record rec;
        bool cand;
        (void)exact_point(fp, a.zgrid, x, y, z, pos, pos, rec, cand);
      }
    }
    return;
  }
  const uint32_t slot = (blockIdx.x - a.seed_blocks) * blockDim.x + threadIdx.x;  // one lane per visit slot; whole warps take the same path
  uint32_t chunk = 0;
  int cls = kChunkSkip;
  ChunkInfo info{0u, 0u, 0.0f};
  if (slot < a.n_chunks) {
    chunk = __ldg(a.order + slot);
    cls = kChunkProcess;
    if (fp.use_prefilter) {  // with a double T (pcp_cull) the gate is taken after T: no chunk culling
      const float4 a0 = __ldg(a.aabb + 2 * slot), a1 = __ldg(a.aabb + 2 * slot + 1);
      cls = classify_chunk(fp, a0, a1, kCount, info);
      if (kCount && cls == kChunkAllCandidates) ncand += (unsigned int)a1.w;
    }
  }
  // one global atomic per CTA (same-address atomics serialise in L2: one per warp was 4.9 k of them at C2)
  __shared__ unsigned int s_cnt[8], s_base;
  const unsigned int m = __ballot_sync(0xffffffffu, cls == kChunkProcess);
  const int warp = threadIdx.x >> 5;
  if (lane == 0) s_cnt[warp] = (unsigned int)__popc(m);
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int tot = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) { const unsigned int c = s_cnt[w]; s_cnt[w] = tot; tot += c; }
    s_base = tot ? atomicAdd(&a.counters->n_active_blocks, tot) : 0u;
  }
  __syncthreads();
  if (cls == kChunkProcess)
    a.active[s_base + s_cnt[warp] + __popc(m & ((1u << lane) - 1u))] = make_uint4(chunk, info.rect, info.org, __float_as_uint(info.d2lo));
  if (kCount) {
    for (int o = 16; o > 0; o >>= 1) ncand += __shfl_down_sync(0xffffffffu, ncand, o);
    if (lane == 0 && ncand) atomicAdd(&a.counters->n_candidates, ncand);
  }
}
template <bool kCount>
__global__ void __launch_bounds__(256) classify_kernel(const __grid_constant__ ScanArgs a, const __grid_constant__ FrameParams fp) {
  classify_body<kCount>(a, fp);
}
// Several frames per launch (small clouds, pcp_submit_frames_device): blockIdx.y picks the frame's arguments.  The frames are on
// different slots (own z-grids, lists, record buffers), so their CTAs are as independent as CTAs of different launches would be.
template <bool kCount, int K>
__global__ void __launch_bounds__(256) classify_multi_kernel(const __grid_constant__ FramePack<ScanArgs, K> p) {
  classify_body<kCount>(p.a[blockIdx.y], p.fp[blockIdx.y]);
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <bool kCount>
__global__ void __launch_bounds__(kScanThreads, PCP_SCAN_CTAS) scan_kernel(const ScanArgs a, const __grid_constant__ FrameParams fp) {
  __shared__ float4 cand_queue[kWarpsPerCta][kCandCap];
  __shared__ Record rec_queue[kWarpsPerCta][kQueueCap];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const unsigned int lt_mask = (1u << lane) - 1u;
  float4* cq = cand_queue[warp];
  Record* rq = rec_queue[warp];
  int cn = 0, rn = 0;
  unsigned int ncand = 0;
  const bool prefilter_on = fp.use_prefilter != 0;
  const bool depth_culling = fp.depth_culling != 0;

  // stage B on the last `cnt` (<= 32) entries of the candidate queue, then record compaction / flush
  auto stage_b = [&](int cnt) {
    Record rec;
    bool cand = false, emit = false;
    cn -= cnt;
    if (lane < cnt) {
      const float4 c = cq[cn + lane];
      const uint32_t slot = __float_as_uint(c.w);
      emit = exact_point(fp, a.zgrid, c.x, c.y, c.z, slot, slot, rec, cand);
    }
    if (kCount) ncand += cand ? 1u : 0u;
    const unsigned int m = __ballot_sync(0xffffffffu, emit);
    if (m) {
      if (emit) rq[rn + __popc(m & lt_mask)] = rec;
      rn += __popc(m);
      __syncwarp();
      if (rn >= 32) {
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(&a.counters->n_records, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        rn -= 32;
        const uint4* src = reinterpret_cast<const uint4*>(rq + rn);
        uint4* dst = reinterpret_cast<uint4*>(a.records + base);
        dst[lane] = src[lane];            // 64 x 16 B = the last 32 records, fully coalesced
        dst[lane + 32] = src[lane + 32];
        __syncwarp();
      }
    }
  };
  auto drain = [&]() {
    __syncwarp();
#pragma unroll 1
    while (cn >= 32) stage_b(32);
  };

  // stage A for kUnroll points per lane: float transform (exact, PCL order), pre-filter, compaction into the queue
  auto stage_a = [&](const bool (&valid)[kUnroll], const float4 (&p)[kUnroll], const uint32_t (&idx)[kUnroll]) {
    float x[kUnroll], y[kUnroll], z[kUnroll];
    bool push[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; k++) {
      x[k] = p[k].x; y[k] = p[k].y; z[k] = p[k].z;
      if (fp.use_w2c) transform_point(fp.w2c, p[k].x, p[k].y, p[k].z, x[k], y[k], z[k]);  // pcl::transformPointCloud (cpp:521)
    }
    if (prefilter_on) {
      PreA pa[kUnroll];
      unsigned int cur[kUnroll];
#pragma unroll
      for (int k = 0; k < kUnroll; k++) pa[k] = prefilter_a<kCount>(fp.pre, x[k], y[k], z[k]);
#pragma unroll
      for (int k = 0; k < kUnroll; k++) {  // only lanes with a known cell touch the grid (no hot spot on cell 0)
        cur[k] = 0x7f7fffffu;
        if ((pa[k].flags & 1u) && !(z[k] < 0.0f)) cur[k] = PCP_ZLOAD(reinterpret_cast<const unsigned int*>(a.zgrid) + 2 * pa[k].cell + 1);
      }
#pragma unroll
      for (int k = 0; k < kUnroll; k++) {
        const unsigned int f = prefilter_b<kCount>(fp.pre, pa[k], cur[k], depth_culling);
        const bool front = valid[k] && !(z[k] < 0.0f);  // view_culling.cpp:76 (exact: double(z) < 0 <=> z < 0)
        push[k] = front && (f & 1u);
        if (kCount) ncand += (front && (f & 2u)) ? 1u : 0u;
      }
    } else {
#pragma unroll
      for (int k = 0; k < kUnroll; k++) push[k] = valid[k];
    }
#pragma unroll
    for (int k = 0; k < kUnroll; k++) {
      const unsigned int m = __ballot_sync(0xffffffffu, push[k]);
      if (push[k]) cq[cn + __popc(m & lt_mask)] = make_float4(x[k], y[k], z[k], __uint_as_float(idx[k]));
      cn += __popc(m);
    }
  };

  const uint32_t chunk = 32u * kUnroll;
  const unsigned long long stride = (unsigned long long)gridDim.x * kWarpsPerCta * chunk;
  for (unsigned long long base64 = (unsigned long long)(blockIdx.x * kWarpsPerCta + warp) * chunk; base64 < a.n; base64 += stride) {
    const uint32_t base = (uint32_t)base64;
    float4 p[kUnroll];
    bool valid[kUnroll];
    uint32_t idx[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; k++) {
      idx[k] = base + k * 32 + lane;
      valid[k] = idx[k] < a.n;
      p[k] = valid[k] ? __ldcs(a.cloud + idx[k]) : make_float4(0.f, 0.f, -1.f, 0.f);
    }
    stage_a(valid, p, idx);
    drain();
  }

  __syncwarp();
  if (cn > 0) stage_b(cn);  // tail of the candidate queue
  // flush the tail of the record queue
  if (rn > 0) {
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(&a.counters->n_records, (unsigned int)rn);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (lane < rn) a.records[base + lane] = rq[lane];
  }
  if (kCount) {  // candidate count: warp reduce, one atomic per warp
    for (int o = 16; o > 0; o >>= 1) ncand += __shfl_down_sync(0xffffffffu, ncand, o);
    if (lane == 0 && ncand) atomicAdd(&a.counters->n_candidates, ncand);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// scan_binned_kernel (pre-binned cloud): three compaction stages per warp, each running with all 32 lanes busy.
//   A0  per chunk: the chunk's cell rectangle (ChunkInfo, from classify_kernel) gives lim2 = max over its cells of
//       (current minimum + tol)^2, read once by the warp (one lane per cell).  A chunk whose distance bound exceeds lim2 is
//       skipped without loading a point.  Otherwise its 128 points are loaded, transformed in float (exact, PCL order) and
//       tested with one compare each (z < 0, ||p||^2 > lim2): ~30 instructions per 32 points instead of the ~110 of the
//       projection pre-filter, for the bulk of the occluded points.
//   A1  the survivors run the FP32 projection pre-filter of prefilter_a/b against their own cell, kA1 x 32 at a time.
//   B   the exact FP64 body (exact_point) and the record queue, as in scan_kernel.
// ---------------------------------------------------------------------------------------------------------------------
#ifndef PCP_A1W
#define PCP_A1W 2
#endif
static constexpr int kA1 = PCP_A1W;
#ifndef PCP_ENT_DIST
#define PCP_ENT_DIST 8
#endif
static constexpr int kEntDist = PCP_ENT_DIST;  // prefetch distance in tickets (one round of the CTA's warps)
static constexpr int kEntCap = 320;  // list entries staged per CTA (PCP_ENT_SMEM): 5 KB; C2 has <= 264 per CTA
static constexpr int kQ0Cap = 32 * (kGroups + 1);  // < 32 left over + one chunk
static constexpr int kQ1Cap = 32 * (kA1 + 1);      // < 32 left over + one A1 batch
#ifndef PCP_RQ_FLUSH
#define PCP_RQ_FLUSH 32
#endif
static constexpr int kRqFlush = PCP_RQ_FLUSH;   // records per global atomicAdd on n_records (same-address atomics serialise in L2)
static constexpr int kRqCap = kRqFlush + 32;    // < kRqFlush left over + one stage-B batch
#ifdef PCP_SCAN_V1  // the round-1 kernel, kept for one A/B measurement against scan_ring_kernel
static constexpr size_t kBinnedSmem = (size_t)kWarpsPerCta * ((kQ0Cap + kQ1Cap) * sizeof(float4) + kRqCap * sizeof(Record));

template <bool kCount>
__global__ void __launch_bounds__(kScanThreads, PCP_SCAN_CTAS) scan_binned_kernel(const ScanArgs a, const __grid_constant__ FrameParams fp) {
  extern __shared__ __align__(16) unsigned char binned_smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const unsigned int lt_mask = (1u << lane) - 1u;
  float4* q0 = reinterpret_cast<float4*>(binned_smem) + (size_t)warp * kQ0Cap;
  float4* q1 = reinterpret_cast<float4*>(binned_smem) + (size_t)kWarpsPerCta * kQ0Cap + (size_t)warp * kQ1Cap;
  Record* rq = reinterpret_cast<Record*>(binned_smem + (size_t)kWarpsPerCta * (kQ0Cap + kQ1Cap) * sizeof(float4)) + (size_t)warp * kRqCap;
  int n0 = 0, n1 = 0, rn = 0;
  unsigned int ncand = 0;
#ifdef PCP_STATS
  unsigned int st_skipped = 0, st_a0 = 0, st_a1 = 0;  // chunks skipped whole, points kept by A0, points kept by A1 (per warp)
#endif
  const bool prefilter_on = fp.use_prefilter != 0;
  const bool depth_culling = fp.depth_culling != 0;
  const unsigned int* zhi = reinterpret_cast<const unsigned int*>(a.zgrid) + 1;

  auto stage_b = [&](int cnt) {
    Record rec;
    bool cand = false, emit = false;
    n1 -= cnt;
    if (lane < cnt) {
      const float4 c = q1[n1 + lane];
      const uint32_t slot = __float_as_uint(c.w);
      // key index = sorted position: no look-up of the original index (a scattered 4-byte read per point, DRAM latency on the
      // critical path of this stage); ties on fl32(dist) then go to the smallest sorted position, which is as deterministic,
      // and nothing ever reads the index word of the grid except the seed pass (the reference never reads its index_map)
      emit = exact_point(fp, a.zgrid, c.x, c.y, c.z, slot, slot, rec, cand);
    }
    if (kCount) ncand += cand ? 1u : 0u;
    const unsigned int m = __ballot_sync(0xffffffffu, emit);
    if (m) {
      if (emit) rq[rn + __popc(m & lt_mask)] = rec;
      rn += __popc(m);
      __syncwarp();
      if (rn >= kRqFlush) {
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(&a.counters->n_records, (unsigned int)kRqFlush);
        base = __shfl_sync(0xffffffffu, base, 0);
        rn -= kRqFlush;
        const uint4* src = reinterpret_cast<const uint4*>(rq + rn);
        uint4* dst = reinterpret_cast<uint4*>(a.records + base);
#pragma unroll
        for (int t = 0; t < kRqFlush / 16; t++) dst[lane + 32 * t] = src[lane + 32 * t];
        __syncwarp();
      }
    }
  };

  // A1 on the last min(n0, 32 kA1) entries of q0
  auto stage_a1 = [&]() {
    const int cnt = n0 < 32 * kA1 ? n0 : 32 * kA1;
    n0 -= cnt;
    float x[kA1], y[kA1], z[kA1], w[kA1];
    bool ok[kA1], push[kA1];
    unsigned int cur[kA1];
#pragma unroll
    for (int k = 0; k < kA1; k++) {
      const int i = k * 32 + lane;
      ok[k] = i < cnt;
      const float4 c = q0[n0 + (ok[k] ? i : 0)];
      x[k] = c.x; y[k] = c.y; z[k] = c.z; w[k] = c.w;
      cur[k] = kZBoundInit;  // FLT_MAX: always an upper bound (cell unknown, or pre-filter off)
    }
    if (prefilter_on) {
      PreA pa[kA1];
#pragma unroll
      for (int k = 0; k < kA1; k++) pa[k] = prefilter_a<kCount>(fp.pre, x[k], y[k], z[k]);
#pragma unroll
      for (int k = 0; k < kA1; k++)
        if (ok[k] && (pa[k].flags & 1u)) cur[k] = PCP_ZLOAD(zhi + 2 * pa[k].cell);
#pragma unroll
      for (int k = 0; k < kA1; k++) {
        const unsigned int f = prefilter_b<kCount>(fp.pre, pa[k], cur[k], depth_culling);
        push[k] = ok[k] && (f & 1u);
        if (kCount) ncand += (ok[k] && (f & 2u)) ? 1u : 0u;
      }
    } else {
#pragma unroll
      for (int k = 0; k < kA1; k++) push[k] = ok[k];
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < kA1; k++) {
      const unsigned int m = __ballot_sync(0xffffffffu, push[k]);
      if (push[k]) q1[n1 + __popc(m & lt_mask)] = make_float4(x[k], y[k], z[k], w[k]);
      n1 += __popc(m);
#ifdef PCP_STATS
      st_a1 += __popc(m);
#endif
    }
    __syncwarp();
#pragma unroll 1
    while (n1 >= 32) stage_b(32);
  };

  // Work distribution.  classify_kernel lists the active chunks in a pseudo-random order (chunks in flight are spread over
  // the whole map, like a shuffled cloud at chunk granularity).  CTA c owns items c, c + G, c + 2G, ...; its warps take them
  // one at a time from a shared-memory ticket (chunk costs differ by orders of magnitude: skipped whole vs full of FP64
  // work).  The next item's list entry is loaded and its points are pulled into L2 while the current one is processed.
  const uint32_t n_active = a.counters->n_active_blocks;
  __shared__ unsigned int s_ticket;
#ifndef PCP_NO_ENT_SMEM
  // The CTA's share of the list (items c, c + G, ...) is staged in shared memory by one bulk load, so that taking an item is
  // one shared-memory atomic plus one broadcast LDS (it was ATOMS -> LDG of the entry from L2 -> prefetch address, a chain of
  // ~1 us per chunk on the warp's critical path).  The points of item j + 8 are pulled into L2 by the warp that takes item j:
  // the prefetch distance is one round of the CTA's warps (~2 us at C2, DRAM latency is < 1 us) and costs no registers.
  __shared__ uint4 s_ent[kEntCap];
  const uint32_t n_mine = n_active > blockIdx.x ? (n_active - blockIdx.x + gridDim.x - 1u) / gridDim.x : 0u;
  for (uint32_t j = threadIdx.x; j < n_mine && j < (uint32_t)kEntCap; j += blockDim.x) s_ent[j] = __ldg(a.active + blockIdx.x + gridDim.x * j);
  if (threadIdx.x == 0) s_ticket = 0u;
  __syncthreads();
  auto entry = [&](uint32_t j) -> uint4 { return j < (uint32_t)kEntCap ? s_ent[j] : __ldg(a.active + blockIdx.x + gridDim.x * j); };
  auto pull = [&](uint32_t j) {
    if (j < n_mine && lane < kChunk / 8) prefetch_l2(a.cloud + (size_t)entry(j).x * kChunk + lane * 8);  // the chunk in lines of 128 B
  };
  for (uint32_t j = (uint32_t)warp; j < (uint32_t)kEntDist; j += kWarpsPerCta) pull(j);
#else
  if (threadIdx.x == 0) s_ticket = 0u;
  __syncthreads();
  auto take = [&]() -> uint32_t {
    uint32_t j = 0;
    if (lane == 0) j = atomicAdd(&s_ticket, 1u);
    j = __shfl_sync(0xffffffffu, j, 0);
    return blockIdx.x + gridDim.x * j;
  };
#endif
  const uint4 kNoEnt = make_uint4(0u, 0u, 0u, 0u);
  // one lane per cell of the entry's rectangle: (minimum + tol)^2 rounded up, as bits (0 for the other lanes / no rectangle)
  auto rect_bits = [&](const uint4& e) -> unsigned int {
    const int w = (int)(e.y & 63u), h = (int)((e.y >> 6) & 63u);
    unsigned int bits = 0u;
    if (lane < w * h) {
      // lane / w without the integer-division sequence: (lane + 0.5) / w lies >= 0.5 / 32 away from every integer and the
      // approximate quotient is off by < 2^-16, so the truncation is exact for lane < 32, w <= 32
      const int cy = __float2int_rz(__fdividef((float)lane + 0.5f, (float)w)), cx = lane - cy * w;
      const int cell = ((int)(e.z >> 16) + cy) * fp.Wc + (int)(e.z & 0xffffu) + cx;
      const float lim = __fadd_ru(__uint_as_float(PCP_ZLOAD0(zhi + 2 * cell)), fp.pre.tol_up);  // >= minimum + tol >= 0
      bits = __float_as_uint(__fmul_ru(lim, lim));
    }
    return bits;
  };
  // The next item's list entry is loaded (and its points pulled into L2) while the current one is processed.  Deeper
  // pipelining (z-grid cells of the next item, record blocks reserved ahead, stage A1's z word reused by stage B) was
  // measured slower: the kernel sits at its 64-register budget and every extra live value costs more than the latency it hides.
#ifndef PCP_NO_ENT_SMEM
  (void)kNoEnt;
  for (;;) {
    uint32_t tj = 0;
    if (lane == 0) tj = atomicAdd(&s_ticket, 1u);
    tj = __shfl_sync(0xffffffffu, tj, 0);
    if (tj >= n_mine) break;
    pull(tj + (uint32_t)kEntDist);
    const uint4 ent = entry(tj);
#else
  uint32_t item = take();
  uint4 ent = item < n_active ? __ldg(a.active + item) : kNoEnt;
  while (item < n_active) {
    const uint32_t nitem = take();
    uint4 nent = kNoEnt;
    if (nitem < n_active) {
      nent = __ldg(a.active + nitem);
      if (lane < kChunk / 8) prefetch_l2(a.cloud + (size_t)nent.x * kChunk + lane * 8);  // the chunk in lines of 128 B
    }
#endif
    const uint32_t base = ent.x * kChunk;
    const unsigned int zb = rect_bits(ent);
    float lim2 = __int_as_float(0x7f800000);  // +inf: nothing is dropped by the distance test
    bool skip = false;
    if (ent.y != 0u) {
      lim2 = __uint_as_float(__reduce_max_sync(0xffffffffu, zb));  // non-negative floats order like their bit patterns
      skip = __uint_as_float(ent.w) > lim2;  // the whole chunk is farther than every cell it can fall into
    }
    if (skip) {
      if (kCount && lane == 0) ncand += ent.y >> 12;  // offered only for all-candidate chunks (classify_chunk)
#ifdef PCP_STATS
      st_skipped++;
#endif
    } else {
      float4 p[kGroups];
      bool valid[kGroups];
#pragma unroll
      for (int k = 0; k < kGroups; k++) {
        const uint32_t i = base + k * 32 + lane;  // sorted position: the accumulator slot
        valid[k] = i < a.n;
        p[k] = valid[k] ? __ldcs(a.cloud + i) : make_float4(0.f, 0.f, -1.f, 0.f);
      }
#pragma unroll
      for (int k = 0; k < kGroups; k++) {
        float x = p[k].x, y = p[k].y, z = p[k].z;
        if (fp.use_w2c) transform_point(fp.w2c, p[k].x, p[k].y, p[k].z, x, y, z);  // pcl::transformPointCloud (cpp:521)
        const bool front = valid[k] && !(z < 0.0f);  // view_culling.cpp:76 (exact: double(z) < 0 <=> z < 0)
        const float dlo = fmaf(z, z, fmaf(y, y, x * x)) * fp.pre.dscale;  // <= the exact squared norm; NaN compares false
        const bool far = dlo > lim2;
        if (kCount) ncand += (front && far) ? 1u : 0u;
        const bool keep = front && !far;
        const unsigned int m = __ballot_sync(0xffffffffu, keep);
        if (keep) q0[n0 + __popc(m & lt_mask)] = make_float4(x, y, z, __uint_as_float(base + k * 32 + lane));
        n0 += __popc(m);
#ifdef PCP_STATS
        st_a0 += __popc(m);
#endif
      }
      __syncwarp();
#pragma unroll 1
      while (n0 >= 32) stage_a1();
    }
#ifdef PCP_NO_ENT_SMEM
    item = nitem;
    ent = nent;
#endif
  }

  __syncwarp();
  if (n0 > 0) stage_a1();
  if (n1 > 0) stage_b(n1);
#ifdef PCP_STATS
  if (lane == 0) {
    atomicAdd(&a.counters->stats[0], st_skipped);
    atomicAdd(&a.counters->stats[1], st_a0);
    atomicAdd(&a.counters->stats[2], st_a1);
  }
#endif
  if (rn > 0) {
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(&a.counters->n_records, (unsigned int)rn);
    base = __shfl_sync(0xffffffffu, base, 0);
    for (int i = lane; i < rn; i += 32) a.records[base + i] = rq[i];
  }
  if (kCount) {
    for (int o = 16; o > 0; o >>= 1) ncand += __shfl_down_sync(0xffffffffu, ncand, o);
    if (lane == 0 && ncand) atomicAdd(&a.counters->n_candidates, ncand);
  }
}

#endif  // PCP_SCAN_V1

// ---------------------------------------------------------------------------------------------------------------------
// scan_ring_kernel (pre-binned cloud): pass 1 as a warp-specialised pipeline.
//
//   scout warp (the CTA's last warp)
//     walks the CTA's share of the active-chunk list 32 entries at a time, ONE LANE PER ENTRY: the lane reads the current
//     minima of the entry's z-grid cell rectangle (ld.cg, L2) and decides whether the whole chunk can be skipped (its
//     distance bound exceeds (minimum + tol)^2 of every cell it can fall into: the reference's pass 2, view_culling.cpp:157,
//     rejects all of its points and none can lower a minimum).  The ids of the chunks that survive go, in one store for the
//     whole batch, into a small shared-memory work ring {chunk, ticket number}.  The scout stays at most kAhead chunks ahead
//     of the workers, so that it tests against minima that are as fresh as the old one-chunk-at-a-time kernel saw.
//   worker warps (the other seven)
//     take the ring's entries in order (one shared-memory ticket per chunk) and fetch the chunk THEMSELVES with one bulk
//     asynchronous copy (cp.async.bulk.shared.global, 1 KB = 64 float4, completion on the warp's own mbarrier) into one of the
//     warp's two buffers; the copy of the warp's next chunk is in flight while it works on the current one.  A skipped chunk
//     is never read from HBM and nothing is prefetched on speculation.  Then the stages that decide: exact float transform
//     (PCL order), FP32 projection pre-filter against the point's own cell (prefilter_a/b), compaction, and the exact FP64
//     body (exact_point) 32 dense entries at a time; records leave through a warp-private queue, 32 records (1 KB, coalesced)
//     per global atomicAdd.
//
// The split takes the per-entry work (ticket, list entry, rectangle test: a chain of dependent shared / L2 accesses that cost
// a whole warp ~100 instructions per entry, 47 % of them for chunks that are then skipped) off the warps that do the
// arithmetic and runs it 32 entries wide.  Every mbarrier is waited on only by the warp that armed it, which also tracks its
// phase: no cross-warp phase bookkeeping (bulk copies complete out of order; a shared ring of slots whose consumers wait by
// parity alone can be passed by a consumer that is a whole lap ahead).  Work distribution between CTAs: entries c, c + G, ...
// of the first (100 - kPoolPct) % of the list are CTA c's own; the rest is a common pool the scouts draw from in batches with
// one global atomic each, which evens out the end of the launch (chunk costs differ by orders of magnitude).
// ---------------------------------------------------------------------------------------------------------------------
#ifndef PCP_AHEAD
#define PCP_AHEAD 16
#endif
#ifndef PCP_POLL_NS
#define PCP_POLL_NS 100
#endif
#ifndef PCP_POOL_PCT
#define PCP_POOL_PCT 30
#endif
#ifndef PCP_POOL_BATCH
#define PCP_POOL_BATCH 16
#endif
#ifndef PCP_CHUNKS_PER_CTA
#define PCP_CHUNKS_PER_CTA 128
#endif
#ifndef PCP_RING_THREADS
#define PCP_RING_THREADS 256
#endif
#ifndef PCP_RING_CTAS
#define PCP_RING_CTAS 4
#endif
static constexpr int kRingThreads = PCP_RING_THREADS;
static constexpr int kWorkers = kRingThreads / 32 - 1;  // warps 0 .. kWorkers-1 work, warp kWorkers scouts
static constexpr int kWorkCap = 128;                    // work-ring entries (power of two, >= kAhead + 32 + kWorkers)
static constexpr int kAhead = PCP_AHEAD;                // the scout tests at most this many chunks ahead of the tickets taken
static constexpr int kPoolPct = PCP_POOL_PCT;           // share of the active list handed out dynamically
static constexpr int kPoolBatch = PCP_POOL_BATCH;       // pool entries per global atomic (<= 32)
static constexpr uint32_t kWorkFree = 0xffffffffu;
static constexpr size_t kBufBytes = (size_t)kWorkers * 2 * kChunk * sizeof(float4);  // two chunk buffers per worker warp
static constexpr int kRingQ1Cap = 32 * (kGroups + 1);  // < 32 left over + one chunk
static constexpr size_t kRingSmem = kBufBytes + (size_t)kWorkers * (kRingQ1Cap * sizeof(float4) + kRqCap * sizeof(Record));
static_assert(kGroups >= 1 && kGroups <= 4, "scan_ring_kernel: kGroups points per worker lane");
static_assert(kPoolBatch >= 1 && kPoolBatch <= 32, "pool batch: one lane per entry");
static_assert((kWorkCap & (kWorkCap - 1)) == 0 && kWorkCap >= kAhead + 32 + kWorkers, "work ring too small");

__device__ __forceinline__ uint2 lds_volatile_v2(const uint2* p) {
  uint2 v;
  asm volatile("ld.volatile.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void sts_volatile_v2(uint2* p, uint2 v) {
  asm volatile("st.volatile.shared.v2.u32 [%0], {%1, %2};" ::"r"(smem_u32(p)), "r"(v.x), "r"(v.y) : "memory");
}

#ifdef PCP_TIMELINE  // tools/ variant only (tools/scan_timeline.py): when each CTA of the last scan_ring launch reached its milestones
__device__ unsigned long long g_timeline[2048][8];
__device__ __forceinline__ void tl_stamp(int k) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  if (blockIdx.x < 2048u) g_timeline[blockIdx.x][k] = t;
}
#define TL(k) tl_stamp(k)
#define TLV(k, v) do { if (blockIdx.x < 2048u) g_timeline[blockIdx.x][k] = (unsigned long long)(v); } while (0)
#else
#define TL(k) ((void)0)
#define TLV(k, v) ((void)0)
#endif

template <bool kCount>
__device__ __forceinline__ void scan_ring_body(const ScanArgs& a, const FrameParams& fp) {
  extern __shared__ __align__(128) unsigned char ring_smem[];
  __shared__ __align__(8) uint64_t s_bar[kWorkers][2];  // completion of the bulk copy into buffer b of worker w
  __shared__ __align__(8) uint2 s_work[kWorkCap];       // {chunk id, ticket number} or {_, kWorkFree}
  __shared__ volatile unsigned int s_total;             // number of work entries of this CTA, 0xffffffff until the scout is done
  __shared__ unsigned int s_ticket;
  __shared__ uint4 s_ent[kEntCap];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const unsigned int lt_mask = (1u << lane) - 1u;
  unsigned int ncand = 0;
  const bool prefilter_on = fp.use_prefilter != 0;
  const bool depth_culling = fp.depth_culling != 0;
  const unsigned int* zhi = reinterpret_cast<const unsigned int*>(a.zgrid) + 1;

  if (threadIdx.x < kWorkCap) s_work[threadIdx.x] = make_uint2(0u, kWorkFree);
  if (threadIdx.x == 0) {
    TL(0);
    for (int w = 0; w < kWorkers; w++) { mbar_init(&s_bar[w][0], 1u); mbar_init(&s_bar[w][1], 1u); }
    s_ticket = 0u;
    s_total = 0xffffffffu;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  pdl_wait();  // everything above overlaps the tail of classify_kernel (programmatic dependent launch); its list is read from here on
  // this CTA may have been resident since before classify_kernel wrote its list: read that data from L2, never through L1 / the
  // read-only path (ld.global.nc promises data that is constant for the lifetime of THIS kernel, which started earlier)
  if (threadIdx.x == 0) TL(1);
  const uint32_t n_active = __ldcg(&a.counters->n_active_blocks);
  const uint32_t G = gridDim.x;
  // static share: entries c + G j, j < n_static (the same count for every CTA); pool: entries [n_static * G, n_active)
  const uint32_t n_static = (uint32_t)(((unsigned long long)n_active * (100 - kPoolPct) / 100) / G);
  const uint32_t pool0 = n_static * G;
  for (uint32_t j = threadIdx.x; j < n_static && j < (uint32_t)kEntCap; j += blockDim.x) s_ent[j] = __ldcg(a.active + blockIdx.x + G * j);
  __syncthreads();

  if (warp == kWorkers) {
    // ------------------------------------------------------------ scout ---------------------------------------------------
    uint32_t count = 0;  // work entries published so far
    auto offer = [&](bool valid, const uint4& ent) {
      // stay close to the workers: a test against minima that are kAhead chunks old skips less
      while ((int)(count - *(volatile unsigned int*)&s_ticket) > kAhead) __nanosleep(64);
      bool skip = false;
      // The rectangles hold <= 32 cells.  Every lane walks ITS entry's cells eight at a time with independent loads (one L2 round
      // trip per eight cells for the whole batch of 32 entries): this loop is what bounds how fast a scout can feed seven workers.
      const int w = (int)(ent.y & 63u), wh = valid ? w * (int)((ent.y >> 6) & 63u) : 0;
      const int wh_max = __reduce_max_sync(0xffffffffu, wh);
      if (wh > 0) {
        const unsigned int* org = zhi + 2 * ((int)(ent.z >> 16) * fp.Wc + (int)(ent.z & 0xffffu));
        // c / w == (c * inv_w) >> 16 for c < 32, w <= 32 with inv_w = ceil(65536 / w) (+1 at most from the approximate division:
        // the quotient c / w lies >= 1/32 below the next integer, the error of the product is < 2 * 32 / 65536)
        const unsigned int inv_w = (unsigned int)__float2int_ru(__fdividef(65536.0f, (float)w));
        unsigned int m = 0u;  // max over the rectangle of (minimum + tol)^2 rounded up, as bits (non-negative floats order like their bits)
        for (int c0 = 0; c0 < wh_max; c0 += 8) {
          unsigned int zb[8];
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const int c = c0 + k;
            const int cy = (int)(((unsigned int)c * inv_w) >> 16), cx = c - cy * w;
            zb[k] = c < wh ? PCP_ZLOAD0(org + 2 * (cy * fp.Wc + cx)) : 0u;
          }
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const float lim = __fadd_ru(__uint_as_float(zb[k]), fp.pre.tol_up);
            if (c0 + k < wh) m = max(m, __float_as_uint(__fmul_ru(lim, lim)));
          }
        }
        skip = __uint_as_float(ent.w) > __uint_as_float(m);  // the whole chunk is farther than every cell it can fall into
        if (kCount && skip) ncand += ent.y >> 12;             // offered only for all-candidate chunks (classify_chunk)
      }
      const bool go = valid && !skip;
      const unsigned int gm = __ballot_sync(0xffffffffu, go);
      if (go) {
        const uint32_t t = count + __popc(gm & lt_mask);
        uint2* e = &s_work[t & (kWorkCap - 1)];
        while (lds_volatile_v2(e).y != kWorkFree) __nanosleep(32);  // the worker that held the entry's previous ticket hands it back
        sts_volatile_v2(e, make_uint2(ent.x, t));
#ifndef PCP_NO_SCOUT_PREFETCH
        // the chunk WILL be read (it passed the test): start it on its way from HBM to L2 now, so that the worker's bulk copy, a few
        // tickets later, finds it there.  One instruction covers one 128-byte line of every surviving lane's chunk.
        const float4* src = a.cloud + (size_t)ent.x * kChunk;
#pragma unroll
        for (int k = 0; k < kChunk / 8; k++)
          if ((uint32_t)(k * 8) < a.n - ent.x * (uint32_t)kChunk) prefetch_l2(src + k * 8);  // not beyond the cloud's last point
#endif
      }
      count += __popc(gm);
      __syncwarp();
    };
    for (uint32_t j0 = 0; j0 < n_static; j0 += 32) {
      const uint32_t j = j0 + lane;
      const bool valid = j < n_static;
      uint4 ent = make_uint4(0u, 0u, 0u, 0u);
      if (valid) ent = j < (uint32_t)kEntCap ? s_ent[j] : __ldcg(a.active + blockIdx.x + G * j);
      offer(valid, ent);
    }
    if (lane == 0) { TL(2); TLV(6, count); }
    for (;;) {  // the common pool
      uint32_t i0 = 0;
      if (lane == 0) i0 = pool0 + atomicAdd(&a.counters->pool_next, (unsigned int)kPoolBatch);
      i0 = __shfl_sync(0xffffffffu, i0, 0);
      if (i0 >= n_active) break;
      const uint32_t i = i0 + lane;
      const bool valid = lane < kPoolBatch && i < n_active;
      uint4 ent = make_uint4(0u, 0u, 0u, 0u);
      if (valid) ent = __ldcg(a.active + i);
      offer(valid, ent);
    }
    if (lane == 0) {
      TL(3); TLV(7, count);
      s_total = count;  // tickets >= count find no work
      if (count) atomicAdd(&a.counters->stats[1], count);  // chunks fetched (1 KB bulk copies) by this CTA: the kernel's cloud traffic
    }
  } else {
    // ------------------------------------------------------------ workers -------------------------------------------------
    float4* buf = reinterpret_cast<float4*>(ring_smem) + (size_t)warp * 2 * kChunk;
    float4* q1 = reinterpret_cast<float4*>(ring_smem + kBufBytes) + (size_t)warp * kRingQ1Cap;
    Record* rq = reinterpret_cast<Record*>(ring_smem + kBufBytes + (size_t)kWorkers * kRingQ1Cap * sizeof(float4)) + (size_t)warp * kRqCap;
    int n1 = 0, rn = 0;

    auto stage_b = [&](int cnt) {
      Record rec;
      bool cand = false, emit = false;
      n1 -= cnt;
      if (lane < cnt) {
        const float4 c = q1[n1 + lane];
        const uint32_t slot = __float_as_uint(c.w);
        // key index = sorted position: ties on fl32(dist) go to the smallest sorted position (deterministic; nothing reads the
        // index word of the grid except the seed pass, and the reference never reads its index_map)
        emit = exact_point(fp, a.zgrid, c.x, c.y, c.z, slot, slot, rec, cand);
      }
      if (kCount) ncand += cand ? 1u : 0u;
      const unsigned int m = __ballot_sync(0xffffffffu, emit);
      if (m) {
        if (emit) rq[rn + __popc(m & lt_mask)] = rec;
        rn += __popc(m);
        __syncwarp();
        if (rn >= kRqFlush) {
          unsigned int base = 0;
          if (lane == 0) base = atomicAdd(&a.counters->n_records, (unsigned int)kRqFlush);
          base = __shfl_sync(0xffffffffu, base, 0);
          rn -= kRqFlush;
          const uint4* src = reinterpret_cast<const uint4*>(rq + rn);
          uint4* dst = reinterpret_cast<uint4*>(a.records + base);
#pragma unroll
          for (int t = 0; t < kRqFlush / 16; t++) dst[lane + 32 * t] = src[lane + 32 * t];
          __syncwarp();
        }
      }
    };

    // Takes the next ticket, waits for its work entry and starts the bulk copy of that chunk into buffer b.
    // `base` / `cnt`: the chunk's first point (sorted position) and point count; returns false when no work is left.
    auto fetch = [&](int b, uint32_t& base, uint32_t& cnt) -> bool {
      uint32_t chunk = kWorkFree;
      if (lane == 0) {
        const uint32_t t = atomicAdd(&s_ticket, 1u);
        uint2* e = &s_work[t & (kWorkCap - 1)];
        for (;;) {
          const uint2 v = lds_volatile_v2(e);
          if (v.y == t) { chunk = v.x; sts_volatile_v2(e, make_uint2(0u, kWorkFree)); break; }
          if (t >= s_total) break;  // the scout is done and never issued this ticket
          __nanosleep(PCP_POLL_NS);
        }
        if (chunk != kWorkFree) {
          const uint32_t b0 = chunk * kChunk;
          const uint32_t c = a.n - b0 < (uint32_t)kChunk ? a.n - b0 : (uint32_t)kChunk;
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the warp's reads of this buffer precede the async-proxy write
          mbar_arrive_expect_tx(&s_bar[warp][b], c * 16u);
          bulk_g2s(buf + (size_t)b * kChunk, a.cloud + b0, c * 16u, &s_bar[warp][b]);
        }
      }
      chunk = __shfl_sync(0xffffffffu, chunk, 0);
      if (chunk == kWorkFree) return false;
      base = chunk * kChunk;
      cnt = a.n - base < (uint32_t)kChunk ? a.n - base : (uint32_t)kChunk;
      return true;
    };

    uint32_t base[2] = {0u, 0u}, cnt[2] = {0u, 0u}, phase[2] = {0u, 0u};
    bool have[2];
    have[0] = fetch(0, base[0], cnt[0]);
    have[1] = have[0] ? fetch(1, base[1], cnt[1]) : false;
    // one chunk out of buffer B (a compile-time index: base / cnt / phase / have stay in registers)
#ifdef PCP_TIMELINE
    bool tl_first = true;
#endif
    auto step = [&](auto B) {
      constexpr int b = decltype(B)::value;
      mbar_wait(&s_bar[warp][b], phase[b]);
#ifdef PCP_TIMELINE
      if (tl_first && warp == 0 && lane == 0) TL(4);
      tl_first = false;
#endif
      phase[b] ^= 1u;
      float4 p[kGroups];
#pragma unroll
      for (int k = 0; k < kGroups; k++) p[k] = buf[(size_t)b * kChunk + k * 32 + lane];
      const uint32_t cbase = base[b], ccnt = cnt[b];
      __syncwarp();  // every lane has its points: the buffer can take the warp's next chunk while this one is processed
      have[b] = fetch(b, base[b], cnt[b]);
      float x[kGroups], y[kGroups], z[kGroups];
      bool ok[kGroups], push[kGroups];
      unsigned int cur[kGroups];
#pragma unroll
      for (int k = 0; k < kGroups; k++) {
        x[k] = p[k].x; y[k] = p[k].y; z[k] = p[k].z;
        if (fp.use_w2c) transform_point(fp.w2c, p[k].x, p[k].y, p[k].z, x[k], y[k], z[k]);  // pcl::transformPointCloud (cpp:521)
        ok[k] = (uint32_t)(k * 32 + lane) < ccnt && !(z[k] < 0.0f);  // view_culling.cpp:76 (exact: double(z) < 0 <=> z < 0)
        cur[k] = kZBoundInit;  // FLT_MAX: always an upper bound (cell unknown, or pre-filter off)
      }
      if (prefilter_on) {
        PreA pa[kGroups];
#pragma unroll
        for (int k = 0; k < kGroups; k++) pa[k] = prefilter_a<kCount>(fp.pre, x[k], y[k], z[k]);
#pragma unroll
        for (int k = 0; k < kGroups; k++)
          if (ok[k] && (pa[k].flags & 1u)) cur[k] = PCP_ZLOAD(zhi + 2 * pa[k].cell);
#pragma unroll
        for (int k = 0; k < kGroups; k++) {
          const unsigned int f = prefilter_b<kCount>(fp.pre, pa[k], cur[k], depth_culling);
          push[k] = ok[k] && (f & 1u);
          if (kCount) ncand += (ok[k] && (f & 2u)) ? 1u : 0u;
        }
      } else {
#pragma unroll
        for (int k = 0; k < kGroups; k++) push[k] = ok[k];
      }
#pragma unroll
      for (int k = 0; k < kGroups; k++) {
        const unsigned int m = __ballot_sync(0xffffffffu, push[k]);
        if (push[k]) q1[n1 + __popc(m & lt_mask)] = make_float4(x[k], y[k], z[k], __uint_as_float(cbase + k * 32 + lane));
        n1 += __popc(m);
      }
      __syncwarp();
#pragma unroll 1
      while (n1 >= 32) stage_b(32);
    };
#pragma unroll 1
    while (have[0] || have[1]) {
      if (have[0]) step(std::integral_constant<int, 0>{});
      if (have[1]) step(std::integral_constant<int, 1>{});
    }
    __syncwarp();
    if (n1 > 0) stage_b(n1);
    if (warp == 0 && lane == 0) TL(5);
    if (rn > 0) {
      unsigned int rbase = 0;
      if (lane == 0) rbase = atomicAdd(&a.counters->n_records, (unsigned int)rn);
      rbase = __shfl_sync(0xffffffffu, rbase, 0);
      for (int i = lane; i < rn; i += 32) a.records[rbase + i] = rq[i];
    }
  }
  if (kCount) {
    for (int o = 16; o > 0; o >>= 1) ncand += __shfl_down_sync(0xffffffffu, ncand, o);
    if (lane == 0 && ncand) atomicAdd(&a.counters->n_candidates, ncand);
  }
}
template <bool kCount>
__global__ void __launch_bounds__(kRingThreads, PCP_RING_CTAS) scan_ring_kernel(const __grid_constant__ ScanArgs a, const __grid_constant__ FrameParams fp) {
  scan_ring_body<kCount>(a, fp);
}
template <bool kCount, int K>
__global__ void __launch_bounds__(kRingThreads, PCP_RING_CTAS) scan_ring_multi_kernel(const __grid_constant__ FramePack<ScanArgs, K> p) {
  scan_ring_body<kCount>(p.a[blockIdx.y], p.fp[blockIdx.y]);
}

// Launch with the programmatic-stream-serialization attribute: the kernel's CTAs may start while the previous kernel of the
// stream drains; the kernel itself waits (griddepcontrol.wait) before it touches anything its predecessor wrote.
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// grids of the pre-binned pass 1: one persistent CTA per resident slot for a large cloud; a small cloud takes fewer CTAs
// (>= PCP_CHUNKS_PER_CTA chunks each), so that the kernels of the frames in flight on the other slots find free SMs instead of
// queueing behind a grid that fills the GPU (measured at C1, 15.6 k chunks: 64 / 128 per CTA)
static void pass1_grids(const ScanArgs& a, int sm_count, int* cgrid, int* rgrid) {
  int g = sm_count * PCP_RING_CTAS;
  const int by_work = (int)(a.n_chunks / (unsigned)PCP_CHUNKS_PER_CTA);
  if (by_work < g) g = by_work > sm_count / 2 ? by_work : sm_count / 2;
  *rgrid = g;
  *cgrid = a.seed_blocks + (int)((a.n_chunks + 255) / 256);
}
// the opt-in to > 48 KB of dynamic shared memory is a per-device attribute: one flag per device, not per process
static int configure_pass1(int device) {
  static bool configured[64] = {};
  if (device >= 0 && device < 64 && configured[device]) return 0;
#ifdef PCP_SCAN_V1
  if (cudaFuncSetAttribute(scan_binned_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBinnedSmem) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(scan_binned_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBinnedSmem) != cudaSuccess) return 1;
#endif
  if (cudaFuncSetAttribute(scan_ring_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmem) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(scan_ring_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmem) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(scan_ring_multi_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmem) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(scan_ring_multi_kernel<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmem) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(scan_ring_multi_kernel<false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmem) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(scan_ring_multi_kernel<true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmem) != cudaSuccess) return 1;
  if (device >= 0 && device < 64) configured[device] = true;
  return 0;
}

template <int K>
static int launch_scan_group_k(const ScanArgs* a, const FrameParams* fp, int sm_count, bool count_candidates, cudaStream_t s) {
  FramePack<ScanArgs, K> p;
  for (int k = 0; k < K; k++) { p.a[k] = a[k]; p.fp[k] = fp[k]; }
  int cgrid, rgrid;
  pass1_grids(a[0], sm_count, &cgrid, &rgrid);
  cudaError_t e;
  if (count_candidates) e = launch_pdl(classify_multi_kernel<true, K>, dim3(cgrid, K), dim3(256), 0, s, p);
  else e = launch_pdl(classify_multi_kernel<false, K>, dim3(cgrid, K), dim3(256), 0, s, p);
  if (e != cudaSuccess) return 1;
  if (count_candidates) e = launch_pdl(scan_ring_multi_kernel<true, K>, dim3(rgrid, K), dim3(kRingThreads), kRingSmem, s, p);
  else e = launch_pdl(scan_ring_multi_kernel<false, K>, dim3(rgrid, K), dim3(kRingThreads), kRingSmem, s, p);
  return e == cudaSuccess ? 0 : 1;
}
int launch_scan_group(const ScanArgs* a, const FrameParams* fp, int k, int sm_count, bool count_candidates, int device, cudaStream_t s) {
  if (!a[0].aabb || configure_pass1(device) != 0) return 1;
  if (k == 2) return launch_scan_group_k<2>(a, fp, sm_count, count_candidates, s);
  if (k == 4) return launch_scan_group_k<4>(a, fp, sm_count, count_candidates, s);
  return 1;
}

int launch_scan(const ScanArgs& a, const FrameParams& fp, int grid, int sm_count, bool count_candidates, int device, cudaStream_t s) {
  if (a.aabb) {
    int cgrid, rgrid;
    pass1_grids(a, sm_count, &cgrid, &rgrid);
    (void)rgrid;
    if (configure_pass1(device) != 0) return 1;
    cudaError_t e;
    if (count_candidates) e = launch_pdl(classify_kernel<true>, dim3(cgrid), dim3(256), 0, s, a, fp);
    else e = launch_pdl(classify_kernel<false>, dim3(cgrid), dim3(256), 0, s, a, fp);
    if (e != cudaSuccess) return 1;
#ifdef PCP_SCAN_V1
    if (count_candidates) scan_binned_kernel<true><<<grid, kScanThreads, kBinnedSmem, s>>>(a, fp);
    else scan_binned_kernel<false><<<grid, kScanThreads, kBinnedSmem, s>>>(a, fp);
#else
#ifdef PCP_NO_PDL_SCAN
    if (count_candidates) scan_ring_kernel<true><<<rgrid, kRingThreads, kRingSmem, s>>>(a, fp);
    else scan_ring_kernel<false><<<rgrid, kRingThreads, kRingSmem, s>>>(a, fp);
#else
    if (count_candidates) e = launch_pdl(scan_ring_kernel<true>, dim3(rgrid), dim3(kRingThreads), kRingSmem, s, a, fp);
    else e = launch_pdl(scan_ring_kernel<false>, dim3(rgrid), dim3(kRingThreads), kRingSmem, s, a, fp);
    if (e != cudaSuccess) return 1;
#endif
#endif
  } else {
    if (count_candidates) scan_kernel<true><<<grid, kScanThreads, 0, s>>>(a, fp);
    else scan_kernel<false><<<grid, kScanThreads, 0, s>>>(a, fp);
  }
  return 0;
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

// PCP_SAMPLE_BILINEAR (north_star (c); the reference samples the nearest pixel): 8.8 fixed-point bilinear value of the three
// channels at the double projection of the record's float point, operation by operation as oracle/pcp_oracle.cpp bilinear_q.
// Kept out of line: it is the rare mode and must not cost the nearest path registers.
__device__ __noinline__ void bilinear_sample(const uint8_t* __restrict__ bgr, size_t pitch, const FrameParams& fp, const Record* rp, int pix,
                                             uint32_t& qr, uint32_t& qg, uint32_t& qb) {
  const float rx = rp->x, ry = rp->y, rz = rp->z;
  double u, v;
  project(fp.cam, (double)rx, (double)ry, (double)rz, u, v);  // the doubles whose truncation is `pix`
  const int y0 = pix / fp.imgW, x0 = pix - y0 * fp.imgW;
  const int x1 = x0 + 1 < fp.imgW ? x0 + 1 : fp.imgW - 1, y1 = y0 + 1 < fp.imgH ? y0 + 1 : fp.imgH - 1;
  double fx = __dsub_rn(u, (double)x0), fy = __dsub_rn(v, (double)y0);
  if (fx < 0.0) fx = 0.0;  // u in (-1, 0) truncates to column 0: no extrapolation
  if (fy < 0.0) fy = 0.0;
  const double gx = __dsub_rn(1.0, fx), gy = __dsub_rn(1.0, fy);
  const uint8_t* r0 = bgr + (size_t)y0 * pitch;
  const uint8_t* r1 = bgr + (size_t)y1 * pitch;
  uint32_t q[3];
#pragma unroll
  for (int ch = 0; ch < 3; ch++) {
    const double p00 = (double)__ldg(r0 + (size_t)x0 * 3 + ch), p10 = (double)__ldg(r0 + (size_t)x1 * 3 + ch);
    const double p01 = (double)__ldg(r1 + (size_t)x0 * 3 + ch), p11 = (double)__ldg(r1 + (size_t)x1 * 3 + ch);
    const double top = __dadd_rn(__dmul_rn(p00, gx), __dmul_rn(p10, fx));
    const double bot = __dadd_rn(__dmul_rn(p01, gx), __dmul_rn(p11, fx));
    const double val = __dadd_rn(__dmul_rn(top, gy), __dmul_rn(bot, fy));
    q[ch] = (uint32_t)__double2uint_rz(__dadd_rn(__dmul_rn(val, 256.0), 0.5));
  }
  qb = q[0]; qg = q[1]; qr = q[2];
}

// kRes records per thread, processed phase by phase so that the kRes independent chains of dependent loads (record ->
// z-grid cell -> pixel -> accumulator) are in flight together: the kernel is latency-bound, not bandwidth-bound.  Only the
// fields every record needs (index, dist, cell, pixel) stay in registers; x,y,z are re-read for the rare paths that use
// them (TOP5 scores, re-projection after a double T), which keeps the kernel at 3 CTAs per SM.
// measured at C2 (serialized resolve / frame with three in flight): 4 records x 3 CTAs per SM x 6 CTAs per SM launched 20.9 / 38.0 us;
// 2 x 4 x 4 (one resident wave, 64 registers) 17.2 / 35.6 us; 1 x 4 x 12 24.1 / 41.4; 8 x 3 x 3 22.9 / 39.9
#ifndef PCP_RES
#define PCP_RES 2
#endif
#ifndef PCP_RES_CTAS
#define PCP_RES_CTAS 4
#endif
static constexpr int kRes = PCP_RES;

__device__ __forceinline__ void resolve_body(const ResolveArgs& a, const FrameParams& fp) {
  const unsigned int nrec = a.counters->n_records;
  unsigned int nvis = 0, ncol = 0;
  const unsigned int tile = blockDim.x * kRes;
  for (unsigned int base = blockIdx.x * tile; base < nrec; base += gridDim.x * tile) {
    uint32_t idx[kRes];
    int cell[kRes], pix[kRes];
    double dist[kRes];
    bool live[kRes];
#pragma unroll
    for (int k = 0; k < kRes; k++) {
      const unsigned int r = base + k * blockDim.x + threadIdx.x;
      live[k] = r < nrec;
      idx[k] = 0; cell[k] = 0; pix[k] = -1; dist[k] = 0.0;
      if (live[k]) {
        const uint4* rp = reinterpret_cast<const uint4*>(a.records + r);
        const uint4 w0 = __ldcs(rp), w1 = __ldcs(rp + 1);
        idx[k] = w0.w;
        dist[k] = __hiloint2double((int)w1.y, (int)w1.x);
        cell[k] = (int)w1.z;
        pix[k] = (int)w1.w;
      }
    }
    // The z-grid word, the pixel and the accumulator are all addressed from the record alone, so their loads are issued
    // together (speculatively for records that turn out to be occluded): two dependent latencies per record, not four.
    unsigned int hi[kRes];
    uint32_t rgb[kRes], mv[kRes];
    uint4 s[kRes];
    const bool gather = a.bgr != nullptr;
    // SUM accumulators are updated with no-return atomics (RED.ADD.U32): integer sums commute, so frames in flight on different
    // slots need no ordering between their resolve kernels and the accumulator is never read here.  TOP5 inserts are
    // read-modify-write on sorted lists: those frames stay ordered (the caller chains their resolve kernels) and keep plain stores.
    const bool ordered = a.accum_mode != 0;
#pragma unroll
    for (int k = 0; k < kRes; k++) {
      hi[k] = 0u; rgb[k] = 0u; mv[k] = 0u; s[k] = make_uint4(0u, 0u, 0u, 0u);
      if (!live[k]) continue;
      if (fp.depth_culling) hi[k] = __ldg(reinterpret_cast<const unsigned int*>(a.zgrid) + 2 * cell[k] + 1);
      if (gather && pix[k] >= 0) {
        const int pv = pix[k] / fp.imgW, pu = pix[k] - pv * fp.imgW;
        const uint8_t* px = a.bgr + (size_t)pv * a.pitch + (size_t)pu * 3;
        rgb[k] = (uint32_t)__ldg(px + 2) | ((uint32_t)__ldg(px + 1) << 8) | ((uint32_t)__ldg(px) << 16);  // r | g<<8 | b<<16 (cpp:760-762)
        if (a.mask) mv[k] = __ldg(a.mask + (size_t)pv * a.mask_pitch + pu);  // cpp:804
        if (ordered) s[k] = a.acc[idx[k]];
      }
    }
    if (fp.depth_culling) {
#pragma unroll
      for (int k = 0; k < kRes; k++)
        live[k] = live[k] && !(dist[k] > __dadd_rn((double)__uint_as_float(hi[k]), fp.tol));  // view_culling.cpp:157
    }
#pragma unroll
    for (int k = 0; k < kRes; k++) {
      if (!live[k]) continue;
      nvis++;
      const uint32_t orig = (a.vis_bits || a.pix_out) ? (a.perm ? __ldg(a.perm + idx[k]) : idx[k]) : 0u;  // caller's index
      if (a.vis_bits) atomicOr(a.vis_bits + (orig >> 5), 1u << (orig & 31));
      if (pix[k] == -2) {  // a double T was applied for culling: generateColorMap projects the float point itself
        const Record* rp = a.records + (base + k * blockDim.x + threadIdx.x);
        const float rx = rp->x, ry = rp->y, rz = rp->z;
        double u, v;
        project(fp.cam, (double)rx, (double)ry, (double)rz, u, v);
        pix[k] = color_pixel((double)rz, u, v, fp.imgW, fp.imgH);
        if (gather && pix[k] >= 0) {  // rare path (pcp_cull with T never gathers; kept for completeness)
          const int pv = pix[k] / fp.imgW, pu = pix[k] - pv * fp.imgW;
          const uint8_t* px = a.bgr + (size_t)pv * a.pitch + (size_t)pu * 3;
          rgb[k] = (uint32_t)__ldg(px + 2) | ((uint32_t)__ldg(px + 1) << 8) | ((uint32_t)__ldg(px) << 16);
          if (a.mask) mv[k] = __ldg(a.mask + (size_t)pv * a.mask_pitch + pu);
          if (ordered) s[k] = a.acc[idx[k]];
        }
      }
      live[k] = pix[k] >= 0;
      if (!live[k]) continue;
      ncol++;
      if (a.pix_out) a.pix_out[orig] = pix[k];
    }
    if (!gather) continue;
#pragma unroll
    for (int k = 0; k < kRes; k++) {
      if (!live[k]) continue;
      uint32_t ar, ag, ab;
      if (a.sample_mode == 0) {
        ar = rgb[k] & 255u; ag = (rgb[k] >> 8) & 255u; ab = rgb[k] >> 16;
      } else {
        bilinear_sample(a.bgr, a.pitch, fp, a.records + (base + k * blockDim.x + threadIdx.x), pix[k], ar, ag, ab);
        rgb[k] = min((ar + 128u) >> 8, 255u) | (min((ag + 128u) >> 8, 255u) << 8) | (min((ab + 128u) >> 8, 255u) << 16);  // TOP5 lists
      }
      if (ordered) {  // a point appears at most once per frame and TOP5 frames are ordered: plain read-modify-write
        s[k].x += ar; s[k].y += ag; s[k].z += ab; s[k].w += 1u;
        a.acc[idx[k]] = s[k];
      } else {
        // {sum_r, sum_g} and {sum_b, count} as two 64-bit words: one no-return RED.ADD.64 each adds both halves at once (a half
        // cannot carry into its neighbour before the 32-bit sum itself would have wrapped)
        unsigned long long* w = reinterpret_cast<unsigned long long*>(a.acc + idx[k]);
        atomicAdd(w, (unsigned long long)ar | ((unsigned long long)ag << 32));
        atomicAdd(w + 1, (unsigned long long)ab | (1ull << 32));
      }
      if (a.accum_mode != 0) {
        const Record* rp = a.records + (base + k * blockDim.x + threadIdx.x);
        const float rx = rp->x, ry = rp->y, rz = rp->z;
        const float o = orientation_score(rx, ry, rz, fp.cam_pos[0], fp.cam_pos[1], fp.cam_pos[2]);
        const float d = distance_score(rx, ry, rz);
        top_insert(a, idx[k], final_score(o, d), rgb[k], (uint32_t)fp.frame_index);
      }
      if (a.mask) {
        unsigned int* mw = reinterpret_cast<unsigned int*>(a.mask_acc + idx[k]);
        if (ordered) {
          uint2 m = a.mask_acc[idx[k]];
          m.x += (mv[k] == 255u) ? 1u : 0u;                                  // cpp:806
          m.y = mv[k] > m.y ? mv[k] : m.y;
          a.mask_acc[idx[k]] = m;
        } else {
          if (mv[k] == 255u) atomicAdd(mw, 1u);                              // cpp:806
          if (mv[k]) atomicMax(mw + 1, mv[k]);
        }
      }
    }
  }
  // counts: warp shuffle, then shared memory, then ONE pair of global atomics per CTA (same-address atomics serialise in
  // L2 at about one per cycle: one pair per warp was 50 k of them, i.e. most of this kernel's 20 us)
  __shared__ unsigned int cta_counts[2];
  if (threadIdx.x < 2) cta_counts[threadIdx.x] = 0u;
  __syncthreads();
  for (int o = 16; o > 0; o >>= 1) {
    nvis += __shfl_down_sync(0xffffffffu, nvis, o);
    ncol += __shfl_down_sync(0xffffffffu, ncol, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (nvis) atomicAdd(&cta_counts[0], nvis);
    if (ncol) atomicAdd(&cta_counts[1], ncol);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (cta_counts[0]) atomicAdd(&a.counters->n_visible, cta_counts[0]);
    if (cta_counts[1]) atomicAdd(&a.counters->n_colored, cta_counts[1]);
  }
}
__global__ void __launch_bounds__(256, PCP_RES_CTAS) resolve_kernel(const __grid_constant__ ResolveArgs a, const __grid_constant__ FrameParams fp) {
  resolve_body(a, fp);
}
template <int K>
__global__ void __launch_bounds__(256, PCP_RES_CTAS) resolve_multi_kernel(const __grid_constant__ FramePack<ResolveArgs, K> p) {
  resolve_body(p.a[blockIdx.y], p.fp[blockIdx.y]);
}

void launch_resolve(const ResolveArgs& a, const FrameParams& fp, int grid, cudaStream_t s) {
  resolve_kernel<<<grid, 256, 0, s>>>(a, fp);
}
template <int K>
static int launch_resolve_group_k(const ResolveArgs* a, const FrameParams* fp, int grid, cudaStream_t s) {
  FramePack<ResolveArgs, K> p;
  for (int k = 0; k < K; k++) { p.a[k] = a[k]; p.fp[k] = fp[k]; }
  resolve_multi_kernel<K><<<dim3(grid, K), 256, 0, s>>>(p);
  return 0;
}
int launch_resolve_group(const ResolveArgs* a, const FrameParams* fp, int k, int grid, cudaStream_t s) {
  if (k == 2) return launch_resolve_group_k<2>(a, fp, grid, s);
  if (k == 4) return launch_resolve_group_k<4>(a, fp, grid, s);
  return 1;
}

// ---------------------------------------------------------------------------------------------------------------------
// Cloud ingest: AoS records (pcl::PointXYZI: 32 B, intensity at byte 16) -> float4 SoA.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void aos_to_soa_kernel(const uint8_t* __restrict__ aos, size_t n, size_t stride, size_t ioff, float4* __restrict__ out) {
  // records of a multiple of 16 bytes on a 16-byte aligned base (pcl::PointXYZI: 32 B): two 16-byte loads per record instead of four
  // 4-byte ones (every sector of the record is fetched once either way; this quarters the load instructions and their replays)
  const bool vec = (stride & 15) == 0 && (reinterpret_cast<uintptr_t>(aos) & 15) == 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint8_t* rec = aos + i * stride;
    if (vec) {
      const float4 a = __ldcs(reinterpret_cast<const float4*>(rec));
      const float4 b = __ldcs(reinterpret_cast<const float4*>(rec + (ioff & ~(size_t)15)));
      const int k = (int)((ioff & 15) >> 2);
      const float inten = k == 0 ? b.x : k == 1 ? b.y : k == 2 ? b.z : b.w;
      out[i] = make_float4(a.x, a.y, a.z, inten);
    } else {
      const float* p = reinterpret_cast<const float*>(rec);
      const float inten = *reinterpret_cast<const float*>(rec + ioff);
      out[i] = make_float4(p[0], p[1], p[2], inten);
    }
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
// `fixed88`: the sums hold 8.8 fixed-point samples (PCP_SAMPLE_BILINEAR): the integer mean is rounded to 8 bits.
// The outputs are in the caller's order, the sums in sorted (Morton) order: a permutation between them.  Three ways were measured at
// 1e7 points (tools/bench_finalize.py): scattering the 3-byte colours through perm (331 us, partial-sector writes); one
// kernel that walks the outputs and gathers the 16-byte sums through inv (143 us: every gather costs a 32-byte DRAM sector and the
// colours leave as byte stores); and the two kernels below, which are what runs:
//   pack:   sorted order, fully coalesced: 16-byte sums -> 8-byte {r | g << 8 | b << 16, count};
//   unpack: output order, four points per thread: one 16-byte load of inv, four independent 8-byte gathers from the packed array (80 MB
//           at 1e7 points: it was just written and is served from L2), 12 + 16 bytes of vector stores.
// The same pair is the middle of pcp_exchange_finalize (pack the rank's slice, all-gather, unpack).
__device__ __forceinline__ uint2 pack_mean(const uint4 s, int fixed88) {
  uint32_t r = 0, g = 0, b = 0;
  if (s.w) { r = s.x / s.w; g = s.y / s.w; b = s.z / s.w; }
  if (fixed88) { r = min((r + 128u) >> 8, 255u); g = min((g + 128u) >> 8, 255u); b = min((b + 128u) >> 8, 255u); }
  return make_uint2(r | (g << 8) | (b << 16), s.w);
}
__device__ __forceinline__ void store_one(uint8_t* rgb, uint32_t* count, uint32_t i, const uint2 v) {
  if (rgb) { rgb[3 * (size_t)i] = (uint8_t)(v.x & 255u); rgb[3 * (size_t)i + 1] = (uint8_t)((v.x >> 8) & 255u); rgb[3 * (size_t)i + 2] = (uint8_t)(v.x >> 16); }
  if (count) count[i] = v.y;
}
__global__ void __launch_bounds__(256) pack_sum_kernel(const uint4* __restrict__ acc, uint32_t j0, uint32_t j1, uint2* __restrict__ packed,
                                                       int fixed88) {
  for (uint32_t j = j0 + blockIdx.x * blockDim.x + threadIdx.x; j < j1; j += gridDim.x * blockDim.x) packed[j] = pack_mean(__ldcs(acc + j), fixed88);
}
// Outputs in the caller's order for 16-byte aligned output pointers.  A CTA takes tiles of 1024 points: every thread fetches four
// points (one 16-byte load of inv, four independent gathers), the counts leave as one 16-byte store per thread and the 3072 colour
// bytes of the tile go through shared memory and leave as 192 16-byte stores (byte or 12-byte-strided word stores reach L2 as partial
// sectors: measured 212 us against 162 us for the byte form at 1e7 points).  kFromAcc: gather the 16-byte sums themselves and take the
// mean here (the one-kernel form); otherwise gather the packed 8-byte values of pack_sum_kernel.
template <bool kFromAcc>
__global__ void __launch_bounds__(256) finalize_tiles_kernel(const void* __restrict__ src, const uint32_t* __restrict__ inv, uint32_t n,
                                                             uint8_t* __restrict__ rgb, uint32_t* __restrict__ count, int fixed88) {
  __shared__ __align__(16) uint32_t s_rgb[768];
  const uint4* acc = reinterpret_cast<const uint4*>(src);
  const uint2* packed = reinterpret_cast<const uint2*>(src);
  const uint32_t n_tiles = (n + 1023u) / 1024u;
  const uint32_t t = threadIdx.x;
  for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const uint32_t base = tile * 1024u, i = base + 4u * t;
    uint2 v[4] = {make_uint2(0, 0), make_uint2(0, 0), make_uint2(0, 0), make_uint2(0, 0)};
    if (i + 3u < n) {
      uint4 j = make_uint4(i, i + 1, i + 2, i + 3);
      if (inv) j = __ldcs(reinterpret_cast<const uint4*>(inv) + (i >> 2));
      if (kFromAcc) {
        const uint4 sa = acc[j.x], sb = acc[j.y], sc = acc[j.z], sd = acc[j.w];
        v[0] = pack_mean(sa, fixed88); v[1] = pack_mean(sb, fixed88); v[2] = pack_mean(sc, fixed88); v[3] = pack_mean(sd, fixed88);
      } else {
        v[0] = packed[j.x]; v[1] = packed[j.y]; v[2] = packed[j.z]; v[3] = packed[j.w];
      }
      if (count) reinterpret_cast<uint4*>(count)[i >> 2] = make_uint4(v[0].y, v[1].y, v[2].y, v[3].y);
    } else {
      for (uint32_t k = 0; k < 4u && i + k < n; k++) {
        const uint32_t j = inv ? inv[i + k] : i + k;
        v[k] = kFromAcc ? pack_mean(acc[j], fixed88) : packed[j];
        if (count) count[i + k] = v[k].y;
      }
    }
    if (rgb) {
      // 12 bytes r g b r | g b r g | b r g b
      s_rgb[3 * t] = (v[0].x & 0xffffffu) | (v[1].x << 24);
      s_rgb[3 * t + 1] = ((v[1].x >> 8) & 0xffffu) | (v[2].x << 16);
      s_rgb[3 * t + 2] = ((v[2].x >> 16) & 0xffu) | (v[3].x << 8);
      __syncthreads();
      const uint32_t bytes = 3u * min(1024u, n - base);
      uint8_t* out = rgb + 3 * (size_t)base;
      if (t < 192u) {
        if (16u * t + 16u <= bytes) reinterpret_cast<uint4*>(out)[t] = reinterpret_cast<const uint4*>(s_rgb)[t];
        else for (uint32_t b = 16u * t; b < bytes; b++) out[b] = reinterpret_cast<const uint8_t*>(s_rgb)[b];
      }
      __syncthreads();
    }
  }
}
// any output alignment (a caller's sub-range): one point per thread, byte stores
__global__ void unpack_colours_bytes_kernel(const uint2* __restrict__ packed, const uint32_t* __restrict__ inv, uint32_t n,
                                            uint8_t* __restrict__ rgb, uint32_t* __restrict__ count) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) store_one(rgb, count, i, packed[inv ? __ldg(inv + i) : i]);
}
__global__ void finalize_sum_kernel(const uint4* __restrict__ acc, const uint32_t* __restrict__ inv, uint32_t n, uint8_t* __restrict__ rgb,
                                    uint32_t* __restrict__ count, int fixed88) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    store_one(rgb, count, i, pack_mean(acc[inv ? __ldg(inv + i) : i], fixed88));
}
static int grid_for(size_t items, int per_cta, int cap) {
  const size_t g = (items + per_cta - 1) / per_cta;
  return (int)(g < (size_t)cap ? (g ? g : 1) : (size_t)cap);
}
static bool quad_aligned(const uint8_t* rgb, const uint32_t* count, const uint32_t* inv) {
  return ((uintptr_t)rgb % 16 == 0) && ((uintptr_t)count % 16 == 0) && ((uintptr_t)inv % 16 == 0);
}
void launch_finalize_sum(const uint4* acc, const uint32_t* inv, uint32_t n, uint8_t* rgb, uint32_t* count, int fixed88, cudaStream_t s) {
  if (n == 0 || (!rgb && !count)) return;
  if (quad_aligned(rgb, count, inv)) finalize_tiles_kernel<true><<<grid_for(n, 1024, 148 * 16), 256, 0, s>>>(acc, inv, n, rgb, count, fixed88);
  else finalize_sum_kernel<<<grid_for(n, 256, 148 * 16), 256, 0, s>>>(acc, inv, n, rgb, count, fixed88);
}
void launch_pack_sum(const uint4* acc, uint32_t j0, uint32_t j1, uint2* packed, int fixed88, cudaStream_t s) {
  if (j1 <= j0) return;
  pack_sum_kernel<<<grid_for(j1 - j0, 256, 148 * 16), 256, 0, s>>>(acc, j0, j1, packed, fixed88);
}
void launch_unpack_colours(const uint2* packed, const uint32_t* inv, uint32_t n, uint8_t* rgb, uint32_t* count, cudaStream_t s) {
  if (n == 0 || (!rgb && !count)) return;
  if (quad_aligned(rgb, count, inv)) finalize_tiles_kernel<false><<<grid_for(n, 1024, 148 * 16), 256, 0, s>>>(packed, inv, n, rgb, count, 0);
  else unpack_colours_bytes_kernel<<<grid_for(n, 256, 148 * 16), 256, 0, s>>>(packed, inv, n, rgb, count);
}

// smoothColors (cpp:612-629): slots are already sorted by (score desc, frame asc); weighted mean in float, truncation.
__global__ void finalize_top_kernel(const float* __restrict__ score, const uint32_t* __restrict__ rgbf, const uint32_t* __restrict__ frame,
                                    const uint32_t* __restrict__ perm, uint32_t n, int top_m, uint8_t* __restrict__ rgb,
                                    uint32_t* __restrict__ count) {
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const uint32_t i = perm ? perm[j] : j;  // outputs in the caller's point order
    float total = 0.f, r = 0.f, g = 0.f, b = 0.f;
    uint32_t m = 0;
    for (int k = 0; k < top_m; k++) {
      const size_t o = (size_t)k * n + j;
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
void launch_finalize_top(const float* score, const uint32_t* rgbf, const uint32_t* frame, const uint32_t* perm, uint32_t n, int top_m,
                         uint8_t* rgb, uint32_t* count, cudaStream_t s) {
  if (n == 0) return;
  const int grid = (n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16;
  finalize_top_kernel<<<grid, 256, 0, s>>>(score, rgbf, frame, perm, n, top_m, rgb, count);
}

// Per-point arrays kept in sorted order (pre-binned cloud) -> the caller's order: dst[perm[j]] = src[j], `w` words each.
__global__ void unpermute_kernel(const uint32_t* __restrict__ src, const uint32_t* __restrict__ perm, uint32_t n, int w, uint32_t* __restrict__ dst) {
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const size_t o = (size_t)perm[j] * w, i = (size_t)j * w;
    for (int k = 0; k < w; k++) dst[o + k] = src[i + k];
  }
}
void launch_unpermute(const uint32_t* src, const uint32_t* perm, uint32_t n, int elem_words, uint32_t* dst, cudaStream_t s) {
  if (n == 0) return;
  const int grid = (n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16;
  unpermute_kernel<<<grid, 256, 0, s>>>(src, perm, n, elem_words, dst);
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

// ---------------------------------------------------------------------------------------------------------------------
// removePointsWithNoColor (include/PointCloudProcessor.hpp:238-252) on the device: bit i = the finalized colour of point i is not
// (0,0,0); the ascending index list comes from launch_compact_bits; the survivors are gathered as pcl::PointXYZRGB payloads
// {x, y, z, b | g << 8 | r << 16 | 255 << 24}, 16 bytes each, in the caller's point order.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void coloured_bits_kernel(const uint8_t* __restrict__ rgb, uint32_t n, uint32_t* __restrict__ bits) {
  const uint32_t words = (n + 31) / 32;
  for (uint32_t w = blockIdx.x * blockDim.x + threadIdx.x; w < words; w += gridDim.x * blockDim.x) {
    uint32_t m = 0;
    for (uint32_t b = 0; b < 32; b++) {
      const uint32_t i = w * 32 + b;
      if (i < n && (rgb[3 * (size_t)i] | rgb[3 * (size_t)i + 1] | rgb[3 * (size_t)i + 2])) m |= 1u << b;
    }
    bits[w] = m;
  }
}
__global__ void gather_coloured_kernel(const int32_t* __restrict__ idx, uint32_t m, const float4* __restrict__ cloud, const uint32_t* __restrict__ inv,
                                       const uint8_t* __restrict__ rgb, uint4* __restrict__ out) {
  for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < m; k += gridDim.x * blockDim.x) {
    const uint32_t i = (uint32_t)idx[k];
    const float4 p = cloud[inv ? inv[i] : i];  // the device copy of the cloud is in sorted order
    const uint32_t r = rgb[3 * (size_t)i], g = rgb[3 * (size_t)i + 1], b = rgb[3 * (size_t)i + 2];
    out[k] = make_uint4(__float_as_uint(p.x), __float_as_uint(p.y), __float_as_uint(p.z), b | (g << 8) | (r << 16) | (255u << 24));
  }
}
// generateSegmentMap (cpp:769-816) on the device: bit i = point i received a pixel in the last frame; then, for the compacted indices,
// the PointXYZRGBMask record {camera-frame position (pcl::transformPointCloud, cpp:521), the image's b g r, 255, the mask value; mask
// 255 paints the point red (cpp:806-811)} of 20 bytes, in the map's point order.
__global__ void pixel_bits_kernel(const int32_t* __restrict__ pix, uint32_t n, uint32_t* __restrict__ bits) {
  const uint32_t words = (n + 31) / 32;
  for (uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; base < words * 32; base += gridDim.x * blockDim.x) {
    const uint32_t i = base + (threadIdx.x & 31u);
    const unsigned int m = __ballot_sync(0xffffffffu, i < n && pix[i] >= 0);
    if ((threadIdx.x & 31u) == 0) bits[base >> 5] = m;
  }
}
__global__ void gather_mask_cloud_kernel(const int32_t* __restrict__ idx, uint32_t m, const float4* __restrict__ cloud, const uint32_t* __restrict__ inv,
                                         const int32_t* __restrict__ pix, const uint8_t* __restrict__ bgr, size_t pitch,
                                         const uint8_t* __restrict__ mask, size_t mask_pitch, int W, const Xf34 w2c, uint32_t* __restrict__ out) {
  for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < m; k += gridDim.x * blockDim.x) {
    const uint32_t i = (uint32_t)idx[k];
    const float4 p = cloud[inv ? inv[i] : i];  // the device copy of the cloud is in sorted order
    float x, y, z;
    transform_point(w2c, p.x, p.y, p.z, x, y, z);
    const int q = pix[i], py = q / W, px = q - py * W;
    const uint8_t* c = bgr + (size_t)py * pitch + 3 * (size_t)px;
    uint32_t b = c[0], g = c[1], r = c[2];
    const uint32_t mk = mask[(size_t)py * mask_pitch + (size_t)px];
    if (mk == 255u) { r = 255u; g = 0u; b = 0u; }
    uint32_t* o = out + 5 * (size_t)k;
    o[0] = __float_as_uint(x); o[1] = __float_as_uint(y); o[2] = __float_as_uint(z);
    o[3] = b | (g << 8) | (r << 16) | (255u << 24);
    o[4] = mk;  // uint16 mask + two bytes of padding
  }
}
void launch_pixel_bits(const int32_t* pix, uint32_t n, uint32_t* bits, cudaStream_t s) {
  if (n == 0) return;
  const uint32_t threads = ((n + 31) / 32) * 32;
  const int grid = (threads + 255) / 256 < 148 * 16 ? (threads + 255) / 256 : 148 * 16;
  pixel_bits_kernel<<<grid, 256, 0, s>>>(pix, n, bits);
}
void launch_gather_mask_cloud(const int32_t* idx, uint32_t m, const float4* cloud, const uint32_t* inv, const int32_t* pix, const uint8_t* bgr,
                              size_t pitch, const uint8_t* mask, size_t mask_pitch, int W, const Xf34& w2c, uint32_t* out, cudaStream_t s) {
  if (m == 0) return;
  const int grid = (m + 255) / 256 < 148 * 8 ? (m + 255) / 256 : 148 * 8;
  gather_mask_cloud_kernel<<<grid, 256, 0, s>>>(idx, m, cloud, inv, pix, bgr, pitch, mask, mask_pitch, W, w2c, out);
}
void launch_coloured_bits(const uint8_t* rgb, uint32_t n, uint32_t* bits, cudaStream_t s) {
  if (n == 0) return;
  const uint32_t words = (n + 31) / 32;
  const int grid = (words + 255) / 256 < 148 * 16 ? (words + 255) / 256 : 148 * 16;
  coloured_bits_kernel<<<grid, 256, 0, s>>>(rgb, n, bits);
}
void launch_gather_coloured(const int32_t* idx, uint32_t m, const float4* cloud, const uint32_t* inv, const uint8_t* rgb, uint4* out, cudaStream_t s) {
  if (m == 0) return;
  const int grid = (m + 255) / 256 < 148 * 16 ? (m + 255) / 256 : 148 * 16;
  gather_coloured_kernel<<<grid, 256, 0, s>>>(idx, m, cloud, inv, rgb, out);
}

}  // namespace pcp

int scan_ctas_per_sm() {
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pcp::scan_kernel<false>, pcp::kScanThreads, 0) != cudaSuccess || occ < 1) occ = 2;
  return occ;
}

#ifdef PCP_TIMELINE
extern "C" int pcp_debug_timeline(unsigned long long* out, size_t ctas) {
  if (!out || ctas > 2048) return 1;
  return cudaMemcpyFromSymbol(out, pcp::g_timeline, ctas * 8 * sizeof(unsigned long long)) == cudaSuccess ? 0 : 2;
}
#endif
