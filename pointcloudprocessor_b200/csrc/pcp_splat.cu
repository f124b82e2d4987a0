// pcp_splat.cu — distance / 3-D point masks of a frame (SURVEY.md 8f-4 "mask splatting"): generate_distance_masks of
// scripts/genNormAndDistanceMask.py:233-265 as two kernels over the outputs of the frame's resolve kernel.
//
// The script walks the frame's culled cloud in file order and lets every point that lands inside the image overwrite
// distance_mask[y, x] and points_3d_mask[y, x]: the last point of the file wins.  The culled cloud is written in ascending
// point index (ViewCulling::cull, view_culling.cpp:176-185), so "last" = the largest original index: a 64-bit atomicMax on
// (original index + 1) << 32 | sorted position picks it deterministically, whatever order the threads run in; the low word
// tells the second kernel where the winner lives in the (pre-binned, i.e. Morton-sorted) device cloud.
#include "pcp_internal.cuh"

namespace pcp {

// pix: per ORIGINAL point index, the resolve kernel's pixel v*W+u of a visible point inside the image, else -1.
__global__ void splat_winner_kernel(const int32_t* __restrict__ pix, const uint32_t* __restrict__ perm, uint32_t n,
                                    unsigned long long* __restrict__ winner) {
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const uint32_t orig = perm ? __ldg(perm + j) : j;
    const int p = __ldg(pix + orig);
    if (p >= 0) atomicMax(winner + p, ((unsigned long long)(orig + 1u) << 32) | (unsigned long long)j);
  }
}

// One thread per pixel: the winner's camera-frame float point (pcl::transformPointCloud, cpp:196) and its range
// np.linalg.norm(xyz) in double rounded to float32 (script :256,:259); zeros / -1 where no point landed (:251-252).
__global__ void splat_fill_kernel(const unsigned long long* __restrict__ winner, int npix, const float4* __restrict__ cloud,
                                  const __grid_constant__ FrameParams fp, float* __restrict__ dist, float* __restrict__ xyz,
                                  int32_t* __restrict__ index, unsigned int* __restrict__ n_filled) {
  unsigned int filled = 0;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += gridDim.x * blockDim.x) {
    const unsigned long long key = winner[p];
    float d = 0.0f, x = 0.0f, y = 0.0f, z = 0.0f;
    int32_t id = -1;
    if (key) {
      const float4 c = __ldg(cloud + (uint32_t)key);
      x = c.x; y = c.y; z = c.z;
      if (fp.use_w2c) transform_point(fp.w2c, c.x, c.y, c.z, x, y, z);
      d = __double2float_rn(norm3((double)x, (double)y, (double)z));
      id = (int32_t)((uint32_t)(key >> 32) - 1u);
      filled++;
    }
    dist[p] = d;
    if (xyz) { xyz[3 * (size_t)p] = x; xyz[3 * (size_t)p + 1] = y; xyz[3 * (size_t)p + 2] = z; }
    if (index) index[p] = id;
  }
  for (int o = 16; o > 0; o >>= 1) filled += __shfl_down_sync(0xffffffffu, filled, o);
  if ((threadIdx.x & 31) == 0 && filled) atomicAdd(n_filled, filled);
}

void launch_splat(const int32_t* pix, const uint32_t* perm, uint32_t n, const float4* cloud, const FrameParams& fp, unsigned long long* winner,
                  float* dist, float* xyz, int32_t* index, unsigned int* n_filled, int sm_count, cudaStream_t s) {
  const int npix = fp.imgW * fp.imgH;
  if (npix <= 0) return;
  if (n) {
    const uint32_t blocks = (n + 255u) / 256u;
    const uint32_t cap = (uint32_t)sm_count * 8u;
    splat_winner_kernel<<<blocks < cap ? blocks : cap, 256, 0, s>>>(pix, perm, n, winner);
  }
  const int blocks = (npix + 255) / 256, cap = sm_count * 8;
  splat_fill_kernel<<<blocks < cap ? blocks : cap, 256, 0, s>>>(winner, npix, cloud, fp, dist, xyz, index, n_filled);
}

}  // namespace pcp
