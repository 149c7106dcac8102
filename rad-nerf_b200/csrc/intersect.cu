// intersect.cu -- ray/AABB, ray/sphere, morton codes and packbits for sm_100a.
// Replaces models/csrc/intersection.cu and raymarching.cu:35-161 of the reference.
#include <stdarg.h>

#include "aabb.cuh"

namespace rnb {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---- ray / box -----------------------------------------------------------------------
// One thread per ray looping over the voxels in index order; the first `max_hits` hits are
// kept (a legal outcome of the reference's atomicAdd race, intersection.cu:49-55), then the
// max_hits rows are ordered by ascending t1 like the reference's torch::sort (:95-97), which
// places unfilled (-1) rows first.  For the path's only configuration (1 voxel, 1 hit)
// this is a single slab test and three coalesced stores.
template <bool kSphere>
__global__ void __launch_bounds__(256) ray_prim_intersect_kernel(
    const float* __restrict__ rays_o, const float* __restrict__ rays_d,
    const float* __restrict__ centers, const float* __restrict__ extents, int64_t n_rays,
    int64_t n_prims, int max_hits, int32_t* __restrict__ hit_cnt, float* __restrict__ hits_t,
    int64_t* __restrict__ hits_idx) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rays) return;
  const float ox = rays_o[3 * r], oy = rays_o[3 * r + 1], oz = rays_o[3 * r + 2];
  const float dx = rays_d[3 * r], dy = rays_d[3 * r + 1], dz = rays_d[3 * r + 2];
  const float ix = __fdiv_rn(1.0f, dx), iy = __fdiv_rn(1.0f, dy), iz = __fdiv_rn(1.0f, dz);
  float* ht = hits_t + r * max_hits * 2;
  int64_t* hi = hits_idx + r * max_hits;
  for (int k = 0; k < max_hits; k++) { ht[2 * k] = -1.0f; ht[2 * k + 1] = -1.0f; hi[k] = -1; }
  int cnt = 0;
  for (int64_t v = 0; v < n_prims; v++) {
    float2 t;
    if (kSphere) {
      // intersection.cu:103-121
      const float cox = ox - centers[3 * v], coy = oy - centers[3 * v + 1], coz = oz - centers[3 * v + 2];
      const float rad = extents[v];
      const float a = dx * dx + dy * dy + dz * dz;
      const float half_b = dx * cox + dy * coy + dz * coz;
      const float c = cox * cox + coy * coy + coz * coz - rad * rad;
      const float disc = half_b * half_b - a * c;
      if (disc < 0) t = make_float2(-1.0f, -1.0f);
      else { const float sq = sqrtf(disc); t = make_float2((-half_b - sq) / a, (-half_b + sq) / a); }
    } else {
      t = slab_test(ox, oy, oz, ix, iy, iz, centers[3 * v], centers[3 * v + 1], centers[3 * v + 2],
                    extents[3 * v], extents[3 * v + 1], extents[3 * v + 2]);
    }
    if (t.y > 0) {
      if (cnt < max_hits) { ht[2 * cnt] = fmaxf(t.x, 0.0f); ht[2 * cnt + 1] = t.y; hi[cnt] = v; }
      cnt++;
    }
  }
  hit_cnt[r] = cnt;
  // ascending insertion sort on t1 over the max_hits rows (stable)
  for (int i = 1; i < max_hits; i++) {
    const float k0 = ht[2 * i], k1 = ht[2 * i + 1]; const int64_t kv = hi[i];
    int j = i - 1;
    while (j >= 0 && ht[2 * j] > k0) { ht[2 * j + 2] = ht[2 * j]; ht[2 * j + 3] = ht[2 * j + 1]; hi[j + 1] = hi[j]; j--; }
    ht[2 * j + 2] = k0; ht[2 * j + 3] = k1; hi[j + 1] = kv;
  }
}

// ---- morton ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ __forceinline__ uint32_t compact_bits(uint32_t x) {
  x = x & 0x49249249u;
  x = (x | (x >> 2)) & 0xc30c30c3u;
  x = (x | (x >> 4)) & 0x0f00f00fu;
  x = (x | (x >> 8)) & 0xff0000ffu;
  x = (x | (x >> 16)) & 0x0000ffffu;
  return x;
}

__global__ void __launch_bounds__(256) morton3D_kernel(const int32_t* __restrict__ coords, int64_t n,
                                                       int32_t* __restrict__ indices) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t x = coords[3 * i], y = coords[3 * i + 1], z = coords[3 * i + 2];
    indices[i] = (int32_t)(expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2));
  }
}
__global__ void __launch_bounds__(256) morton3D_invert_kernel(const int32_t* __restrict__ indices, int64_t n,
                                                              int32_t* __restrict__ coords) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t ind = indices[i];   // arithmetic shifts of the signed value, as raymarching.cu:97-100
    coords[3 * i] = (int32_t)compact_bits((uint32_t)(ind >> 0));
    coords[3 * i + 1] = (int32_t)compact_bits((uint32_t)(ind >> 1));
    coords[3 * i + 2] = (int32_t)compact_bits((uint32_t)(ind >> 2));
  }
}

// ---- packbits --------------------------------------------------------------------------
// Warp w owns 256 consecutive cells (1 KiB of fp32) per iteration: lane l loads cell
// base+32*j+l for j=0..7 (eight fully coalesced 128 B requests in flight), and
// __ballot_sync of (v > thr) IS the little-endian 32-bit word holding bitfield bytes
// 4j..4j+3 of that slab (bit i of byte n <-> cell 8n+i, raymarching.cu:135-138).  Lanes 0..7
// then store one word each = one 32 B sector.  HBM-bound: 4.125 B per cell.
template <typename T>
__device__ __forceinline__ float load_cell(const T* p);
template <>
__device__ __forceinline__ float load_cell<float>(const float* p) { return ld_stream_f32(p); }
template <>
__device__ __forceinline__ float load_cell<__half>(const __half* p) { return __half2float(*p); }

template <typename T, bool kDevThr>
__global__ void __launch_bounds__(256) packbits_kernel(const T* __restrict__ grid, int64_t n_bytes, float thr,
                                                       const float* __restrict__ mean_density,
                                                       uint8_t* __restrict__ bitfield) {
  if (kDevThr) thr = fminf(*mean_density, thr);
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t n_slabs = n_bytes >> 5;  // 32 bytes (256 cells) per slab
  for (int64_t s = warp; s < n_slabs; s += n_warps) {
    const T* g = grid + s * 256 + lane;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) v[j] = load_cell<T>(g + 32 * j);
    uint32_t mine = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const uint32_t w = __ballot_sync(kFullMask, v[j] > thr);
      if (lane == j) mine = w;
    }
    if (lane < 8) reinterpret_cast<uint32_t*>(bitfield)[s * 8 + lane] = mine;
  }
  // tail (n_bytes not a multiple of 32): one thread per byte
  const int64_t tail0 = n_slabs << 5;
  for (int64_t n = tail0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < n_bytes;
       n += (int64_t)gridDim.x * blockDim.x) {
    uint8_t bits = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) bits |= (load_cell<T>(grid + 8 * n + i) > thr) ? (uint8_t)(1u << i) : 0;
    bitfield[n] = bits;
  }
}

}  // namespace rnb

using namespace rnb;

extern "C" {

int rnb_abi_version(void) { return RNB_ABI_VERSION; }
const char* rnb_last_error(void) { return rnb::g_err; }
int rnb_device_arch(void) {
  int dev = 0, major = 0, minor = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); set_error("no CUDA device"); return RNB_ERR_LAUNCH; }
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  return major * 10 + minor;
}

int rnb_ray_aabb_intersect(const float* rays_o, const float* rays_d, const float* centers,
                           const float* half_sizes, int64_t n_rays, int64_t n_voxels, int max_hits,
                           int32_t* hit_cnt, float* hits_t, int64_t* hits_voxel_idx, rnb_stream_t stream) {
  RNB_REQUIRE(n_rays >= 0 && n_voxels >= 0 && max_hits >= 1, "rnb_ray_aabb_intersect");
  if (n_rays == 0) return RNB_OK;
  RNB_REQUIRE(rays_o && rays_d && centers && half_sizes && hit_cnt && hits_t && hits_voxel_idx, "rnb_ray_aabb_intersect");
  ray_prim_intersect_kernel<false><<<ceil_div(n_rays, 256), 256, 0, as_stream(stream)>>>(
      rays_o, rays_d, centers, half_sizes, n_rays, n_voxels, max_hits, hit_cnt, hits_t, hits_voxel_idx);
  return check_launch("rnb_ray_aabb_intersect");
}

int rnb_ray_sphere_intersect(const float* rays_o, const float* rays_d, const float* centers,
                             const float* radii, int64_t n_rays, int64_t n_spheres, int max_hits,
                             int32_t* hit_cnt, float* hits_t, int64_t* hits_sphere_idx, rnb_stream_t stream) {
  RNB_REQUIRE(n_rays >= 0 && n_spheres >= 0 && max_hits >= 1, "rnb_ray_sphere_intersect");
  if (n_rays == 0) return RNB_OK;
  RNB_REQUIRE(rays_o && rays_d && centers && radii && hit_cnt && hits_t && hits_sphere_idx, "rnb_ray_sphere_intersect");
  ray_prim_intersect_kernel<true><<<ceil_div(n_rays, 256), 256, 0, as_stream(stream)>>>(
      rays_o, rays_d, centers, radii, n_rays, n_spheres, max_hits, hit_cnt, hits_t, hits_sphere_idx);
  return check_launch("rnb_ray_sphere_intersect");
}

static inline int elementwise_grid(int64_t n, int per_block) {
  int64_t b = (n + per_block - 1) / per_block;
  const int64_t cap = (int64_t)kNumSMs * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

int rnb_morton3D(const int32_t* coords, int64_t n, int32_t* indices, rnb_stream_t stream) {
  RNB_REQUIRE(n >= 0, "rnb_morton3D");
  if (n == 0) return RNB_OK;
  RNB_REQUIRE(coords && indices, "rnb_morton3D");
  morton3D_kernel<<<elementwise_grid(n, 256), 256, 0, as_stream(stream)>>>(coords, n, indices);
  return check_launch("rnb_morton3D");
}

int rnb_morton3D_invert(const int32_t* indices, int64_t n, int32_t* coords, rnb_stream_t stream) {
  RNB_REQUIRE(n >= 0, "rnb_morton3D_invert");
  if (n == 0) return RNB_OK;
  RNB_REQUIRE(indices && coords, "rnb_morton3D_invert");
  morton3D_invert_kernel<<<elementwise_grid(n, 256), 256, 0, as_stream(stream)>>>(indices, n, coords);
  return check_launch("rnb_morton3D_invert");
}

int rnb_packbits(const void* density_grid, int grid_is_half, int64_t n_bytes, float threshold,
                 uint8_t* density_bitfield, rnb_stream_t stream) {
  RNB_REQUIRE(n_bytes >= 0, "rnb_packbits");
  if (n_bytes == 0) return RNB_OK;
  RNB_REQUIRE(density_grid && density_bitfield, "rnb_packbits");
  RNB_REQUIRE(((uintptr_t)density_bitfield & 3) == 0, "rnb_packbits");
  // one warp per 256-cell slab; 8 warps per block
  const int grid = elementwise_grid(n_bytes / 32 * 32, 256 * 8 / 8);
  if (grid_is_half)
    packbits_kernel<__half, false><<<grid, 256, 0, as_stream(stream)>>>((const __half*)density_grid, n_bytes, threshold, nullptr, density_bitfield);
  else
    packbits_kernel<float, false><<<grid, 256, 0, as_stream(stream)>>>((const float*)density_grid, n_bytes, threshold, nullptr, density_bitfield);
  return check_launch("rnb_packbits");
}

int rnb_packbits_devthr(const float* density_grid, int64_t n_bytes, const float* mean_density,
                        float threshold_cap, uint8_t* density_bitfield, rnb_stream_t stream) {
  RNB_REQUIRE(n_bytes >= 0, "rnb_packbits_devthr");
  if (n_bytes == 0) return RNB_OK;
  RNB_REQUIRE(density_grid && density_bitfield && mean_density, "rnb_packbits_devthr");
  RNB_REQUIRE(((uintptr_t)density_bitfield & 3) == 0, "rnb_packbits_devthr");
  const int grid = elementwise_grid(n_bytes / 32 * 32, 256);
  packbits_kernel<float, true><<<grid, 256, 0, as_stream(stream)>>>(density_grid, n_bytes, threshold_cap, mean_density, density_bitfield);
  return check_launch("rnb_packbits_devthr");
}

}  // extern "C"
