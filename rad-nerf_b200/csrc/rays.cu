// rays.cu -- training-batch assembly on the device: pixel sampling + get_rays + target-colour gather in one kernel.
//
// Replaces, per step, BaseDataset.__getitem__ (datasets/base.py:22-29: two np.random.choice calls and a fancy-indexed
// gather on the host), the host->device copy of the sample, `self.directions[pix_idxs]`, `self.poses[img_idxs]` and
// get_rays (datasets/ray_utils.py:46-70: a batched 1x3 @ 3x3 matmul + expand) -- train_ml.py:183-186.
// Thread = ray.  Indices are either supplied (parity tests, or a host sampler) or drawn from a counter-based generator.
#include "common.cuh"

namespace rnb {

__device__ __forceinline__ uint64_t splitmix(uint64_t seed, uint64_t ctr) {
  uint64_t z = seed + (ctr + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256) batch_rays_kernel(const float* __restrict__ directions, const float* __restrict__ poses,
                                                         const float* __restrict__ images, const int64_t* __restrict__ img_idxs,
                                                         const int64_t* __restrict__ pix_idxs, int64_t n_img, int64_t n_pix, int64_t B,
                                                         uint64_t seed, float* __restrict__ rays_o, float* __restrict__ rays_d,
                                                         float* __restrict__ rgb, int64_t* __restrict__ img_out, int64_t* __restrict__ pix_out) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  // np.random.choice(n, B): iid uniform with replacement (multiply-shift of a 64-bit draw: bias < 2^-32)
  const int64_t im = img_idxs ? img_idxs[b] : (int64_t)(((splitmix(seed, 2 * (uint64_t)b) >> 32) * (uint64_t)n_img) >> 32);
  const int64_t px = pix_idxs ? pix_idxs[b] : (int64_t)(((splitmix(seed, 2 * (uint64_t)b + 1) >> 32) * (uint64_t)n_pix) >> 32);
  const float dx = directions[3 * px], dy = directions[3 * px + 1], dz = directions[3 * px + 2];
  const float* c = poses + 12 * im;     // (3,4) row-major: [R | t]
#pragma unroll
  for (int i = 0; i < 3; i++) {
    // rays_d = directions @ R^T: d_i = sum_j dir_j * R[i][j]   (ray_utils.py:62-66), accumulated in index order
    rays_d[3 * b + i] = fmaf(dz, c[4 * i + 2], fmaf(dy, c[4 * i + 1], dx * c[4 * i]));
    rays_o[3 * b + i] = c[4 * i + 3];                                                   // ray_utils.py:68
  }
  if (images != nullptr && rgb != nullptr) {
    const float* p = images + 3 * (im * n_pix + px);
    rgb[3 * b] = p[0]; rgb[3 * b + 1] = p[1]; rgb[3 * b + 2] = p[2];
  }
  if (img_out) img_out[b] = im;
  if (pix_out) pix_out[b] = px;
}

}  // namespace rnb

using namespace rnb;

extern "C" int rnb_batch_rays(const float* directions, const float* poses, const float* images, const int64_t* img_idxs,
                              const int64_t* pix_idxs, int64_t n_img, int64_t n_pix, int64_t B, uint64_t seed, float* rays_o,
                              float* rays_d, float* rgb, int64_t* img_idxs_out, int64_t* pix_idxs_out, rnb_stream_t stream) {
  const char* what = "rnb_batch_rays";
  RNB_REQUIRE(B >= 0 && n_img >= 1 && n_pix >= 1, what);
  if (B == 0) return RNB_OK;
  RNB_REQUIRE(directions && poses && rays_o && rays_d, what);
  RNB_REQUIRE((images == nullptr) == (rgb == nullptr), what);
  batch_rays_kernel<<<ceil_div(B, 256), 256, 0, as_stream(stream)>>>(directions, poses, images, img_idxs, pix_idxs, n_img, n_pix, B, seed,
                                                                   rays_o, rays_d, rgb, img_idxs_out, pix_idxs_out);
  return check_launch(what);
}
