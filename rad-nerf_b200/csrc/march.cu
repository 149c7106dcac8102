// march.cu -- training / test-time occupancy-grid marchers (replaces raymarching.cu:166-454).
#include "march.cuh"

namespace rnb {

// ---- training: phase 1 (count) ---------------------------------------------------------
// One thread per ray; 64-thread CTAs so that a training batch of 8192 rays spreads over
// 128 of the 148 SMs (the loop is latency-bound: one dependent bitfield load per step).
// Writes the per-ray count into rays_a[r][2]; the scan kernel below completes the rows.
__global__ void __launch_bounds__(64) march_train_count_kernel(
    const float* __restrict__ rays_o, const float* __restrict__ rays_d, const float* __restrict__ hits_t,
    const float* __restrict__ noise, MarchParams p, int64_t n_rays, int64_t* __restrict__ rays_a) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rays) return;
  const Ray ray = load_ray(rays_o, rays_d, r);
  float t1 = __ldg(hits_t + 2 * r);
  const float t2 = __ldg(hits_t + 2 * r + 1);
  if (t1 >= 0) t1 = __fmaf_rn(calc_dt(t1, p), __ldg(noise + r), t1);  // raymarching.cu:195-198
  float t = t1; int n = 0; Probe s;
  while (0 <= t && t < t2 && n < p.max_samples) {
    if (march_probe(ray, p, t, s)) { t = __fadd_rn(t, s.dt); n++; }
  }
  rays_a[3 * r + 2] = n;
}

// ---- training: exclusive scan of the counts -> rays_a rows + counter --------------------
// Single CTA of 1024 threads, 4 rays per thread per sweep with a running carry; for the
// training batch (8192 rays) that is two sweeps.  Deterministic ascending-ray CSR (Q2).
__global__ void __launch_bounds__(1024) march_train_scan_kernel(int64_t n_rays, int64_t* __restrict__ rays_a,
                                                                int32_t* __restrict__ counter) {
  __shared__ int warp_tot[32];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < n_rays; base += 4096) {
    int c[4]; int mine = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int64_t r = base + (int64_t)tid * 4 + k;
      c[k] = r < n_rays ? (int)rays_a[3 * r + 2] : 0;
      mine += c[k];
    }
    const int incl = warp_incl_scan_i(mine, lane);
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      const int wi = warp_incl_scan_i(warp_tot[lane], lane);
      warp_tot[lane] = wi;  // inclusive over warps
    }
    __syncthreads();
    const int carry = carry_s;
    int off = carry + (wid ? warp_tot[wid - 1] : 0) + incl - mine;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int64_t r = base + (int64_t)tid * 4 + k;
      if (r < n_rays) { rays_a[3 * r] = r; rays_a[3 * r + 1] = off; }
      off += c[k];
    }
    __syncthreads();
    if (tid == 0) carry_s = carry + warp_tot[31];
    __syncthreads();
  }
  if (tid == 0) { counter[0] = carry_s; counter[1] = (int32_t)n_rays; }
}

// ---- training: phase 2 (write) -----------------------------------------------------------
__global__ void __launch_bounds__(64) march_train_write_kernel(
    const float* __restrict__ rays_o, const float* __restrict__ rays_d, const float* __restrict__ hits_t,
    const float* __restrict__ noise, MarchParams p, int64_t n_rays, const int64_t* __restrict__ rays_a,
    int64_t cap, float* __restrict__ xyzs, float* __restrict__ dirs, float* __restrict__ deltas,
    float* __restrict__ ts) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rays) return;
  const int64_t start = rays_a[3 * r + 1];
  const int n_samples = (int)rays_a[3 * r + 2];
  if (n_samples == 0) return;
  const Ray ray = load_ray(rays_o, rays_d, r);
  float t1 = __ldg(hits_t + 2 * r);
  const float t2 = __ldg(hits_t + 2 * r + 1);
  if (t1 >= 0) t1 = __fmaf_rn(calc_dt(t1, p), __ldg(noise + r), t1);
  float t = t1; int k = 0; Probe s;
  while (t < t2 && k < n_samples) {
    if (march_probe(ray, p, t, s)) {
      const int64_t o = start + k;
      if (o < cap) {
        xyzs[3 * o] = s.x; xyzs[3 * o + 1] = s.y; xyzs[3 * o + 2] = s.z;
        dirs[3 * o] = ray.dx; dirs[3 * o + 1] = ray.dy; dirs[3 * o + 2] = ray.dz;
        ts[o] = t; deltas[o] = s.dt;
      }
      t = __fadd_rn(t, s.dt); k++;
    }
  }
}

// ---- test time -------------------------------------------------------------------------------
// One thread per alive ray; the (n_alive, N_samples) slabs are written completely (zeros
// behind N_eff) so the caller never memsets (raymarching.cu:421-426 does).
__global__ void __launch_bounds__(128) march_test_kernel(
    const float* __restrict__ rays_o, const float* __restrict__ rays_d, float* __restrict__ hits_t,
    const int64_t* __restrict__ alive, int64_t n_alive, MarchParams p, int N_samples,
    float* __restrict__ xyzs, float* __restrict__ dirs, float* __restrict__ deltas, float* __restrict__ ts,
    int32_t* __restrict__ n_eff) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_alive) return;
  const int64_t r = alive[n];
  const Ray ray = load_ray(rays_o, rays_d, r);
  float t = hits_t[2 * r];
  const float t2 = hits_t[2 * r + 1];
  int k = 0; Probe s;
  float t_resume = t;
  const int64_t o0 = n * N_samples;
  while (t < t2 && k < N_samples) {
    if (march_probe(ray, p, t, s)) {
      const int64_t o = o0 + k;
      xyzs[3 * o] = s.x; xyzs[3 * o + 1] = s.y; xyzs[3 * o + 2] = s.z;
      dirs[3 * o] = ray.dx; dirs[3 * o + 1] = ray.dy; dirs[3 * o + 2] = ray.dz;
      ts[o] = t; deltas[o] = s.dt;
      t = __fadd_rn(t, s.dt); k++;
      t_resume = t;  // only a taken sample moves the resume point (raymarching.cu:389-391)
    }
  }
  if (k > 0) hits_t[2 * r] = t_resume;
  n_eff[n] = k;
  for (int j = k; j < N_samples; j++) {
    const int64_t o = o0 + j;
    xyzs[3 * o] = 0.f; xyzs[3 * o + 1] = 0.f; xyzs[3 * o + 2] = 0.f;
    dirs[3 * o] = 0.f; dirs[3 * o + 1] = 0.f; dirs[3 * o + 2] = 0.f;
    ts[o] = 0.f; deltas[o] = 0.f;
  }
}

}  // namespace rnb

using namespace rnb;

extern "C" {

int rnb_raymarching_train_count(const float* rays_o, const float* rays_d, const float* hits_t,
                                const uint8_t* density_bitfield, int cascades, float scale,
                                float exp_step_factor, const float* noise, int grid_size, int max_samples,
                                int64_t n_rays, int64_t* rays_a, int32_t* counter, rnb_stream_t stream) {
  const char* what = "rnb_raymarching_train_count";
  RNB_REQUIRE(n_rays >= 0 && cascades >= 1 && grid_size >= 1 && grid_size <= 1024 && max_samples >= 1, what);
  RNB_REQUIRE(counter != nullptr, what);
  RNB_REQUIRE(n_rays == 0 || (rays_o && rays_d && hits_t && density_bitfield && noise && rays_a), what);
  const MarchParams p = make_march_params(density_bitfield, cascades, scale, exp_step_factor, grid_size, max_samples, scale);
  if (n_rays > 0)
    march_train_count_kernel<<<ceil_div(n_rays, 64), 64, 0, as_stream(stream)>>>(rays_o, rays_d, hits_t, noise, p, n_rays, rays_a);
  march_train_scan_kernel<<<1, 1024, 0, as_stream(stream)>>>(n_rays, rays_a, counter);
  return check_launch(what);
}

int rnb_raymarching_train_write(const float* rays_o, const float* rays_d, const float* hits_t,
                                const uint8_t* density_bitfield, int cascades, float scale,
                                float exp_step_factor, const float* noise, int grid_size, int max_samples,
                                int64_t n_rays, const int64_t* rays_a, int64_t cap, float* xyzs, float* dirs,
                                float* deltas, float* ts, rnb_stream_t stream) {
  const char* what = "rnb_raymarching_train_write";
  RNB_REQUIRE(n_rays >= 0 && cap >= 0 && cascades >= 1 && grid_size >= 1 && grid_size <= 1024 && max_samples >= 1, what);
  if (n_rays == 0 || cap == 0) return RNB_OK;
  RNB_REQUIRE(rays_o && rays_d && hits_t && density_bitfield && noise && rays_a && xyzs && dirs && deltas && ts, what);
  const MarchParams p = make_march_params(density_bitfield, cascades, scale, exp_step_factor, grid_size, max_samples, scale);
  march_train_write_kernel<<<ceil_div(n_rays, 64), 64, 0, as_stream(stream)>>>(rays_o, rays_d, hits_t, noise, p, n_rays, rays_a, cap, xyzs, dirs, deltas, ts);
  return check_launch(what);
}

int rnb_raymarching_test(const float* rays_o, const float* rays_d, float* hits_t, const int64_t* alive_indices,
                         int64_t n_alive, const uint8_t* density_bitfield, int cascades, float scale,
                         float exp_step_factor, int grid_size, int max_samples, int N_samples, float* xyzs,
                         float* dirs, float* deltas, float* ts, int32_t* N_eff_samples, rnb_stream_t stream) {
  const char* what = "rnb_raymarching_test";
  RNB_REQUIRE(n_alive >= 0 && cascades >= 1 && grid_size >= 1 && grid_size <= 1024 && max_samples >= 1 && N_samples >= 1, what);
  if (n_alive == 0) return RNB_OK;
  RNB_REQUIRE(rays_o && rays_d && hits_t && alive_indices && density_bitfield && xyzs && dirs && deltas && ts && N_eff_samples, what);
  // hazard Q1: the reference's test kernel hands `cascades` to calc_dt where `scale` is expected
  const MarchParams p = make_march_params(density_bitfield, cascades, scale, exp_step_factor, grid_size, max_samples, (float)cascades);
  march_test_kernel<<<ceil_div(n_alive, 128), 128, 0, as_stream(stream)>>>(rays_o, rays_d, hits_t, alive_indices, n_alive, p, N_samples, xyzs, dirs, deltas, ts, N_eff_samples);
  return check_launch(what);
}

}  // extern "C"
