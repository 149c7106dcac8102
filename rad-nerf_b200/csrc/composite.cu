// composite.cu -- front-to-back volume compositing with early termination
// (replaces models/csrc/volumerendering.cu of the reference).
//
// Training kernels are WARP-PER-RAY: the 32 lanes load 32 consecutive samples of the ray's
// CSR segment (coalesced), evaluate alpha = 1 - __expf(-sigma*delta) in parallel, and the
// transmittance chain T_k = T_{k-1} * (1 - alpha_k) is replayed with one SHFL + FMUL per
// sample in exactly the reference's serial order -- so T, the weights ws and the
// termination index (total_samples) are bit-identical to the reference's one-thread-per-ray
// loop (volumerendering.cu:27-44), while the per-ray sums are warp tree reductions (fp32
// re-association only).  HBM-bound: 28 B/sample forward, 48 B/sample backward.
#include "composite.cuh"

namespace rnb {

__global__ void __launch_bounds__(256) composite_train_fw_kernel(
    const float* __restrict__ sigmas, const float* __restrict__ rgbs, const float* __restrict__ deltas,
    const float* __restrict__ ts, const int64_t* __restrict__ rays_a, float T_threshold, int64_t n_rays,
    int64_t* __restrict__ total_samples, float* __restrict__ opacity, float* __restrict__ depth,
    float* __restrict__ rgb, float* __restrict__ ws) {
  const int lane = threadIdx.x & 31;
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (n >= n_rays) return;
  const int64_t ray_idx = rays_a[3 * n], start = rays_a[3 * n + 1];
  const int N = (int)rays_a[3 * n + 2];
  const RayAccum r = composite_fw_segment<false>(sigmas, rgbs, deltas, ts, start, N, T_threshold, ws);
  if (lane == 0) {
    rgb[3 * ray_idx] = r.R; rgb[3 * ray_idx + 1] = r.G; rgb[3 * ray_idx + 2] = r.B;
    depth[ray_idx] = r.D; opacity[ray_idx] = r.O; total_samples[ray_idx] = r.total;
  }
}

__global__ void __launch_bounds__(256) composite_train_bw_kernel(
    const float* __restrict__ dL_dopacity, const float* __restrict__ dL_ddepth, const float* __restrict__ dL_drgb,
    const float* __restrict__ dL_dws, const float* __restrict__ sigmas, const float* __restrict__ rgbs,
    const float* __restrict__ ws, const float* __restrict__ deltas, const float* __restrict__ ts,
    const int64_t* __restrict__ rays_a, const float* __restrict__ opacity, const float* __restrict__ depth,
    const float* __restrict__ rgb, float T_threshold, int64_t n_rays, float* __restrict__ dL_dsigmas,
    float* __restrict__ dL_drgbs) {
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (n >= n_rays) return;
  const int64_t ray_idx = rays_a[3 * n], start = rays_a[3 * n + 1];
  const int N = (int)rays_a[3 * n + 2];
  if (N == 0) return;
  RayGrad q;
  q.Rf = rgb[3 * ray_idx]; q.Gf = rgb[3 * ray_idx + 1]; q.Bf = rgb[3 * ray_idx + 2];
  q.Of = opacity[ray_idx]; q.Df = depth[ray_idx];
  q.gr = dL_drgb[3 * ray_idx]; q.gg = dL_drgb[3 * ray_idx + 1]; q.gb = dL_drgb[3 * ray_idx + 2];
  q.gO = dL_dopacity[ray_idx]; q.gD = dL_ddepth[ray_idx];
  composite_bw_segment<false>(q, dL_dws, sigmas, rgbs, ws, deltas, ts, start, N, T_threshold, dL_dsigmas, dL_drgbs);
}

// Test-time in-place compositing (volumerendering.cu:206-250): one thread per alive ray
// (N_eff <= 64 samples); per-ray state is kept in registers and written back once.
__global__ void __launch_bounds__(128) composite_test_fw_kernel(
    const float* __restrict__ sigmas, const float* __restrict__ rgbs, const float* __restrict__ deltas,
    const float* __restrict__ ts, int64_t* __restrict__ alive, float T_threshold,
    const int32_t* __restrict__ n_eff, int64_t n_alive, int N_samples, float* __restrict__ opacity,
    float* __restrict__ depth, float* __restrict__ rgb) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_alive) return;
  const int ne = n_eff[n];
  if (ne == 0) { alive[n] = -1; return; }
  const int64_t r = alive[n];
  float O = opacity[r], D = depth[r], R = rgb[3 * r], G = rgb[3 * r + 1], B = rgb[3 * r + 2];
  float T = 1.0f - O;
  const int64_t o0 = n * N_samples;
  for (int s = 0; s < ne; s++) {
    const int64_t o = o0 + s;
    const float a = 1.0f - __expf(-sigmas[o] * deltas[o]);
    const float w = a * T;
    R = fmaf(w, rgbs[3 * o], R); G = fmaf(w, rgbs[3 * o + 1], G); B = fmaf(w, rgbs[3 * o + 2], B);
    D = fmaf(w, ts[o], D); O += w;
    T *= 1.0f - a;
    if (T <= T_threshold) { alive[n] = -1; break; }
  }
  opacity[r] = O; depth[r] = D; rgb[3 * r] = R; rgb[3 * r + 1] = G; rgb[3 * r + 2] = B;
}

}  // namespace rnb

using namespace rnb;

extern "C" {

int rnb_composite_train_fw(const float* sigmas, const float* rgbs, const float* deltas, const float* ts,
                           const int64_t* rays_a, float T_threshold, int64_t n_rays, int64_t N,
                           int64_t* total_samples, float* opacity, float* depth, float* rgb, float* ws,
                           rnb_stream_t stream) {
  const char* what = "rnb_composite_train_fw";
  RNB_REQUIRE(n_rays >= 0 && N >= 0, what);
  if (n_rays == 0) return RNB_OK;
  RNB_REQUIRE(rays_a && total_samples && opacity && depth && rgb, what);
  RNB_REQUIRE(N == 0 || (sigmas && rgbs && deltas && ts && ws), what);
  composite_train_fw_kernel<<<ceil_div(n_rays * 32, 256), 256, 0, as_stream(stream)>>>(
      sigmas, rgbs, deltas, ts, rays_a, T_threshold, n_rays, total_samples, opacity, depth, rgb, ws);
  return check_launch(what);
}

int rnb_composite_train_bw(const float* dL_dopacity, const float* dL_ddepth, const float* dL_drgb,
                           const float* dL_dws, const float* sigmas, const float* rgbs, const float* ws,
                           const float* deltas, const float* ts, const int64_t* rays_a, const float* opacity,
                           const float* depth, const float* rgb, float T_threshold, int64_t n_rays, int64_t N,
                           float* dL_dsigmas, float* dL_drgbs, rnb_stream_t stream) {
  const char* what = "rnb_composite_train_bw";
  RNB_REQUIRE(n_rays >= 0 && N >= 0, what);
  if (n_rays == 0 || N == 0) return RNB_OK;
  RNB_REQUIRE(dL_dopacity && dL_ddepth && dL_drgb && sigmas && rgbs && deltas && ts && rays_a && opacity && depth && rgb && dL_dsigmas && dL_drgbs, what);
  RNB_REQUIRE(dL_dws == nullptr || ws != nullptr, what);
  composite_train_bw_kernel<<<ceil_div(n_rays * 32, 256), 256, 0, as_stream(stream)>>>(
      dL_dopacity, dL_ddepth, dL_drgb, dL_dws, sigmas, rgbs, ws, deltas, ts, rays_a, opacity, depth, rgb,
      T_threshold, n_rays, dL_dsigmas, dL_drgbs);
  return check_launch(what);
}

int rnb_composite_test_fw(const float* sigmas, const float* rgbs, const float* deltas, const float* ts,
                          const float* hits_t, int64_t* alive_indices, float T_threshold,
                          const int32_t* N_eff_samples, int64_t n_alive, int N_samples, float* opacity,
                          float* depth, float* rgb, rnb_stream_t stream) {
  const char* what = "rnb_composite_test_fw";
  (void)hits_t;  // unused by the reference kernel as well (volumerendering.cu:212)
  RNB_REQUIRE(n_alive >= 0 && N_samples >= 1, what);
  if (n_alive == 0) return RNB_OK;
  RNB_REQUIRE(sigmas && rgbs && deltas && ts && alive_indices && N_eff_samples && opacity && depth && rgb, what);
  composite_test_fw_kernel<<<ceil_div(n_alive, 128), 128, 0, as_stream(stream)>>>(
      sigmas, rgbs, deltas, ts, alive_indices, T_threshold, N_eff_samples, n_alive, N_samples, opacity, depth, rgb);
  return check_launch(what);
}

}  // extern "C"
