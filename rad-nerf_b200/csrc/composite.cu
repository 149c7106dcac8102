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
#include "common.cuh"

namespace rnb {

// Serial transmittance over the 32 lanes.  On entry T_in = transmittance before lane 0's
// sample (warp-uniform); returns this lane's transmittance BEFORE its sample and updates
// T_in to the value after lane 31's sample.
__device__ __forceinline__ float serial_T(float one_minus_a, float& T_in, int lane) {
  float run = T_in, mine = T_in;
#pragma unroll
  for (int j = 0; j < 32; j++) {
    const float om = __shfl_sync(kFullMask, one_minus_a, j);
    if (lane == j) mine = run;
    run = __fmul_rn(run, om);
  }
  T_in = run;
  return mine;
}

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
  float T = 1.0f, R = 0.f, G = 0.f, B = 0.f, D = 0.f, O = 0.f;
  int total = N;
  int base = 0;
  for (; base < N; base += 32) {
    const int k = base + lane;
    const bool valid = k < N;
    const int64_t s = start + k;
    float a = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, tt = 0.f;
    if (valid) {
      a = 1.0f - __expf(-ld_stream_f32(sigmas + s) * ld_stream_f32(deltas + s));
      cr = ld_stream_f32(rgbs + 3 * s); cg = ld_stream_f32(rgbs + 3 * s + 1); cb = ld_stream_f32(rgbs + 3 * s + 2);
      tt = ld_stream_f32(ts + s);
    }
    const float om = 1.0f - a;
    const float Tb = serial_T(om, T, lane);
    const float Ta = __fmul_rn(Tb, om);
    const unsigned term = __ballot_sync(kFullMask, valid && Ta <= T_threshold);
    const int first = term ? (__ffs(term) - 1) : 32;
    const float w = (valid && lane <= first) ? __fmul_rn(a, Tb) : 0.f;
    if (valid) ws[s] = w;
    R = fmaf(w, cr, R); G = fmaf(w, cg, G); B = fmaf(w, cb, B); D = fmaf(w, tt, D); O += w;
    if (term) { total = base + first; base += 32; break; }
  }
  // samples behind the terminating chunk keep ws = 0 (reference: zero-initialised, Q8)
  for (int k = base + lane; k < N; k += 32) ws[start + k] = 0.f;
  R = warp_sum(R); G = warp_sum(G); B = warp_sum(B); D = warp_sum(D); O = warp_sum(O);
  if (lane == 0) {
    rgb[3 * ray_idx] = R; rgb[3 * ray_idx + 1] = G; rgb[3 * ray_idx + 2] = B;
    depth[ray_idx] = D; opacity[ray_idx] = O; total_samples[ray_idx] = total;
  }
}

// Backward (volumerendering.cu:109-151).  Per sample, with T = transmittance AFTER the sample
// and r,g,b,d the inclusive prefix sums of w*c / w*t:
//   dL_drgbs = dL_drgb * w
//   dL_dsigma = delta * ( sum_c dL_drgb_c * (c*T - (C - c_prefix)) + dL_dopacity * (1 - O)
//                         + dL_ddepth * (t*T - (D - d)) + T*dL_dws - (sum(dL_dws*ws) - prefix) )
__global__ void __launch_bounds__(256) composite_train_bw_kernel(
    const float* __restrict__ dL_dopacity, const float* __restrict__ dL_ddepth, const float* __restrict__ dL_drgb,
    const float* __restrict__ dL_dws, const float* __restrict__ sigmas, const float* __restrict__ rgbs,
    const float* __restrict__ ws, const float* __restrict__ deltas, const float* __restrict__ ts,
    const int64_t* __restrict__ rays_a, const float* __restrict__ opacity, const float* __restrict__ depth,
    const float* __restrict__ rgb, float T_threshold, int64_t n_rays, float* __restrict__ dL_dsigmas,
    float* __restrict__ dL_drgbs) {
  const int lane = threadIdx.x & 31;
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (n >= n_rays) return;
  const int64_t ray_idx = rays_a[3 * n], start = rays_a[3 * n + 1];
  const int N = (int)rays_a[3 * n + 2];
  if (N == 0) return;
  const float Rf = rgb[3 * ray_idx], Gf = rgb[3 * ray_idx + 1], Bf = rgb[3 * ray_idx + 2];
  const float Of = opacity[ray_idx], Df = depth[ray_idx];
  const float gr = dL_drgb[3 * ray_idx], gg = dL_drgb[3 * ray_idx + 1], gb = dL_drgb[3 * ray_idx + 2];
  const float gO = dL_dopacity[ray_idx], gD = dL_ddepth[ray_idx];
  float ws_total = 0.f;
  if (dL_dws) {
    for (int k = lane; k < N; k += 32) ws_total = fmaf(dL_dws[start + k], ws[start + k], ws_total);
    ws_total = warp_sum(ws_total);
  }
  float T = 1.0f, r = 0.f, g = 0.f, b = 0.f, d = 0.f, wsrun = 0.f;  // carries (warp-uniform)
  int base = 0;
  for (; base < N; base += 32) {
    const int k = base + lane;
    const bool valid = k < N;
    const int64_t s = start + k;
    float a = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, tt = 0.f, dl = 0.f, gw = 0.f, wv = 0.f;
    if (valid) {
      dl = ld_stream_f32(deltas + s);
      a = 1.0f - __expf(-ld_stream_f32(sigmas + s) * dl);
      cr = ld_stream_f32(rgbs + 3 * s); cg = ld_stream_f32(rgbs + 3 * s + 1); cb = ld_stream_f32(rgbs + 3 * s + 2);
      tt = ld_stream_f32(ts + s);
      if (dL_dws) { gw = ld_stream_f32(dL_dws + s); wv = ld_stream_f32(ws + s); }
    }
    const float om = 1.0f - a;
    const float Tb = serial_T(om, T, lane);
    const float Ta = __fmul_rn(Tb, om);
    const unsigned term = __ballot_sync(kFullMask, valid && Ta <= T_threshold);
    const int first = term ? (__ffs(term) - 1) : 32;
    const bool live = valid && lane <= first;
    const float w = live ? __fmul_rn(a, Tb) : 0.f;
    const float pr = r + warp_incl_scan(w * cr, lane), pg = g + warp_incl_scan(w * cg, lane);
    const float pb = b + warp_incl_scan(w * cb, lane), pd = d + warp_incl_scan(w * tt, lane);
    const float pw = wsrun + warp_incl_scan(live ? gw * wv : 0.f, lane);
    if (valid) {
      float gs = 0.f, g0 = 0.f, g1 = 0.f, g2 = 0.f;
      if (live) {
        g0 = gr * w; g1 = gg * w; g2 = gb * w;
        gs = dl * (gr * (cr * Ta - (Rf - pr)) + gg * (cg * Ta - (Gf - pg)) + gb * (cb * Ta - (Bf - pb)) +
                   gO * (1.0f - Of) + gD * (tt * Ta - (Df - pd)) + Ta * gw - (ws_total - pw));
      }
      dL_dsigmas[s] = gs;
      dL_drgbs[3 * s] = g0; dL_drgbs[3 * s + 1] = g1; dL_drgbs[3 * s + 2] = g2;
    }
    if (term) { base += 32; break; }
    r = __shfl_sync(kFullMask, pr, 31); g = __shfl_sync(kFullMask, pg, 31); b = __shfl_sync(kFullMask, pb, 31);
    d = __shfl_sync(kFullMask, pd, 31); wsrun = __shfl_sync(kFullMask, pw, 31);
  }
  for (int k = base + lane; k < N; k += 32) {
    const int64_t s = start + k;
    dL_dsigmas[s] = 0.f; dL_drgbs[3 * s] = 0.f; dL_drgbs[3 * s + 1] = 0.f; dL_drgbs[3 * s + 2] = 0.f;
  }
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
