// composite.cuh -- warp-per-ray front-to-back compositing over one CSR segment (device functions
// shared by composite.cu, the operator-level kernels, and engine.cu, the fused training step).
// See composite.cu for the design notes (bit-identical transmittance chain, warp reductions).
#pragma once
#include "common.cuh"

namespace rnb {

// sample colour accessor: kHalf4 = false: (N,3) f32 rows; true: (N,4) f16 rows (the fused field's output)
template <bool kHalf4>
__device__ __forceinline__ void load_rgb(const void* __restrict__ rgbs, int64_t s, float& r, float& g, float& b) {
  if (kHalf4) {
    const uint2 q = __ldg(reinterpret_cast<const uint2*>(rgbs) + s);
    const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&q.x));
    const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&q.y));
    r = lo.x; g = lo.y; b = hi.x;
  } else {
    const float* p = reinterpret_cast<const float*>(rgbs) + 3 * s;
    r = ld_stream_f32(p); g = ld_stream_f32(p + 1); b = ld_stream_f32(p + 2);
  }
}

// Serial transmittance over the 32 lanes.  On entry T_in = transmittance before lane 0's
// sample (warp-uniform); returns this lane's transmittance BEFORE its sample and updates
// T_in to the value after lane 31's sample.
// Every lane replays the chain ((T_in * om_0) * om_1) * ... itself, stopping at its own index, from a
// shared-memory copy of the 32 factors read as eight broadcast 16-byte loads: the loop-carried dependency is the
// 32 multiplies alone (~130 cycles per chunk).  The earlier form -- one SHFL per factor feeding one running
// product -- compiled to SHFL -> FMUL pairs through a single register, ~1000 cycles per chunk, and the longest
// ray of the batch (up to 32 chunks) set the duration of the whole compositing kernel.
__device__ __forceinline__ float serial_T(float one_minus_a, float& T_in, int lane) {
  __shared__ __align__(16) float s_om[32][32];      // [warp of the CTA][lane]
  float* row = s_om[threadIdx.x >> 5];
  __syncwarp();                                     // the previous chunk's reads are complete
  row[lane] = one_minus_a;
  __syncwarp();
  float mine = T_in;
#pragma unroll
  for (int q = 0; q < 8; q++) {
    const float4 v = reinterpret_cast<const float4*>(row)[q];
    if (4 * q + 0 < lane) mine = __fmul_rn(mine, v.x);
    if (4 * q + 1 < lane) mine = __fmul_rn(mine, v.y);
    if (4 * q + 2 < lane) mine = __fmul_rn(mine, v.z);
    if (4 * q + 3 < lane) mine = __fmul_rn(mine, v.w);
  }
  T_in = __shfl_sync(kFullMask, __fmul_rn(mine, one_minus_a), 31);
  return mine;
}

// One lane's sample of a chunk (zeros past the end of the segment).
struct SampleIn { float sigma, dl, r, g, b, t; };
template <bool kHalf4>
__device__ __forceinline__ SampleIn load_sample(const float* __restrict__ sigmas, const void* __restrict__ rgbs,
                                                const float* __restrict__ deltas, const float* __restrict__ ts,
                                                int64_t start, int k, int N) {
  SampleIn v = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (k < N) {
    const int64_t s = start + k;
    v.sigma = ld_stream_f32(sigmas + s); v.dl = ld_stream_f32(deltas + s);
    load_rgb<kHalf4>(rgbs, s, v.r, v.g, v.b);
    v.t = ld_stream_f32(ts + s);
  }
  return v;
}

struct RayAccum { float R, G, B, D, O; int total; };

// Forward over the segment [start, start+N) (volumerendering.cu:23-44).  Writes ws for every sample
// of the segment (0 behind the terminating sample, Q8); returns warp-uniform sums.
template <bool kHalf4>
__device__ __forceinline__ RayAccum composite_fw_segment(const float* __restrict__ sigmas, const void* __restrict__ rgbs,
                                                         const float* __restrict__ deltas, const float* __restrict__ ts,
                                                         int64_t start, int N, float T_threshold, float* __restrict__ ws) {
  const int lane = threadIdx.x & 31;
  float T = 1.0f, R = 0.f, G = 0.f, B = 0.f, D = 0.f, O = 0.f;
  int total = N;
  int base = 0;
  // The chunk loop is software-pipelined: the loads of chunk c+1 are issued before the serial chain of chunk c,
  // so a long ray pays one memory latency, not one per 32 samples (the longest ray sets the kernel's duration).
  SampleIn cur = load_sample<kHalf4>(sigmas, rgbs, deltas, ts, start, lane, N);
  for (; base < N; base += 32) {
    const SampleIn nxt = load_sample<kHalf4>(sigmas, rgbs, deltas, ts, start, base + 32 + lane, N);
    const int k = base + lane;
    const bool valid = k < N;
    const int64_t s = start + k;
    const float a = valid ? 1.0f - __expf(-cur.sigma * cur.dl) : 0.f;
    const float cr = cur.r, cg = cur.g, cb = cur.b, tt = cur.t;
    const float om = 1.0f - a;
    const float Tb = serial_T(om, T, lane);
    const float Ta = __fmul_rn(Tb, om);
    const unsigned term = __ballot_sync(kFullMask, valid && Ta <= T_threshold);
    const int first = term ? (__ffs(term) - 1) : 32;
    const float w = (valid && lane <= first) ? __fmul_rn(a, Tb) : 0.f;
    if (valid) ws[s] = w;
    R = fmaf(w, cr, R); G = fmaf(w, cg, G); B = fmaf(w, cb, B); D = fmaf(w, tt, D); O += w;
    if (term) { total = base + first; base += 32; break; }
    cur = nxt;
  }
  // samples behind the terminating chunk keep ws = 0 (reference: zero-initialised, Q8)
  for (int k = base + lane; k < N; k += 32) ws[start + k] = 0.f;
  RayAccum r;
  r.R = warp_sum(R); r.G = warp_sum(G); r.B = warp_sum(B); r.D = warp_sum(D); r.O = warp_sum(O); r.total = total;
  return r;
}

struct RayGrad { float Rf, Gf, Bf, Of, Df, gr, gg, gb, gO, gD; };

// Backward (volumerendering.cu:109-151).  Per sample, with T = transmittance AFTER the sample
// and r,g,b,d the inclusive prefix sums of w*c / w*t:
//   dL_drgbs = dL_drgb * w
//   dL_dsigma = delta * ( sum_c dL_drgb_c * (c*T - (C - c_prefix)) + dL_dopacity * (1 - O)
//                         + dL_ddepth * (t*T - (D - d)) + T*dL_dws - (sum(dL_dws*ws) - prefix) )
template <bool kHalf4>
__device__ __forceinline__ void composite_bw_segment(const RayGrad q, const float* __restrict__ dL_dws,
                                                     const float* __restrict__ sigmas, const void* __restrict__ rgbs,
                                                     const float* __restrict__ ws, const float* __restrict__ deltas,
                                                     const float* __restrict__ ts, int64_t start, int N, float T_threshold,
                                                     float* __restrict__ dL_dsigmas, float* __restrict__ dL_drgbs) {
  const int lane = threadIdx.x & 31;
  const float Rf = q.Rf, Gf = q.Gf, Bf = q.Bf, Of = q.Of, Df = q.Df, gr = q.gr, gg = q.gg, gb = q.gb, gO = q.gO, gD = q.gD;
  float ws_total = 0.f;
  if (dL_dws) {
    for (int k = lane; k < N; k += 32) ws_total = fmaf(dL_dws[start + k], ws[start + k], ws_total);
    ws_total = warp_sum(ws_total);
  }
  float T = 1.0f, r = 0.f, g = 0.f, b = 0.f, d = 0.f, wsrun = 0.f;  // carries (warp-uniform)
  int base = 0;
  // software-pipelined like the forward loop: chunk c+1 is in flight while chunk c's chain and scans run
  SampleIn cur = load_sample<kHalf4>(sigmas, rgbs, deltas, ts, start, lane, N);
  float cgw = 0.f, cwv = 0.f;
  if (dL_dws && lane < N) { cgw = ld_stream_f32(dL_dws + start + lane); cwv = ld_stream_f32(ws + start + lane); }
  for (; base < N; base += 32) {
    const int kn = base + 32 + lane;
    const SampleIn nxt = load_sample<kHalf4>(sigmas, rgbs, deltas, ts, start, kn, N);
    float ngw = 0.f, nwv = 0.f;
    if (dL_dws && kn < N) { ngw = ld_stream_f32(dL_dws + start + kn); nwv = ld_stream_f32(ws + start + kn); }
    const int k = base + lane;
    const bool valid = k < N;
    const int64_t s = start + k;
    const float dl = cur.dl, cr = cur.r, cg = cur.g, cb = cur.b, tt = cur.t, gw = cgw, wv = cwv;
    const float a = valid ? 1.0f - __expf(-cur.sigma * dl) : 0.f;
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
    cur = nxt; cgw = ngw; cwv = nwv;
  }
  for (int k = base + lane; k < N; k += 32) {
    const int64_t s = start + k;
    dL_dsigmas[s] = 0.f; dL_drgbs[3 * s] = 0.f; dL_drgbs[3 * s + 1] = 0.f; dL_drgbs[3 * s + 2] = 0.f;
  }
}

}  // namespace rnb
