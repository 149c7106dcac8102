// engine.cu -- one Rad-NeRF training step as a short chain of kernels with no host synchronisation.
//
// Replaces everything below the data loader in NeRFSystem.training_step (train_ml.py:172-200):
//   Ray_Gate (networks.py:1087-1093)                 -> gate_in + rnb_mlp_fwd + gate_softmax
//   ml_render's K-loop (ml_rendering.py:47-68)       -> eng_march (all K passes, AABB inline) -> eng_scan
//                                                       -> eng_expand -> rnb_field_fwd -> eng_composite_fw
//   background / gated blend / NeRFLoss (losses.py)  -> eng_blend_loss (also emits every gradient seed)
//   autograd backward                                -> eng_composite_bw -> rnb_field_bwd -> rnb_mlp_bwd (gate)
//   FusedAdam                                        -> eng_adam (also refreshes the fp16 copy, zeroes grads)
// The reference runs ~100 launches and K host round trips (`counter[0]`) per step; here sample counts
// live in device memory (`seg`), every kernel sizes its work from them, and the whole step can be
// captured in a CUDA graph.  Samples of the K passes are packed into one list (128-aligned pass bases).
#include "aabb.cuh"
#include "composite.cuh"
#include "march_warp.cuh"
#include "coarse.cuh"
#include "adam.cuh"

namespace rnb {

static inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

struct Carver {
  int64_t off = 0;
  int64_t take(int64_t bytes) { const int64_t o = off; off = align_up(off + bytes, 256); return o; }
};

// ---- gate -----------------------------------------------------------------------------------------
// gate input = cat(rays_o, rays_d) (raw, Q18), padded to 16 columns with ones like tinycudann
__global__ void __launch_bounds__(256) eng_gate_in_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                                          int n_rays, __half* __restrict__ gate_in) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rays) return;
  const float ox = rays_o[3 * r], oy = rays_o[3 * r + 1], oz = rays_o[3 * r + 2];
  const float dx = rays_d[3 * r], dy = rays_d[3 * r + 1], dz = rays_d[3 * r + 2];
  const __half2 one = __floats2half2_rn(1.f, 1.f);
  const __half2 a = __floats2half2_rn(ox, oy), b = __floats2half2_rn(oz, dx), c = __floats2half2_rn(dy, dz);
  uint4 q0, q1;
  q0.x = *reinterpret_cast<const uint32_t*>(&a); q0.y = *reinterpret_cast<const uint32_t*>(&b);
  q0.z = *reinterpret_cast<const uint32_t*>(&c); q0.w = *reinterpret_cast<const uint32_t*>(&one);
  q1.x = q1.y = q1.z = q1.w = *reinterpret_cast<const uint32_t*>(&one);
  reinterpret_cast<uint4*>(gate_in + (int64_t)r * 16)[0] = q0;
  reinterpret_cast<uint4*>(gate_in + (int64_t)r * 16)[1] = q1;
}

// softmax over the first K logits (fp32), importance[k] += sum over rays (networks.py:1090-1092)
__global__ void __launch_bounds__(256) eng_gate_softmax_kernel(const __half* __restrict__ logits, int n_rays, int K,
                                                               float* __restrict__ prob, float* __restrict__ importance) {
  __shared__ float red[8][RNB_MAX_K];
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float p[RNB_MAX_K];
#pragma unroll
  for (int k = 0; k < RNB_MAX_K; k++) p[k] = 0.f;
  if (r < n_rays) {
    float m = -INFINITY;
#pragma unroll
    for (int k = 0; k < RNB_MAX_K; k++) if (k < K) { p[k] = __half2float(logits[(int64_t)r * 16 + k]); m = fmaxf(m, p[k]); }
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < RNB_MAX_K; k++) if (k < K) { p[k] = expf(p[k] - m); sum += p[k]; }
#pragma unroll
    for (int k = 0; k < RNB_MAX_K; k++) if (k < K) { p[k] = p[k] / sum; prob[(int64_t)r * K + k] = p[k]; }
  }
#pragma unroll
  for (int k = 0; k < RNB_MAX_K; k++) {
    if (k < K) {
      const float s = warp_sum(p[k]);
      if (lane == 0) red[warp][k] = s;
    }
  }
  __syncthreads();
  if (threadIdx.x < K) {
    float s = 0.f;
    for (int w = 0; w < 8; w++) s += red[w][threadIdx.x];
    atomicAdd(importance + threadIdx.x, s);
  }
}

// ---- marching of all K passes -------------------------------------------------------------------------
struct BitfieldPtrs { const uint8_t* p[RNB_MAX_K]; };

template <bool kConstDt, bool kOneCascade>
__global__ void __launch_bounds__(256) eng_march_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                                        const float* __restrict__ noise, MarchParams p, BitfieldPtrs bf,
                                                        float cx, float cy, float cz, float hx, float hy, float hz, float near_d,
                                                        int n_rays, int32_t* __restrict__ counts, float* __restrict__ t_scratch,
                                                        const uint64_t* __restrict__ coarse, uint64_t noise_seed,
                                                        const int32_t* __restrict__ march_calls, uint64_t noise_slot_salt) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int pass = blockIdx.y;
  if (r >= n_rays) return;
  const Ray ray = load_ray(rays_o, rays_d, r);
  // scene-box interval (intersection.cu:40-55) + NEAR_DISTANCE clamp (ml_rendering.py:48-50)
  const float2 hit = slab_test(ray.ox, ray.oy, ray.oz, ray.ix, ray.iy, ray.iz, cx, cy, cz, hx, hy, hz);
  float t1 = -1.f, t2 = -1.f;
  if (hit.y > 0) { t1 = fmaxf(hit.x, 0.0f); t2 = hit.y; }
  if (t1 >= 0 && t1 < near_d) t1 = near_d;
  // empty-space test against the dilated coarse grid (coarse.cuh): skip rays that cannot meet an occupied cell, stop the
  // others behind the last block they can meet
  if (coarse != nullptr && t1 >= 0 && t1 < t2) t2 = coarse_last_hit(ray, t1, t2, coarse + (int64_t)pass * p.cascades * kCoarseWords, p.cascades, p.scale);
  // perturbation (raymarching.cu:195-198): the caller's noise, or -- none given -- the library's counter generator keyed by
  // (seed, march call, pass, ray): one fresh draw per sub-NeRF call like the reference's rand_like (custom_functions.py:83)
  if (t1 >= 0) {
    const float u = noise != nullptr ? __ldg(noise + (int64_t)pass * n_rays + r)
                                     : counter_uniform(noise_seed, ((uint64_t)(uint32_t)*march_calls * gridDim.y + pass) * (uint64_t)n_rays + r + noise_slot_salt);
    t1 = __fmaf_rn(calc_dt(t1, p), u, t1);
  }
  float t_next;
  const int64_t row = (int64_t)pass * n_rays + r;
  const int n = t2 < 0 ? 0 : march_ray_warp<kConstDt, kOneCascade>(ray, p, bf.p[pass], t1, t2, p.max_samples, t_scratch + row * p.max_samples, t_next);
  if ((threadIdx.x & 31) == 0) counts[row] = n;
}

// Exclusive scan of the per-ray counts of every pass (single CTA).  Pass bases are rounded up to 128
// so that a 128-sample tile never straddles two passes.  seg = [base_0..base_{K-1}, end, n_0..n_{K-1},
// overflow, total].
__global__ void __launch_bounds__(1024) eng_scan_kernel(const int32_t* __restrict__ counts, int n_rays, int K, int64_t capacity,
                                                        int32_t* __restrict__ starts, int32_t* __restrict__ seg, int32_t* __restrict__ march_calls) {
  __shared__ int warp_tot[32];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  int base = 0, total = 0;
  for (int pass = 0; pass < K; pass++) {
    if (tid == 0) carry_s = 0;
    __syncthreads();
    const int32_t* c_in = counts + (int64_t)pass * n_rays;
    int32_t* s_out = starts + (int64_t)pass * n_rays;
    for (int b0 = 0; b0 < n_rays; b0 += 4096) {
      int c[4]; int mine = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int r = b0 + tid * 4 + k;
        c[k] = r < n_rays ? c_in[r] : 0;
        mine += c[k];
      }
      const int incl = warp_incl_scan_i(mine, lane);
      if (lane == 31) warp_tot[wid] = incl;
      __syncthreads();
      if (wid == 0) warp_tot[lane] = warp_incl_scan_i(warp_tot[lane], lane);
      __syncthreads();
      const int carry = carry_s;
      int off = base + carry + (wid ? warp_tot[wid - 1] : 0) + incl - mine;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int r = b0 + tid * 4 + k;
        if (r < n_rays) s_out[r] = off;
        off += c[k];
      }
      __syncthreads();
      if (tid == 0) carry_s = carry + warp_tot[31];
      __syncthreads();
    }
    const int n_pass = carry_s;
    if (tid == 0) { seg[pass] = base; seg[K + 1 + pass] = n_pass; }
    total += n_pass;
    base = (base + n_pass + 127) & ~127;
    __syncthreads();
  }
  if (tid == 0) {
    seg[K] = base;
    seg[2 * K + 1] = (int64_t)base > capacity ? 1 : 0;
    seg[2 * K + 2] = total;
    *march_calls += 1;          // the marching kernel in front of this one has drawn its noise: next call, next stream of draws
  }
}

// per-ray sample rows -> packed list (ray id, t, dt); warp per (pass, ray)
__global__ void __launch_bounds__(256) eng_expand_kernel(const int32_t* __restrict__ counts, const int32_t* __restrict__ starts,
                                                         const float* __restrict__ t_scratch, MarchParams p, int n_rays,
                                                         int64_t capacity, int32_t* __restrict__ ray_id, float* __restrict__ ts,
                                                         float* __restrict__ deltas) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= n_rays) return;
  const int64_t row = (int64_t)blockIdx.y * n_rays + r;
  const int n = counts[row];
  const int64_t start = starts[row];
  const float* src = t_scratch + row * p.max_samples;
  for (int k = lane; k < n; k += 32) {
    const int64_t o = start + k;
    if (o >= capacity) break;
    const float t = ld_stream_f32(src + k);
    ray_id[o] = r; ts[o] = t; deltas[o] = calc_dt(t, p);
  }
}

// ---- compositing of all K passes ----------------------------------------------------------------------
__global__ void __launch_bounds__(256) eng_composite_fw_kernel(const float* __restrict__ sigmas, const __half* __restrict__ rgb4,
                                                               const float* __restrict__ deltas, const float* __restrict__ ts,
                                                               const int32_t* __restrict__ counts, const int32_t* __restrict__ starts,
                                                               float T_threshold, int n_rays, int64_t capacity,
                                                               float* __restrict__ ws, float* __restrict__ p_opacity,
                                                               float* __restrict__ p_depth, float* __restrict__ p_rgb,
                                                               int32_t* __restrict__ p_total) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rays) return;
  const int64_t row = (int64_t)blockIdx.y * n_rays + r;
  const int64_t start = starts[row];
  int N = counts[row];
  if (start + N > capacity) N = start < capacity ? (int)(capacity - start) : 0;   // overflow: the tail is dropped (flagged in seg)
  const RayAccum a = composite_fw_segment<true>(sigmas, rgb4, deltas, ts, start, N, T_threshold, ws);
  if ((threadIdx.x & 31) == 0) {
    p_rgb[3 * row] = a.R; p_rgb[3 * row + 1] = a.G; p_rgb[3 * row + 2] = a.B;
    p_depth[row] = a.D; p_opacity[row] = a.O; p_total[row] = a.total;
  }
}

__global__ void __launch_bounds__(256) eng_composite_bw_kernel(const float* __restrict__ sigmas, const __half* __restrict__ rgb4,
                                                               const float* __restrict__ deltas, const float* __restrict__ ts,
                                                               const int32_t* __restrict__ counts, const int32_t* __restrict__ starts,
                                                               float T_threshold, int n_rays, int64_t capacity,
                                                               const float* __restrict__ p_opacity, const float* __restrict__ p_depth,
                                                               const float* __restrict__ p_rgb, const float* __restrict__ d_opacity,
                                                               const float* __restrict__ d_depth, const float* __restrict__ d_rgb,
                                                               float* __restrict__ dsig, float* __restrict__ drgb) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rays) return;
  const int64_t row = (int64_t)blockIdx.y * n_rays + r;
  const int64_t start = starts[row];
  int N = counts[row];
  if (start + N > capacity) N = start < capacity ? (int)(capacity - start) : 0;
  if (N == 0) return;
  RayGrad q;
  q.Rf = p_rgb[3 * row]; q.Gf = p_rgb[3 * row + 1]; q.Bf = p_rgb[3 * row + 2];
  q.Of = p_opacity[row]; q.Df = p_depth[row];
  q.gr = d_rgb[3 * row]; q.gg = d_rgb[3 * row + 1]; q.gb = d_rgb[3 * row + 2];
  q.gO = d_opacity[row]; q.gD = d_depth[row];
  composite_bw_segment<true>(q, nullptr, sigmas, rgb4, nullptr, deltas, ts, start, N, T_threshold, dsig, drgb);
}


// ---- background + gated blend + NeRFLoss + every gradient seed ----------------------------------------
// One thread per ray.  Forward: rgb_i += bg_i (1 - opacity_i) (ml_rendering.py:192-200); rgb = sum_i g_i rgb_i,
// opacity = sum_i g_i opacity_i, depth[:, i] = depth_i (ml_rendering.py:64-68).  Loss (losses.py:52-74, summed
// term means as train_ml.py:192): rgb MSE, opacity entropy, CV^2 of the importance, depth-mutual.  Backward
// seeds: dL/d(rgb_i, depth_i, opacity_i) before the background blend, and dL/dlogits of the gate (softmax
// backward, scaled to fp16 for the gate MLP's backward).
struct BlendArgs {
  int n_rays, K;
  float inv_rays_global;
  float lambda_opacity, lambda_cv, lambda_dm, gs;
  float bg[RNB_MAX_K][3];
  const float* target; const float* prob; const float* importance;
  const float* p_opacity; const float* p_depth; const float* p_rgb;
  float* rgb; float* opacity; float* depth;
  float* d_opacity; float* d_depth; float* d_rgb; __half* dlogits;
  float* loss;
  int with_backward;
  // step guard (RNB_GRAD_TAIL): the marcher's overflow flag and aligned sample total, the step's flag word, the status block
  const int32_t* seg; int32_t* flags; int32_t* status;
};

__global__ void __launch_bounds__(256) eng_blend_loss_kernel(BlendArgs A) {
  __shared__ float red[8][4];
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int K = A.K, R = A.n_rays;
  if (A.flags != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    // samples beyond sample_capacity were dropped by every kernel of this step: the optimizer must not apply it
    if (A.seg[2 * K + 1]) atomicOr(A.flags, RNB_FLAG_CAPACITY);
    atomicMax(A.status + 2, A.seg[K]);
  }
  bool saturated = false;
  // CV^2 of the (global) importance and its derivative (losses.py:69-71; unbiased variance)
  float cv = 0.f, dcv[RNB_MAX_K];
#pragma unroll
  for (int k = 0; k < RNB_MAX_K; k++) dcv[k] = 0.f;
  if (K > 1 && A.lambda_cv > 0.f) {
    float mean = 0.f;
    for (int k = 0; k < K; k++) mean += A.importance[k];
    mean /= K;
    float var = 0.f;
    for (int k = 0; k < K; k++) { const float d = A.importance[k] - mean; var += d * d; }
    var /= (K - 1);
    const float den = mean * mean + 1e-10f;
    cv = A.lambda_cv * var / den;
#pragma unroll
    for (int k = 0; k < RNB_MAX_K; k++)
      if (k < K) dcv[k] = A.lambda_cv * (2.f * (A.importance[k] - mean) / (K - 1) / den - var * 2.f * mean / K / (den * den));
  }
  float l_rgb = 0.f, l_op = 0.f, l_dm = 0.f;
  if (r < R) {
    float g[RNB_MAX_K], o_i[RNB_MAX_K], d_i[RNB_MAX_K], c_i[RNB_MAX_K][3];
    float rgb[3] = {0.f, 0.f, 0.f}, op = 0.f, fused = 0.f;
#pragma unroll
    for (int k = 0; k < RNB_MAX_K; k++) {
      if (k < K) {
        const int64_t row = (int64_t)k * R + r;
        g[k] = K > 1 ? A.prob[(int64_t)r * K + k] : 1.f;
        o_i[k] = A.p_opacity[row]; d_i[k] = A.p_depth[row];
#pragma unroll
        for (int c = 0; c < 3; c++) {
          c_i[k][c] = A.p_rgb[3 * row + c] + A.bg[k][c] * (1.f - o_i[k]);
          rgb[c] = fmaf(c_i[k][c], g[k], rgb[c]);
        }
        op = fmaf(o_i[k], g[k], op);
        fused = fmaf(d_i[k], g[k], fused);
        A.depth[(int64_t)r * K + k] = d_i[k];
      }
    }
    A.opacity[r] = op;
    float dl_rgb[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
      A.rgb[3 * r + c] = rgb[c];
      const float e = rgb[c] - A.target[3 * r + c];
      l_rgb += e * e;
      dl_rgb[c] = 2.f * e * A.inv_rays_global * (1.f / 3.f);
    }
    l_rgb *= A.inv_rays_global * (1.f / 3.f);
    const float o1 = op + 1e-10f, lo = logf(o1);
    l_op = A.lambda_opacity * (-o1 * lo) * A.inv_rays_global;
    const float dl_op = A.lambda_opacity * (-(lo + 1.f)) * A.inv_rays_global;
    float dg[RNB_MAX_K], gdot = 0.f;
#pragma unroll
    for (int k = 0; k < RNB_MAX_K; k++) {
      if (k < K) {
        const int64_t row = (int64_t)k * R + r;
        float dl_d = 0.f;
        if (K > 1 && A.lambda_dm > 0.f) {
          const float e = d_i[k] - fused;
          l_dm += A.lambda_dm * e * e * A.inv_rays_global / K;
          dl_d = 2.f * A.lambda_dm * e * A.inv_rays_global / K;
        }
        if (A.with_backward) {
          float d_o = g[k] * dl_op;
#pragma unroll
          for (int c = 0; c < 3; c++) {
            A.d_rgb[3 * row + c] = g[k] * dl_rgb[c];
            d_o -= g[k] * dl_rgb[c] * A.bg[k][c];
          }
          A.d_opacity[row] = d_o; A.d_depth[row] = dl_d;
        }
        dg[k] = dl_rgb[0] * c_i[k][0] + dl_rgb[1] * c_i[k][1] + dl_rgb[2] * c_i[k][2] + dl_op * o_i[k] + dcv[k];
        gdot = fmaf(g[k], dg[k], gdot);
      }
    }
    if (A.with_backward && K > 1) {
      __half2 out[8];
#pragma unroll
      for (int k = 0; k < 16; k += 2) {
        const float u0 = k < K ? g[k] * (dg[k] - gdot) * A.gs : 0.f, u1 = k + 1 < K ? g[k + 1] * (dg[k + 1] - gdot) * A.gs : 0.f;
        saturated |= !(fabsf(u0) <= 65504.f) || !(fabsf(u1) <= 65504.f);      // also true for nan
        const float v0 = fminf(fmaxf(u0, -65504.f), 65504.f), v1 = fminf(fmaxf(u1, -65504.f), 65504.f);
        out[k >> 1] = __floats2half2_rn(v0, v1);
      }
      uint4* dst = reinterpret_cast<uint4*>(A.dlogits + (int64_t)r * 16);
      dst[0] = *reinterpret_cast<const uint4*>(&out[0]);
      dst[1] = *reinterpret_cast<const uint4*>(&out[4]);
    }
  }
  if (A.flags != nullptr && __any_sync(kFullMask, saturated) && lane == 0) atomicOr(A.flags, RNB_FLAG_NONFINITE);
  l_rgb = warp_sum(l_rgb); l_op = warp_sum(l_op); l_dm = warp_sum(l_dm);
  if (lane == 0) { red[warp][0] = l_rgb; red[warp][1] = l_op; red[warp][2] = l_dm; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, b = 0.f, c = 0.f;
    for (int w = 0; w < 8; w++) { a += red[w][0]; b += red[w][1]; c += red[w][2]; }
    const float cvv = blockIdx.x == 0 ? cv : 0.f;
    atomicAdd(A.loss + 1, a); atomicAdd(A.loss + 2, b); atomicAdd(A.loss + 4, c);
    if (blockIdx.x == 0) atomicAdd(A.loss + 3, cvv);
    atomicAdd(A.loss + 0, a + b + c + cvv);
  }
}

// ---- Adam (torch.optim.Adam, amsgrad = False, weight_decay = 0; train_ml.py:143 uses FusedAdam(eps=1e-15)) ----
__global__ void __launch_bounds__(256) eng_adam_kernel(float* __restrict__ p, __half* __restrict__ p16, float* __restrict__ g,
                                                       float* __restrict__ m, float* __restrict__ v, int64_t n, float lr, int64_t host_step,
                                                       float beta1, float beta2, float eps, float grad_scale,
                                                       const int32_t* __restrict__ flags, int32_t* __restrict__ status) {
  __shared__ AdamStep sh;
  adam_prologue(&sh, flags != nullptr ? *(volatile const int32_t*)flags : 0, status, host_step, lr, beta1, beta2);
  const float step_size = sh.step_size, inv_sqrt_bc2 = sh.inv_sqrt_bc2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t n4 = n >> 2;
  if (sh.skip) {      // a non-finite gradient or dropped samples: clear the gradients, leave parameters and moments alone
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) g[i] = 0.f;
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i], gg = reinterpret_cast<float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    float* P = &pp.x; float* G = &gg.x; float* M = &mm.x; float* V = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const float gr = G[k] * grad_scale;
      M[k] = beta1 * M[k] + (1.f - beta1) * gr;
      V[k] = beta2 * V[k] + (1.f - beta2) * gr * gr;
      P[k] -= step_size * M[k] / (sqrtf(V[k]) * inv_sqrt_bc2 + eps);
    }
    reinterpret_cast<float4*>(p)[i] = pp; reinterpret_cast<float4*>(m)[i] = mm; reinterpret_cast<float4*>(v)[i] = vv;
    reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const __half2 a = __floats2half2_rn(pp.x, pp.y), b = __floats2half2_rn(pp.z, pp.w);
    reinterpret_cast<uint2*>(p16)[i] = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
  }
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float gr = g[i] * grad_scale;
    const float mi = beta1 * m[i] + (1.f - beta1) * gr, vi = beta2 * v[i] + (1.f - beta2) * gr * gr;
    const float pi = p[i] - step_size * mi / (sqrtf(vi) * inv_sqrt_bc2 + eps);
    p[i] = pi; m[i] = mi; v[i] = vi; g[i] = 0.f; p16[i] = __float2half_rn(pi);
  }
}

// inf / nan anywhere in grads[lo, hi) -> flag (the dense networks' block: a few ten thousand values)
__global__ void __launch_bounds__(256) eng_grad_check_kernel(const float* __restrict__ g, int64_t lo, int64_t hi, int32_t* __restrict__ flags) {
  bool bad = false;
  for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (int64_t)gridDim.x * blockDim.x)
    bad |= !(fabsf(g[i]) <= 3.0e38f);
  if (__any_sync(kFullMask, bad) && (threadIdx.x & 31) == 0) atomicOr(flags, RNB_FLAG_NONFINITE);
}

}  // namespace rnb

using namespace rnb;

// ---- optional per-stage device timing (bench.py's roofline leg; off by default) -------------------------
#include <vector>
namespace {
enum Stage { ST_GATE = 0, ST_MARCH, ST_SCAN, ST_EXPAND, ST_FIELD_FWD, ST_COMP_FW, ST_BLEND, ST_COMP_BW, ST_FIELD_BW, ST_GATE_BW, ST_ADAM, ST_BARRIER, ST_COUNT };
struct ProfRec { int stage; cudaEvent_t a, b; };
std::vector<ProfRec> g_prof;
bool g_prof_on = false;
struct StageTimer {
  cudaStream_t st; cudaEvent_t b = nullptr;
  StageTimer(int stage, cudaStream_t s) : st(s) {
    if (!g_prof_on) return;
    ProfRec r; r.stage = stage;
    cudaEventCreate(&r.a); cudaEventCreate(&r.b);
    cudaEventRecord(r.a, st);
    b = r.b; g_prof.push_back(r);
  }
  ~StageTimer() { if (b) cudaEventRecord(b, st); }
};
}  // namespace

// the empty-space test of coarse.cuh is written for cascades of 128^3 cells (RNB_NO_CULL=1 turns it off)
static bool coarse_enabled(const rnb_engine_config* c) {
  static const bool off = getenv("RNB_NO_CULL") != nullptr && atoi(getenv("RNB_NO_CULL")) != 0;
  return !off && c->grid_size == 2 * kCoarse;
}

static int check_cfg(const rnb_engine_config* c, const char* what) {
  RNB_REQUIRE(c != nullptr, what);
  RNB_REQUIRE(c->K >= 1 && c->K <= RNB_MAX_K && c->n_rays >= 1 && c->max_samples >= 1 && c->max_samples <= 65536, what);
  RNB_REQUIRE(c->grid_size >= 1 && c->grid_size <= 1024 && c->cascades >= 1, what);
  RNB_REQUIRE(c->sample_capacity >= 128 && c->sample_capacity % 128 == 0 && c->sample_capacity < (1ll << 31), what);
  RNB_REQUIRE(c->loss_scale > 0.f && c->n_params > 0, what);
  return RNB_OK;
}

extern "C" {

int rnb_engine_layout_of(const rnb_engine_config* c, rnb_engine_layout* L) {
  const char* what = "rnb_engine_layout_of";
  if (int rc = check_cfg(c, what)) return rc;
  RNB_REQUIRE(L != nullptr, what);
  const int64_t R = c->n_rays, K = c->K, S = c->sample_capacity, M = c->max_samples;
  Carver w;
  L->gate_in = w.take(R * 16 * 2); L->gate_hidden = w.take(4 * R * 64 * 2); L->gate_logits = w.take(R * 16 * 2);
  L->gate_dlogits = w.take(R * 16 * 2); L->gate_prob = w.take(R * K * 4); L->dgate = w.take(R * K * 4);
  L->importance = w.take(RNB_MAX_K * 4);
  L->t_scratch = w.take(K * R * M * 4);
  L->counts = w.take(K * R * 4); L->starts = w.take(K * R * 4);
  L->seg = w.take((2 * RNB_MAX_K + 4) * 4);
  L->ray_id = w.take(S * 4); L->ts = w.take(S * 4); L->deltas = w.take(S * 4);
  L->coarse = w.take((int64_t)K * c->cascades * kCoarseWords * 8);      // dilated 64^3-block occupancy of every sub-NeRF (coarse.cuh)
  L->enc = w.take(S * 32 * 2); L->rgb_in = w.take(S * 32 * 2); L->geo_h = w.take(S * 64 * 2);
  L->rgb_h1 = w.take(S * 64 * 2); L->rgb_h2 = w.take(S * 64 * 2); L->sraw = w.take(S * 2);
  L->sigmas = w.take(S * 4); L->rgb4 = w.take(S * 4 * 2); L->ws = w.take(S * 4);
  L->dsigmas = w.take(S * 4); L->drgbs = w.take(S * 3 * 4);
  L->pass_opacity = w.take(K * R * 4); L->pass_depth = w.take(K * R * 4); L->pass_rgb = w.take(K * R * 3 * 4);
  L->pass_total = w.take(K * R * 4);
  L->d_pass_opacity = w.take(K * R * 4); L->d_pass_depth = w.take(K * R * 4); L->d_pass_rgb = w.take(K * R * 3 * 4);
  L->rgb = w.take(R * 3 * 4); L->opacity = w.take(R * 4); L->depth = w.take(R * K * 4);
  L->loss = w.take(8 * 4);
  L->status = w.take(16 * 4);
  L->march_slot_bytes = L->enc - L->t_scratch;     // [t_scratch, counts, starts, seg, ray_id, ts, deltas] are contiguous
  L->march_slot1 = w.take(L->march_slot_bytes);
  L->total_bytes = w.off;
  return RNB_OK;
}

static const rnb_mlp_desc kGateDesc = {16, 16, 4, 0};

int rnb_gate_fwd(const void* gate_w_f16, const float* rays_o, const float* rays_d, int64_t n_rays, int K, void* gate_in_f16,
                 void* gate_hidden_f16, void* gate_logits_f16, float* prob, float* importance, rnb_stream_t stream) {
  const char* what = "rnb_gate_fwd";
  RNB_REQUIRE(gate_w_f16 && rays_o && rays_d && gate_in_f16 && gate_logits_f16 && prob && importance, what);
  RNB_REQUIRE(n_rays >= 1 && K >= 2 && K <= RNB_MAX_K, what);
  cudaStream_t st = as_stream(stream);
  const int R = (int)n_rays;
  eng_gate_in_kernel<<<ceil_div(R, 256), 256, 0, st>>>(rays_o, rays_d, R, (__half*)gate_in_f16);
  if (int rc = rnb_mlp_fwd(&kGateDesc, gate_w_f16, gate_in_f16, R, gate_logits_f16, gate_hidden_f16, stream)) return rc;
  RNB_CUDA(cudaMemsetAsync(importance, 0, K * 4, st), what);
  eng_gate_softmax_kernel<<<ceil_div(R, 256), 256, 0, st>>>((const __half*)gate_logits_f16, R, K, prob, importance);
  return check_launch(what);
}

int rnb_engine_gate(const rnb_engine_config* c, void* workspace, const void* params_f16, const float* rays_o,
                    const float* rays_d, rnb_stream_t stream) {
  const char* what = "rnb_engine_gate";
  if (int rc = check_cfg(c, what)) return rc;
  RNB_REQUIRE(workspace && params_f16 && rays_o && rays_d && c->K > 1, what);
  rnb_engine_layout L;
  rnb_engine_layout_of(c, &L);
  char* ws = (char*)workspace;
  StageTimer tm(ST_GATE, as_stream(stream));
  return rnb_gate_fwd((const __half*)params_f16 + c->gate_off, rays_o, rays_d, c->n_rays, c->K, ws + L.gate_in, ws + L.gate_hidden,
                      ws + L.gate_logits, (float*)(ws + L.gate_prob), (float*)(ws + L.importance), stream);
}

int rnb_engine_march(const rnb_engine_config* c, void* workspace, const float* rays_o, const float* rays_d, const float* noise,
                     const uint8_t* const* bitfields, int slot, rnb_stream_t stream) {
  const char* what = "rnb_engine_march";
  if (int rc = check_cfg(c, what)) return rc;
  RNB_REQUIRE(workspace && rays_o && rays_d && bitfields && (slot == 0 || slot == 1), what);
  rnb_engine_layout L;
  rnb_engine_layout_of(c, &L);
  char* ws = (char*)workspace + (slot == 1 ? L.march_slot1 - L.t_scratch : 0);
  cudaStream_t st = as_stream(stream);
  const int R = c->n_rays, K = c->K;
  const int64_t S = c->sample_capacity;
  int32_t* counts = (int32_t*)(ws + L.counts); int32_t* starts = (int32_t*)(ws + L.starts); int32_t* seg = (int32_t*)(ws + L.seg);
  float* t_scratch = (float*)(ws + L.t_scratch);
  int32_t* ray_id = (int32_t*)(ws + L.ray_id); float* ts = (float*)(ws + L.ts); float* deltas = (float*)(ws + L.deltas);
  // calls so far INTO THIS SLOT (status[7] / status[9]): the two slots are marched on different streams, possibly at the same
  // time, so each keeps its own count; (slot, count, pass, ray) keys the noise draw
  int32_t* march_calls = (int32_t*)((char*)workspace + L.status) + (slot == 0 ? 7 : 9);
  const MarchParams mp = make_march_params(nullptr, c->cascades, c->scale, c->exp_step_factor, c->grid_size, c->max_samples, c->scale);
  BitfieldPtrs bf;
  for (int i = 0; i < K; i++) { RNB_REQUIRE(bitfields[i] != nullptr, what); bf.p[i] = bitfields[i]; }
  const dim3 ray_grid(ceil_div((int64_t)R * 32, 256), K);
  {
  StageTimer tm(ST_MARCH, st);
  // conservative coarse occupancy of every cascade, rebuilt from the bitfields on every call (one small kernel)
  uint64_t* coarse = nullptr;
  if (coarse_enabled(c)) {
    coarse = (uint64_t*)(ws + L.coarse);
    if (int rc = coarse_build(bitfields, K, c->cascades, coarse, st)) return rc;
  }
#define RNB_ENG_MARCH(CONST_DT, ONE)                                                                                              \
  eng_march_kernel<CONST_DT, ONE><<<ray_grid, 256, 0, st>>>(rays_o, rays_d, noise, mp, bf, c->center[0], c->center[1], c->center[2], \
                                                            c->half_size[0], c->half_size[1], c->half_size[2], c->near_distance, R,  \
                                                            counts, t_scratch, coarse, c->noise_seed, march_calls, slot ? (1ull << 62) : 0ull)
  if (c->exp_step_factor == 0.0f && c->cascades == 1) RNB_ENG_MARCH(true, true);     // bounded scenes (train_ml.py:99-101)
  else if (c->exp_step_factor == 0.0f) RNB_ENG_MARCH(true, false);
  else RNB_ENG_MARCH(false, false);
#undef RNB_ENG_MARCH
  }
  { StageTimer tm(ST_SCAN, st); eng_scan_kernel<<<1, 1024, 0, st>>>(counts, R, K, S, starts, seg, march_calls); }
  { StageTimer tm(ST_EXPAND, st); eng_expand_kernel<<<ray_grid, 256, 0, st>>>(counts, starts, t_scratch, mp, R, S, ray_id, ts, deltas); }
  return check_launch(what);
}

int rnb_engine_gate_bwd(const rnb_engine_config* c, void* workspace, const void* params_f16, float* grads_f32, rnb_stream_t stream) {
  const char* what = "rnb_engine_gate_bwd";
  if (int rc = check_cfg(c, what)) return rc;
  RNB_REQUIRE(workspace && params_f16 && grads_f32 && c->K > 1, what);
  rnb_engine_layout L;
  rnb_engine_layout_of(c, &L);
  char* ws = (char*)workspace;
  StageTimer tm(ST_GATE_BW, as_stream(stream));
  return rnb_mlp_bwd(&kGateDesc, (const __half*)params_f16 + c->gate_off, ws + L.gate_in, ws + L.gate_hidden, ws + L.gate_logits,
                     ws + L.gate_dlogits, c->n_rays, nullptr, grads_f32 + c->gate_off, 1.0f / c->loss_scale, stream);
}

int rnb_engine_fwd_bwd_ex(const rnb_engine_config* c, void* workspace, const void* params_f16, float* grads_f32,
                       const float* rays_o, const float* rays_d, const float* target_rgb, const float* noise,
                       const uint8_t* const* bitfields, const float* bg_host, int64_t n_rays_global,
                       const float* importance, int with_backward, int march_slot, void* gate_ready_event,
                          void* blend_done_event, int skip_gate_bwd, rnb_stream_t stream) {
  const char* what = "rnb_engine_fwd_bwd";
  if (int rc = check_cfg(c, what)) return rc;
  RNB_REQUIRE(workspace && params_f16 && rays_o && rays_d && target_rgb && bg_host, what);
  RNB_REQUIRE(n_rays_global >= c->n_rays && (c->K == 1 || importance != nullptr), what);
  RNB_REQUIRE(!with_backward || grads_f32 != nullptr, what);
  rnb_engine_layout L;
  rnb_engine_layout_of(c, &L);
  char* ws = (char*)workspace;
  cudaStream_t st = as_stream(stream);
  const int R = c->n_rays, K = c->K;
  const int64_t S = c->sample_capacity;
  const __half* p16 = (const __half*)params_f16;
  float* sigmas = (float*)(ws + L.sigmas); __half* rgb4 = (__half*)(ws + L.rgb4); float* wsamp = (float*)(ws + L.ws);

  RNB_REQUIRE(march_slot >= -1 && march_slot <= 1, what);
  // the marcher's outputs live in one of two slots (see rnb_engine_march)
  const int64_t slot_off = march_slot == 1 ? L.march_slot1 - L.t_scratch : 0;
  int32_t* counts = (int32_t*)(ws + slot_off + L.counts); int32_t* starts = (int32_t*)(ws + slot_off + L.starts);
  int32_t* seg = (int32_t*)(ws + slot_off + L.seg);
  int32_t* ray_id = (int32_t*)(ws + slot_off + L.ray_id); float* ts = (float*)(ws + slot_off + L.ts); float* deltas = (float*)(ws + slot_off + L.deltas);
  RNB_CUDA(cudaMemsetAsync(ws + L.loss, 0, 8 * 4, st), what);
  // step guard: this step's flag word lives behind the gradients (RNB_GRAD_TAIL); cleared here, ORed into by the blend kernel,
  // the field backward and rnb_engine_grad_check, read by the Adam kernels
  int32_t* flags = with_backward ? reinterpret_cast<int32_t*>(grads_f32 + c->n_params) : nullptr;
  if (flags) RNB_CUDA(cudaMemsetAsync(flags, 0, 4, st), what);
  // 1. march all K passes (AABB inline), 2. scan, 3. expand -- unless the batch was marched ahead of time
  const MarchParams mp = make_march_params(nullptr, c->cascades, c->scale, c->exp_step_factor, c->grid_size, c->max_samples, c->scale);
  const dim3 ray_grid(ceil_div((int64_t)R * 32, 256), K);
  if (march_slot < 0) {
    RNB_REQUIRE(bitfields, what);
    if (int rc = rnb_engine_march(c, workspace, rays_o, rays_d, noise, bitfields, 0, stream)) return rc;
  }
  // 4. field forward
  rnb_field_params P;
  P.K = K; P.grid = c->grid;
  for (int k = 0; k < 3; k++) { P.xyz_min[k] = c->xyz_min[k]; P.xyz_max[k] = c->xyz_max[k]; }
  P.table_f16 = p16 + c->table_off;
  for (int i = 0; i < K; i++) { P.geo_w_f16[i] = p16 + c->geo_off[i]; P.rgb_w_f16[i] = p16 + c->rgb_off[i]; }
  rnb_samples Sm = {rays_o, rays_d, ray_id, ts, seg, S};
  rnb_field_saved Sv = {ws + L.enc, ws + L.rgb_in, ws + L.geo_h, ws + L.rgb_h1, ws + L.rgb_h2, ws + L.sraw};
  { StageTimer tm(ST_FIELD_FWD, st); if (int rc = rnb_field_fwd(&P, &Sm, with_backward ? &Sv : nullptr, sigmas, rgb4, stream)) return rc; }
  // 5. composite, 6. blend + loss + gradient seeds
  float* p_op = (float*)(ws + L.pass_opacity); float* p_de = (float*)(ws + L.pass_depth); float* p_rgb = (float*)(ws + L.pass_rgb);
  BlendArgs B;
  B.n_rays = R; B.K = K; B.inv_rays_global = 1.0f / (float)n_rays_global;
  B.lambda_opacity = c->lambda_opacity; B.lambda_cv = c->lambda_cv; B.lambda_dm = c->lambda_dm; B.gs = c->loss_scale;
  for (int i = 0; i < K; i++) for (int k = 0; k < 3; k++) B.bg[i][k] = bg_host[3 * i + k];
  B.target = target_rgb; B.prob = (const float*)(ws + L.gate_prob); B.importance = importance;
  B.p_opacity = p_op; B.p_depth = p_de; B.p_rgb = p_rgb;
  B.rgb = (float*)(ws + L.rgb); B.opacity = (float*)(ws + L.opacity); B.depth = (float*)(ws + L.depth);
  B.d_opacity = (float*)(ws + L.d_pass_opacity); B.d_depth = (float*)(ws + L.d_pass_depth); B.d_rgb = (float*)(ws + L.d_pass_rgb);
  B.dlogits = (__half*)(ws + L.gate_dlogits); B.loss = (float*)(ws + L.loss); B.with_backward = with_backward;
  B.seg = seg; B.flags = flags; B.status = (int32_t*)(ws + L.status);
  float* dsig = (float*)(ws + L.dsigmas); float* drgb = (float*)(ws + L.drgbs);
  { StageTimer tm(ST_COMP_FW, st);
  eng_composite_fw_kernel<<<ray_grid, 256, 0, st>>>(sigmas, rgb4, deltas, ts, counts, starts, c->T_threshold, R, S, wsamp, p_op, p_de, p_rgb,
                                                    (int32_t*)(ws + L.pass_total)); }
  // the gate forward may have been issued on another stream (it is needed from here on only)
  if (gate_ready_event != nullptr) RNB_CUDA(cudaStreamWaitEvent(st, (cudaEvent_t)gate_ready_event, 0), what);
  { StageTimer tm(ST_BLEND, st); eng_blend_loss_kernel<<<ceil_div(R, 256), 256, 0, st>>>(B); }
  if (int rc = check_launch(what)) return rc;
  // ... and the gate backward may run on another stream next to the compositing backward: it needs the blend only
  if (blend_done_event != nullptr) RNB_CUDA(cudaEventRecord((cudaEvent_t)blend_done_event, st), what);
  if (!with_backward) return RNB_OK;
  // 7. backward: composite, field, gate
  { StageTimer tm(ST_COMP_BW, st);
  eng_composite_bw_kernel<<<ray_grid, 256, 0, st>>>(sigmas, rgb4, deltas, ts, counts, starts, c->T_threshold, R, S, p_op, p_de, p_rgb,
                                                    B.d_opacity, B.d_depth, B.d_rgb, dsig, drgb); }
  if (int rc = check_launch(what)) return rc;
  float* gg[RNB_MAX_K]; float* gr[RNB_MAX_K];
  for (int i = 0; i < K; i++) { gg[i] = grads_f32 + c->geo_off[i]; gr[i] = grads_f32 + c->rgb_off[i]; }
  { StageTimer tm(ST_FIELD_BW, st);
  if (int rc = rnb_field_bwd_guarded(&P, &Sm, &Sv, rgb4, dsig, drgb, c->loss_scale, grads_f32 + c->table_off, gg, gr, flags, stream)) return rc; }
  if (K > 1 && !skip_gate_bwd) return rnb_engine_gate_bwd(c, workspace, params_f16, grads_f32, stream);
  return RNB_OK;
}

int rnb_engine_fwd_bwd(const rnb_engine_config* c, void* workspace, const void* params_f16, float* grads_f32, const float* rays_o,
                       const float* rays_d, const float* target_rgb, const float* noise, const uint8_t* const* bitfields,
                       const float* bg_host, int64_t n_rays_global, const float* importance, int with_backward, int march_slot,
                       rnb_stream_t stream) {
  return rnb_engine_fwd_bwd_ex(c, workspace, params_f16, grads_f32, rays_o, rays_d, target_rgb, noise, bitfields, bg_host, n_rays_global,
                               importance, with_backward, march_slot, nullptr, nullptr, 0, stream);
}

static int launch_adam(float* params_f32, void* params_f16, float* grads_f32, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                       float beta1, float beta2, float eps, int64_t step, float grad_scale, const int32_t* flags, int32_t* status,
                       rnb_stream_t stream, const char* what) {
  RNB_REQUIRE(params_f32 && params_f16 && grads_f32 && exp_avg && exp_avg_sq && n > 0, what);
  RNB_REQUIRE(((uintptr_t)params_f32 & 15) == 0 && ((uintptr_t)grads_f32 & 15) == 0 && ((uintptr_t)exp_avg & 15) == 0 &&
              ((uintptr_t)exp_avg_sq & 15) == 0 && ((uintptr_t)params_f16 & 7) == 0, what);
  int64_t blocks = (n / 4 + 255) / 256;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;      // (a single resident wave, 5 blocks per SM, measured 3 % slower)
  if (blocks < 1) blocks = 1;
  StageTimer tm(ST_ADAM, as_stream(stream));
  eng_adam_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(params_f32, (__half*)params_f16, grads_f32, exp_avg, exp_avg_sq, n, lr, step,
                                                              beta1, beta2, eps, grad_scale, flags, status);
  return check_launch(what);
}

int rnb_engine_adam(float* params_f32, void* params_f16, float* grads_f32, float* exp_avg, float* exp_avg_sq, int64_t n,
                    float lr, float beta1, float beta2, float eps, int64_t step, float grad_scale, rnb_stream_t stream) {
  const char* what = "rnb_engine_adam";
  RNB_REQUIRE(step >= 1, what);
  return launch_adam(params_f32, params_f16, grads_f32, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, step, grad_scale, nullptr, nullptr,
                     stream, what);
}

int rnb_engine_adam_guarded(float* params_f32, void* params_f16, float* grads_f32, float* exp_avg, float* exp_avg_sq, int64_t n,
                            float lr, float beta1, float beta2, float eps, float grad_scale, const int32_t* flags, int32_t* status,
                            rnb_stream_t stream) {
  const char* what = "rnb_engine_adam_guarded";
  RNB_REQUIRE(flags != nullptr && status != nullptr, what);
  return launch_adam(params_f32, params_f16, grads_f32, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, 1, grad_scale, flags, status,
                     stream, what);
}

int rnb_engine_grad_check(const float* grads_f32, int64_t lo, int64_t hi, int32_t* flags, rnb_stream_t stream) {
  const char* what = "rnb_engine_grad_check";
  RNB_REQUIRE(grads_f32 && flags && lo >= 0 && hi >= lo, what);
  if (hi == lo) return RNB_OK;
  int64_t blocks = (hi - lo + 1023) / 1024;
  if (blocks > kNumSMs) blocks = kNumSMs;
  eng_grad_check_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(grads_f32, lo, hi, flags);
  return check_launch(what);
}


int rnb_engine_profile(int enable) {
  g_prof_on = enable != 0;
  return RNB_OK;
}

/* ms[ST_COUNT], calls[ST_COUNT]: device time and number of timed stages since the last read (synchronises). */
int rnb_engine_profile_read(double* ms, int32_t* calls, int n) {
  const char* what = "rnb_engine_profile_read";
  RNB_REQUIRE(ms && calls && n >= ST_COUNT, what);
  for (int i = 0; i < n; i++) { ms[i] = 0.0; calls[i] = 0; }
  for (auto& r : g_prof) {
    cudaEventSynchronize(r.b);
    float t = 0.f;
    if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) { ms[r.stage] += t; calls[r.stage]++; }
    cudaEventDestroy(r.a); cudaEventDestroy(r.b);
  }
  g_prof.clear();
  return RNB_OK;
}

/* Spans recorded by the caller around work the library does not launch itself (the host side's collectives, the
 * peer-memory optimizer): begin returns a handle (>= 0) or -1 when profiling is off; end records the closing event. */
int rnb_engine_profile_span_begin(int stage, rnb_stream_t stream) {
  if (!g_prof_on || stage < 0 || stage >= ST_COUNT) return -1;
  ProfRec r; r.stage = stage;
  cudaEventCreate(&r.a); cudaEventCreate(&r.b);
  cudaEventRecord(r.a, as_stream(stream));
  g_prof.push_back(r);
  return (int)g_prof.size() - 1;
}

int rnb_engine_profile_span_end(int handle, rnb_stream_t stream) {
  if (handle < 0 || handle >= (int)g_prof.size()) return RNB_OK;
  cudaEventRecord(g_prof[handle].b, as_stream(stream));
  return RNB_OK;
}

/* Timeline of the recorded stages since the last read: stage id, start and end in milliseconds relative to the first
 * record (events of different streams share the device clock), in issue order.  Returns the records through *n_out
 * (at most n_max), synchronises and clears like rnb_engine_profile_read. */
int rnb_engine_profile_timeline(int32_t* stage, double* start_ms, double* end_ms, int n_max, int32_t* n_out) {
  const char* what = "rnb_engine_profile_timeline";
  RNB_REQUIRE(stage && start_ms && end_ms && n_out && n_max >= 0, what);
  int n = 0;
  cudaEvent_t origin = g_prof.empty() ? nullptr : g_prof.front().a;
  for (auto& r : g_prof) cudaEventSynchronize(r.b);
  for (auto& r : g_prof) {
    float t0 = 0.f, t1 = 0.f;
    if (n < n_max && cudaEventElapsedTime(&t0, origin, r.a) == cudaSuccess && cudaEventElapsedTime(&t1, origin, r.b) == cudaSuccess) {
      stage[n] = r.stage; start_ms[n] = t0; end_ms[n] = t1; n++;
    }
  }
  for (auto& r : g_prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  g_prof.clear();
  *n_out = n;
  return RNB_OK;
}

}  // extern "C"
