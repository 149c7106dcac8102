// march_warp.cuh -- warp-per-ray occupancy-grid marcher (B200-first redesign of
// models/csrc/raymarching.cu:184-235 / :353-401 of the reference).
//
// The reference walks each ray in one thread: every loop iteration probes ONE grid cell with a
// dependent byte load and then either takes a sample (t += dt) or skips to the voxel's exit face
// (do t += dt while t < t_target).  Both branches advance t along the SAME chain
//     c_0 = t_start,  c_{k+1} = c_k + calc_dt(c_k)           (fp32, round-to-nearest)
// so the set of parameters a ray can ever visit does not depend on the occupancy; occupancy only
// decides which chain elements are visited.  Here a warp owns a ray and processes the chain in
// chunks of 32: lane j holds c_j, all 32 cells are probed in parallel (32 independent loads in
// flight instead of one), every empty lane finds the first chain element at or behind its own
// t_target (in closed form for a constant step, else by a 5-step binary search over the warp), and
// the reference's visit order is then recovered as the orbit of the entry lane under "next lane"
// by pointer doubling (5 warp-wide OR reductions).
// Visited+occupied lanes are the samples; their t is stored lane-contiguously.  The result is
// bit-identical to the serial loop (same chain values, same comparisons).
#pragma once
#include "march.cuh"

namespace rnb {

struct CellProbe {
  float x, y, z, dt;
  bool occ;
  float t_target;   // only meaningful when !occ
};

// Same arithmetic as march_probe() (march.cuh), without the data-dependent advance.  kOneCascade (cascades == 1): the
// mip level is 0 whatever the position, so its bound min(2^-1, scale) and the reciprocal (bound0, bound0_inv) are
// computed once per ray by the caller instead of per probe.
template <bool kOneCascade>
__device__ __forceinline__ CellProbe probe_cell(const Ray& r, const MarchParams& p, const uint8_t* __restrict__ bitfield,
                                                float tt, float dt, float bound0, float bound0_inv) {
  CellProbe s;
  s.x = __fmaf_rn(r.dx, tt, r.ox); s.y = __fmaf_rn(r.dy, tt, r.oy); s.z = __fmaf_rn(r.dz, tt, r.oz);
  s.dt = dt;
  int mip = 0;
  float mip_bound = bound0, mip_bound_inv = bound0_inv;
  if (!kOneCascade) {
    const float mx = fmaxf(fabsf(s.x), fmaxf(fabsf(s.y), fabsf(s.z)));
    const int mp = min(p.cascades - 1, max(0, frexp_exponent(mx) + 1));
    const int md = min(p.cascades - 1, max(0, frexp_exponent(fabsf(__fmul_rn(s.dt, p.gs_f)))));
    mip = max(mp, md);
    mip_bound = fminf(__uint_as_float((uint32_t)(mip - 1 + 127) << 23), p.scale);
    // 1 / 2^k = 2^-k exactly: the same bits as the IEEE division, without its dozen instructions per probe
    mip_bound_inv = p.pow2_scale ? __uint_as_float(0x7f000000u - __float_as_uint(mip_bound)) : __fdiv_rn(1.0f, mip_bound);
  }
  const int nx = (int)fmaxf(0.0f, fminf(__fmul_rn(__fmul_rn(0.5f, __fmaf_rn(s.x, mip_bound_inv, 1.0f)), p.gs_f), p.gs_m1));
  const int ny = (int)fmaxf(0.0f, fminf(__fmul_rn(__fmul_rn(0.5f, __fmaf_rn(s.y, mip_bound_inv, 1.0f)), p.gs_f), p.gs_m1));
  const int nz = (int)fmaxf(0.0f, fminf(__fmul_rn(__fmul_rn(0.5f, __fmaf_rn(s.z, mip_bound_inv, 1.0f)), p.gs_f), p.gs_m1));
  const uint32_t idx = (uint32_t)mip * p.grid_size3 + (spread3(nx) | (spread3(ny) << 1) | (spread3(nz) << 2));
  const uint32_t byte = __ldg(bitfield + (idx >> 3));
  s.occ = (byte >> (idx & 7)) & 1u;
  const float tx = __fmul_rn(__fmaf_rn(mip_bound, __fmaf_rn(__fmul_rn(__fadd_rn(__fadd_rn((float)nx, 0.5f), r.hx), p.gs_inv), 2.0f, -1.0f), -s.x), r.ix);
  const float ty = __fmul_rn(__fmaf_rn(mip_bound, __fmaf_rn(__fmul_rn(__fadd_rn(__fadd_rn((float)ny, 0.5f), r.hy), p.gs_inv), 2.0f, -1.0f), -s.y), r.iy);
  const float tz = __fmul_rn(__fmaf_rn(mip_bound, __fmaf_rn(__fmul_rn(__fadd_rn(__fadd_rn((float)nz, 0.5f), r.hz), p.gs_inv), 2.0f, -1.0f), -s.z), r.iz);
  s.t_target = __fadd_rn(tt, fmaxf(0.0f, fminf(tx, fminf(ty, tz))));
  return s;
}

// Marches one ray with the whole warp.  All arguments are warp-uniform.
//   t_start : first chain element (already perturbed for training), t2 : exit parameter
//   budget  : maximum number of samples to take (max_samples, or N_samples at test time)
//   t_out   : receives the t of sample k at t_out[k] (k < return value); may be nullptr
//   t_next  : (out) chain element following the last sample taken (the test-time resume point);
//             untouched when no sample was taken
// Returns the number of samples.
template <bool kConstDt, bool kOneCascade = false>
__device__ __forceinline__ int march_ray_warp(const Ray& ray, const MarchParams& p, const uint8_t* __restrict__ bitfield,
                                              float t_start, float t2, int budget, float* __restrict__ t_out,
                                              float& t_next) {
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  float tc = t_start;
  float pending = -INFINITY;   // chain elements < pending are hopped over (a skip that left the previous chunk)
  int n = 0;
  if (!(tc >= 0.0f) || budget <= 0) return 0;
  const float bound0 = fminf(0.5f, p.scale), bound0_inv = __fdiv_rn(1.0f, bound0);   // mip level 0
  while (true) {
    // ---- chain: lane j <- c_j ------------------------------------------------------------
    float mine = tc;
    bool fast = false;
    float q_step = 0.f;
    if (kConstDt) {
      // Constant step: while the chunk stays inside one binade [2^e, 2^(e+1)) every c_k is a multiple of u = ulp(2^e),
      // so fl(c_k + dt) = c_k + q with q = dt rounded to a multiple of u -- the same q for every k (barring an exact
      // tie |dt - q| = u/2, where round-to-even would depend on c_k).  Then c_j = c_0 + j*q exactly (one FMA, no
      // rounding: the value is a multiple of u below 2^(e+1)) and the 31 dependent additions collapse.
      const float c1 = __fadd_rn(tc, p.dt_min);
      const float q = __fsub_rn(c1, tc);
      const float c31 = __fmaf_rn(31.0f, q, tc);
      const uint32_t eb = __float_as_uint(tc) & 0x7f800000u;
      const float half_ulp = __uint_as_float(eb - (24u << 23));
      fast = eb > (24u << 23) && eb == (__float_as_uint(c1) & 0x7f800000u) && eb == (__float_as_uint(c31) & 0x7f800000u) &&
             fabsf(__fsub_rn(p.dt_min, q)) != half_ulp;
      if (fast) mine = __fmaf_rn((float)lane, q, tc);
      q_step = q;
    }
    if (!kConstDt && (__float_as_uint(p.esf) & 0x007fffffu) == 0u && __fmul_rn(tc, p.esf) >= p.dt_min && tc >= 1e-30f) {
      // Exponential stepping with a power-of-two factor (1/256, train_ml.py:101): t*esf is exact, so while the clamp of
      // calc_dt is inactive over the whole chunk, t + fl(t*esf) = fma(t, esf, t) -- one instruction per chain element
      // instead of four.  The chain is increasing: the clamp is inactive iff it is at both ends.
      float m = tc;
#pragma unroll
      for (int k = 1; k < 32; k++)
        if (k <= lane) m = __fmaf_rn(m, p.esf, m);
      const float c30 = __shfl_sync(kFullMask, m, 30);
      if (__fmul_rn(c30, p.esf) <= p.dt_max) { mine = m; fast = true; }
    }
    if (!fast) {
#pragma unroll
      for (int k = 1; k < 32; k++) {
        const float d = kConstDt ? p.dt_min : calc_dt(mine, p);
        if (k <= lane) mine = __fadd_rn(mine, d);
      }
    }
    const float my_dt = kConstDt ? p.dt_min : calc_dt(mine, p);
    const bool valid = mine < t2;
    const unsigned validmask = __ballot_sync(kFullMask, valid);   // a prefix of the lanes (chain is increasing)
    const int nvalid = __popc(validmask);
    // ---- probe all 32 cells ---------------------------------------------------------------
    CellProbe s;
    s.occ = false; s.t_target = 0.f;
    if (valid) s = probe_cell<kOneCascade>(ray, p, bitfield, mine, my_dt, bound0, bound0_inv);
    const unsigned occmask = __ballot_sync(kFullMask, valid && s.occ);
    // ---- next visited lane of every empty lane: first k > lane with !(c_k < t_target) -------
    int lo = lane + 1, hi = 32;
    if (occmask != validmask) {   // at least one empty valid lane (warp-uniform)
      if (kConstDt && fast) {
        // closed-form chain c_k = tc + k*q (exact, see above): k* = ceil((t_target - tc) / q) up to one unit of rounding in
        // the quotient; one exact comparison on either side settles it
        const float kf = __fmul_rn(__fsub_rn(s.t_target, tc), __frcp_rn(q_step));
        int k = min(max((int)ceilf(kf), lane + 1), 32);
        if (k > lane + 1 && !(__fmaf_rn((float)(k - 1), q_step, tc) < s.t_target)) k--;
        if (k < 32 && __fmaf_rn((float)k, q_step, tc) < s.t_target) k++;
        lo = k;
      } else {
#pragma unroll
        for (int it = 0; it < 5; it++) {
          const int mid = (lo + hi) >> 1;
          const float cm = __shfl_sync(kFullMask, mine, mid & 31);
          if (lo < hi) {
            if (!(cm < s.t_target)) hi = mid; else lo = mid + 1;
          }
        }
      }
    }
    // ---- the reference's visit order as an orbit: lane l is followed by l + 1 when its cell is occupied (a sample was
    // taken) and by its skip target otherwise; lanes past t2 end the walk.  The set of visited lanes is the orbit of the
    // first lane not hopped over, found by pointer doubling (5 rounds: one warp-wide OR and one shuffle each) ----------
    int nx = valid ? (s.occ ? lane + 1 : lo) : 32;           // in [lane+1, 32]; 32 = leaves the chunk
    const int pos = __popc(__ballot_sync(kFullMask, mine < pending));   // first lane not hopped over
    unsigned visited = 0u;
    if (pos < 32) {
      pending = -INFINITY;
      visited = 1u << pos;
#pragma unroll
      for (int it = 0; it < 5; it++) {
        const unsigned to = (((visited >> lane) & 1u) && nx < 32) ? (1u << nx) : 0u;
        visited |= __reduce_or_sync(kFullMask, to);
        const int nn = __shfl_sync(kFullMask, nx, nx & 31);
        nx = nx < 32 ? nn : 32;
      }
      visited &= validmask;
    }
    unsigned samplemask = visited & occmask;
    bool done = false;
    if (n + __popc(samplemask) >= budget) {                  // the budget ends the walk inside this chunk
      unsigned beyond = samplemask;
      for (int k = budget - n; k > 0; k--) beyond &= beyond - 1u;   // clears the first (budget - n) samples
      samplemask &= ~beyond;
      done = true;
    }
    n += __popc(samplemask);
    if (!done && visited != 0u) {
      const int last_visited = 31 - __clz(visited);
      // an empty last lane hops out of the chunk: its target carries over
      if (!((occmask >> last_visited) & 1u)) pending = __shfl_sync(kFullMask, s.t_target, last_visited);
    }
    // ---- emit this chunk's samples -----------------------------------------------------------
    if (samplemask) {
      const int n0 = n - __popc(samplemask);
      if (t_out != nullptr && ((samplemask >> lane) & 1u)) t_out[n0 + __popc(samplemask & lt_mask)] = mine;
      const int last = 31 - __clz(samplemask);
      t_next = __shfl_sync(kFullMask, __fadd_rn(mine, my_dt), last);
    }
    if (done || nvalid < 32) break;
    const float c31 = __shfl_sync(kFullMask, mine, 31), d31 = __shfl_sync(kFullMask, my_dt, 31);
    tc = __fadd_rn(c31, d31);
  }
  return n;
}

}  // namespace rnb
