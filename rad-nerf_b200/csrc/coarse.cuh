// coarse.cuh -- conservative empty-space test for the marchers.
//
// Two thirds of the rays of a typical batch cross the scene box without touching an occupied cell, yet the marcher has
// to walk the whole chain for them (~25 chunks of 32 elements).  A ray can only produce a sample where a chain element
// falls into an occupied cell, i.e. within rounding error (1e-6) of the ray's segment.  coarse_build() reduces every
// cascade of each sub-NeRF's 128^3 bitfield to 64^3 blocks (one block = one bitfield byte, Morton order) DILATED by one
// block.  For every cascade m a warp clips the ray to that cascade's box [-b_m, b_m]^3 (b_m = min(2^(m-1), scale)) and
// tests the clipped segment -- extended by one spacing at both ends -- at points no more than one block of that cascade
// apart (Chebyshev), so every point of the segment is within half a block of a tested point and an occupied cell of
// cascade m near the segment always raises a flag.  A chain element that the reference assigns to cascade m
// (max(mip_from_pos, mip_from_dt), raymarching.cu:19-32) lies inside box m whatever made the level m, so the union over
// cascades covers every cell any element can probe.  No flag -> the ray yields no sample whatever the marcher would do:
// it is skipped.  Flags also bound the marching interval from above: past the last flagged point (+ one spacing) of
// every cascade all probed cells are empty, so the marcher stops there.  Sample counts, offsets and positions are
// unchanged (tests/test_golden.py, test_engine_gpu.py stay bit-exact).
#pragma once
#include "march.cuh"

namespace rnb {

constexpr int kCoarse = 64;                       // blocks per axis for a 128^3 grid
constexpr int kCoarseWords = kCoarse * kCoarse;   // one 64-bit word per (y, z) row

struct CoarseBitfields { const uint8_t* p[RNB_MAX_K]; };

// One warp per (y, z) row of blocks: lane l reads the bitfield bytes of blocks x = l and x = l + 32 of the nine rows
// (y + dy, z + dz) around it, two ballots per row give the row's 64 occupancy bits, OR over the rows and a shift by one
// bit either way complete the dilation.  The whole bitfield (256 KB per sub-NeRF) is read nine times from L2.
static __global__ void __launch_bounds__(256) coarse_build_kernel(CoarseBitfields bf, int cascades, uint64_t* __restrict__ cg) {
  const int p = blockIdx.y / cascades, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int y = row & 63, z = row >> 6;
  const uint8_t* b = bf.p[p] + (int64_t)(blockIdx.y % cascades) * (kCoarse * kCoarse * kCoarse);   // one byte per block
  const uint32_t mx0 = spread3((uint32_t)lane), mx1 = spread3((uint32_t)lane + 32);
  uint32_t v0[9], v1[9];
#pragma unroll
  for (int k = 0; k < 9; k++) {
    const int yy = y + k % 3 - 1, zz = z + k / 3 - 1;
    const bool in = yy >= 0 && yy < kCoarse && zz >= 0 && zz < kCoarse;
    const uint32_t myz = in ? (spread3((uint32_t)yy) << 1) | (spread3((uint32_t)zz) << 2) : 0u;
    v0[k] = in ? __ldg(b + (mx0 | myz)) : 0u;
    v1[k] = in ? __ldg(b + (mx1 | myz)) : 0u;
  }
  uint32_t any0 = 0, any1 = 0;
#pragma unroll
  for (int k = 0; k < 9; k++) { any0 |= v0[k]; any1 |= v1[k]; }
  const uint64_t w = (uint64_t)__ballot_sync(kFullMask, any0 != 0) | ((uint64_t)__ballot_sync(kFullMask, any1 != 0) << 32);
  if (lane == 0) cg[(int64_t)blockIdx.y * kCoarseWords + row] = w | (w << 1) | (w >> 1);
}

// cg: (K, cascades, 4096) words
static inline int coarse_build(const uint8_t* const* bitfields, int K, int cascades, uint64_t* cg, cudaStream_t st) {
  CoarseBitfields bf;
  for (int i = 0; i < K; i++) bf.p[i] = bitfields[i];
  coarse_build_kernel<<<dim3(kCoarseWords / 8, K * cascades), 256, 0, st>>>(bf, cascades, cg);
  return check_launch("coarse_build");
}

// Warp-collective.  cg: this sub-NeRF's dilated grids (cascades, 4096).  Returns the parameter past which the segment
// [t1, t2] cannot meet an occupied cell of any cascade (<= t2), or a negative value when it cannot meet one at all.
__device__ __forceinline__ float coarse_last_hit(const Ray& r, float t1, float t2, const uint64_t* __restrict__ cg, int cascades, float scale) {
  const int lane = threadIdx.x & 31;
  const float dmax = fmaxf(fabsf(r.dx), fmaxf(fabsf(r.dy), fabsf(r.dz)));
  float t_end = -1.0f;
  for (int m = 0; m < cascades; m++, cg += kCoarseWords) {
    const float bound = fminf(__uint_as_float((uint32_t)(m - 1 + 127) << 23), scale);   // min(2^(m-1), scale)
    const float inv_bound = 1.0f / bound;
    const float step = (2.0f / kCoarse) * bound / dmax;                // one block of this cascade, in units of t
    // the ray inside this cascade's box -- grown by one block so that a ray grazing it within rounding error still gets
    // a proper interval -- and one spacing more at both ends.  The outermost cascade's box is the scene box: [t1, t2].
    float a = t1, b = t2;
    if (m < cascades - 1) {
      const float g = bound * (1.0f + 2.0f / kCoarse);
      const float x0 = (-g - r.ox) * r.ix, x1 = (g - r.ox) * r.ix;
      const float y0 = (-g - r.oy) * r.iy, y1 = (g - r.oy) * r.iy;
      const float z0 = (-g - r.oz) * r.iz, z1 = (g - r.oz) * r.iz;
      const float lo = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fminf(z0, z1));
      const float hi = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
      if (lo > hi) continue;                                           // misses the box: no element of this cascade
      if (lo <= hi) { a = fmaxf(t1, lo - step); b = fminf(t2, hi + step); }   // (NaN from 0 * inf: no clipping)
      if (a > b) continue;
    }
    const float span = (b - a) / step;
    if (!(span < 4096.0f)) return t2;                                  // degenerate direction: do not cull
    const int n = (int)span + 2;                                       // points a + j*step, j < n, the last ones clamped to b
    int last = -1;
    for (int base = 0; base < n; base += 32) {
      const int j = base + lane;
      const float t = fminf(__fmaf_rn((float)j, step, a), b);
      const float x = __fmaf_rn(r.dx, t, r.ox), y = __fmaf_rn(r.dy, t, r.oy), z = __fmaf_rn(r.dz, t, r.oz);
      const int bx = min(kCoarse - 1, max(0, (int)(0.5f * (x * inv_bound + 1.0f) * kCoarse)));
      const int by = min(kCoarse - 1, max(0, (int)(0.5f * (y * inv_bound + 1.0f) * kCoarse)));
      const int bz = min(kCoarse - 1, max(0, (int)(0.5f * (z * inv_bound + 1.0f) * kCoarse)));
      const bool hit = j < n && ((__ldg(cg + by + kCoarse * bz) >> bx) & 1ull);
      const unsigned msk = __ballot_sync(kFullMask, hit);
      if (msk) last = base + 31 - __clz(msk);
    }
    if (last >= 0) t_end = fmaxf(t_end, fminf(__fmaf_rn((float)(last + 1), step, a), t2));
  }
  return t_end;
}

}  // namespace rnb
