// coarse.cuh -- conservative empty-space test for single-cascade scenes (scale <= 0.5).
//
// Two thirds of the rays of a typical batch cross the scene box without touching an occupied cell, yet the marcher has
// to walk the whole chain for them (~25 chunks of 32 elements).  A ray can only produce a sample where a chain element
// falls into an occupied cell, i.e. within rounding error (1e-6) of the ray's segment.  coarse_build() reduces each
// sub-NeRF's 128^3 bitfield to 64^3 blocks (one block = one bitfield byte, Morton order) DILATED by one block; a warp
// then tests the segment at points no more than one block apart (Chebyshev), so every point of the segment is within
// half a block of a tested point and an occupied cell near the segment always raises a flag.  No flag -> the ray yields
// no sample whatever the marcher would do: it is skipped.  Flags also bound the marching interval from above: past the
// last flagged point (+ one spacing) every cell is empty, so the marcher stops there.  Sample counts, offsets and
// positions are unchanged (tests/test_golden.py, test_engine_gpu.py stay bit-exact).
#pragma once
#include "march.cuh"

namespace rnb {

constexpr int kCoarse = 64;                       // blocks per axis for a 128^3 grid
constexpr int kCoarseWords = kCoarse * kCoarse;   // one 64-bit word per (y, z) row

struct CoarseBitfields { const uint8_t* p[RNB_MAX_K]; };

// One warp per (y, z) row of blocks: lane l reads the bitfield bytes of blocks x = l and x = l + 32 of the nine rows
// (y + dy, z + dz) around it, two ballots per row give the row's 64 occupancy bits, OR over the rows and a shift by one
// bit either way complete the dilation.  The whole bitfield (256 KB per sub-NeRF) is read nine times from L2.
static __global__ void __launch_bounds__(256) coarse_build_kernel(CoarseBitfields bf, uint64_t* __restrict__ cg) {
  const int p = blockIdx.y, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int y = row & 63, z = row >> 6;
  const uint8_t* b = bf.p[p];
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
  if (lane == 0) cg[(int64_t)p * kCoarseWords + row] = w | (w << 1) | (w >> 1);
}

static inline int coarse_build(const uint8_t* const* bitfields, int K, uint64_t* cg, cudaStream_t st) {
  CoarseBitfields bf;
  for (int i = 0; i < K; i++) bf.p[i] = bitfields[i];
  coarse_build_kernel<<<dim3(kCoarseWords / 8, K), 256, 0, st>>>(bf, cg);
  return check_launch("coarse_build");
}

// Warp-collective.  cg: this sub-NeRF's dilated 64^3 grid; inv_bound = 1 / scale.  Returns the parameter past which the
// segment [t1, t2] cannot meet an occupied cell (<= t2), or a negative value when it cannot meet one at all.
__device__ __forceinline__ float coarse_last_hit(const Ray& r, float t1, float t2, const uint64_t* __restrict__ cg, float inv_bound) {
  const int lane = threadIdx.x & 31;
  const float dmax = fmaxf(fabsf(r.dx), fmaxf(fabsf(r.dy), fabsf(r.dz)));
  const float step = (2.0f / kCoarse) / (inv_bound * dmax);          // one block, in units of t
  const float span = (t2 - t1) / step;
  if (!(span < 2048.0f)) return t2;                                  // degenerate direction: do not cull
  const int n = (int)span + 2;                                       // points t1 + j*step, j < n, the last ones clamped to t2
  int last = -1;
  for (int base = 0; base < n; base += 32) {
    const int j = base + lane;
    const float t = fminf(__fmaf_rn((float)j, step, t1), t2);
    const float x = __fmaf_rn(r.dx, t, r.ox), y = __fmaf_rn(r.dy, t, r.oy), z = __fmaf_rn(r.dz, t, r.oz);
    const int bx = min(kCoarse - 1, max(0, (int)(0.5f * (x * inv_bound + 1.0f) * kCoarse)));
    const int by = min(kCoarse - 1, max(0, (int)(0.5f * (y * inv_bound + 1.0f) * kCoarse)));
    const int bz = min(kCoarse - 1, max(0, (int)(0.5f * (z * inv_bound + 1.0f) * kCoarse)));
    const bool hit = j < n && ((__ldg(cg + by + kCoarse * bz) >> bx) & 1ull);
    const unsigned m = __ballot_sync(kFullMask, hit);
    if (m) last = base + 31 - __clz(m);
  }
  if (last < 0) return -1.0f;
  return fminf(__fmaf_rn((float)(last + 1), step, t1), t2);
}

}  // namespace rnb
