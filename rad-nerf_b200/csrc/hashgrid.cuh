// hashgrid.cuh -- device/host helpers of the multiresolution hash grid (level table, corner
// location, dense/hash index, warp-aggregated 8-byte gradient reductions).  Shared by hashgrid.cu
// (operator level) and field.cu (fused field kernels).
#pragma once
#include "common.cuh"

namespace rnb {

struct GridLevels {
  int n_levels;
  float scale[RNB_MAX_LEVELS];
  uint32_t res[RNB_MAX_LEVELS];
  uint32_t offset[RNB_MAX_LEVELS + 1];
};

__device__ __forceinline__ uint32_t grid_index(uint32_t x, uint32_t y, uint32_t z, uint32_t res, uint32_t size, bool dense) {
  uint32_t idx;
  if (dense) {
    idx = x + y * res + z * res * res;
    if (idx >= size) idx %= size;   // only the +1 corner on the upper boundary can leave the level
  } else {
    idx = x ^ (y * 2654435761u) ^ (z * 805459861u);
    idx = ((size & (size - 1u)) == 0u) ? (idx & (size - 1u)) : (idx % size);
  }
  return idx;
}

// tcnn grid_index(): the dense index is accumulated while stride <= size; the level is hashed iff
// the full dense grid does not fit (size < res^3).
__device__ __forceinline__ bool level_is_dense(uint32_t res, uint32_t size) {
  return (uint64_t)res * res * res <= (uint64_t)size;
}

struct Corner {
  uint32_t cx, cy, cz;
  float wx, wy, wz;
};

__device__ __forceinline__ Corner locate(float x, float y, float z, float scale) {
  Corner c;
  const float px = fmaf(scale, x, 0.5f), py = fmaf(scale, y, 0.5f), pz = fmaf(scale, z, 0.5f);
  const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
  c.cx = (uint32_t)(int)fx; c.cy = (uint32_t)(int)fy; c.cz = (uint32_t)(int)fz;
  c.wx = px - fx; c.wy = py - fy; c.wz = pz - fz;
  return c;
}

__device__ __forceinline__ void red_add_f32x2(float* addr, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

__device__ __forceinline__ void red_add_f32x4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Warp-aggregated scatter.  Consecutive lanes are consecutive samples of one ray, so lanes
// that hit the same coarse entry form RUNS: a segmented warp reduction over runs of equal index
// (5 shuffle steps) leaves the run total in the run's first lane, which issues one reduction.
__device__ __forceinline__ void scatter_aggregated(float* grad_level, uint32_t idx, float g0, float g1, bool active) {
  const int lane = threadIdx.x & 31;
  const uint32_t key = active ? idx : 0xffffffffu;
  const uint32_t prev = __shfl_up_sync(kFullMask, key, 1);
  const unsigned heads = __ballot_sync(kFullMask, lane == 0 || prev != key);
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float v0 = __shfl_down_sync(kFullMask, g0, o), v1 = __shfl_down_sync(kFullMask, g1, o);
    // lane+o belongs to my run iff no run head lies in (lane, lane+o]
    const unsigned between = ((heads >> lane) >> 1) & ((1u << o) - 1u);
    if (lane + o < 32 && between == 0u) { g0 += v0; g1 += v1; }
  }
  if (active && ((heads >> lane) & 1u)) red_add_f32x2(grad_level + 2 * (size_t)idx, g0, g1);
}

inline int to_levels(const rnb_grid_meta* m, GridLevels* g, const char* what) {
  RNB_REQUIRE(m != nullptr, what);
  RNB_REQUIRE(m->n_levels >= 1 && m->n_levels <= RNB_MAX_LEVELS && m->n_features == 2, what);
  g->n_levels = m->n_levels;
  for (int l = 0; l < m->n_levels; l++) {
    g->scale[l] = m->scale[l]; g->res[l] = m->res[l]; g->offset[l] = m->offset[l];
    RNB_REQUIRE(m->offset[l + 1] > m->offset[l] && m->res[l] >= 2, what);
  }
  g->offset[m->n_levels] = m->offset[m->n_levels];
  return RNB_OK;
}

// tcnn SphericalHarmonics degree 4 on the unit vector (x,y,z): 16 real SH basis values.
__device__ __forceinline__ void sh4_eval(float x, float y, float z, float (&o)[16]) {
  const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
  o[0] = 0.28209479177387814f;
  o[1] = -0.48860251190291987f * y;
  o[2] = 0.48860251190291987f * z;
  o[3] = -0.48860251190291987f * x;
  o[4] = 1.0925484305920792f * xy;
  o[5] = -1.0925484305920792f * yz;
  o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
  o[7] = -1.0925484305920792f * xz;
  o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
  o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
  o[10] = 2.8906114426405538f * xy * z;
  o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
  o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
  o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
  o[14] = 1.4453057213202769f * z * (x2 - y2);
  o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
}

}  // namespace rnb

namespace rnb {

// ---- lean per-level corner math (shared by the fused field kernels) ----------------------------------
// The eight corner indices of a cell need only the two candidates per axis: for hashed levels
// (table size a power of two) idx = (x' ^ y' ^ z') & mask with y' = y*2654435761, z' = z*805459861
// precomputed for y, y+1 / z, z+1 (one LOP3 + one AND per corner); for dense levels
// idx = x + y*res + z*res^2 with the y / z terms precomputed (one IADD3 per corner) and a conditional
// subtract for the single +1 corner that can leave the level on the upper boundary.
struct LevelCorners {
  uint32_t idx[8];
  float w[8];
  uint32_t cell;   // cx | cy << 11 | cz << 22: identifies the cell on levels with res < 1024
};

__device__ __forceinline__ LevelCorners level_corners(float x, float y, float z, float scale, uint32_t res, uint32_t size) {
  LevelCorners L;
  const float px = fmaf(scale, x, 0.5f), py = fmaf(scale, y, 0.5f), pz = fmaf(scale, z, 0.5f);
  const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
  const uint32_t cx = (uint32_t)(int)fx, cy = (uint32_t)(int)fy, cz = (uint32_t)(int)fz;
  const float wx1 = px - fx, wy1 = py - fy, wz1 = pz - fz;
  const float wx0 = 1.f - wx1, wy0 = 1.f - wy1, wz0 = 1.f - wz1;
  L.cell = cx | (cy << 11) | (cz << 22);
  // weight of corner k = (x-term) * (y-term) * (z-term), evaluated in the same order as hashgrid_fwd_kernel
  const float xy00 = wx0 * wy0, xy10 = wx1 * wy0, xy01 = wx0 * wy1, xy11 = wx1 * wy1;
  L.w[0] = xy00 * wz0; L.w[1] = xy10 * wz0; L.w[2] = xy01 * wz0; L.w[3] = xy11 * wz0;
  L.w[4] = xy00 * wz1; L.w[5] = xy10 * wz1; L.w[6] = xy01 * wz1; L.w[7] = xy11 * wz1;
  if ((uint64_t)res * res * res <= (uint64_t)size) {   // dense level (warp-uniform)
    const uint32_t y0 = cy * res, y1 = y0 + res, z0 = cz * res * res, z1 = z0 + res * res;
    const uint32_t b00 = cx + y0 + z0, b01 = cx + y1 + z0, b10 = cx + y0 + z1, b11 = cx + y1 + z1;
    L.idx[0] = b00; L.idx[1] = b00 + 1; L.idx[2] = b01; L.idx[3] = b01 + 1;
    L.idx[4] = b10; L.idx[5] = b10 + 1; L.idx[6] = b11; L.idx[7] = b11 + 1;
#pragma unroll
    for (int k = 0; k < 8; k++) if (L.idx[k] >= size) L.idx[k] -= size;
  } else if ((size & (size - 1u)) == 0u) {             // hashed, power-of-two table
    const uint32_t m = size - 1u;
    const uint32_t y0 = cy * 2654435761u, y1 = y0 + 2654435761u, z0 = cz * 805459861u, z1 = z0 + 805459861u;
    const uint32_t x1 = cx + 1u;
    L.idx[0] = (cx ^ y0 ^ z0) & m; L.idx[1] = (x1 ^ y0 ^ z0) & m; L.idx[2] = (cx ^ y1 ^ z0) & m; L.idx[3] = (x1 ^ y1 ^ z0) & m;
    L.idx[4] = (cx ^ y0 ^ z1) & m; L.idx[5] = (x1 ^ y0 ^ z1) & m; L.idx[6] = (cx ^ y1 ^ z1) & m; L.idx[7] = (x1 ^ y1 ^ z1) & m;
  } else {                                             // hashed, general size
    const uint32_t y0 = cy * 2654435761u, y1 = y0 + 2654435761u, z0 = cz * 805459861u, z1 = z0 + 805459861u;
    const uint32_t x1 = cx + 1u;
    L.idx[0] = (cx ^ y0 ^ z0) % size; L.idx[1] = (x1 ^ y0 ^ z0) % size; L.idx[2] = (cx ^ y1 ^ z0) % size; L.idx[3] = (x1 ^ y1 ^ z0) % size;
    L.idx[4] = (cx ^ y0 ^ z1) % size; L.idx[5] = (x1 ^ y0 ^ z1) % size; L.idx[6] = (cx ^ y1 ^ z1) % size; L.idx[7] = (x1 ^ y1 ^ z1) % size;
  }
  return L;
}

// The same for all 8 corners of one level at once.  Lanes in the same CELL share all eight corner entries, so the
// run structure (heads, reach of every shuffle step) is computed once per level from the cell key instead of once
// per corner.  kSteps = 5 combines whole-warp runs; kSteps = 3 cuts runs at 8-lane boundaries (fine levels, where a
// ray spends only a few samples per cell: fewer shuffles for at most a few extra reductions).
template <int kSteps>
__device__ __forceinline__ void scatter_level_aggregated(float* grad_level, const LevelCorners& L, float d0, float d1, bool active) {
  const int lane = threadIdx.x & 31;
  const uint32_t key = active ? L.cell : 0xffffffffu;
  const uint32_t prev = __shfl_up_sync(kFullMask, key, 1);
  const unsigned heads = __ballot_sync(kFullMask, lane == 0 || prev != key || (kSteps < 5 && (lane & ((1 << kSteps) - 1)) == 0));
  const unsigned after = (heads >> lane) >> 1;          // run heads strictly behind this lane
  // Branch-free and corner-parallel: step j of all 16 values (8 corners x 2 features) is issued together -- 16 independent
  // shuffles in flight -- and a lane that must not take its neighbour's partial sum adds 0 * v instead of branching around
  // the add.  (The first form looped corner by corner with `if (take) g += v`: ncu showed the adds stalled on their own
  // shuffle (short scoreboard) behind BSSY/BSYNC pairs, ~2700 cycles per level and warp.)
  float m[kSteps];
#pragma unroll
  for (int j = 0; j < kSteps; j++) m[j] = (lane + (1 << j) < 32 && (after & ((1u << (1 << j)) - 1u)) == 0u) ? 1.f : 0.f;
  const bool emit = active && ((heads >> lane) & 1u);
  float g0[8], g1[8];
#pragma unroll
  for (int k = 0; k < 8; k++) { g0[k] = active ? L.w[k] * d0 : 0.f; g1[k] = active ? L.w[k] * d1 : 0.f; }
#pragma unroll
  for (int j = 0; j < kSteps; j++) {
    float v0[8], v1[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { v0[k] = __shfl_down_sync(kFullMask, g0[k], 1 << j); v1[k] = __shfl_down_sync(kFullMask, g1[k], 1 << j); }
#pragma unroll
    for (int k = 0; k < 8; k++) { g0[k] = fmaf(m[j], v0[k], g0[k]); g1[k] = fmaf(m[j], v1[k], g1[k]); }
  }
  if (emit) {
#pragma unroll
    for (int k = 0; k < 8; k++) red_add_f32x2(grad_level + 2 * (size_t)L.idx[k], g0[k], g1[k]);
  }
}

// The same with the two features of a corner carried as ONE packed fp16 pair through the run reduction: half the shuffles
// and adds (the reduction is what a scatter warp spends its issue slots on).  d is the fp16 pair the activation gradient is
// carried in anyway (loss-scaled, so in fp16's range); it enters divided by 32 so that a 32-lane sum cannot overflow, the
// run total leaves in fp32 times `mult` (= 32 / loss scale).  Rounding: each partial sum is rounded to fp16 (11 bits) -- the
// table gradient agrees with the fp32 reduction to ~4e-4 L2-relative, against a 1e-2 gradient tolerance.
template <int kSteps>
__device__ __forceinline__ void scatter_level_aggregated_h2(float* grad_level, const LevelCorners& L, __half2 d, float mult, bool active) {
  const int lane = threadIdx.x & 31;
  const uint32_t key = active ? L.cell : 0xffffffffu;
  const uint32_t prev = __shfl_up_sync(kFullMask, key, 1);
  const unsigned heads = __ballot_sync(kFullMask, lane == 0 || prev != key || (kSteps < 5 && (lane & ((1 << kSteps) - 1)) == 0));
  const unsigned after = (heads >> lane) >> 1;          // run heads strictly behind this lane
  __half2 m[kSteps];
#pragma unroll
  for (int j = 0; j < kSteps; j++) {
    const float t = (lane + (1 << j) < 32 && (after & ((1u << (1 << j)) - 1u)) == 0u) ? 1.f : 0.f;
    m[j] = __floats2half2_rn(t, t);
  }
  const bool emit = active && ((heads >> lane) & 1u);
  const __half2 ds = active ? __hmul2(d, __floats2half2_rn(1.f / 32.f, 1.f / 32.f)) : __floats2half2_rn(0.f, 0.f);
  __half2 g[8];
#pragma unroll
  for (int k = 0; k < 8; k++) g[k] = __hmul2(__float2half2_rn(L.w[k]), ds);
#pragma unroll
  for (int j = 0; j < kSteps; j++) {
    __half2 v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint32_t u = __shfl_down_sync(kFullMask, *reinterpret_cast<const uint32_t*>(&g[k]), 1 << j);
      v[k] = *reinterpret_cast<const __half2*>(&u);
    }
#pragma unroll
    for (int k = 0; k < 8; k++) g[k] = __hfma2(m[j], v[k], g[k]);
  }
  if (emit) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const float2 f = __half22float2(g[k]);
      red_add_f32x2(grad_level + 2 * (size_t)L.idx[k], f.x * mult, f.y * mult);
    }
  }
}

}  // namespace rnb
