// aabb.cuh -- ray / axis-aligned box slab test in the reference's operation order
// (models/csrc/intersection.cu:12-21); shared by intersect.cu and the fused training step.
#pragma once
#include "common.cuh"

namespace rnb {

__device__ __forceinline__ float2 slab_test(float ox, float oy, float oz, float ix, float iy, float iz,
                                            float cx, float cy, float cz, float hx, float hy, float hz) {
  // same operation order as intersection.cu:12-21: (center -/+ half_size - ray_o) * inv_d
  const float ax = __fmul_rn(__fsub_rn(__fsub_rn(cx, hx), ox), ix), bx = __fmul_rn(__fsub_rn(__fadd_rn(cx, hx), ox), ix);
  const float ay = __fmul_rn(__fsub_rn(__fsub_rn(cy, hy), oy), iy), by = __fmul_rn(__fsub_rn(__fadd_rn(cy, hy), oy), iy);
  const float az = __fmul_rn(__fsub_rn(__fsub_rn(cz, hz), oz), iz), bz = __fmul_rn(__fsub_rn(__fadd_rn(cz, hz), oz), iz);
  const float t1 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
  const float t2 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
  if (t1 > t2) return make_float2(-1.0f, -1.0f);
  return make_float2(t1, t2);
}

}  // namespace rnb
