// march.cuh -- the occupancy-grid DDA step shared by the train / test / fused marchers.
//
// Restates one loop iteration of the reference marcher (models/csrc/raymarching.cu:204-235,
// :245-279, :353-401) with every fp32 operation spelled out as a round-to-nearest intrinsic,
// in the order (and with the FFMA contractions) that nvcc emits for the reference source --
// verified against the SASS of the reference built for sm_100a (DESIGN.md "FMA placement").
// That is what makes per-ray sample counts / offsets bit-exact (BASELINE.json north_star).
#pragma once
#include <string.h>

#include "common.cuh"

namespace rnb {

struct MarchParams {
  const uint8_t* bitfield;
  int cascades;
  int grid_size;
  int max_samples;
  uint32_t grid_size3;
  float gs_f;        // (float)grid_size
  float gs_m1;       // grid_size - 1.0f
  float gs_inv;      // 1.0f / grid_size
  float scale;
  float esf;         // exp_step_factor
  float dt_min;      // SQRT3 / max_samples                         (raymarching.cu:12)
  float dt_max;      // SQRT3*2*dt_scale / grid_size, dt_scale = scale (train) or cascades (test, Q1)
  int pow2_scale;    // scale is a power of two: every mip bound min(2^(mip-1), scale) is one, and 1 / bound is an exponent flip
};

// host side: IEEE fp32 arithmetic identical to what the device would compute
inline MarchParams make_march_params(const uint8_t* bitfield, int cascades, float scale, float esf,
                                     int grid_size, int max_samples, float dt_scale) {
  MarchParams p;
  p.bitfield = bitfield; p.cascades = cascades; p.grid_size = grid_size; p.max_samples = max_samples;
  p.grid_size3 = (uint32_t)grid_size * grid_size * grid_size;
  p.gs_f = (float)grid_size; p.gs_m1 = (float)grid_size - 1.0f; p.gs_inv = 1.0f / (float)grid_size;
  p.scale = scale; p.esf = esf;
  volatile float sqrt3 = 1.73205080757f, two_sqrt3 = 1.73205080757f * 2;
  p.dt_min = sqrt3 / (float)max_samples;
  volatile float num = two_sqrt3 * dt_scale;
  p.dt_max = num / (float)grid_size;
  uint32_t sb; memcpy(&sb, &scale, 4);
  p.pow2_scale = (sb & 0x007fffffu) == 0u && (sb >> 23) > 1u && (sb >> 23) < 253u;
  return p;
}

struct Ray {
  float ox, oy, oz, dx, dy, dz, ix, iy, iz;  // origin, direction, 1/direction
  float hx, hy, hz;                          // 0.5*sign(d): which voxel face the ray leaves through
};

__device__ __forceinline__ Ray load_ray(const float* __restrict__ rays_o, const float* __restrict__ rays_d, int64_t r) {
  Ray y;
  y.ox = __ldg(rays_o + 3 * r); y.oy = __ldg(rays_o + 3 * r + 1); y.oz = __ldg(rays_o + 3 * r + 2);
  y.dx = __ldg(rays_d + 3 * r); y.dy = __ldg(rays_d + 3 * r + 1); y.dz = __ldg(rays_d + 3 * r + 2);
  y.ix = __fdiv_rn(1.0f, y.dx); y.iy = __fdiv_rn(1.0f, y.dy); y.iz = __fdiv_rn(1.0f, y.dz);
  y.hx = copysignf(0.5f, y.dx); y.hy = copysignf(0.5f, y.dy); y.hz = copysignf(0.5f, y.dz);
  return y;
}

// raymarching.cu:11-13: clamp(t*esf, dt_min, dt_max) = fmaxf(dt_min, fminf(t*esf, dt_max))
__device__ __forceinline__ float calc_dt(float t, const MarchParams& p) {
  return fmaxf(p.dt_min, fminf(__fmul_rn(t, p.esf), p.dt_max));
}

// exponent e with |x| = m * 2^e, m in [0.5,1)  (frexpf); 0 for x == 0 / inf / nan
__device__ __forceinline__ int frexp_exponent(float ax) {
  uint32_t b = __float_as_uint(ax) & 0x7fffffffu;
  if (b == 0u || b >= 0x7f800000u) return 0;
  int adj = 0;
  if (b < 0x00800000u) { b = __float_as_uint(__fmul_rn(__uint_as_float(b), 16777216.0f)); adj = 24; }
  return (int)(b >> 23) - 126 - adj;
}

__device__ __forceinline__ uint32_t spread3(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

struct Probe {
  float x, y, z, dt;
};

// Probes the grid at parameter t.  Returns true when the cell is occupied (sample = {x,y,z,t,dt};
// the caller advances t += dt).  Otherwise advances t past the empty voxel and returns false.
__device__ __forceinline__ bool march_probe(const Ray& r, const MarchParams& p, float& t, Probe& s) {
  const float tt = t;
  s.x = __fmaf_rn(r.dx, tt, r.ox); s.y = __fmaf_rn(r.dy, tt, r.oy); s.z = __fmaf_rn(r.dz, tt, r.oz);
  s.dt = calc_dt(tt, p);
  // mip_from_pos / mip_from_dt (raymarching.cu:19-32)
  const float mx = fmaxf(fabsf(s.x), fmaxf(fabsf(s.y), fabsf(s.z)));
  const int mp = min(p.cascades - 1, max(0, frexp_exponent(mx) + 1));
  const int md = min(p.cascades - 1, max(0, frexp_exponent(fabsf(__fmul_rn(s.dt, p.gs_f)))));
  const int mip = max(mp, md);
  // mip_bound = fminf(scalbnf(1, mip-1), scale)
  const float mip_bound = fminf(__uint_as_float((uint32_t)(mip - 1 + 127) << 23), p.scale);
  const float mip_bound_inv = __fdiv_rn(1.0f, mip_bound);
  const int nx = (int)fmaxf(0.0f, fminf(__fmul_rn(__fmul_rn(0.5f, __fmaf_rn(s.x, mip_bound_inv, 1.0f)), p.gs_f), p.gs_m1));
  const int ny = (int)fmaxf(0.0f, fminf(__fmul_rn(__fmul_rn(0.5f, __fmaf_rn(s.y, mip_bound_inv, 1.0f)), p.gs_f), p.gs_m1));
  const int nz = (int)fmaxf(0.0f, fminf(__fmul_rn(__fmul_rn(0.5f, __fmaf_rn(s.z, mip_bound_inv, 1.0f)), p.gs_f), p.gs_m1));
  const uint32_t idx = (uint32_t)mip * p.grid_size3 + (spread3(nx) | (spread3(ny) << 1) | (spread3(nz) << 2));
  const uint32_t byte = __ldg(p.bitfield + (idx >> 3));
  if (byte & (1u << (idx & 7))) return true;
  // skip to the face through which the ray leaves the voxel (raymarching.cu:227-233)
  const float tx = __fmul_rn(__fmaf_rn(mip_bound, __fmaf_rn(__fmul_rn(__fadd_rn(__fadd_rn((float)nx, 0.5f), r.hx), p.gs_inv), 2.0f, -1.0f), -s.x), r.ix);
  const float ty = __fmul_rn(__fmaf_rn(mip_bound, __fmaf_rn(__fmul_rn(__fadd_rn(__fadd_rn((float)ny, 0.5f), r.hy), p.gs_inv), 2.0f, -1.0f), -s.y), r.iy);
  const float tz = __fmul_rn(__fmaf_rn(mip_bound, __fmaf_rn(__fmul_rn(__fadd_rn(__fadd_rn((float)nz, 0.5f), r.hz), p.gs_inv), 2.0f, -1.0f), -s.z), r.iz);
  const float t_target = __fadd_rn(tt, fmaxf(0.0f, fminf(tx, fminf(ty, tz))));
  float tn = tt;
  do {
    tn = __fadd_rn(tn, calc_dt(tn, p));
  } while (tn < t_target);
  t = tn;
  return false;
}

}  // namespace rnb
