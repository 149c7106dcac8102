/*
 * oracle/vren_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked/loaded by the product).
 *
 * Plain-C (CPU) restatement of the reference's `vren` CUDA kernels that lie on the
 * Rad-NeRF hot path.  Every function cites the reference file:line it follows
 * (paths relative to /root/reference/models/csrc/).  fp32 arithmetic is written with
 * explicit fmaf() where nvcc contracts the reference expression into an FFMA
 * (checked against the SASS of the reference compiled for sm_100a), so
 * that integer outputs (sample counts, offsets, bitfields) can be compared bit-exactly
 * with the GPU.  __expf is restated with expf (tolerance, hazard Q10 in SURVEY.md).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SQRT3 1.73205080757f

static inline float clampf(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }
static inline float signf(float x) { return copysignf(1.0f, x); }

/* raymarching.cu:11-13 */
static inline float calc_dt(float t, float exp_step_factor, int max_samples, int grid_size, float scale) {
  return clampf(t * exp_step_factor, SQRT3 / (float)max_samples, SQRT3 * 2 * scale / (float)grid_size);
}

/* raymarching.cu:19-23 */
static inline int mip_from_pos(float x, float y, float z, int cascades) {
  const float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
  int exponent; frexpf(mx, &exponent);
  int m = exponent + 1; if (m < 0) m = 0; if (m > cascades - 1) m = cascades - 1;
  return m;
}

/* raymarching.cu:29-32 */
static inline int mip_from_dt(float dt, int grid_size, int cascades) {
  int exponent; frexpf(dt * (float)grid_size, &exponent);
  int m = exponent; if (m < 0) m = 0; if (m > cascades - 1) m = cascades - 1;
  return m;
}

/* raymarching.cu:35-50 */
static inline uint32_t expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
static inline uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) {
  return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2);
}
/* raymarching.cu:52-60 */
static inline uint32_t morton3D_invert1(uint32_t x) {
  x = x & 0x49249249u;
  x = (x | (x >> 2)) & 0xc30c30c3u;
  x = (x | (x >> 4)) & 0x0f00f00fu;
  x = (x | (x >> 8)) & 0xff0000ffu;
  x = (x | (x >> 16)) & 0x0000ffffu;
  return x;
}

/* raymarching.cu:62-70 */
void orc_morton3D(const int32_t* coords, int64_t n, int32_t* indices) {
  for (int64_t i = 0; i < n; i++)
    indices[i] = (int32_t)morton3D((uint32_t)coords[3 * i], (uint32_t)coords[3 * i + 1], (uint32_t)coords[3 * i + 2]);
}

/* raymarching.cu:90-101 */
void orc_morton3D_invert(const int32_t* indices, int64_t n, int32_t* coords) {
  for (int64_t i = 0; i < n; i++) {
    const int32_t ind = indices[i];
    coords[3 * i + 0] = (int32_t)morton3D_invert1((uint32_t)(ind >> 0));
    coords[3 * i + 1] = (int32_t)morton3D_invert1((uint32_t)(ind >> 1));
    coords[3 * i + 2] = (int32_t)morton3D_invert1((uint32_t)(ind >> 2));
  }
}

/* raymarching.cu:122-141 -- n_bytes = number of bitfield bytes; grid has 8*n_bytes floats */
void orc_packbits(const float* density_grid, int64_t n_bytes, float thr, uint8_t* bitfield) {
  for (int64_t n = 0; n < n_bytes; n++) {
    uint8_t bits = 0;
    for (int i = 0; i < 8; i++) bits |= (density_grid[8 * n + i] > thr) ? (uint8_t)(1u << i) : 0;
    bitfield[n] = bits;
  }
}

/* intersection.cu:5-56 (+ host :59-100 for max_hits == 1, N_voxels == 1: the sort over a
 * length-1 axis is the identity).  hits_t (n,1,2) init -1, hits_idx (n,1) init -1. */
void orc_ray_aabb_intersect(const float* rays_o, const float* rays_d, const float* center,
                            const float* half_size, int64_t n_rays,
                            int32_t* hit_cnt, float* hits_t, int64_t* hits_idx) {
  for (int64_t r = 0; r < n_rays; r++) {
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) {
      const float inv_d = 1.0f / rays_d[3 * r + a];
      const float tmin = (center[a] - half_size[a] - rays_o[3 * r + a]) * inv_d;
      const float tmax = (center[a] + half_size[a] - rays_o[3 * r + a]) * inv_d;
      lo[a] = fminf(tmin, tmax);
      hi[a] = fmaxf(tmin, tmax);
    }
    const float t1 = fmaxf(fmaxf(lo[0], lo[1]), lo[2]);
    const float t2 = fminf(fminf(hi[0], hi[1]), hi[2]);
    hit_cnt[r] = 0; hits_t[2 * r] = -1.0f; hits_t[2 * r + 1] = -1.0f; hits_idx[r] = -1;
    float a0 = t1, a1 = t2;
    if (t1 > t2) { a0 = -1.0f; a1 = -1.0f; }
    if (a1 > 0) {
      hit_cnt[r] = 1; hits_t[2 * r] = fmaxf(a0, 0.0f); hits_t[2 * r + 1] = a1; hits_idx[r] = 0;
    }
  }
}

/* ---- occupancy-grid marching ------------------------------------------------------ */

typedef struct {
  float ox, oy, oz, dx, dy, dz, dx_inv, dy_inv, dz_inv;
  const uint8_t* bitfield;
  int cascades, grid_size, max_samples;
  uint32_t grid_size3;
  float grid_size_inv, scale, exp_step_factor;
  float dt_scale; /* the 5th argument handed to calc_dt: `scale` in the train kernel
                     (raymarching.cu:196,207,231), `cascades` in the test kernel
                     (raymarching.cu:370,399 -- hazard Q1) */
} march_ctx;

/* One iteration of the loop body shared by raymarching.cu:204-235, :245-279 and :353-401.
 * Returns 1 when the cell at t is occupied (*x,*y,*z,*dt describe the sample; the caller
 * advances t by *dt), else 0 after having advanced *t past the empty voxel. */
static inline int march_probe(const march_ctx* c, float* t, float* x, float* y, float* z, float* dt) {
  const float tt = *t;
  *x = fmaf(tt, c->dx, c->ox); *y = fmaf(tt, c->dy, c->oy); *z = fmaf(tt, c->dz, c->oz);
  *dt = calc_dt(tt, c->exp_step_factor, c->max_samples, c->grid_size, c->dt_scale);
  const int mp = mip_from_pos(*x, *y, *z, c->cascades);
  const int md = mip_from_dt(*dt, c->grid_size, c->cascades);
  const int mip = mp > md ? mp : md;
  const float mip_bound = fminf(scalbnf(1.0f, mip - 1), c->scale);
  const float mip_bound_inv = 1 / mip_bound;
  const float gs = (float)c->grid_size;
  const int nx = (int)clampf(0.5f * fmaf(*x, mip_bound_inv, 1.0f) * gs, 0.0f, gs - 1.0f);
  const int ny = (int)clampf(0.5f * fmaf(*y, mip_bound_inv, 1.0f) * gs, 0.0f, gs - 1.0f);
  const int nz = (int)clampf(0.5f * fmaf(*z, mip_bound_inv, 1.0f) * gs, 0.0f, gs - 1.0f);
  const uint32_t idx = (uint32_t)mip * c->grid_size3 + morton3D((uint32_t)nx, (uint32_t)ny, (uint32_t)nz);
  const int occ = c->bitfield[idx / 8] & (1 << (idx % 8));
  if (occ) return 1;
  const float tx = fmaf(fmaf(fmaf(0.5f, signf(c->dx), (float)nx + 0.5f) * c->grid_size_inv, 2.0f, -1.0f), mip_bound, -*x) * c->dx_inv;
  const float ty = fmaf(fmaf(fmaf(0.5f, signf(c->dy), (float)ny + 0.5f) * c->grid_size_inv, 2.0f, -1.0f), mip_bound, -*y) * c->dy_inv;
  const float tz = fmaf(fmaf(fmaf(0.5f, signf(c->dz), (float)nz + 0.5f) * c->grid_size_inv, 2.0f, -1.0f), mip_bound, -*z) * c->dz_inv;
  const float t_target = tt + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
  float tn = tt;
  do {
    tn += calc_dt(tn, c->exp_step_factor, c->max_samples, c->grid_size, c->dt_scale);
  } while (tn < t_target);
  *t = tn;
  return 0;
}

static inline void march_ctx_init(march_ctx* c, const float* o, const float* d, const uint8_t* bitfield,
                                  int cascades, float scale, float esf, int grid_size, int max_samples,
                                  float dt_scale) {
  c->ox = o[0]; c->oy = o[1]; c->oz = o[2];
  c->dx = d[0]; c->dy = d[1]; c->dz = d[2];
  c->dx_inv = 1.0f / c->dx; c->dy_inv = 1.0f / c->dy; c->dz_inv = 1.0f / c->dz;
  c->bitfield = bitfield; c->cascades = cascades; c->grid_size = grid_size; c->max_samples = max_samples;
  c->grid_size3 = (uint32_t)(grid_size * grid_size * grid_size);
  c->grid_size_inv = 1.0f / (float)grid_size; c->scale = scale; c->exp_step_factor = esf;
  c->dt_scale = dt_scale;
}

/* raymarching.cu:184-235 first pass: per-ray sample count (the starting t is perturbed by
 * dt(t1)*noise only when t1 >= 0, :195-198). */
void orc_march_train_count(const float* rays_o, const float* rays_d, const float* hits_t,
                           const uint8_t* bitfield, int cascades, float scale, float esf,
                           const float* noise, int grid_size, int max_samples, int64_t n_rays,
                           int32_t* n_samples) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t r = 0; r < n_rays; r++) {
    march_ctx c; march_ctx_init(&c, rays_o + 3 * r, rays_d + 3 * r, bitfield, cascades, scale, esf, grid_size, max_samples, scale);
    float t1 = hits_t[2 * r], t2 = hits_t[2 * r + 1];
    if (t1 >= 0) t1 = fmaf(calc_dt(t1, esf, max_samples, grid_size, scale), noise[r], t1);
    float t = t1; int N = 0; float x, y, z, dt;
    while (0 <= t && t < t2 && N < max_samples) {
      if (march_probe(&c, &t, &x, &y, &z, &dt)) { t += dt; N++; }
    }
    n_samples[r] = N;
  }
}

/* raymarching.cu:237-279 second pass, with the CANONICAL slot assignment (hazard Q2):
 * rays_a row r = (r, exclusive_sum(n_samples)[r], n_samples[r]).  Returns total samples. */
int64_t orc_march_train_write(const float* rays_o, const float* rays_d, const float* hits_t,
                              const uint8_t* bitfield, int cascades, float scale, float esf,
                              const float* noise, int grid_size, int max_samples, int64_t n_rays,
                              const int32_t* n_samples, int64_t* rays_a,
                              float* xyzs, float* dirs, float* deltas, float* ts) {
  int64_t total = 0;
  for (int64_t r = 0; r < n_rays; r++) {
    rays_a[3 * r] = r; rays_a[3 * r + 1] = total; rays_a[3 * r + 2] = n_samples[r];
    total += n_samples[r];
  }
  if (!xyzs) return total;
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t r = 0; r < n_rays; r++) {
    march_ctx c; march_ctx_init(&c, rays_o + 3 * r, rays_d + 3 * r, bitfield, cascades, scale, esf, grid_size, max_samples, scale);
    float t1 = hits_t[2 * r], t2 = hits_t[2 * r + 1];
    if (t1 >= 0) t1 = fmaf(calc_dt(t1, esf, max_samples, grid_size, scale), noise[r], t1);
    float t = t1; int samples = 0; float x, y, z, dt;
    const int64_t start = rays_a[3 * r + 1]; const int N = n_samples[r];
    while (t < t2 && samples < N) {
      if (march_probe(&c, &t, &x, &y, &z, &dt)) {
        const int64_t s = start + samples;
        xyzs[3 * s] = x; xyzs[3 * s + 1] = y; xyzs[3 * s + 2] = z;
        dirs[3 * s] = c.dx; dirs[3 * s + 1] = c.dy; dirs[3 * s + 2] = c.dz;
        ts[s] = t; deltas[s] = dt;
        t += dt; samples++;
      }
    }
  }
  return total;
}

/* raymarching.cu:335-404.  hits_t (n_rays,2) is mutated ([r][0] <- resume point, :390).
 * Output slabs (n_alive, N_samples[,3]) must be zero-filled by the caller (host :421-426). */
void orc_march_test(const float* rays_o, const float* rays_d, float* hits_t, const int64_t* alive,
                    const uint8_t* bitfield, int cascades, float scale, float esf, int grid_size,
                    int max_samples, int N_samples, int64_t n_alive,
                    float* xyzs, float* dirs, float* deltas, float* ts, int32_t* n_eff) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t n = 0; n < n_alive; n++) {
    const int64_t r = alive[n];
    march_ctx c; march_ctx_init(&c, rays_o + 3 * r, rays_d + 3 * r, bitfield, cascades, scale, esf, grid_size, max_samples, (float)cascades);
    float t = hits_t[2 * r], t2 = hits_t[2 * r + 1];
    int s = 0; float x, y, z, dt;
    while (t < t2 && s < N_samples) {
      if (march_probe(&c, &t, &x, &y, &z, &dt)) {
        const int64_t o = n * N_samples + s;
        xyzs[3 * o] = x; xyzs[3 * o + 1] = y; xyzs[3 * o + 2] = z;
        dirs[3 * o] = c.dx; dirs[3 * o + 1] = c.dy; dirs[3 * o + 2] = c.dz;
        ts[o] = t; deltas[o] = dt;
        t += dt;
        hits_t[2 * r] = t;
        s++;
      }
    }
    n_eff[n] = s;
  }
}

/* ---- volume compositing ------------------------------------------------------------ */

/* volumerendering.cu:6-45.  Outputs opacity,depth (n_rays), rgb (n_rays,3), ws (N) and
 * total_samples (n_rays, int64) must be zero-filled by the caller (host :57-61). */
void orc_composite_train_fw(const float* sigmas, const float* rgbs, const float* deltas, const float* ts,
                            const int64_t* rays_a, float T_threshold, int64_t n_rays,
                            int64_t* total_samples, float* opacity, float* depth, float* rgb, float* ws) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t n = 0; n < n_rays; n++) {
    const int64_t ray_idx = rays_a[3 * n], start_idx = rays_a[3 * n + 1], N_samples = rays_a[3 * n + 2];
    int64_t samples = 0; float T = 1.0f;
    float R = 0, G = 0, B = 0, D = 0, O = 0;
    while (samples < N_samples) {
      const int64_t s = start_idx + samples;
      const float a = 1.0f - expf(-sigmas[s] * deltas[s]);
      const float w = a * T;
      R = fmaf(w, rgbs[3 * s], R); G = fmaf(w, rgbs[3 * s + 1], G); B = fmaf(w, rgbs[3 * s + 2], B);
      D = fmaf(w, ts[s], D);
      O += w;
      ws[s] = w;
      T *= 1.0f - a;
      if (T <= T_threshold) break;
      samples++;
    }
    rgb[3 * ray_idx] = R; rgb[3 * ray_idx + 1] = G; rgb[3 * ray_idx + 2] = B;
    depth[ray_idx] = D; opacity[ray_idx] = O;
    total_samples[ray_idx] = samples;
  }
}

/* volumerendering.cu:87-152 (host :155-203).  dL_dsigmas (N), dL_drgbs (N,3) zero-filled
 * by the caller.  The in-kernel inclusive scan of dL_dws*ws (:119-123) is restated as a
 * running sum; the N_samples==0 out-of-segment read (hazard Q9) is not reproduced. */
void orc_composite_train_bw(const float* dL_dopacity, const float* dL_ddepth, const float* dL_drgb,
                            const float* dL_dws, const float* sigmas, const float* rgbs, const float* ws,
                            const float* deltas, const float* ts, const int64_t* rays_a,
                            const float* opacity, const float* depth, const float* rgb,
                            float T_threshold, int64_t n_rays, float* dL_dsigmas, float* dL_drgbs) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t n = 0; n < n_rays; n++) {
    const int64_t ray_idx = rays_a[3 * n], start_idx = rays_a[3 * n + 1], N_samples = rays_a[3 * n + 2];
    if (N_samples == 0) continue;
    int64_t samples = 0;
    const float R = rgb[3 * ray_idx], G = rgb[3 * ray_idx + 1], B = rgb[3 * ray_idx + 2];
    const float O = opacity[ray_idx], D = depth[ray_idx];
    float T = 1.0f, r = 0, g = 0, b = 0, d = 0;
    float total = 0; /* inclusive scan, sequential like thrust::inclusive_scan(thrust::device) inside a thread */
    for (int64_t k = 0; k < N_samples; k++) total += dL_dws[start_idx + k] * ws[start_idx + k];
    float run = 0;
    const float gr = dL_drgb[3 * ray_idx], gg = dL_drgb[3 * ray_idx + 1], gb = dL_drgb[3 * ray_idx + 2];
    while (samples < N_samples) {
      const int64_t s = start_idx + samples;
      const float a = 1.0f - expf(-sigmas[s] * deltas[s]);
      const float w = a * T;
      r = fmaf(w, rgbs[3 * s], r); g = fmaf(w, rgbs[3 * s + 1], g); b = fmaf(w, rgbs[3 * s + 2], b);
      d = fmaf(w, ts[s], d);
      T *= 1.0f - a;
      run += dL_dws[s] * ws[s];
      dL_drgbs[3 * s] = gr * w; dL_drgbs[3 * s + 1] = gg * w; dL_drgbs[3 * s + 2] = gb * w;
      dL_dsigmas[s] = deltas[s] * (
          gr * (rgbs[3 * s] * T - (R - r)) +
          gg * (rgbs[3 * s + 1] * T - (G - g)) +
          gb * (rgbs[3 * s + 2] * T - (B - b)) +
          dL_dopacity[ray_idx] * (1 - O) +
          dL_ddepth[ray_idx] * (ts[s] * T - (D - d)) +
          T * dL_dws[s] - (total - run));
      if (T <= T_threshold) break;
      samples++;
    }
  }
}

/* volumerendering.cu:206-250: in-place accumulate into opacity/depth/rgb[r], kill rays. */
void orc_composite_test_fw(const float* sigmas, const float* rgbs, const float* deltas, const float* ts,
                           int64_t* alive, float T_threshold, const int32_t* n_eff, int N_samples,
                           int64_t n_alive, float* opacity, float* depth, float* rgb) {
  for (int64_t n = 0; n < n_alive; n++) {
    if (n_eff[n] == 0) { alive[n] = -1; continue; }
    const int64_t r = alive[n];
    int s = 0; float T = 1 - opacity[r];
    float R = rgb[3 * r], G = rgb[3 * r + 1], B = rgb[3 * r + 2], D = depth[r], O = opacity[r];
    while (s < n_eff[n]) {
      const int64_t o = n * N_samples + s;
      const float a = 1.0f - expf(-sigmas[o] * deltas[o]);
      const float w = a * T;
      R = fmaf(w, rgbs[3 * o], R); G = fmaf(w, rgbs[3 * o + 1], G); B = fmaf(w, rgbs[3 * o + 2], B);
      D = fmaf(w, ts[o], D);
      O += w;
      T *= 1.0f - a;
      if (T <= T_threshold) { alive[n] = -1; break; }
      s++;
    }
    rgb[3 * r] = R; rgb[3 * r + 1] = G; rgb[3 * r + 2] = B; depth[r] = D; opacity[r] = O;
  }
}
