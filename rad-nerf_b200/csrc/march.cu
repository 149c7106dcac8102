// march.cu -- training / test-time occupancy-grid marchers (replaces raymarching.cu:166-454).
#include "march_warp.cuh"

namespace rnb {

// ---- training: phase 1 (march once) ------------------------------------------------------
// Warp per ray (march_warp.cuh); 8 rays per 256-thread CTA.  Writes the per-ray count into
// rays_a[r][2] (the scan kernel below completes the rows) and the t of every sample into
// t_scratch[r * max_samples + k], so that phase 2 is a pure streaming expansion (no re-march).
template <bool kConstDt>
__global__ void __launch_bounds__(256) march_train_count_kernel(
    const float* __restrict__ rays_o, const float* __restrict__ rays_d, const float* __restrict__ hits_t,
    const float* __restrict__ noise, MarchParams p, int64_t n_rays, int64_t* __restrict__ rays_a,
    float* __restrict__ t_scratch) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rays) return;
  const Ray ray = load_ray(rays_o, rays_d, r);
  float t1 = __ldg(hits_t + 2 * r);
  const float t2 = __ldg(hits_t + 2 * r + 1);
  if (t1 >= 0) t1 = __fmaf_rn(calc_dt(t1, p), __ldg(noise + r), t1);  // raymarching.cu:195-198
  float t_next;
  const int n = march_ray_warp<kConstDt>(ray, p, p.bitfield, t1, t2, p.max_samples,
                                         t_scratch + r * (int64_t)p.max_samples, t_next);
  if ((threadIdx.x & 31) == 0) rays_a[3 * r + 2] = n;
}

// ---- training: exclusive scan of the counts -> rays_a rows + counter --------------------
// Single CTA of 1024 threads, 4 rays per thread per sweep with a running carry; for the
// training batch (8192 rays) that is two sweeps.  Deterministic ascending-ray CSR (Q2).
__global__ void __launch_bounds__(1024) march_train_scan_kernel(int64_t n_rays, int64_t* __restrict__ rays_a,
                                                                int32_t* __restrict__ counter) {
  __shared__ int warp_tot[32];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < n_rays; base += 4096) {
    int c[4]; int mine = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int64_t r = base + (int64_t)tid * 4 + k;
      c[k] = r < n_rays ? (int)rays_a[3 * r + 2] : 0;
      mine += c[k];
    }
    const int incl = warp_incl_scan_i(mine, lane);
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      const int wi = warp_incl_scan_i(warp_tot[lane], lane);
      warp_tot[lane] = wi;  // inclusive over warps
    }
    __syncthreads();
    const int carry = carry_s;
    int off = carry + (wid ? warp_tot[wid - 1] : 0) + incl - mine;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int64_t r = base + (int64_t)tid * 4 + k;
      if (r < n_rays) { rays_a[3 * r] = r; rays_a[3 * r + 1] = off; }
      off += c[k];
    }
    __syncthreads();
    if (tid == 0) carry_s = carry + warp_tot[31];
    __syncthreads();
  }
  if (tid == 0) { counter[0] = carry_s; counter[1] = (int32_t)n_rays; }
}

// ---- training: phase 2 (expand) ----------------------------------------------------------
// Warp per ray: sample k of ray r is t = t_scratch[r*max_samples + k]; position and step are
// recomputed with the marcher's own expressions (x = fma(d, t, o), dt = calc_dt(t)), so the
// payload is bit-identical to what the reference writes in its second march.  Lane-contiguous
// (coalesced) stores; nothing is memset.
__global__ void __launch_bounds__(256) march_train_write_kernel(
    const float* __restrict__ rays_o, const float* __restrict__ rays_d, MarchParams p, int64_t n_rays,
    const int64_t* __restrict__ rays_a, const float* __restrict__ t_scratch, int64_t cap,
    float* __restrict__ xyzs, float* __restrict__ dirs, float* __restrict__ deltas, float* __restrict__ ts) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= n_rays) return;
  const int64_t start = rays_a[3 * r + 1];
  const int n_samples = (int)rays_a[3 * r + 2];
  if (n_samples == 0) return;
  const Ray ray = load_ray(rays_o, rays_d, r);
  const float* tsrc = t_scratch + r * (int64_t)p.max_samples;
  for (int k = lane; k < n_samples; k += 32) {
    const int64_t o = start + k;
    if (o >= cap) break;
    const float t = ld_stream_f32(tsrc + k);
    xyzs[3 * o] = __fmaf_rn(ray.dx, t, ray.ox); xyzs[3 * o + 1] = __fmaf_rn(ray.dy, t, ray.oy); xyzs[3 * o + 2] = __fmaf_rn(ray.dz, t, ray.oz);
    dirs[3 * o] = ray.dx; dirs[3 * o + 1] = ray.dy; dirs[3 * o + 2] = ray.dz;
    ts[o] = t; deltas[o] = calc_dt(t, p);
  }
}

// ---- test time -------------------------------------------------------------------------------
// Warp per alive ray; the (n_alive, N_samples) slabs are written completely (zeros behind N_eff)
// so the caller never memsets (raymarching.cu:421-426 does).
template <bool kConstDt>
__global__ void __launch_bounds__(256) march_test_kernel(
    const float* __restrict__ rays_o, const float* __restrict__ rays_d, float* __restrict__ hits_t,
    const int64_t* __restrict__ alive, int64_t n_alive, MarchParams p, int N_samples,
    float* __restrict__ xyzs, float* __restrict__ dirs, float* __restrict__ deltas, float* __restrict__ ts,
    int32_t* __restrict__ n_eff) {
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (n >= n_alive) return;
  const int64_t r = alive[n];
  const Ray ray = load_ray(rays_o, rays_d, r);
  const float t = hits_t[2 * r];
  const float t2 = hits_t[2 * r + 1];
  const int64_t o0 = n * N_samples;
  float t_next = t;
  // unlike the training kernel the reference's test loop has no `0 <= t` guard (raymarching.cu:365)
  int k = 0;
  if (t < t2) k = march_ray_warp<kConstDt>(ray, p, p.bitfield, t, t2, N_samples, ts + o0, t_next);
  __syncwarp();
  if (lane == 0) {
    if (k > 0) hits_t[2 * r] = t_next;   // only a taken sample moves the resume point (raymarching.cu:389-391)
    n_eff[n] = k;
  }
  for (int j = lane; j < N_samples; j += 32) {
    const int64_t o = o0 + j;
    if (j < k) {
      const float tj = ts[o];
      xyzs[3 * o] = __fmaf_rn(ray.dx, tj, ray.ox); xyzs[3 * o + 1] = __fmaf_rn(ray.dy, tj, ray.oy); xyzs[3 * o + 2] = __fmaf_rn(ray.dz, tj, ray.oz);
      dirs[3 * o] = ray.dx; dirs[3 * o + 1] = ray.dy; dirs[3 * o + 2] = ray.dz;
      deltas[o] = calc_dt(tj, p);
    } else {
      xyzs[3 * o] = 0.f; xyzs[3 * o + 1] = 0.f; xyzs[3 * o + 2] = 0.f;
      dirs[3 * o] = 0.f; dirs[3 * o + 1] = 0.f; dirs[3 * o + 2] = 0.f;
      ts[o] = 0.f; deltas[o] = 0.f;
    }
  }
}

}  // namespace rnb

using namespace rnb;

extern "C" {

int rnb_raymarching_train_count(const float* rays_o, const float* rays_d, const float* hits_t,
                                const uint8_t* density_bitfield, int cascades, float scale,
                                float exp_step_factor, const float* noise, int grid_size, int max_samples,
                                int64_t n_rays, int64_t* rays_a, int32_t* counter, float* t_scratch,
                                rnb_stream_t stream) {
  const char* what = "rnb_raymarching_train_count";
  RNB_REQUIRE(n_rays >= 0 && cascades >= 1 && grid_size >= 1 && grid_size <= 1024 && max_samples >= 1, what);
  RNB_REQUIRE(counter != nullptr, what);
  RNB_REQUIRE(n_rays == 0 || (rays_o && rays_d && hits_t && density_bitfield && noise && rays_a && t_scratch), what);
  const MarchParams p = make_march_params(density_bitfield, cascades, scale, exp_step_factor, grid_size, max_samples, scale);
  if (n_rays > 0) {
    const int grid = ceil_div(n_rays * 32, 256);
    if (exp_step_factor == 0.0f)
      march_train_count_kernel<true><<<grid, 256, 0, as_stream(stream)>>>(rays_o, rays_d, hits_t, noise, p, n_rays, rays_a, t_scratch);
    else
      march_train_count_kernel<false><<<grid, 256, 0, as_stream(stream)>>>(rays_o, rays_d, hits_t, noise, p, n_rays, rays_a, t_scratch);
  }
  march_train_scan_kernel<<<1, 1024, 0, as_stream(stream)>>>(n_rays, rays_a, counter);
  return check_launch(what);
}

int rnb_raymarching_train_write(const float* rays_o, const float* rays_d, float scale, float exp_step_factor,
                                int grid_size, int max_samples, int64_t n_rays, const int64_t* rays_a,
                                const float* t_scratch, int64_t cap, float* xyzs, float* dirs, float* deltas,
                                float* ts, rnb_stream_t stream) {
  const char* what = "rnb_raymarching_train_write";
  RNB_REQUIRE(n_rays >= 0 && cap >= 0 && grid_size >= 1 && grid_size <= 1024 && max_samples >= 1, what);
  if (n_rays == 0 || cap == 0) return RNB_OK;
  RNB_REQUIRE(rays_o && rays_d && rays_a && t_scratch && xyzs && dirs && deltas && ts, what);
  const MarchParams p = make_march_params(nullptr, 1, scale, exp_step_factor, grid_size, max_samples, scale);
  march_train_write_kernel<<<ceil_div(n_rays * 32, 256), 256, 0, as_stream(stream)>>>(rays_o, rays_d, p, n_rays, rays_a, t_scratch, cap, xyzs, dirs, deltas, ts);
  return check_launch(what);
}

// ---- no-allocation variant of the training marcher ----------------------------------------------------------------
// (No counterpart in the reference: added for callers that cannot call cudaMalloc --
// code under CUDA-graph capture, tight serving loops, harnesses that forbid it.)  Same kernels as
// rnb_raymarching_train_count + rnb_raymarching_train_write; the per-ray sample-t rows live in g_noalloc_t_scratch below
// instead of a caller buffer.  The array is file-scope device memory: every GPU that loads the library holds its 32 MiB
// until exit.  In HOST code the array's name is only a stand-in for the symbol, not a device address: its address is taken
// once per device, with that device current, through cudaGetSymbolAddress, and that pointer is what the kernels get.
}  // extern "C"
namespace rnb {
static __device__ __align__(256) float g_noalloc_t_scratch[RNB_NOALLOC_MARCH_MAX_FLOATS];
}
#include <mutex>
namespace {
constexpr int kMaxDevices = 64;
float* g_noalloc_ptr[kMaxDevices] = {};
std::mutex g_noalloc_mutex;
// the device address of the scratch array on the CURRENT device (set-up on the first call per device)
int noalloc_scratch(float** out, const char* what) {
  int dev = -1;
  RNB_CUDA(cudaGetDevice(&dev), what);
  RNB_REQUIRE(dev >= 0 && dev < kMaxDevices, what);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    RNB_CUDA(cudaGetSymbolAddress((void**)&p, rnb::g_noalloc_t_scratch), what);
    RNB_REQUIRE(p != nullptr, what);
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return RNB_OK;
}
}  // namespace
extern "C" {

int rnb_raymarching_train_noalloc(const float* rays_o, const float* rays_d, const float* hits_t, const uint8_t* density_bitfield,
                                  int cascades, float scale, float exp_step_factor, const float* noise, int grid_size,
                                  int max_samples, int64_t n_rays, int64_t* rays_a, int32_t* counter, int64_t cap, float* xyzs,
                                  float* dirs, float* deltas, float* ts, rnb_stream_t stream) {
  const char* what = "rnb_raymarching_train_noalloc";
  RNB_REQUIRE(n_rays >= 0 && cap >= 0 && cascades >= 1 && grid_size >= 1 && grid_size <= 1024 && max_samples >= 1, what);
  RNB_REQUIRE(counter != nullptr, what);
  if (n_rays * (int64_t)max_samples > (int64_t)RNB_NOALLOC_MARCH_MAX_FLOATS) {
    set_error("%s: n_rays * max_samples = %lld exceeds the built-in scratch of %lld floats (RNB_NOALLOC_MARCH_MAX_FLOATS); "
              "use rnb_raymarching_train_count / _write with a caller-provided t_scratch",
              what, (long long)(n_rays * (int64_t)max_samples), (long long)RNB_NOALLOC_MARCH_MAX_FLOATS);
    return RNB_ERR_BADARG;
  }
  RNB_REQUIRE(n_rays == 0 || (rays_o && rays_d && hits_t && density_bitfield && noise && rays_a), what);
  RNB_REQUIRE(n_rays == 0 || cap == 0 || (xyzs && dirs && deltas && ts), what);
  float* t_scratch = nullptr;
  if (int rc = noalloc_scratch(&t_scratch, what)) return rc;
  cudaStream_t st = as_stream(stream);
  const MarchParams p = make_march_params(density_bitfield, cascades, scale, exp_step_factor, grid_size, max_samples, scale);
  // nothing of the scratch has to start at zero: a row's first N_samples entries are written by the count kernel and
  // only those are read by the write kernel, so the previous call's contents are never seen
  if (n_rays > 0) {
    const int grid = ceil_div(n_rays * 32, 256);
    if (exp_step_factor == 0.0f)
      march_train_count_kernel<true><<<grid, 256, 0, st>>>(rays_o, rays_d, hits_t, noise, p, n_rays, rays_a, t_scratch);
    else
      march_train_count_kernel<false><<<grid, 256, 0, st>>>(rays_o, rays_d, hits_t, noise, p, n_rays, rays_a, t_scratch);
  }
  march_train_scan_kernel<<<1, 1024, 0, st>>>(n_rays, rays_a, counter);
  if (n_rays > 0 && cap > 0)
    march_train_write_kernel<<<ceil_div(n_rays * 32, 256), 256, 0, st>>>(rays_o, rays_d, p, n_rays, rays_a, t_scratch, cap, xyzs, dirs, deltas, ts);
  return check_launch(what);
}

int rnb_raymarching_test(const float* rays_o, const float* rays_d, float* hits_t, const int64_t* alive_indices,
                         int64_t n_alive, const uint8_t* density_bitfield, int cascades, float scale,
                         float exp_step_factor, int grid_size, int max_samples, int N_samples, float* xyzs,
                         float* dirs, float* deltas, float* ts, int32_t* N_eff_samples, rnb_stream_t stream) {
  const char* what = "rnb_raymarching_test";
  RNB_REQUIRE(n_alive >= 0 && cascades >= 1 && grid_size >= 1 && grid_size <= 1024 && max_samples >= 1 && N_samples >= 1, what);
  if (n_alive == 0) return RNB_OK;
  RNB_REQUIRE(rays_o && rays_d && hits_t && alive_indices && density_bitfield && xyzs && dirs && deltas && ts && N_eff_samples, what);
  // hazard Q1: the reference's test kernel hands `cascades` to calc_dt where `scale` is expected
  const MarchParams p = make_march_params(density_bitfield, cascades, scale, exp_step_factor, grid_size, max_samples, (float)cascades);
  const int grid = ceil_div(n_alive * 32, 256);
  if (exp_step_factor == 0.0f)
    march_test_kernel<true><<<grid, 256, 0, as_stream(stream)>>>(rays_o, rays_d, hits_t, alive_indices, n_alive, p, N_samples, xyzs, dirs, deltas, ts, N_eff_samples);
  else
    march_test_kernel<false><<<grid, 256, 0, as_stream(stream)>>>(rays_o, rays_d, hits_t, alive_indices, n_alive, p, N_samples, xyzs, dirs, deltas, ts, N_eff_samples);
  return check_launch(what);
}

}  // extern "C"
