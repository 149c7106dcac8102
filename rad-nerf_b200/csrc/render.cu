// render.cu -- full-frame (test-time) rendering of all K sub-NeRFs with the gated blend.
//
// Replaces ml_render(test_time=True): K x __render_rays_test (models/ml_rendering.py:81-155) -- the
// alive-ray loop around vren.raymarching_test / MNGP.forward / vren.composite_test_fw -- plus the
// background and gate-weighted blend (ml_rendering.py:64-77).  The reference runs, per sub-NeRF and per
// iteration, two vren kernels, ~10 ATen kernels (boolean-mask compaction and scatter), the tcnn kernels
// and two host synchronisations, on zero-filled (N_alive, N_samples) slabs.  Here ALL K sub-NeRFs advance
// together, one round = 4 launches and one 4*K-byte read-back:
//   inf_march      warp per alive (pass, ray): resumes the chain at t_cur, takes up to NS samples
//                  (march_warp.cuh), reserves a dense range of the pass's sample region with one atomicAdd
//                  and writes (ray, t) -- no slabs, no masks, no memset;
//   rnb_field_fwd  the fused tcgen05 field on the packed list (all passes, nothing saved);
//   inf_composite  warp per alive (pass, ray): resumes T = 1 - opacity, composites its samples with the
//                  reference's recurrence (volumerendering.cu:223-249), appends survivors to the next list.
// NS follows the reference's schedule clamp(N_rays / N_alive, min, 64) (ml_rendering.py:115) with a floor of
// 8 (Q13: the schedule only matters for the 1024-scheduled-samples cut-off).
#include <stdlib.h>

#include <vector>

#include "aabb.cuh"
#include "composite.cuh"
#include "march_warp.cuh"
#include "coarse.cuh"

namespace rnb {

struct BitfieldPtrsR { const uint8_t* p[RNB_MAX_K]; };
// The plan of one round, in device memory: written by inf_plan_kernel from the previous round's survivors, read by the
// round's kernels.  The host launches rounds without knowing how many rays are alive (no read-back per round).
struct PassCounts {
  int32_t n[RNB_MAX_K]; int32_t base[RNB_MAX_K + 1];   // alive items / sample-region base per pass
  int32_t NS;               // sample budget per item in this round (ml_rendering.py:115)
  int32_t scheduled;        // budgets handed out so far (the reference's `samples`, ml_rendering.py:111-116)
  int32_t done;             // no ray alive or the 1024 scheduled samples are spent: every later launch returns at once
  int32_t rounds;           // rounds planned so far
  long long total_samples;  // samples evaluated (statistics)
};

struct InfState {
  float* t_cur; float* t2; float* opacity; float* depth; float* rgb;   // (K*R), rgb (K*R,3)
  int32_t* item_start; int32_t* item_n;                                  // (K*R)
  int32_t* alive[2];                                                     // (K*R) each, ping-pong
  int32_t* alive_cnt[2];                                                 // (K) each
  int32_t* sample_cnt;                                                   // (K) samples reserved in this round
};

// all (pass, ray) items alive; scene-box interval + NEAR_DISTANCE clamp (ml_rendering.py:48-50)
__global__ void __launch_bounds__(256) inf_init_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d, int R, int K,
                                                       float cx, float cy, float cz, float hx, float hy, float hz, float near_d, InfState S,
                                                       PassCounts* __restrict__ plan) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  const float ox = rays_o[3 * r], oy = rays_o[3 * r + 1], oz = rays_o[3 * r + 2];
  const float ix = __fdiv_rn(1.0f, rays_d[3 * r]), iy = __fdiv_rn(1.0f, rays_d[3 * r + 1]), iz = __fdiv_rn(1.0f, rays_d[3 * r + 2]);
  const float2 hit = slab_test(ox, oy, oz, ix, iy, iz, cx, cy, cz, hx, hy, hz);
  float t1 = -1.f, t2 = -1.f;
  if (hit.y > 0) { t1 = fmaxf(hit.x, 0.0f); t2 = hit.y; }
  if (t1 >= 0 && t1 < near_d) t1 = near_d;
  for (int p = 0; p < K; p++) {
    const int64_t i = (int64_t)p * R + r;
    S.t_cur[i] = t1; S.t2[i] = t2; S.opacity[i] = 0.f; S.depth[i] = 0.f;
    S.rgb[3 * i] = 0.f; S.rgb[3 * i + 1] = 0.f; S.rgb[3 * i + 2] = 0.f;
    S.alive[0][i] = r;
    if (r == 0) { S.alive_cnt[0][p] = R; S.sample_cnt[p] = 0; }
  }
  if (r == 0) { plan->NS = 0; plan->scheduled = 0; plan->done = 0; plan->rounds = 0; plan->total_samples = 0; }
}

// One thread plans the round: N_samples = max(min(N_rays // N_alive, 64), min_samples) (ml_rendering.py:115) with a floor set
// by the workspace the caller provided (samples per alive item and round that always fit: 8..64; a warp computes 32 chain
// elements at a time, so small budgets re-do marching work round after round), shrunk until the round's samples fit.
__global__ void inf_plan_kernel(PassCounts* __restrict__ C, InfState S, int K, int R, int in, int max_samples, int min_samples,
                                int ns_floor, int64_t capacity) {
  if (threadIdx.x != 0 || C->done) return;
  int64_t alive = 0;
  for (int p = 0; p < K; p++) { C->total_samples += S.sample_cnt[p]; S.sample_cnt[p] = 0; alive += S.alive_cnt[in][p]; }
  if (C->scheduled >= max_samples || alive == 0) { C->done = 1; return; }
  int64_t NS = ((int64_t)K * R) / alive;
  if (NS > 64) NS = 64;
  if (NS < min_samples) NS = min_samples;
  if (NS < ns_floor) NS = ns_floor;
  while (NS > 1) {
    int64_t need = 0;
    for (int p = 0; p < K; p++) need += ((int64_t)S.alive_cnt[in][p] * NS + 127) / 128 * 128;
    if (need <= capacity) break;
    NS--;
  }
  int64_t base = 0;
  for (int p = 0; p < K; p++) {
    C->n[p] = S.alive_cnt[in][p];
    C->base[p] = (int32_t)base; base += ((int64_t)C->n[p] * NS + 127) / 128 * 128;
    S.alive_cnt[in ^ 1][p] = 0;
  }
  C->base[K] = (int32_t)base;
  C->NS = (int32_t)NS; C->scheduled += (int32_t)NS; C->rounds += 1;
}

// flat alive-item index -> (pass, position in the pass's list)
__device__ __forceinline__ bool locate_item(const PassCounts& C, int K, int64_t w, int& pass, int& j) {
  int64_t acc = 0;
  for (int p = 0; p < K; p++) {
    if (w < acc + C.n[p]) { pass = p; j = (int)(w - acc); return true; }
    acc += C.n[p];
  }
  return false;
}

template <bool kConstDt, bool kOneCascade>
__global__ void __launch_bounds__(256) inf_march_kernel(const float* __restrict__ rays_o, const float* __restrict__ rays_d, MarchParams mp,
                                                        BitfieldPtrsR bf, const PassCounts* __restrict__ plan, int K, int R, int in,
                                                        int64_t capacity, InfState S, int32_t* __restrict__ ray_id,
                                                        float* __restrict__ ts, const uint64_t* __restrict__ coarse_grid) {
  __shared__ float tbuf[8][64];
  __shared__ PassCounts C;
  if (plan->done) return;
  if (threadIdx.x == 0) C = *plan;
  __syncthreads();
  const int NS = C.NS;
  const uint64_t* coarse = C.rounds == 1 ? coarse_grid : nullptr;      // the empty-space test runs in the first round only
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  int64_t total = 0;
  for (int p = 0; p < K; p++) total += C.n[p];
  for (int64_t w = (int64_t)blockIdx.x * 8 + wib; w < total; w += (int64_t)gridDim.x * 8) {
    int pass, j;
    locate_item(C, K, w, pass, j);
    const int r = S.alive[in][(int64_t)pass * R + j];
    const int64_t item = (int64_t)pass * R + r;
    const Ray ray = load_ray(rays_o, rays_d, r);
    const float t = S.t_cur[item];
    float t2 = S.t2[item];
    if (coarse != nullptr && t < t2) {
      // first round only (coarse.cuh): rays that cannot meet an occupied cell end here, the others stop marching behind
      // the last block they can meet -- in this and in every later round
      t2 = coarse_last_hit(ray, t, t2, coarse + (int64_t)pass * mp.cascades * kCoarseWords, mp.cascades, mp.scale);
      if (lane == 0) S.t2[item] = t2;
    }
    float t_next = t;
    int n = 0;
    if (t < t2) n = march_ray_warp<kConstDt, kOneCascade>(ray, mp, bf.p[pass], t, t2, NS, tbuf[wib], t_next);   // raymarching.cu:365
    __syncwarp();
    int start = 0;
    if (lane == 0 && n > 0) start = atomicAdd(S.sample_cnt + pass, n);
    start = __shfl_sync(kFullMask, start, 0);
    const int64_t o0 = (int64_t)C.base[pass] + start;
    for (int k = lane; k < n; k += 32) {
      if (o0 + k < capacity) { ray_id[o0 + k] = r; ts[o0 + k] = tbuf[wib][k]; }
    }
    if (lane == 0) {
      S.item_start[item] = (int)o0; S.item_n[item] = n;
      if (n > 0) S.t_cur[item] = t_next;            // only a taken sample moves the resume point (raymarching.cu:389-391)
    }
    __syncwarp();
  }
}

// seg for rnb_field_fwd: pass bases (host-planned upper bounds) and the sample counts actually reserved
__global__ void inf_seg_kernel(const PassCounts* __restrict__ C, int K, const int32_t* __restrict__ sample_cnt, int32_t* __restrict__ seg) {
  const int p = threadIdx.x;
  const bool done = C->done != 0;       // finished frame: an empty sample list (the field kernel then has nothing to do)
  if (p < K) { seg[p] = done ? 0 : C->base[p]; seg[K + 1 + p] = done ? 0 : sample_cnt[p]; }
  if (p == 0) seg[K] = done ? 0 : C->base[K];
}

// Resume-able compositing of one round (volumerendering.cu:223-249), warp per alive item.
__global__ void __launch_bounds__(256) inf_composite_kernel(const float* __restrict__ sigmas, const __half* __restrict__ rgb4,
                                                            const float* __restrict__ ts, MarchParams mp, const PassCounts* __restrict__ plan,
                                                            int K, int R, int in, float T_threshold, int64_t capacity, InfState S) {
  __shared__ PassCounts C;
  if (plan->done) return;
  if (threadIdx.x == 0) C = *plan;
  __syncthreads();
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  int64_t total = 0;
  for (int p = 0; p < K; p++) total += C.n[p];
  for (int64_t w = (int64_t)blockIdx.x * 8 + wib; w < total; w += (int64_t)gridDim.x * 8) {
    int pass, j;
    locate_item(C, K, w, pass, j);
    const int r = S.alive[in][(int64_t)pass * R + j];
    const int64_t item = (int64_t)pass * R + r;
    int n = S.item_n[item];
    const int64_t start = S.item_start[item];
    if (start + n > capacity) n = start < capacity ? (int)(capacity - start) : 0;
    if (n == 0) continue;                                   // no sample taken: the ray is finished (Q11)
    float T = 1.0f - S.opacity[item];
    float Rs = 0.f, Gs = 0.f, Bs = 0.f, Ds = 0.f, Os = 0.f;
    bool dead = false;
    for (int base = 0; base < n; base += 32) {
      const int k = base + lane;
      const bool valid = k < n;
      float a = 0.f, cr = 0.f, cg = 0.f, cb = 0.f, tt = 0.f;
      if (valid) {
        const int64_t s = start + k;
        tt = ts[s];
        a = 1.0f - __expf(-sigmas[s] * calc_dt(tt, mp));
        load_rgb<true>(rgb4, s, cr, cg, cb);
      }
      const float om = 1.0f - a;
      const float Tb = serial_T(om, T, lane);
      const float Ta = __fmul_rn(Tb, om);
      const unsigned term = __ballot_sync(kFullMask, valid && Ta <= T_threshold);
      const int first = term ? (__ffs(term) - 1) : 32;
      const float wgt = (valid && lane <= first) ? __fmul_rn(a, Tb) : 0.f;
      Rs = fmaf(wgt, cr, Rs); Gs = fmaf(wgt, cg, Gs); Bs = fmaf(wgt, cb, Bs); Ds = fmaf(wgt, tt, Ds); Os += wgt;
      if (term) { dead = true; break; }
    }
    Rs = warp_sum(Rs); Gs = warp_sum(Gs); Bs = warp_sum(Bs); Ds = warp_sum(Ds); Os = warp_sum(Os);
    if (lane == 0) {
      S.opacity[item] += Os; S.depth[item] += Ds;
      S.rgb[3 * item] += Rs; S.rgb[3 * item + 1] += Gs; S.rgb[3 * item + 2] += Bs;
      if (!dead) S.alive[in ^ 1][(int64_t)pass * R + atomicAdd(S.alive_cnt[in ^ 1] + pass, 1)] = r;
    }
  }
}

struct BgR { float c[RNB_MAX_K][3]; };

// background per sub-NeRF + gate-weighted blend (ml_rendering.py:149-153, 64-68)
__global__ void __launch_bounds__(256) inf_finish_kernel(InfState S, const float* __restrict__ prob, BgR bg, int R, int K,
                                                         float* __restrict__ rgb, float* __restrict__ depth, float* __restrict__ opacity,
                                                         float* __restrict__ pass_rgb) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  float c[3] = {0.f, 0.f, 0.f}, o = 0.f;
  for (int p = 0; p < K; p++) {
    const int64_t i = (int64_t)p * R + r;
    const float g = K > 1 ? prob[(int64_t)r * K + p] : 1.f;
    const float oi = S.opacity[i];
    for (int k = 0; k < 3; k++) {
      const float ci = S.rgb[3 * i + k] + bg.c[p][k] * (1.f - oi);
      if (pass_rgb) pass_rgb[3 * i + k] = ci;
      c[k] = fmaf(ci, g, c[k]);
    }
    o = fmaf(oi, g, o);
    depth[(int64_t)r * K + p] = S.depth[i];
  }
  rgb[3 * r] = c[0]; rgb[3 * r + 1] = c[1]; rgb[3 * r + 2] = c[2];
  opacity[r] = o;
}

}  // namespace rnb

using namespace rnb;

extern "C" {

int rnb_render_layout_of(const rnb_engine_config* c, rnb_render_layout* L) {
  const char* what = "rnb_render_layout_of";
  RNB_REQUIRE(c && L && c->K >= 1 && c->K <= RNB_MAX_K && c->n_rays >= 1, what);
  const int64_t R = c->n_rays, K = c->K, S = c->sample_capacity;
  RNB_REQUIRE(S % 128 == 0 && S >= (K * R * 8 + 127) / 128 * 128 + 128 * K, what);   // room for the first round at NS = 8
  int64_t off = 0;
  auto take = [&](int64_t bytes) { const int64_t o = off; off = (off + bytes + 255) / 256 * 256; return o; };
  L->gate_in = take(R * 16 * 2); L->gate_logits = take(R * 16 * 2); L->gate_prob = take(R * K * 4); L->importance = take(RNB_MAX_K * 4);
  L->t_cur = take(K * R * 4); L->t2 = take(K * R * 4); L->acc_opacity = take(K * R * 4); L->acc_depth = take(K * R * 4);
  L->acc_rgb = take(K * R * 12); L->item_start = take(K * R * 4); L->item_n = take(K * R * 4);
  L->alive0 = take(K * R * 4); L->alive1 = take(K * R * 4);
  L->counters = take(512);      // alive counts (2 x K), sample counts (K); at byte 256: the round plan
  L->seg = take((2 * RNB_MAX_K + 4) * 4);
  L->ray_id = take(S * 4); L->ts = take(S * 4); L->sigmas = take(S * 4); L->rgb4 = take(S * 8);
  L->coarse = take((int64_t)K * c->cascades * kCoarseWords * 8);
  L->total_bytes = off;
  return RNB_OK;
}

int rnb_render(const rnb_engine_config* c, void* workspace, const void* params_f16, const float* rays_o, const float* rays_d,
               const uint8_t* const* bitfields, const float* bg_host, float* rgb, float* depth, float* opacity, float* pass_rgb,
               int64_t* stats_host, rnb_stream_t stream) {
  const char* what = "rnb_render";
  rnb_render_layout L;
  if (int rc = rnb_render_layout_of(c, &L)) return rc;
  RNB_REQUIRE(workspace && params_f16 && rays_o && rays_d && bitfields && bg_host && rgb && depth && opacity, what);
  char* ws = (char*)workspace;
  cudaStream_t st = as_stream(stream);
  const int R = c->n_rays, K = c->K;
  const int64_t S = c->sample_capacity;
  const __half* p16 = (const __half*)params_f16;
  if (K > 1) {
    if (int rc = rnb_gate_fwd(p16 + c->gate_off, rays_o, rays_d, R, K, ws + L.gate_in, nullptr, ws + L.gate_logits,
                              (float*)(ws + L.gate_prob), (float*)(ws + L.importance), stream)) return rc;
  }
  InfState St;
  St.t_cur = (float*)(ws + L.t_cur); St.t2 = (float*)(ws + L.t2); St.opacity = (float*)(ws + L.acc_opacity);
  St.depth = (float*)(ws + L.acc_depth); St.rgb = (float*)(ws + L.acc_rgb);
  St.item_start = (int32_t*)(ws + L.item_start); St.item_n = (int32_t*)(ws + L.item_n);
  St.alive[0] = (int32_t*)(ws + L.alive0); St.alive[1] = (int32_t*)(ws + L.alive1);
  int32_t* counters = (int32_t*)(ws + L.counters);
  St.alive_cnt[0] = counters; St.alive_cnt[1] = counters + RNB_MAX_K; St.sample_cnt = counters + 2 * RNB_MAX_K;
  int32_t* seg = (int32_t*)(ws + L.seg);
  int32_t* ray_id = (int32_t*)(ws + L.ray_id); float* ts = (float*)(ws + L.ts);
  float* sigmas = (float*)(ws + L.sigmas); __half* rgb4 = (__half*)(ws + L.rgb4);
  // hazard Q1: the reference's test kernel hands `cascades` to calc_dt where `scale` is expected
  const MarchParams mp = make_march_params(nullptr, c->cascades, c->scale, c->exp_step_factor, c->grid_size, c->max_samples, (float)c->cascades);
  BitfieldPtrsR bf;
  for (int i = 0; i < K; i++) { RNB_REQUIRE(bitfields[i] != nullptr, what); bf.p[i] = bitfields[i]; }
  uint64_t* coarse = nullptr;
  if (c->grid_size == 2 * kCoarse && !(getenv("RNB_NO_CULL") && atoi(getenv("RNB_NO_CULL")))) {
    coarse = (uint64_t*)(ws + L.coarse);
    if (int rc = coarse_build(bitfields, K, c->cascades, coarse, st)) return rc;
  }
  PassCounts* plan = reinterpret_cast<PassCounts*>(ws + L.counters + 256);
  static_assert(sizeof(PassCounts) <= 256 && 3 * RNB_MAX_K * 4 <= 256, "round plan fits the counters region");
  inf_init_kernel<<<ceil_div(R, 256), 256, 0, st>>>(rays_o, rays_d, R, K, c->center[0], c->center[1], c->center[2], c->half_size[0],
                                                    c->half_size[1], c->half_size[2], c->near_distance, St, plan);
  rnb_field_params P;
  P.K = K; P.grid = c->grid;
  for (int k = 0; k < 3; k++) { P.xyz_min[k] = c->xyz_min[k]; P.xyz_max[k] = c->xyz_max[k]; }
  P.table_f16 = p16 + c->table_off;
  for (int i = 0; i < K; i++) { P.geo_w_f16[i] = p16 + c->geo_off[i]; P.rgb_w_f16[i] = p16 + c->rgb_off[i]; }
  const int min_samples = c->exp_step_factor == 0.0f ? 1 : 4;   // ml_rendering.py:108
  int64_t ns_floor = (S - 128 * (int64_t)K) / ((int64_t)K * R);
  ns_floor = ns_floor < 8 ? 8 : ns_floor > 64 ? 64 : ns_floor;
  // RNB_RENDER_PROF=1: per-stage device time of this call on stderr (CUDA events; diagnostics only)
  const bool prof = getenv("RNB_RENDER_PROF") != nullptr;
  std::vector<cudaEvent_t> pev;
  auto mark = [&]() { if (prof) { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); pev.push_back(e); } };
  // The reference's loop (ml_rendering.py:108-141) reads the number of alive rays back twice per iteration and sub-NeRF.
  // Here a round is planned ON THE DEVICE (inf_plan_kernel) and the host launches rounds without waiting: kBatch rounds
  // at a time, with an asynchronous copy of the plan's `done` flag behind every batch that is looked at only after the NEXT
  // batch has been enqueued -- the GPU never idles on the host, and a finished frame costs at most 2 * kBatch - 1 empty
  // rounds (four kernels each that return at once) before the host notices.
  struct Pinned { int32_t* p = nullptr; cudaEvent_t ev[2] = {nullptr, nullptr}; };
  static thread_local Pinned pin;       // 1 KB of pinned memory and two events per calling thread, created once
  if (pin.p == nullptr) {
    RNB_CUDA(cudaHostAlloc((void**)&pin.p, 1024, cudaHostAllocDefault), what);
    RNB_CUDA(cudaEventCreateWithFlags(&pin.ev[0], cudaEventDisableTiming), what);
    RNB_CUDA(cudaEventCreateWithFlags(&pin.ev[1], cudaEventDisableTiming), what);
  }
  const int kBatch = 4;
  const int max_rounds = c->max_samples;      // a round schedules at least one sample: the loop below ends on `done` long before
  const int grid = kNumSMs * 8;
  rnb_samples Sm = {rays_o, rays_d, ray_id, ts, seg, S};
  int launched = 0, batches = 0;
  bool finished = false;
  while (!finished && launched < max_rounds + 2 * kBatch) {
    for (int b = 0; b < kBatch; b++, launched++) {
      const int in = launched & 1;
      mark();
      inf_plan_kernel<<<1, 32, 0, st>>>(plan, St, K, R, in, c->max_samples, min_samples, (int)ns_floor, S);
#define RNB_INF_MARCH(CONST_DT, ONE) \
  inf_march_kernel<CONST_DT, ONE><<<grid, 256, 0, st>>>(rays_o, rays_d, mp, bf, plan, K, R, in, S, St, ray_id, ts, coarse)
      if (c->exp_step_factor == 0.0f && c->cascades == 1) RNB_INF_MARCH(true, true);
      else if (c->exp_step_factor == 0.0f) RNB_INF_MARCH(true, false);
      else RNB_INF_MARCH(false, false);
#undef RNB_INF_MARCH
      inf_seg_kernel<<<1, 32, 0, st>>>(plan, K, St.sample_cnt, seg);
      if (int rc = check_launch(what)) return rc;
      mark();
      if (int rc = rnb_field_fwd(&P, &Sm, nullptr, sigmas, rgb4, stream)) return rc;
      mark();
      inf_composite_kernel<<<grid, 256, 0, st>>>(sigmas, rgb4, ts, mp, plan, K, R, in, c->T_threshold, S, St);
      if (int rc = check_launch(what)) return rc;
      mark();
    }
    RNB_CUDA(cudaMemcpyAsync(pin.p + 16 * (batches & 1), &plan->done, 4, cudaMemcpyDeviceToHost, st), what);
    RNB_CUDA(cudaEventRecord(pin.ev[batches & 1], st), what);
    if (batches > 0) {      // the PREVIOUS batch's flag: that batch has long finished
      RNB_CUDA(cudaEventSynchronize(pin.ev[(batches - 1) & 1]), what);
      finished = pin.p[16 * ((batches - 1) & 1)] != 0;
    }
    batches++;
  }
  // the plan of the round that found nothing alive carries the totals
  inf_plan_kernel<<<1, 32, 0, st>>>(plan, St, K, R, launched & 1, c->max_samples, min_samples, (int)ns_floor, S);
  static_assert(256 + sizeof(PassCounts) <= 1024, "pinned block");
  RNB_CUDA(cudaMemcpyAsync(pin.p + 64, plan, sizeof(PassCounts), cudaMemcpyDeviceToHost, st), what);
  BgR bg;
  for (int i = 0; i < K; i++) for (int k = 0; k < 3; k++) bg.c[i][k] = bg_host[3 * i + k];
  inf_finish_kernel<<<ceil_div(R, 256), 256, 0, st>>>(St, (const float*)(ws + L.gate_prob), bg, R, K, rgb, depth, opacity, pass_rgb);
  RNB_CUDA(cudaStreamSynchronize(st), what);      // one synchronisation per frame: the statistics (and the pinned words are free again)
  const PassCounts* fin = reinterpret_cast<const PassCounts*>(pin.p + 64);
  const int rounds = fin->rounds;
  const int64_t total_samples = fin->total_samples;
  if (prof) {
    float t_march = 0.f, t_field = 0.f, t_comp = 0.f;
    for (size_t i = 0; i + 3 < pev.size(); i += 4) {
      float a = 0.f, b2 = 0.f, c2 = 0.f;
      cudaEventElapsedTime(&a, pev[i], pev[i + 1]); cudaEventElapsedTime(&b2, pev[i + 1], pev[i + 2]); cudaEventElapsedTime(&c2, pev[i + 2], pev[i + 3]);
      t_march += a; t_field += b2; t_comp += c2;
    }
    fprintf(stderr, "[rnb_render] rays %d K %d rounds %d (launched %d) samples %lld: plan+march %.3f ms, field %.3f ms, composite %.3f ms\n", R, K,
            rounds, launched, (long long)total_samples, t_march, t_field, t_comp);
    for (cudaEvent_t e : pev) cudaEventDestroy(e);
  }
  if (stats_host) { stats_host[0] = total_samples; stats_host[1] = rounds; }
  return check_launch(what);
}

}  // extern "C"
