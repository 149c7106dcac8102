// peer.cu -- the data-parallel exchange of the training step as ONE kernel over NVLink peer memory.
//
// The reference's multi-GPU step is DDP: bucketed NCCL all-reduce of every gradient, then the same Adam step on every
// rank (train_ml.py:143, 309-318).  Here rank r OWNS shard r of the flat parameter vector.  After the backward pass,
// one kernel per rank does, for its shard only,
//     g   = sum over ranks p = 0..W-1 of grads_p[shard]      (plain peer loads over NVLink, fixed order: the same bits
//                                                             on every run and independent of the reading rank)
//     Adam on the fp32 master / moments of the shard          (local, 1/W of the optimizer work)
//     fp16 compute copy of the shard -> ALL ranks' buffers    (peer stores)
// i.e. reduce-scatter + optimizer + all-gather with no staging buffer and half the all-gather payload (fp16, 2 B per
// parameter instead of the 4 B an all-reduce returns).  Ordering between ranks comes from the two small collectives the
// host side issues around it (parallel.py): every rank's backward precedes the kernel, and nobody reads parameters or
// re-zeroes gradients before every rank's kernel is done.  The kernel itself never waits on another GPU.
//
// Buffers that peers touch (gradients: read; fp16 parameters: written) are allocated with rnb_peer_alloc (cudaMalloc +
// CUDA IPC handle) and mapped into the other processes with rnb_peer_open.
#include <string.h>

#include "adam.cuh"

namespace rnb {

struct PeerPtrs {
  const float* grads[RNB_MAX_PEERS];
  __half* p16[RNB_MAX_PEERS];
  int32_t* flags[RNB_MAX_PEERS];      // every rank's barrier words (RNB_PEER_FLAG_WORDS), or all NULL: the caller orders the ranks
};

// ---- flag barrier over peer memory ------------------------------------------------------------------------------------
// Every rank owns 2 x RNB_MAX_PEERS counters in its own (peer-mapped) memory: ARRIVE[p] is written by rank p when its backward
// of step v is complete, DONE[p] when its optimizer kernel of step v has stored its shard everywhere.  Counters only grow
// (v = optimizer step), so nothing is ever reset.  A spin is bounded (~seconds): on a timeout status[6] is set and the kernel
// goes on, so a lost peer shows up as a reported error instead of a hung GPU.
constexpr int kFlagArrive = 0, kFlagDone = RNB_MAX_PEERS;
__device__ __forceinline__ int32_t ld_acquire_sys(const int32_t* p) {
  int32_t v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(int32_t* p, int32_t v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ bool spin_until(const int32_t* flag, int32_t value) {
  for (int it = 0; it < (1 << 24); it++) {
    if (ld_acquire_sys(flag) >= value) return true;
    __nanosleep(it < 64 ? 20 : 200);
  }
  return false;
}
// one thread per peer: "my backward of step `value` is done" into every rank's ARRIVE[rank]
__global__ void peer_signal_kernel(PeerPtrs P, int world, int rank, int slot, int32_t value) {
  const int p = threadIdx.x;
  if (p < world) { __threadfence_system(); st_release_sys(P.flags[p] + slot + rank, value); }
}
// one thread per peer: wait until every rank's counter in MY words has reached `value`
__global__ void peer_wait_kernel(const int32_t* mine, int world, int slot, int32_t value, int32_t* status) {
  const int p = threadIdx.x;
  if (p < world && !spin_until(mine + slot + p, value) && status != nullptr) status[6] = 1;
}

// The W gradient vectors of one float4 index: every load is issued before any is used (the remote ones cross NVLink).
template <int W>
__device__ __forceinline__ void load_peers(const PeerPtrs& P, int world, int64_t i4, float4 (&g)[W]) {
#pragma unroll
  for (int p = 0; p < W; p++)
    if (p < world) g[p] = __ldcs(reinterpret_cast<const float4*>(P.grads[p]) + i4);
}
// ... and their sum in rank order from zero: the same bits on every rank and on every run.
template <int W>
__device__ __forceinline__ float4 add_peers(int world, const float4 (&g)[W]) {
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int p = 0; p < W; p++)
    if (p < world) { acc.x += g[p].x; acc.y += g[p].y; acc.z += g[p].z; acc.w += g[p].w; }
  return acc;
}

__device__ __forceinline__ float4 adam4(float4 gg, float* __restrict__ p32, float* __restrict__ m, float* __restrict__ v, int64_t i,
                                        float step_size, float beta1, float beta2, float eps, float inv_sqrt_bc2, float grad_scale) {
  float4 pp = reinterpret_cast<float4*>(p32)[i], mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
  float* Pp = &pp.x; float* G = &gg.x; float* M = &mm.x; float* V = &vv.x;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const float gr = G[k] * grad_scale;
    M[k] = beta1 * M[k] + (1.f - beta1) * gr;
    V[k] = beta2 * V[k] + (1.f - beta2) * gr * gr;
    Pp[k] -= step_size * M[k] / (sqrtf(V[k]) * inv_sqrt_bc2 + eps);
  }
  reinterpret_cast<float4*>(p32)[i] = pp; reinterpret_cast<float4*>(m)[i] = mm; reinterpret_cast<float4*>(v)[i] = vv;
  return pp;
}
__device__ __forceinline__ uint2 to_half4(float4 pp) {
  const __half2 a = __floats2half2_rn(pp.x, pp.y), b = __floats2half2_rn(pp.z, pp.w);
  return make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}
// All indices in units of 4 parameters.
//   [lo4, hi4)      this rank's shard: reduce over ranks, Adam, fp16 result to EVERY rank
//   [rep_lo4, rep_hi4)  replicated block (the gate network): every rank reduces and updates its own replica, so that the
//                   next gate forward needs no remote data
//   zero_buf[0, zero_n4)  the local gradient buffer of the PREVIOUS step (all peers finished reading it before this
//                   kernel was allowed to start): cleared here for the next step
// A remote load takes microseconds, so the loop is software-pipelined: the gradients of a thread's NEXT index are in flight
// while the current one goes through Adam, and the first ones are issued before the (store-only) clearing of the old
// gradient buffer, which then hides their latency.  (Unpipelined, the N = 2 kernel took 96 us against 66 us with every
// pointer local and 38 us of HBM traffic.)
template <int W>
__global__ void __launch_bounds__(256) adam_sharded_kernel(PeerPtrs P, int world, int rank, float* __restrict__ p32, float* __restrict__ m,
                                                           float* __restrict__ v, int64_t lo4, int64_t hi4, int64_t rep_lo4, int64_t rep_hi4,
                                                           float* __restrict__ zero_buf, int64_t zero_n4, float lr, int64_t host_step,
                                                           float beta1, float beta2, float eps, float grad_scale, int64_t flag_index,
                                                           int32_t* __restrict__ status, int32_t barrier_value) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float4 g[W];
  int64_t i = lo4 + t0;
  // in-kernel barrier (flag words given): nothing of a peer is read before its ARRIVE counter shows this step
  if (P.flags[rank] != nullptr) {
    if (threadIdx.x < world && !spin_until(P.flags[rank] + kFlagArrive + threadIdx.x, barrier_value) && status != nullptr) status[6] = 1;
    __syncthreads();
  }
  if (i < hi4) load_peers<W>(P, world, i, g);
  // step guard: OR of every rank's flag word (peer loads) -> all ranks skip or apply the same step
  int flag_bits = 0;
  if (flag_index >= 0 && threadIdx.x == 0) {
#pragma unroll
    for (int p = 0; p < W; p++)
      if (p < world) flag_bits |= *reinterpret_cast<const volatile int32_t*>(P.grads[p] + flag_index);
  }
  __shared__ AdamStep sh;
  adam_prologue(&sh, flag_bits, status, host_step, lr, beta1, beta2);
  const float step_size = sh.step_size, inv_sqrt_bc2 = sh.inv_sqrt_bc2;
  if (zero_buf != nullptr)
    for (int64_t z = t0; z < zero_n4; z += stride) __stcs(reinterpret_cast<float4*>(zero_buf) + z, make_float4(0.f, 0.f, 0.f, 0.f));
  // skipped step: nothing is applied.  The gradient buffer the peers are reading now is cleared by the NEXT step's kernel
  // like any other (zero_buf alternation), so there is nothing else to do.
  // "my optimizer kernel of this step is done" into every rank's DONE[rank], by the last block to get here: every block
  // fences its peer stores first, so a rank that sees the counter also sees the parameters
  auto signal_done = [&]() {
    if (P.flags[rank] == nullptr) return;
    __threadfence_system();
    __syncthreads();
    __shared__ int last;
    if (threadIdx.x == 0) last = atomicAdd(status + 5, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (last) {
      if (threadIdx.x == 0) status[5] = 0;
      if (threadIdx.x < world) st_release_sys(P.flags[threadIdx.x] + kFlagDone + rank, barrier_value);
    }
  };
  if (sh.skip) { signal_done(); return; }
  for (; i < hi4; i += stride) {
    float4 gn[W];
    const bool more = i + stride < hi4;
    if (more) load_peers<W>(P, world, i + stride, gn);
    const float4 pp = adam4(add_peers<W>(world, g), p32, m, v, i, step_size, beta1, beta2, eps, inv_sqrt_bc2, grad_scale);
    const uint2 q = to_half4(pp);
#pragma unroll
    for (int p = 0; p < W; p++)
      if (p < world) reinterpret_cast<uint2*>(P.p16[p])[i] = q;
    if (more) {
#pragma unroll
      for (int p = 0; p < W; p++) g[p] = gn[p];
    }
  }
  for (int64_t r = rep_lo4 + t0; r < rep_hi4; r += stride) {
    load_peers<W>(P, world, r, g);
    const float4 pp = adam4(add_peers<W>(world, g), p32, m, v, r, step_size, beta1, beta2, eps, inv_sqrt_bc2, grad_scale);
    reinterpret_cast<uint2*>(P.p16[rank])[r] = to_half4(pp);
  }
  signal_done();
}

}  // namespace rnb

using namespace rnb;

extern "C" {

int rnb_peer_alloc(int64_t bytes, void** dev_ptr, void* ipc_handle_64) {
  const char* what = "rnb_peer_alloc";
  RNB_REQUIRE(bytes > 0 && dev_ptr && ipc_handle_64, what);
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  RNB_CUDA(cudaMalloc(&p, (size_t)bytes), what);
  RNB_CUDA(cudaMemset(p, 0, (size_t)bytes), what);
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) { cudaFree(p); RNB_CUDA(e, what); }
  memcpy(ipc_handle_64, &h, 64);
  *dev_ptr = p;
  return RNB_OK;
}

int rnb_peer_open(const void* ipc_handle_64, void** dev_ptr) {
  const char* what = "rnb_peer_open";
  RNB_REQUIRE(ipc_handle_64 && dev_ptr, what);
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle_64, 64);
  RNB_CUDA(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess), what);
  return RNB_OK;
}

int rnb_peer_close(void* dev_ptr) {
  RNB_CUDA(cudaIpcCloseMemHandle(dev_ptr), "rnb_peer_close");
  return RNB_OK;
}

int rnb_peer_free(void* dev_ptr) {
  RNB_CUDA(cudaFree(dev_ptr), "rnb_peer_free");
  return RNB_OK;
}

/* this rank's shard [lo, hi) of n parameters split over `world` ranks in multiples of 4 */
int rnb_adam_shard_bounds(int64_t n, int rank, int world, int64_t* lo, int64_t* hi) {
  const char* what = "rnb_adam_shard_bounds";
  RNB_REQUIRE(n > 0 && n % 4 == 0 && world >= 1 && rank >= 0 && rank < world && lo && hi, what);
  const int64_t n4 = n / 4, per = (n4 + world - 1) / world;
  const int64_t a = per * rank < n4 ? per * rank : n4, b = per * (rank + 1) < n4 ? per * (rank + 1) : n4;
  *lo = a * 4; *hi = b * 4;
  return RNB_OK;
}

static int launch_adam_sharded(int rank, int world, const float* const* grads_peers, void* const* params_f16_peers, float* params_f32,
                               float* exp_avg, float* exp_avg_sq, int64_t n, int64_t n_sharded, float* zero_grads, float lr, float beta1,
                               float beta2, float eps, int64_t step, float grad_scale, int64_t flag_index, int32_t* status,
                               void* const* flag_peers, int32_t barrier_value, rnb_stream_t stream, const char* what) {
  RNB_REQUIRE(world >= 1 && world <= RNB_MAX_PEERS && rank >= 0 && rank < world && grads_peers && params_f16_peers, what);
  RNB_REQUIRE(params_f32 && exp_avg && exp_avg_sq && n > 0 && n % 4 == 0 && step >= 1, what);
  RNB_REQUIRE(n_sharded > 0 && n_sharded <= n && n_sharded % 4 == 0 && ((uintptr_t)zero_grads & 15) == 0, what);
  PeerPtrs P;
  for (int p = 0; p < RNB_MAX_PEERS; p++) { P.grads[p] = nullptr; P.p16[p] = nullptr; P.flags[p] = nullptr; }
  for (int p = 0; p < world; p++) {
    RNB_REQUIRE(grads_peers[p] && params_f16_peers[p], what);
    RNB_REQUIRE(((uintptr_t)grads_peers[p] & 15) == 0 && ((uintptr_t)params_f16_peers[p] & 7) == 0, what);
    P.grads[p] = grads_peers[p]; P.p16[p] = (__half*)params_f16_peers[p];
    if (flag_peers != nullptr) { RNB_REQUIRE(flag_peers[p] != nullptr && status != nullptr && barrier_value >= 1, what); P.flags[p] = (int32_t*)flag_peers[p]; }
  }
  int64_t lo, hi;
  if (int rc = rnb_adam_shard_bounds(n_sharded, rank, world, &lo, &hi)) return rc;
  // instantiation width: the smallest of 2 / 4 / 8 peers that holds `world` (RNB_PEER_MIN_WIDTH: test hook that runs the
  // wider instantiations on a box with fewer GPUs, tests/test_multi_gpu.py)
  int width = world <= 2 ? 2 : world <= 4 ? 4 : RNB_MAX_PEERS;
  if (const char* e = getenv("RNB_PEER_MIN_WIDTH")) { const int w = atoi(e); if (w > width && w <= RNB_MAX_PEERS) width = w > 4 ? RNB_MAX_PEERS : 4; }
  // one resident wave of grid-stride blocks (the pipelined loop wants every thread to start at once)
  int per_sm = 0;
  cudaError_t oe = width == 2 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, adam_sharded_kernel<2>, 256, 0)
                 : width == 4 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, adam_sharded_kernel<4>, 256, 0)
                              : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, adam_sharded_kernel<RNB_MAX_PEERS>, 256, 0);
  if (oe != cudaSuccess || per_sm < 1) per_sm = 2;
  int64_t blocks = ((zero_grads ? n : hi - lo) / 4 + 255) / 256;
  if (blocks > (int64_t)kNumSMs * per_sm) blocks = (int64_t)kNumSMs * per_sm;
  if (blocks < 1) blocks = 1;
  // flag barrier: announce this rank's backward, let the kernel wait for everybody's, and hold the stream behind it until
  // every rank's kernel has stored its shard here (three launches, no collective)
  if (flag_peers != nullptr) peer_signal_kernel<<<1, 32, 0, as_stream(stream)>>>(P, world, rank, kFlagArrive, barrier_value);
#define RNB_LAUNCH_ADAM_SHARDED(W)                                                                                                   \
  adam_sharded_kernel<W><<<(int)blocks, 256, 0, as_stream(stream)>>>(P, world, rank, params_f32, exp_avg, exp_avg_sq, lo / 4, hi / 4,   \
                                                                    n_sharded / 4, n / 4, zero_grads, n / 4, lr, step, beta1, beta2,    \
                                                                    eps, grad_scale, flag_index, status, barrier_value)
  if (width == 2) RNB_LAUNCH_ADAM_SHARDED(2);
  else if (width == 4) RNB_LAUNCH_ADAM_SHARDED(4);
  else RNB_LAUNCH_ADAM_SHARDED(RNB_MAX_PEERS);
#undef RNB_LAUNCH_ADAM_SHARDED
  if (flag_peers != nullptr) peer_wait_kernel<<<1, 32, 0, as_stream(stream)>>>(P.flags[rank], world, kFlagDone, barrier_value, status);
  return check_launch(what);
}

int rnb_engine_adam_sharded(int rank, int world, const float* const* grads_peers, void* const* params_f16_peers, float* params_f32,
                            float* exp_avg, float* exp_avg_sq, int64_t n, int64_t n_sharded, float* zero_grads, float lr, float beta1,
                            float beta2, float eps, int64_t step, float grad_scale, rnb_stream_t stream) {
  return launch_adam_sharded(rank, world, grads_peers, params_f16_peers, params_f32, exp_avg, exp_avg_sq, n, n_sharded, zero_grads, lr,
                             beta1, beta2, eps, step, grad_scale, -1, nullptr, nullptr, 0, stream, "rnb_engine_adam_sharded");
}

int rnb_engine_adam_sharded_guarded(int rank, int world, const float* const* grads_peers, void* const* params_f16_peers,
                                    float* params_f32, float* exp_avg, float* exp_avg_sq, int64_t n, int64_t n_sharded,
                                    float* zero_grads, float lr, float beta1, float beta2, float eps, float grad_scale,
                                    int32_t* status, void* const* flag_peers, int32_t barrier_value, rnb_stream_t stream) {
  const char* what = "rnb_engine_adam_sharded_guarded";
  RNB_REQUIRE(status != nullptr, what);
  return launch_adam_sharded(rank, world, grads_peers, params_f16_peers, params_f32, exp_avg, exp_avg_sq, n, n_sharded, zero_grads, lr,
                             beta1, beta2, eps, 1, grad_scale, n, status, flag_peers, barrier_value, stream, what);
}

}  // extern "C"
