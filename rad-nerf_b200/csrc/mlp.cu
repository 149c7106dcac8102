// mlp.cu -- 64-wide bias-free fp16 MLPs (tinycudann FullyFusedMLP semantics) on tensor cores.
//
// Replaces tcnn.Network{FullyFusedMLP} as instantiated by the reference
// (models/networks.py:269-289 geo_net_i / rgb_net_i, :1075-1085 Ray_Gate.encoder):
// weights are row-major (out,in) matrices [first | hidden... | last], fp16 compute copy of an
// fp32 master; ReLU between layers; output activation None or Sigmoid; fp32 accumulation;
// every layer output rounded to fp16.
//
// This file is the operator-level (unfused) implementation: one CTA = 4 warps = a tile of
// 128 rows, 32 rows per warp.  The layer chain runs ENTIRELY IN REGISTERS: the fp32
// accumulator fragments of layer l (mma.m16n8k16) are ReLU'd, packed to fp16 and are, by
// construction of the fragment layouts, exactly the A fragments of layer l+1 -- activations
// never touch shared memory in the forward pass; only the (<= 28 KB) weights live in smem.
// Backward: dgrad chains the same way through W^T; wgrad stages dZ_l and the layer inputs of
// the 128-row tile in smem, reads them with ldmatrix.trans and reduces the per-tile
// dW slabs into the fp32 gradient with 8-byte vector reductions.
// (The tcgen05/TMEM version of the fused field kernel lives in field_tc.cu.)
#include "mlp_frag.cuh"

namespace rnb {

// ---- forward ------------------------------------------------------------------------------
template <int IN_W, int OUT_W, int N_HIDDEN, int ACT>
__global__ void __launch_bounds__(128) mlp_fwd_kernel(const __half* __restrict__ weights, const __half* __restrict__ X,
                                                      int64_t n, __half* __restrict__ Y, __half* __restrict__ hidden) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __half* w_in = reinterpret_cast<__half*>(smem_raw);                  // [64][IN_W+8]
  __half* w_hid = w_in + 64 * (IN_W + 8);                              // (N_HIDDEN-1) x [64][72]
  __half* w_out = w_hid + (N_HIDDEN - 1) * 64 * 72;                    // [OUT_W][72]
  load_weight_smem(w_in, weights, 64, IN_W);
  for (int l = 0; l < N_HIDDEN - 1; l++) load_weight_smem(w_hid + l * 64 * 72, weights + 64 * IN_W + l * 64 * 64, 64, 64);
  load_weight_smem(w_out, weights + 64 * IN_W + (N_HIDDEN - 1) * 64 * 64, OUT_W, 64);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int64_t n_tiles = (n + kTileRows - 1) / kTileRows;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t row0 = tile * kTileRows + warp * 32;
    if (row0 >= n) continue;
    uint32_t a[2][4][4];
    float acc[2][8][4];
    load_afrag_global<IN_W / 16>(X, IN_W, row0, n, a, g, t);
    warp_layer<IN_W / 16, 8>(a, w_in, acc, g, t);
    acc_to_afrag<true>(acc, a);
    if (hidden) store_afrag_global<4>(hidden, 64, row0, n, a, g, t);
#pragma unroll
    for (int l = 0; l < N_HIDDEN - 1; l++) {
      warp_layer<4, 8>(a, w_hid + l * 64 * 72, acc, g, t);
      acc_to_afrag<true>(acc, a);
      if (hidden) store_afrag_global<4>(hidden + (int64_t)(l + 1) * n * 64, 64, row0, n, a, g, t);
    }
    warp_layer<4, OUT_W / 8>(a, w_out, acc, g, t);
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
      const int64_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
#pragma unroll
      for (int nt = 0; nt < OUT_W / 8; nt++) {
        float v0 = acc[mt][nt][0], v1 = acc[mt][nt][1], v2 = acc[mt][nt][2], v3 = acc[mt][nt][3];
        if (ACT == 1) {
          v0 = 1.f / (1.f + __expf(-v0)); v1 = 1.f / (1.f + __expf(-v1));
          v2 = 1.f / (1.f + __expf(-v2)); v3 = 1.f / (1.f + __expf(-v3));
        }
        const int c = nt * 8 + 2 * t;
        if (r_lo < n) *reinterpret_cast<uint32_t*>(Y + r_lo * OUT_W + c) = pack_h2(v0, v1);
        if (r_hi < n) *reinterpret_cast<uint32_t*>(Y + r_hi * OUT_W + c) = pack_h2(v2, v3);
      }
    }
  }
}

// ---- backward -----------------------------------------------------------------------------
template <int IN_W, int OUT_W, int N_HIDDEN, int ACT>
struct MlpBwdSmem {
  static constexpr int wt_out = 64 * (OUT_W + 8);               // W_L^T : [64][OUT_W+8]
  static constexpr int wt_hid = (N_HIDDEN - 1) * 64 * 72;       // W_l^T : [64][72]
  static constexpr int wt_in = IN_W * 72;                       // W_0^T : [IN_W][72]
  static constexpr int dz_hid = N_HIDDEN * kTileRows * 72;      // dZ_0..dZ_{L-1}
  static constexpr int dz_out = kTileRows * (OUT_W + 8);        // dZ_L
  static constexpr int in_x = kTileRows * (IN_W + 8);           // In_0 = X
  static constexpr int in_hid = N_HIDDEN * kTileRows * 72;      // In_l = H_l
  static constexpr int total_halves = wt_out + wt_hid + wt_in + dz_hid + dz_out + in_x + in_hid;
  static constexpr int bytes = total_halves * 2;
};

template <int IN_W, int OUT_W, int N_HIDDEN, int ACT>
__global__ void __launch_bounds__(128) mlp_bwd_kernel(const __half* __restrict__ weights, const __half* __restrict__ X,
                                                      const __half* __restrict__ hidden, const __half* __restrict__ Y,
                                                      const __half* __restrict__ dY, int64_t n, __half* __restrict__ dX,
                                                      float* __restrict__ gw, float gscale) {
  using S = MlpBwdSmem<IN_W, OUT_W, N_HIDDEN, ACT>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __half* wt_out = reinterpret_cast<__half*>(smem_raw);
  __half* wt_hid = wt_out + S::wt_out;
  __half* wt_in = wt_hid + S::wt_hid;
  __half* dz_hid = wt_in + S::wt_in;
  __half* dz_out = dz_hid + S::dz_hid;
  __half* in_x = dz_out + S::dz_out;
  __half* in_hid = in_x + S::in_x;
  const __half* W0 = weights;
  const __half* WL = weights + 64 * IN_W + (N_HIDDEN - 1) * 64 * 64;
  load_weight_smem_T(wt_out, WL, OUT_W, 64);
  for (int l = 0; l < N_HIDDEN - 1; l++) load_weight_smem_T(wt_hid + l * 64 * 72, weights + 64 * IN_W + l * 64 * 64, 64, 64);
  if (dX) load_weight_smem_T(wt_in, W0, 64, IN_W);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int64_t n_tiles = (n + kTileRows - 1) / kTileRows;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t row0 = tile * kTileRows + warp * 32;
    uint32_t a[2][4][4], h[2][4][4];
    float acc[2][8][4];
    // ---- phase A: dgrad chain on this warp's 32 rows (rows >= n contribute zeros) ----
    load_afrag_global<OUT_W / 16>(dY, OUT_W, row0, n, a, g, t);
    if (ACT == 1) {  // sigmoid': dZ = dY * y * (1 - y)
      load_afrag_global<OUT_W / 16>(Y, OUT_W, row0, n, h, g, t);
#pragma unroll
      for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int kc = 0; kc < OUT_W / 16; kc++)
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const float2 d = unpack_h2(a[mt][kc][j]), y = unpack_h2(h[mt][kc][j]);
            a[mt][kc][j] = pack_h2(d.x * y.x * (1.f - y.x), d.y * y.y * (1.f - y.y));
          }
    }
    store_afrag_smem<OUT_W / 16>(dz_out, OUT_W + 8, warp * 32, a, g, t);
    // X tile -> smem (In_0)
    load_afrag_global<IN_W / 16>(X, IN_W, row0, n, h, g, t);
    store_afrag_smem<IN_W / 16>(in_x, IN_W + 8, warp * 32, h, g, t);
#pragma unroll
    for (int l = N_HIDDEN; l >= 1; l--) {
      if (l == N_HIDDEN) warp_layer<OUT_W / 16, 8>(a, wt_out, acc, g, t);
      else warp_layer<4, 8>(a, wt_hid + (l - 1) * 64 * 72, acc, g, t);
      load_afrag_global<4>(hidden + (int64_t)(l - 1) * n * 64, 64, row0, n, h, g, t);
      store_afrag_smem<4>(in_hid + (l - 1) * kTileRows * 72, 72, warp * 32, h, g, t);
#pragma unroll
      for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int kc = 0; kc < 4; kc++)
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const float2 hv = unpack_h2(h[mt][kc][j]);
            const float* c = acc[mt][2 * kc + (j >> 1)];
            const float d0 = c[(j & 1) * 2], d1 = c[(j & 1) * 2 + 1];
            a[mt][kc][j] = pack_h2(hv.x > 0.f ? d0 : 0.f, hv.y > 0.f ? d1 : 0.f);
          }
      store_afrag_smem<4>(dz_hid + (l - 1) * kTileRows * 72, 72, warp * 32, a, g, t);
    }
    if (dX) {
      warp_layer<4, IN_W / 8>(a, wt_in, acc, g, t);
#pragma unroll
      for (int mt = 0; mt < 2; mt++) {
        const int64_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
#pragma unroll
        for (int nt = 0; nt < IN_W / 8; nt++) {
          const int c = nt * 8 + 2 * t;
          if (r_lo < n) *reinterpret_cast<uint32_t*>(dX + r_lo * IN_W + c) = pack_h2(acc[mt][nt][0], acc[mt][nt][1]);
          if (r_hi < n) *reinterpret_cast<uint32_t*>(dX + r_hi * IN_W + c) = pack_h2(acc[mt][nt][2], acc[mt][nt][3]);
        }
      }
    }
    __syncthreads();
    // ---- phase B: weight gradients of the tile ----
    if (gw) {
      tile_wgrad<64, IN_W>(dz_hid, in_x, gw, gscale, warp, lane);
#pragma unroll
      for (int l = 1; l < N_HIDDEN; l++)
        tile_wgrad<64, 64>(dz_hid + l * kTileRows * 72, in_hid + (l - 1) * kTileRows * 72,
                           gw + 64 * IN_W + (l - 1) * 64 * 64, gscale, warp, lane);
      tile_wgrad<OUT_W, 64>(dz_out, in_hid + (N_HIDDEN - 1) * kTileRows * 72,
                            gw + 64 * IN_W + (N_HIDDEN - 1) * 64 * 64, gscale, warp, lane);
    }
    __syncthreads();
  }
}

template <int IN_W, int OUT_W, int N_HIDDEN, int ACT>
static int launch_fwd(const __half* w, const __half* x, int64_t n, __half* y, __half* hidden, cudaStream_t st) {
  const int smem = (64 * (IN_W + 8) + (N_HIDDEN - 1) * 64 * 72 + OUT_W * 72) * 2;
  auto k = mlp_fwd_kernel<IN_W, OUT_W, N_HIDDEN, ACT>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int64_t tiles = (n + kTileRows - 1) / kTileRows;
  const int grid = (int)(tiles < (int64_t)kNumSMs * 8 ? tiles : (int64_t)kNumSMs * 8);
  k<<<grid, 128, smem, st>>>(w, x, n, y, hidden);
  return check_launch("rnb_mlp_fwd");
}

template <int IN_W, int OUT_W, int N_HIDDEN, int ACT>
static int launch_bwd(const __half* w, const __half* x, const __half* hidden, const __half* y, const __half* dy,
                      int64_t n, __half* dx, float* gw, float gscale, cudaStream_t st) {
  using S = MlpBwdSmem<IN_W, OUT_W, N_HIDDEN, ACT>;
  auto k = mlp_bwd_kernel<IN_W, OUT_W, N_HIDDEN, ACT>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, S::bytes);
  int64_t tiles = (n + kTileRows - 1) / kTileRows;
  const int per_sm = S::bytes > 110 * 1024 ? 1 : 2;
  const int grid = (int)(tiles < (int64_t)kNumSMs * per_sm ? tiles : (int64_t)kNumSMs * per_sm);
  k<<<grid, 128, S::bytes, st>>>(w, x, hidden, y, dy, n, dx, gw, gscale);
  return check_launch("rnb_mlp_bwd");
}

}  // namespace rnb

using namespace rnb;

#define RNB_MLP_DISPATCH(CALL)                                                                         \
  if (d->in_width == 32 && d->out_width == 32 && d->n_hidden == 1 && d->out_activation == 0) { return CALL(32, 32, 1, 0); } \
  if (d->in_width == 32 && d->out_width == 16 && d->n_hidden == 2 && d->out_activation == 1) { return CALL(32, 16, 2, 1); } \
  if (d->in_width == 32 && d->out_width == 16 && d->n_hidden == 2 && d->out_activation == 0) { return CALL(32, 16, 2, 0); } \
  if (d->in_width == 16 && d->out_width == 16 && d->n_hidden == 4 && d->out_activation == 0) { return CALL(16, 16, 4, 0); } \
  if (d->in_width == 32 && d->out_width == 16 && d->n_hidden == 1 && d->out_activation == 0) { return CALL(32, 16, 1, 0); } \
  if (d->in_width == 16 && d->out_width == 16 && d->n_hidden == 2 && d->out_activation == 0) { return CALL(16, 16, 2, 0); }

extern "C" {

int rnb_mlp_fwd(const rnb_mlp_desc* d, const void* weights_f16, const void* in_f16, int64_t n, void* out_f16,
                void* hidden_f16, rnb_stream_t stream) {
  const char* what = "rnb_mlp_fwd";
  RNB_REQUIRE(d != nullptr && n >= 0, what);
  if (n == 0) return RNB_OK;
  RNB_REQUIRE(weights_f16 && in_f16 && out_f16, what);
#define FWD(I, O, H, A) launch_fwd<I, O, H, A>((const __half*)weights_f16, (const __half*)in_f16, n, (__half*)out_f16, (__half*)hidden_f16, as_stream(stream))
  RNB_MLP_DISPATCH(FWD)
#undef FWD
  set_error("%s: unsupported MLP shape in=%d out=%d hidden=%d act=%d", what, d->in_width, d->out_width, d->n_hidden, d->out_activation);
  return RNB_ERR_BADARG;
}

int rnb_mlp_bwd(const rnb_mlp_desc* d, const void* weights_f16, const void* in_f16, const void* hidden_f16,
                const void* out_f16, const void* dL_dout_f16, int64_t n, void* dL_din_f16, float* grad_weights,
                float grad_scale, rnb_stream_t stream) {
  const char* what = "rnb_mlp_bwd";
  RNB_REQUIRE(d != nullptr && n >= 0, what);
  if (n == 0) return RNB_OK;
  RNB_REQUIRE(weights_f16 && in_f16 && hidden_f16 && dL_dout_f16, what);
  RNB_REQUIRE(d->out_activation == 0 || out_f16 != nullptr, what);
#define BWD(I, O, H, A) launch_bwd<I, O, H, A>((const __half*)weights_f16, (const __half*)in_f16, (const __half*)hidden_f16, (const __half*)out_f16, (const __half*)dL_dout_f16, n, (__half*)dL_din_f16, grad_weights, grad_scale, as_stream(stream))
  RNB_MLP_DISPATCH(BWD)
#undef BWD
  set_error("%s: unsupported MLP shape in=%d out=%d hidden=%d act=%d", what, d->in_width, d->out_width, d->n_hidden, d->out_activation);
  return RNB_ERR_BADARG;
}

}  // extern "C"
