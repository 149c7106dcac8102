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
#include "common.cuh"

namespace rnb {

constexpr int kMlpWidth = 64;
constexpr int kTileRows = 128;

__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  const __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_h2(uint32_t v) {
  return __half22float2(*reinterpret_cast<const __half2*>(&v));
}

__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem_ptr) {
  const uint32_t addr = (uint32_t)__cvta_generic_to_shared(smem_ptr);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}

// Copies a row-major (rows, cols) fp16 matrix from global into smem with row stride cols+8
// (the +8 halves skew makes the 8 rows of a B fragment / ldmatrix land in distinct banks).
__device__ __forceinline__ void load_weight_smem(__half* dst, const __half* __restrict__ src, int rows, int cols) {
  const int stride = cols + 8;
  const int vec_per_row = cols / 8;
  for (int i = threadIdx.x; i < rows * vec_per_row; i += blockDim.x) {
    const int r = i / vec_per_row, v = i % vec_per_row;
    *reinterpret_cast<uint4*>(dst + r * stride + v * 8) = __ldg(reinterpret_cast<const uint4*>(src + r * cols) + v);
  }
}
// Same, transposed: dst[c][r] = src[r][c], dst row stride rows+8.
__device__ __forceinline__ void load_weight_smem_T(__half* dst, const __half* __restrict__ src, int rows, int cols) {
  const int stride = rows + 8;
  for (int i = threadIdx.x; i < rows * cols; i += blockDim.x) {
    const int r = i / cols, c = i % cols;
    dst[c * stride + r] = src[i];
  }
}

// One dense layer on a warp's 32 rows: acc[mt][nt] (+)= A[mt] (32 x K, fragments) * W^T,
// W in smem as [N][K+8].  KC = K/16 k-chunks, NT = N/8 n-tiles.
template <int KC, int NT>
__device__ __forceinline__ void warp_layer(const uint32_t (&a)[2][4][4], const __half* __restrict__ w_smem,
                                           float (&acc)[2][8][4], int g, int t) {
  constexpr int stride = KC * 16 + 8;
#pragma unroll
  for (int nt = 0; nt < NT; nt++) {
#pragma unroll
    for (int mt = 0; mt < 2; mt++) { acc[mt][nt][0] = acc[mt][nt][1] = acc[mt][nt][2] = acc[mt][nt][3] = 0.f; }
#pragma unroll
    for (int kc = 0; kc < KC; kc++) {
      const __half* wp = w_smem + (nt * 8 + g) * stride + kc * 16 + 2 * t;
      const uint32_t b0 = *reinterpret_cast<const uint32_t*>(wp);
      const uint32_t b1 = *reinterpret_cast<const uint32_t*>(wp + 8);
      mma16816(acc[0][nt], a[0][kc], b0, b1);
      mma16816(acc[1][nt], a[1][kc], b0, b1);
    }
  }
}

// accumulator fragments (32 x 64 fp32) -> A fragments of the next layer (fp16), optional ReLU
template <bool kRelu>
__device__ __forceinline__ void acc_to_afrag(const float (&acc)[2][8][4], uint32_t (&a)[2][4][4]) {
#pragma unroll
  for (int mt = 0; mt < 2; mt++)
#pragma unroll
    for (int kc = 0; kc < 4; kc++) {
      const float* lo = acc[mt][2 * kc];
      const float* hi = acc[mt][2 * kc + 1];
      if (kRelu) {
        a[mt][kc][0] = pack_h2(fmaxf(lo[0], 0.f), fmaxf(lo[1], 0.f));
        a[mt][kc][1] = pack_h2(fmaxf(lo[2], 0.f), fmaxf(lo[3], 0.f));
        a[mt][kc][2] = pack_h2(fmaxf(hi[0], 0.f), fmaxf(hi[1], 0.f));
        a[mt][kc][3] = pack_h2(fmaxf(hi[2], 0.f), fmaxf(hi[3], 0.f));
      } else {
        a[mt][kc][0] = pack_h2(lo[0], lo[1]); a[mt][kc][1] = pack_h2(lo[2], lo[3]);
        a[mt][kc][2] = pack_h2(hi[0], hi[1]); a[mt][kc][3] = pack_h2(hi[2], hi[3]);
      }
    }
}

// A fragments of a warp's 32 rows from a row-major global matrix X (n, W) fp16.
template <int KC>
__device__ __forceinline__ void load_afrag_global(const __half* __restrict__ X, int W, int64_t row0, int64_t n,
                                                  uint32_t (&a)[2][4][4], int g, int t) {
#pragma unroll
  for (int mt = 0; mt < 2; mt++) {
    const int64_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
#pragma unroll
    for (int kc = 0; kc < KC; kc++) {
      const int c = kc * 16 + 2 * t;
      a[mt][kc][0] = r_lo < n ? __ldg(reinterpret_cast<const uint32_t*>(X + r_lo * W + c)) : 0u;
      a[mt][kc][1] = r_hi < n ? __ldg(reinterpret_cast<const uint32_t*>(X + r_hi * W + c)) : 0u;
      a[mt][kc][2] = r_lo < n ? __ldg(reinterpret_cast<const uint32_t*>(X + r_lo * W + c + 8)) : 0u;
      a[mt][kc][3] = r_hi < n ? __ldg(reinterpret_cast<const uint32_t*>(X + r_hi * W + c + 8)) : 0u;
    }
  }
}

// stores A-layout fragments (fp16 pairs) of a warp's 32 rows to a row-major global matrix (n, W)
template <int KC>
__device__ __forceinline__ void store_afrag_global(__half* __restrict__ Y, int W, int64_t row0, int64_t n,
                                                   const uint32_t (&a)[2][4][4], int g, int t) {
#pragma unroll
  for (int mt = 0; mt < 2; mt++) {
    const int64_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
#pragma unroll
    for (int kc = 0; kc < KC; kc++) {
      const int c = kc * 16 + 2 * t;
      if (r_lo < n) { *reinterpret_cast<uint32_t*>(Y + r_lo * W + c) = a[mt][kc][0]; *reinterpret_cast<uint32_t*>(Y + r_lo * W + c + 8) = a[mt][kc][2]; }
      if (r_hi < n) { *reinterpret_cast<uint32_t*>(Y + r_hi * W + c) = a[mt][kc][1]; *reinterpret_cast<uint32_t*>(Y + r_hi * W + c + 8) = a[mt][kc][3]; }
    }
  }
}

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
template <int KC>
__device__ __forceinline__ void store_afrag_smem(__half* tile, int stride, int row0, const uint32_t (&a)[2][4][4], int g, int t) {
#pragma unroll
  for (int mt = 0; mt < 2; mt++) {
    const int r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
#pragma unroll
    for (int kc = 0; kc < KC; kc++) {
      const int c = kc * 16 + 2 * t;
      *reinterpret_cast<uint32_t*>(tile + r_lo * stride + c) = a[mt][kc][0];
      *reinterpret_cast<uint32_t*>(tile + r_hi * stride + c) = a[mt][kc][1];
      *reinterpret_cast<uint32_t*>(tile + r_lo * stride + c + 8) = a[mt][kc][2];
      *reinterpret_cast<uint32_t*>(tile + r_hi * stride + c + 8) = a[mt][kc][3];
    }
  }
}

__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

// dW[O][I] (+)= dZ^T (O x 128) * In (128 x I) for the CTA's 128-row tile.  Warp w owns the
// 16-row slab o in [16w, 16w+16) (idle when 16w >= O).  dZ tile: [128][O+8]; In tile: [128][I+8].
template <int O, int I>
__device__ __forceinline__ void tile_wgrad(const __half* dz, const __half* in, float* __restrict__ gw, float gscale,
                                           int warp, int lane) {
  if (warp * 16 >= O) return;
  constexpr int sdz = O + 8, sin_ = I + 8;
  float acc[I / 8][4];
#pragma unroll
  for (int nt = 0; nt < I / 8; nt++) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
  const int q = lane >> 3, r = lane & 7;
#pragma unroll 2
  for (int k0 = 0; k0 < kTileRows; k0 += 16) {
    uint32_t a[4];
    // A[m=o][k=row] = dz[row][o]: matrices (k lo, m lo), (k lo, m hi), (k hi, m lo), (k hi, m hi)
    ldmatrix_x4_trans(a, dz + (k0 + (q >> 1) * 8 + r) * sdz + warp * 16 + (q & 1) * 8);
#pragma unroll
    for (int np = 0; np < I / 16; np++) {
      uint32_t b[4];
      // B[k=row][n=i] = in[row][i]: matrices (k lo, n0), (k hi, n0), (k lo, n0+8), (k hi, n0+8)
      ldmatrix_x4_trans(b, in + (k0 + (q & 1) * 8 + r) * sin_ + np * 16 + (q >> 1) * 8);
      mma16816(acc[2 * np], a, b[0], b[1]);
      mma16816(acc[2 * np + 1], a, b[2], b[3]);
    }
  }
  const int g = lane >> 2, t = lane & 3;
  const int o_lo = warp * 16 + g, o_hi = o_lo + 8;
#pragma unroll
  for (int nt = 0; nt < I / 8; nt++) {
    const int c = nt * 8 + 2 * t;
    red_add_v2(gw + o_lo * I + c, acc[nt][0] * gscale, acc[nt][1] * gscale);
    red_add_v2(gw + o_hi * I + c, acc[nt][2] * gscale, acc[nt][3] * gscale);
  }
}

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
