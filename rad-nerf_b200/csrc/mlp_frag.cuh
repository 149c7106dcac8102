// mlp_frag.cuh -- warp-level building blocks of the 64-wide fp16 MLPs (mma.sync.m16n8k16 fragments):
// weight staging, one dense layer on a warp's 32 rows, accumulator->A-fragment chaining, fragment
// loads/stores and the per-tile weight-gradient product.  Shared by mlp.cu (operator level) and
// field.cu (fused field kernels).
#pragma once
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


}  // namespace rnb
