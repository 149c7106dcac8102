// field.cu -- the fused field of the K sub-NeRFs (forward and backward) on a packed sample list.
//
// Replaces, in ONE kernel per direction, the reference's per-sub-NeRF chain
//   xyz normalise -> tcnn HashGrid -> geo_net_i (32->64->17) -> TruncExp -> dir normalise ->
//   tcnn SH4 -> cat -> rgb_net_i (32->64->64->3, sigmoid)        (models/networks.py:291-328)
// and its backward (tcnn FullyFusedMLP backward + kernel_grid_backward), for ALL K passes of
// a training step: samples of the K passes are packed back to back (128-aligned), a 128-sample
// tile belongs to exactly one pass and picks that pass's weights.
//
// Mapping (sm_100a): CTA = 4 warps = one 128-sample tile; a warp owns 32 samples end to end.
//   phase A (lane = sample): x = o + t*d, normalise, 16 levels x 8 gathers from the L2-resident
//     fp16 table (independent loads, dozens in flight per thread), SH4 of the unit direction;
//     the 48 fp16 inputs of the two MLPs go to a warp-private smem tile, never to HBM;
//   phase B (lane = mma fragment): the five dense layers chained in registers -- the fp32
//     accumulator fragments of a layer are, after ReLU + fp16 rounding, exactly the A fragments
//     of the next one (mma.sync.m16n8k16; weights of the pass staged once per CTA in smem).
// The geo output neurons are permuted while staging W (feature rows first, sigma row at 16) so
// that the geometry feature lands directly in the K-chunk of the rgb net's input.
// Rounding points equal the unfused operators' (fp16 features / activations / outputs, fp32
// accumulation), so both paths agree to fp16 resolution.
#include "hashgrid.cuh"
#include "mlp_frag.cuh"

namespace rnb {

constexpr int kStageStride = 56;   // halves per row of the warp-private input tile: 32 enc | 16 sh | 8 pad

struct FieldArgs {
  GridLevels g;
  float xmin[3], xinv_den[3];   // xyz_min and (xyz_max - xyz_min)
  const __half2* table;
  const __half* geo_w[RNB_MAX_K];
  const __half* rgb_w[RNB_MAX_K];
  int K;
  const float* rays_o; const float* rays_d;
  const int32_t* ray_id; const float* ts; const int32_t* seg;
  int64_t capacity;
};

// sample -> normalised position in [0,1]^3 (networks.py:300-301) and ray direction
struct SamplePoint { float x, y, z, dx, dy, dz; };

__device__ __forceinline__ SamplePoint load_sample(const FieldArgs& A, int64_t s) {
  SamplePoint p;
  const int r = __ldg(A.ray_id + s);
  const float t = __ldg(A.ts + s);
  const float ox = __ldg(A.rays_o + 3 * r), oy = __ldg(A.rays_o + 3 * r + 1), oz = __ldg(A.rays_o + 3 * r + 2);
  p.dx = __ldg(A.rays_d + 3 * r); p.dy = __ldg(A.rays_d + 3 * r + 1); p.dz = __ldg(A.rays_d + 3 * r + 2);
  // the marcher's own expression (one FMA per axis) -- identical to the xyzs the operator path stores
  const float wx = __fmaf_rn(p.dx, t, ox), wy = __fmaf_rn(p.dy, t, oy), wz = __fmaf_rn(p.dz, t, oz);
  p.x = fminf(fmaxf(__fdiv_rn(wx - A.xmin[0], A.xinv_den[0]), 0.f), 1.f);
  p.y = fminf(fmaxf(__fdiv_rn(wy - A.xmin[1], A.xinv_den[1]), 0.f), 1.f);
  p.z = fminf(fmaxf(__fdiv_rn(wz - A.xmin[2], A.xinv_den[2]), 0.f), 1.f);
  return p;
}

// one level of the multiresolution grid -> packed half2 (same arithmetic as hashgrid_fwd_kernel)
__device__ __forceinline__ uint32_t encode_level(const FieldArgs& A, int l, float x, float y, float z) {
  const uint32_t res = A.g.res[l], off = A.g.offset[l], size = A.g.offset[l + 1] - off;
  const bool dense = level_is_dense(res, size);
  const Corner c = locate(x, y, z, A.g.scale[l]);
  const __half2* t = A.table + off;
  __half2 v[8];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const uint32_t idx = grid_index(c.cx + (k & 1), c.cy + ((k >> 1) & 1), c.cz + (k >> 2), res, size, dense);
    v[k] = __ldg(t + idx);
  }
  float a0 = 0.f, a1 = 0.f;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const float w = ((k & 1) ? c.wx : 1.f - c.wx) * (((k >> 1) & 1) ? c.wy : 1.f - c.wy) * ((k >> 2) ? c.wz : 1.f - c.wz);
    const float2 f = __half22float2(v[k]);
    a0 = fmaf(w, f.x, a0); a1 = fmaf(w, f.y, a1);
  }
  return pack_h2(a0, a1);
}

// geo_net output neuron held by smem row j: features (neurons 1..16) first, sigma (neuron 0) at 16
__device__ __forceinline__ int geo_out_perm(int j) { return j < 16 ? j + 1 : (j == 16 ? 0 : j); }

// W (rows, cols) row-major in global -> smem [rows][cols+8]; row j of smem = row perm(j) of W
template <bool kPerm>
__device__ __forceinline__ void stage_weight(__half* dst, const __half* __restrict__ src, int rows, int cols) {
  const int stride = cols + 8, vec_per_row = cols / 8;
  for (int i = threadIdx.x; i < rows * vec_per_row; i += blockDim.x) {
    const int r = i / vec_per_row, v = i % vec_per_row;
    const int sr = kPerm ? geo_out_perm(r) : r;
    *reinterpret_cast<uint4*>(dst + r * stride + v * 8) = __ldg(reinterpret_cast<const uint4*>(src + sr * cols) + v);
  }
}

struct FwdSmem {
  static constexpr int g1 = 0;                      // [64][40]
  static constexpr int g2 = g1 + 64 * 40;           // [32][72]  (rows permuted)
  static constexpr int r1 = g2 + 32 * 72;           // [64][40]
  static constexpr int r2 = r1 + 64 * 40;           // [64][72]
  static constexpr int r3 = r2 + 64 * 72;           // [16][72]
  static constexpr int stage = r3 + 16 * 72;        // 4 x [32][56]
  static constexpr int total = stage + 4 * 32 * kStageStride;
};

// pass of a tile: the i with seg[i] <= tile_base < seg[i+1]
__device__ __forceinline__ int pass_of(const int32_t* __restrict__ seg, int K, int64_t tile_base) {
  int i = 0;
  while (i + 1 < K && tile_base >= (int64_t)__ldg(seg + i + 1)) i++;
  return i;
}

template <bool kSave>
__global__ void __launch_bounds__(128) field_fwd_kernel(FieldArgs A, __half* __restrict__ s_enc, __half* __restrict__ s_rgb_in,
                                                        __half* __restrict__ s_geo_h, __half* __restrict__ s_rgb_h1,
                                                        __half* __restrict__ s_rgb_h2, __half* __restrict__ s_sraw,
                                                        float* __restrict__ sigmas, __half* __restrict__ rgb4) {
  __shared__ __align__(16) __half sm[FwdSmem::total];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  __half* stage = sm + FwdSmem::stage + warp * 32 * kStageStride;
  const int K = A.K;
  int64_t end = __ldg(A.seg + K);
  if (end > A.capacity) end = A.capacity;
  const int64_t n_tiles = end / kTileRows;
  int cur_pass = -1;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t tile_base = tile * kTileRows;
    const int pass = pass_of(A.seg, K, tile_base);
    const int64_t pass_end = (int64_t)__ldg(A.seg + pass) + __ldg(A.seg + K + 1 + pass);
    if (tile_base >= pass_end) continue;          // CTA-uniform: the tile is alignment padding
    if (pass != cur_pass) {                       // CTA-uniform
      __syncthreads();
      const __half* gw = A.geo_w[pass];
      const __half* rw = A.rgb_w[pass];
      stage_weight<false>(sm + FwdSmem::g1, gw, 64, 32);
      stage_weight<true>(sm + FwdSmem::g2, gw + 64 * 32, 32, 64);
      stage_weight<false>(sm + FwdSmem::r1, rw, 64, 32);
      stage_weight<false>(sm + FwdSmem::r2, rw + 64 * 32, 64, 64);
      stage_weight<false>(sm + FwdSmem::r3, rw + 64 * 32 + 64 * 64, 16, 64);
      __syncthreads();
      cur_pass = pass;
    }
    const int64_t row0 = tile_base + warp * 32;
    if (row0 >= pass_end) continue;               // warp-uniform; no block barrier below
    // ---- phase A: lane = sample ------------------------------------------------------------
    {
      const int64_t s = row0 + lane;
      const bool valid = s < pass_end;
      uint4* srow = reinterpret_cast<uint4*>(stage + lane * kStageStride);
      if (valid) {
        const SamplePoint p = load_sample(A, s);
#pragma unroll
        for (int l0 = 0; l0 < 16; l0 += 4) {
          uint4 q;
          q.x = encode_level(A, l0, p.x, p.y, p.z); q.y = encode_level(A, l0 + 1, p.x, p.y, p.z);
          q.z = encode_level(A, l0 + 2, p.x, p.y, p.z); q.w = encode_level(A, l0 + 3, p.x, p.y, p.z);
          srow[l0 >> 2] = q;
          if (kSave) reinterpret_cast<uint4*>(s_enc + s * 32)[l0 >> 2] = q;
        }
        // direction: d/|d| -> (d+1)/2 -> (tcnn) 2u-1 -> SH4   (networks.py:322-323)
        const float nrm = sqrtf(p.dx * p.dx + p.dy * p.dy + p.dz * p.dz);
        const float ux = (__fdiv_rn(p.dx, nrm) + 1.f) * 0.5f, uy = (__fdiv_rn(p.dy, nrm) + 1.f) * 0.5f, uz = (__fdiv_rn(p.dz, nrm) + 1.f) * 0.5f;
        float o[16];
        sh4_eval(ux * 2.f - 1.f, uy * 2.f - 1.f, uz * 2.f - 1.f, o);
        uint4 q0 = make_uint4(pack_h2(o[0], o[1]), pack_h2(o[2], o[3]), pack_h2(o[4], o[5]), pack_h2(o[6], o[7]));
        uint4 q1 = make_uint4(pack_h2(o[8], o[9]), pack_h2(o[10], o[11]), pack_h2(o[12], o[13]), pack_h2(o[14], o[15]));
        srow[4] = q0; srow[5] = q1;
        if (kSave) { reinterpret_cast<uint4*>(s_rgb_in + s * 32)[0] = q0; reinterpret_cast<uint4*>(s_rgb_in + s * 32)[1] = q1; }
      } else {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int q = 0; q < 6; q++) srow[q] = z;
      }
    }
    __syncwarp();
    // ---- phase B: lane = fragment ------------------------------------------------------------
    uint32_t a[2][4][4];
    float acc[2][8][4];
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
      for (int kc = 0; kc < 2; kc++) {
        const __half* lo = stage + (mt * 16 + g) * kStageStride + kc * 16 + 2 * t;
        const __half* hi = lo + 8 * kStageStride;
        a[mt][kc][0] = *reinterpret_cast<const uint32_t*>(lo); a[mt][kc][1] = *reinterpret_cast<const uint32_t*>(hi);
        a[mt][kc][2] = *reinterpret_cast<const uint32_t*>(lo + 8); a[mt][kc][3] = *reinterpret_cast<const uint32_t*>(hi + 8);
      }
    // geo layer 1: 32 -> 64, ReLU
    warp_layer<2, 8>(a, sm + FwdSmem::g1, acc, g, t);
    acc_to_afrag<true>(acc, a);
    if (kSave) store_afrag_global<4>(s_geo_h, 64, row0, pass_end, a, g, t);
    // geo layer 2: 64 -> 32 (cols 0..15 = geometry feature, col 16 = raw sigma)
    warp_layer<4, 4>(a, sm + FwdSmem::g2, acc, g, t);
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
      // rgb input K-chunk 1 = geometry feature (fp16-rounded geo output), chunk 0 = SH from smem
      a[mt][1][0] = pack_h2(acc[mt][0][0], acc[mt][0][1]); a[mt][1][1] = pack_h2(acc[mt][0][2], acc[mt][0][3]);
      a[mt][1][2] = pack_h2(acc[mt][1][0], acc[mt][1][1]); a[mt][1][3] = pack_h2(acc[mt][1][2], acc[mt][1][3]);
      const __half* lo = stage + (mt * 16 + g) * kStageStride + 32 + 2 * t;
      const __half* hi = lo + 8 * kStageStride;
      a[mt][0][0] = *reinterpret_cast<const uint32_t*>(lo); a[mt][0][1] = *reinterpret_cast<const uint32_t*>(hi);
      a[mt][0][2] = *reinterpret_cast<const uint32_t*>(lo + 8); a[mt][0][3] = *reinterpret_cast<const uint32_t*>(hi + 8);
      const int64_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
      if (kSave) {
        const int c = 16 + 2 * t;
        if (r_lo < pass_end) { *reinterpret_cast<uint32_t*>(s_rgb_in + r_lo * 32 + c) = a[mt][1][0]; *reinterpret_cast<uint32_t*>(s_rgb_in + r_lo * 32 + c + 8) = a[mt][1][2]; }
        if (r_hi < pass_end) { *reinterpret_cast<uint32_t*>(s_rgb_in + r_hi * 32 + c) = a[mt][1][1]; *reinterpret_cast<uint32_t*>(s_rgb_in + r_hi * 32 + c + 8) = a[mt][1][3]; }
      }
      if (t == 0) {   // column 16 lives in n-tile 2, lanes with t == 0
        const __half raw_lo = __float2half_rn(acc[mt][2][0]), raw_hi = __float2half_rn(acc[mt][2][2]);
        if (r_lo < pass_end) { sigmas[r_lo] = expf(__half2float(raw_lo)); if (kSave) s_sraw[r_lo] = raw_lo; }
        if (r_hi < pass_end) { sigmas[r_hi] = expf(__half2float(raw_hi)); if (kSave) s_sraw[r_hi] = raw_hi; }
      }
    }
    // rgb layer 1: 32 -> 64, ReLU ; layer 2: 64 -> 64, ReLU ; layer 3: 64 -> 16, sigmoid
    warp_layer<2, 8>(a, sm + FwdSmem::r1, acc, g, t);
    acc_to_afrag<true>(acc, a);
    if (kSave) store_afrag_global<4>(s_rgb_h1, 64, row0, pass_end, a, g, t);
    warp_layer<4, 8>(a, sm + FwdSmem::r2, acc, g, t);
    acc_to_afrag<true>(acc, a);
    if (kSave) store_afrag_global<4>(s_rgb_h2, 64, row0, pass_end, a, g, t);
    warp_layer<4, 2>(a, sm + FwdSmem::r3, acc, g, t);
    if (t < 2) {   // columns 0..3 live in n-tile 0, lanes t = 0 (cols 0,1) and t = 1 (cols 2,3)
#pragma unroll
      for (int mt = 0; mt < 2; mt++) {
        const int64_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
        const float v0 = 1.f / (1.f + __expf(-acc[mt][0][0])), v1 = 1.f / (1.f + __expf(-acc[mt][0][1]));
        const float v2 = 1.f / (1.f + __expf(-acc[mt][0][2])), v3 = 1.f / (1.f + __expf(-acc[mt][0][3]));
        if (r_lo < pass_end) *reinterpret_cast<uint32_t*>(rgb4 + r_lo * 4 + 2 * t) = pack_h2(v0, v1);
        if (r_hi < pass_end) *reinterpret_cast<uint32_t*>(rgb4 + r_hi * 4 + 2 * t) = pack_h2(v2, v3);
      }
    }
    __syncwarp();   // the staging tile is rewritten by the next iteration's phase A
  }
}


// ================================ backward ====================================================
// Per 128-sample tile the CTA walks the five layers from the rgb output back to the encoding.
// For layer l it (1) parks dZ_l (registers -> smem) next to the layer's saved input In_l
// (global -> smem), (2) lets warp w accumulate its 16-row slab of dW_l = dZ_l^T In_l over the
// tile's 128 rows on the tensor cores into fp32 accumulators that stay in shared memory for the
// CTA's whole life (no atomics: a slab has one owner; one flush of 11 264 values per pass and CTA
// at the end instead of one per tile), and (3) chains dZ_{l-1} = (dZ_l W_l) * relu'(In_l) in
// registers.  The encoding gradient is transposed through smem (lane = sample again) and
// scattered into the fp32 table gradient with 8-byte vector reductions, aggregated per warp on
// the dense coarse levels.
struct BwdSmem {
  // transposed weights for the dgrad chain: [I][O+8] halves
  static constexpr int wt_r3 = 0;                        // [64][24]
  static constexpr int wt_r2 = wt_r3 + 64 * 24;          // [64][72]
  static constexpr int wt_r1 = wt_r2 + 64 * 72;          // [16][72]  (input columns 16..31 = geometry feature)
  static constexpr int wt_g2 = wt_r1 + 16 * 72;          // [64][40]  (output neurons permuted)
  static constexpr int wt_g1 = wt_g2 + 64 * 40;          // [32][72]
  static constexpr int dz = wt_g1 + 32 * 72;             // [128][<=72]
  static constexpr int in = dz + 128 * 72;               // [128][<=72]
  static constexpr int halves = in + 128 * 72;
  // fp32 weight-gradient accumulators [O][I+4]
  static constexpr int acc_g1 = 0;                       // [64][36]
  static constexpr int acc_g2 = acc_g1 + 64 * 36;        // [32][68]
  static constexpr int acc_r1 = acc_g2 + 32 * 68;        // [64][36]
  static constexpr int acc_r2 = acc_r1 + 64 * 36;        // [64][68]
  static constexpr int acc_r3 = acc_r2 + 64 * 68;        // [16][68]
  static constexpr int floats = acc_r3 + 16 * 68;
  static constexpr int bytes = halves * 2 + floats * 4;
};

struct GradPtrs { float* geo[RNB_MAX_K]; float* rgb[RNB_MAX_K]; };

// transposed staging, dst[c * (rows+8) + r] = W[perm(r)][c0 + c]
template <bool kPerm>
__device__ __forceinline__ void stage_weight_T(__half* dst, const __half* __restrict__ src, int rows, int cols, int c0, int ncols) {
  const int stride = rows + 8;
  for (int i = threadIdx.x; i < rows * ncols; i += blockDim.x) {
    const int r = i / ncols, c = i % ncols;
    const int sr = kPerm ? geo_out_perm(r) : r;
    dst[c * stride + r] = src[sr * cols + c0 + c];
  }
}

// dW slab of warp `warp` (16 output rows) over the tile's 128 rows, accumulated into smem fp32.
template <int O, int I>
__device__ __forceinline__ void tile_wgrad_acc(const __half* dz, const __half* in, float* accum, int warp, int lane) {
  if (warp * 16 >= O) return;
  constexpr int sdz = O + 8, sin_ = I + 8, sacc = I + 4;
  float acc[I / 8][4];
#pragma unroll
  for (int nt = 0; nt < I / 8; nt++) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
  const int q = lane >> 3, r = lane & 7;
#pragma unroll 2
  for (int k0 = 0; k0 < kTileRows; k0 += 16) {
    uint32_t a[4];
    ldmatrix_x4_trans(a, dz + (k0 + (q >> 1) * 8 + r) * sdz + warp * 16 + (q & 1) * 8);
#pragma unroll
    for (int np = 0; np < I / 16; np++) {
      uint32_t b[4];
      ldmatrix_x4_trans(b, in + (k0 + (q & 1) * 8 + r) * sin_ + np * 16 + (q >> 1) * 8);
      mma16816(acc[2 * np], a, b[0], b[1]);
      mma16816(acc[2 * np + 1], a, b[2], b[3]);
    }
  }
  const int g = lane >> 2, t = lane & 3;
  float* p_lo = accum + (warp * 16 + g) * sacc + 2 * t;
  float* p_hi = p_lo + 8 * sacc;
#pragma unroll
  for (int nt = 0; nt < I / 8; nt++) {
    float2 v = *reinterpret_cast<float2*>(p_lo + nt * 8);
    v.x += acc[nt][0]; v.y += acc[nt][1];
    *reinterpret_cast<float2*>(p_lo + nt * 8) = v;
    float2 u = *reinterpret_cast<float2*>(p_hi + nt * 8);
    u.x += acc[nt][2]; u.y += acc[nt][3];
    *reinterpret_cast<float2*>(p_hi + nt * 8) = u;
  }
}

// the warp's 32 rows of a saved activation (cap, W) f16 -> smem tile rows [warp*32, +32), stride W+8
template <int W>
__device__ __forceinline__ void load_in_tile(__half* tile, const __half* __restrict__ X, int64_t row0, int64_t n_end, int warp, int lane) {
  constexpr int vpr = W / 8;
#pragma unroll
  for (int j = 0; j < vpr; j++) {
    const int idx = lane + 32 * j;
    const int r = idx / vpr, v = idx % vpr;
    const int64_t s = row0 + r;
    uint4 q = make_uint4(0u, 0u, 0u, 0u);
    if (s < n_end) q = __ldg(reinterpret_cast<const uint4*>(X + s * W) + v);
    *reinterpret_cast<uint4*>(tile + (warp * 32 + r) * (W + 8) + v * 8) = q;
  }
}

// dZ = dH (accumulators, 32 x 64) masked by relu'(h) with h read from the smem input tile
__device__ __forceinline__ void relu_mask_to_afrag(const float (&acc)[2][8][4], const __half* in_tile72, int warp, int g, int t,
                                                   uint32_t (&a)[2][4][4]) {
#pragma unroll
  for (int mt = 0; mt < 2; mt++)
#pragma unroll
    for (int kc = 0; kc < 4; kc++)
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int row = warp * 32 + mt * 16 + g + (j & 1) * 8;
        const int col = kc * 16 + 2 * t + (j >> 1) * 8;
        const float2 hv = unpack_h2(*reinterpret_cast<const uint32_t*>(in_tile72 + row * 72 + col));
        const float* c = acc[mt][2 * kc + (j >> 1)];
        const float d0 = c[(j & 1) * 2], d1 = c[(j & 1) * 2 + 1];
        a[mt][kc][j] = pack_h2(hv.x > 0.f ? d0 : 0.f, hv.y > 0.f ? d1 : 0.f);
      }
}

__device__ __forceinline__ float sat_h(float v) { return fminf(fmaxf(v, -65504.f), 65504.f); }

template <int O, int I>
__device__ __forceinline__ void flush_accum(float* accum, float* __restrict__ gw, float inv_scale, bool perm) {
  constexpr int sacc = I + 4;
  for (int i = threadIdx.x; i < O * I / 2; i += blockDim.x) {
    const int o = i / (I / 2), c = (i % (I / 2)) * 2;
    float2 v = *reinterpret_cast<float2*>(accum + o * sacc + c);
    *reinterpret_cast<float2*>(accum + o * sacc + c) = make_float2(0.f, 0.f);
    if (v.x != 0.f || v.y != 0.f) red_add_v2(gw + (perm ? geo_out_perm(o) : o) * I + c, v.x * inv_scale, v.y * inv_scale);
  }
}

__global__ void __launch_bounds__(128) field_bwd_kernel(FieldArgs A, const __half* __restrict__ s_enc, const __half* __restrict__ s_rgb_in,
                                                        const __half* __restrict__ s_geo_h, const __half* __restrict__ s_rgb_h1,
                                                        const __half* __restrict__ s_rgb_h2, const __half* __restrict__ s_sraw,
                                                        const __half* __restrict__ rgb4, const float* __restrict__ dsig,
                                                        const float* __restrict__ drgb, float gs, float* __restrict__ grad_table,
                                                        GradPtrs G, int n_aggregate_levels) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __half* sm = reinterpret_cast<__half*>(smem_raw);
  float* accum = reinterpret_cast<float*>(smem_raw + BwdSmem::halves * 2);
  __half* dzbuf = sm + BwdSmem::dz;
  __half* inbuf = sm + BwdSmem::in;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const float inv_gs = 1.0f / gs;
  for (int i = threadIdx.x; i < BwdSmem::floats; i += blockDim.x) accum[i] = 0.f;
  const int K = A.K;
  int64_t end = __ldg(A.seg + K);
  if (end > A.capacity) end = A.capacity;
  const int64_t n_tiles = end / kTileRows;
  int cur_pass = -1;
  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t tile_base = tile * kTileRows;
    const int pass = pass_of(A.seg, K, tile_base);
    const int64_t pass_end = (int64_t)__ldg(A.seg + pass) + __ldg(A.seg + K + 1 + pass);
    if (tile_base >= pass_end) continue;          // CTA-uniform
    if (pass != cur_pass) {                       // CTA-uniform: flush the finished pass, stage the new weights
      __syncthreads();
      if (cur_pass >= 0) {
        float* gg = G.geo[cur_pass]; float* gr = G.rgb[cur_pass];
        flush_accum<64, 32>(accum + BwdSmem::acc_g1, gg, inv_gs, false);
        flush_accum<32, 64>(accum + BwdSmem::acc_g2, gg + 64 * 32, inv_gs, true);
        flush_accum<64, 32>(accum + BwdSmem::acc_r1, gr, inv_gs, false);
        flush_accum<64, 64>(accum + BwdSmem::acc_r2, gr + 64 * 32, inv_gs, false);
        flush_accum<16, 64>(accum + BwdSmem::acc_r3, gr + 64 * 32 + 64 * 64, inv_gs, false);
      }
      const __half* gw = A.geo_w[pass];
      const __half* rw = A.rgb_w[pass];
      stage_weight_T<false>(sm + BwdSmem::wt_r3, rw + 64 * 32 + 64 * 64, 16, 64, 0, 64);
      stage_weight_T<false>(sm + BwdSmem::wt_r2, rw + 64 * 32, 64, 64, 0, 64);
      stage_weight_T<false>(sm + BwdSmem::wt_r1, rw, 64, 32, 16, 16);
      stage_weight_T<true>(sm + BwdSmem::wt_g2, gw + 64 * 32, 32, 64, 0, 64);
      stage_weight_T<false>(sm + BwdSmem::wt_g1, gw, 64, 32, 0, 32);
      __syncthreads();
      cur_pass = pass;
    }
    const int64_t row0 = tile_base + warp * 32;
    uint32_t a[2][4][4];
    float acc[2][8][4];
    // ---- rgb layer 3 (64 -> 16, sigmoid): dZ = dL/drgb * gs * y (1 - y) -------------------------
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
      uint32_t v_lo = 0u, v_hi = 0u;
      if (t < 2) {
        const int64_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
        if (r_lo < pass_end) {
          const float2 y = unpack_h2(__ldg(reinterpret_cast<const uint32_t*>(rgb4 + r_lo * 4 + 2 * t)));
          const float d0 = __ldg(drgb + 3 * r_lo + 2 * t), d1 = t == 0 ? __ldg(drgb + 3 * r_lo + 1) : 0.f;
          v_lo = pack_h2(sat_h(d0 * gs * y.x * (1.f - y.x)), sat_h(d1 * gs * y.y * (1.f - y.y)));
        }
        if (r_hi < pass_end) {
          const float2 y = unpack_h2(__ldg(reinterpret_cast<const uint32_t*>(rgb4 + r_hi * 4 + 2 * t)));
          const float d0 = __ldg(drgb + 3 * r_hi + 2 * t), d1 = t == 0 ? __ldg(drgb + 3 * r_hi + 1) : 0.f;
          v_hi = pack_h2(sat_h(d0 * gs * y.x * (1.f - y.x)), sat_h(d1 * gs * y.y * (1.f - y.y)));
        }
      }
      a[mt][0][0] = v_lo; a[mt][0][1] = v_hi; a[mt][0][2] = 0u; a[mt][0][3] = 0u;
    }
    store_afrag_smem<1>(dzbuf, 24, warp * 32, a, g, t);
    load_in_tile<64>(inbuf, s_rgb_h2, row0, pass_end, warp, lane);
    __syncthreads();
    tile_wgrad_acc<16, 64>(dzbuf, inbuf, accum + BwdSmem::acc_r3, warp, lane);
    warp_layer<1, 8>(a, sm + BwdSmem::wt_r3, acc, g, t);
    relu_mask_to_afrag(acc, inbuf, warp, g, t, a);
    __syncthreads();
    // ---- rgb layer 2 (64 -> 64) ----------------------------------------------------------------
    store_afrag_smem<4>(dzbuf, 72, warp * 32, a, g, t);
    load_in_tile<64>(inbuf, s_rgb_h1, row0, pass_end, warp, lane);
    __syncthreads();
    tile_wgrad_acc<64, 64>(dzbuf, inbuf, accum + BwdSmem::acc_r2, warp, lane);
    warp_layer<4, 8>(a, sm + BwdSmem::wt_r2, acc, g, t);
    relu_mask_to_afrag(acc, inbuf, warp, g, t, a);
    __syncthreads();
    // ---- rgb layer 1 (32 -> 64): input = [SH | geometry feature] ----------------------------------
    store_afrag_smem<4>(dzbuf, 72, warp * 32, a, g, t);
    load_in_tile<32>(inbuf, s_rgb_in, row0, pass_end, warp, lane);
    __syncthreads();
    tile_wgrad_acc<64, 32>(dzbuf, inbuf, accum + BwdSmem::acc_r1, warp, lane);
    warp_layer<4, 2>(a, sm + BwdSmem::wt_r1, acc, g, t);       // d(feature): 32 x 16
    // ---- geo layer 2 (64 -> 32): dZ = [d feature (cols 0..15) | d raw sigma (col 16) | 0] --------
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
      a[mt][0][0] = pack_h2(acc[mt][0][0], acc[mt][0][1]); a[mt][0][1] = pack_h2(acc[mt][0][2], acc[mt][0][3]);
      a[mt][0][2] = pack_h2(acc[mt][1][0], acc[mt][1][1]); a[mt][0][3] = pack_h2(acc[mt][1][2], acc[mt][1][3]);
      uint32_t v_lo = 0u, v_hi = 0u;
      if (t == 0) {   // TruncExp backward: g * exp(clamp(x, -15, 15))  (custom_functions.py:170-173)
        const int64_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
        if (r_lo < pass_end) {
          const float raw = __half2float(s_sraw[r_lo]);
          v_lo = pack_h2(sat_h(__ldg(dsig + r_lo) * gs * expf(fminf(fmaxf(raw, -15.f), 15.f))), 0.f);
        }
        if (r_hi < pass_end) {
          const float raw = __half2float(s_sraw[r_hi]);
          v_hi = pack_h2(sat_h(__ldg(dsig + r_hi) * gs * expf(fminf(fmaxf(raw, -15.f), 15.f))), 0.f);
        }
      }
      a[mt][1][0] = v_lo; a[mt][1][1] = v_hi; a[mt][1][2] = 0u; a[mt][1][3] = 0u;
    }
    __syncthreads();
    store_afrag_smem<2>(dzbuf, 40, warp * 32, a, g, t);
    load_in_tile<64>(inbuf, s_geo_h, row0, pass_end, warp, lane);
    __syncthreads();
    tile_wgrad_acc<32, 64>(dzbuf, inbuf, accum + BwdSmem::acc_g2, warp, lane);
    warp_layer<2, 8>(a, sm + BwdSmem::wt_g2, acc, g, t);
    relu_mask_to_afrag(acc, inbuf, warp, g, t, a);
    __syncthreads();
    // ---- geo layer 1 (32 -> 64) ----------------------------------------------------------------
    store_afrag_smem<4>(dzbuf, 72, warp * 32, a, g, t);
    load_in_tile<32>(inbuf, s_enc, row0, pass_end, warp, lane);
    __syncthreads();
    tile_wgrad_acc<64, 32>(dzbuf, inbuf, accum + BwdSmem::acc_g1, warp, lane);
    warp_layer<4, 4>(a, sm + BwdSmem::wt_g1, acc, g, t);       // d(encoding): 32 x 32
    __syncthreads();
    // ---- encoding gradient: fragments -> smem (own rows) -> lane = sample -> scatter ---------------
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
      for (int nt = 0; nt < 4; nt++) {
        const int r_lo = warp * 32 + mt * 16 + g;
        *reinterpret_cast<uint32_t*>(dzbuf + r_lo * 40 + nt * 8 + 2 * t) = pack_h2(acc[mt][nt][0], acc[mt][nt][1]);
        *reinterpret_cast<uint32_t*>(dzbuf + (r_lo + 8) * 40 + nt * 8 + 2 * t) = pack_h2(acc[mt][nt][2], acc[mt][nt][3]);
      }
    __syncwarp();
    {
      const int64_t s = row0 + lane;
      const bool active = s < pass_end;
      SamplePoint p; p.x = p.y = p.z = 0.f;
      if (active) p = load_sample(A, s);
      const __half2* drow = reinterpret_cast<const __half2*>(dzbuf + (warp * 32 + lane) * 40);
      for (int l = 0; l < A.g.n_levels; l++) {
        const uint32_t off = A.g.offset[l];
        const LevelCorners L = level_corners(p.x, p.y, p.z, A.g.scale[l], A.g.res[l], A.g.offset[l + 1] - off);
        float2 d = __half22float2(drow[l]);
        d.x *= inv_gs; d.y *= inv_gs;
        float* gl = grad_table + 2 * (size_t)off;
        const bool aggregate = l < n_aggregate_levels;   // warp-uniform
#pragma unroll
        for (int k = 0; k < 8; k++) {
          if (aggregate) scatter_aggregated(gl, L.idx[k], L.w[k] * d.x, L.w[k] * d.y, active);
          else if (active) red_add_f32x2(gl + 2 * (size_t)L.idx[k], L.w[k] * d.x, L.w[k] * d.y);
        }
      }
    }
    __syncthreads();   // dzbuf / inbuf are rewritten by the next tile
  }
  __syncthreads();
  if (cur_pass >= 0) {
    float* gg = G.geo[cur_pass]; float* gr = G.rgb[cur_pass];
    flush_accum<64, 32>(accum + BwdSmem::acc_g1, gg, inv_gs, false);
    flush_accum<32, 64>(accum + BwdSmem::acc_g2, gg + 64 * 32, inv_gs, true);
    flush_accum<64, 32>(accum + BwdSmem::acc_r1, gr, inv_gs, false);
    flush_accum<64, 64>(accum + BwdSmem::acc_r2, gr + 64 * 32, inv_gs, false);
    flush_accum<16, 64>(accum + BwdSmem::acc_r3, gr + 64 * 32 + 64 * 64, inv_gs, false);
  }
}

}  // namespace rnb

using namespace rnb;

static int make_field_args(const rnb_field_params* P, const rnb_samples* S, FieldArgs* A, const char* what) {
  RNB_REQUIRE(P != nullptr && S != nullptr, what);
  RNB_REQUIRE(P->K >= 1 && P->K <= RNB_MAX_K, what);
  RNB_REQUIRE(P->grid.n_levels == 16, what);   // the fused kernels are specialised for L=16, F=2 (32 features)
  if (int rc = to_levels(&P->grid, &A->g, what)) return rc;
  RNB_REQUIRE(P->table_f16 != nullptr, what);
  for (int c = 0; c < 3; c++) { A->xmin[c] = P->xyz_min[c]; A->xinv_den[c] = P->xyz_max[c] - P->xyz_min[c]; }
  A->table = (const __half2*)P->table_f16;
  for (int i = 0; i < P->K; i++) {
    RNB_REQUIRE(P->geo_w_f16[i] && P->rgb_w_f16[i], what);
    A->geo_w[i] = (const __half*)P->geo_w_f16[i]; A->rgb_w[i] = (const __half*)P->rgb_w_f16[i];
  }
  A->K = P->K;
  RNB_REQUIRE(S->rays_o && S->rays_d && S->ray_id && S->ts && S->seg, what);
  RNB_REQUIRE(S->capacity > 0 && S->capacity % kTileRows == 0, what);
  A->rays_o = S->rays_o; A->rays_d = S->rays_d; A->ray_id = S->ray_id; A->ts = S->ts; A->seg = S->seg;
  A->capacity = S->capacity;
  return RNB_OK;
}

static int field_grid(int64_t capacity, int per_sm) {
  const int64_t tiles = capacity / kTileRows;
  const int64_t cap = (int64_t)kNumSMs * per_sm;
  return (int)(tiles < cap ? tiles : cap);
}

int rnb_field_fwd_tc_launch(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved, float* sigmas,
                            void* rgb4_f16, rnb_stream_t stream);
int rnb_field_bwd_tc_launch(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved, const void* rgb4_f16,
                            const float* dL_dsigmas, const float* dL_drgbs, float loss_scale, float* grad_table,
                            float* const* grad_geo, float* const* grad_rgb, rnb_stream_t stream);
static int g_fwd_impl = 1;   // 1 = tcgen05 (field_tc.cu), 0 = mma.sync (this file)
static int g_bwd_impl = 1;

extern "C" {

int rnb_set_field_impl(int fwd_tcgen05, int bwd_tcgen05) {
  // the two implementations keep the saved activations in different layouts (operand tiles vs rows): no mixing
  RNB_REQUIRE((fwd_tcgen05 != 0) == (bwd_tcgen05 != 0), "rnb_set_field_impl");
  g_fwd_impl = fwd_tcgen05 ? 1 : 0;
  g_bwd_impl = bwd_tcgen05 ? 1 : 0;
  return RNB_OK;
}

int rnb_field_fwd(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved, float* sigmas,
                  void* rgb4_f16, rnb_stream_t stream) {
  const char* what = "rnb_field_fwd";
  FieldArgs A;
  if (int rc = make_field_args(P, S, &A, what)) return rc;
  RNB_REQUIRE(sigmas && rgb4_f16, what);
  const bool save = saved != nullptr && saved->enc != nullptr;
  if (save) RNB_REQUIRE(saved->rgb_in && saved->geo_h && saved->rgb_h1 && saved->rgb_h2 && saved->sraw, what);
  if (g_fwd_impl == 1) return rnb_field_fwd_tc_launch(P, S, saved, sigmas, rgb4_f16, stream);
  const int grid = field_grid(S->capacity, 4);
  if (save) {
    RNB_REQUIRE(saved->rgb_in && saved->geo_h && saved->rgb_h1 && saved->rgb_h2 && saved->sraw, what);
    field_fwd_kernel<true><<<grid, 128, 0, as_stream(stream)>>>(A, (__half*)saved->enc, (__half*)saved->rgb_in, (__half*)saved->geo_h,
                                                                (__half*)saved->rgb_h1, (__half*)saved->rgb_h2, (__half*)saved->sraw,
                                                                sigmas, (__half*)rgb4_f16);
  } else {
    field_fwd_kernel<false><<<grid, 128, 0, as_stream(stream)>>>(A, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, sigmas, (__half*)rgb4_f16);
  }
  return check_launch(what);
}

int rnb_field_bwd(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved, const void* rgb4_f16,
                  const float* dL_dsigmas, const float* dL_drgbs, float loss_scale, float* grad_table,
                  float* const* grad_geo, float* const* grad_rgb, rnb_stream_t stream) {
  const char* what = "rnb_field_bwd";
  FieldArgs A;
  if (int rc = make_field_args(P, S, &A, what)) return rc;
  RNB_REQUIRE(saved && saved->enc && saved->rgb_in && saved->geo_h && saved->rgb_h1 && saved->rgb_h2 && saved->sraw, what);
  RNB_REQUIRE(rgb4_f16 && dL_dsigmas && dL_drgbs && grad_table && grad_geo && grad_rgb && loss_scale > 0.f, what);
  for (int i = 0; i < P->K; i++) RNB_REQUIRE(grad_geo[i] && grad_rgb[i], what);
  if (g_bwd_impl == 1)
    return rnb_field_bwd_tc_launch(P, S, saved, rgb4_f16, dL_dsigmas, dL_drgbs, loss_scale, grad_table, grad_geo, grad_rgb, stream);
  GradPtrs G;
  for (int i = 0; i < P->K; i++) { RNB_REQUIRE(grad_geo[i] && grad_rgb[i], what); G.geo[i] = grad_geo[i]; G.rgb[i] = grad_rgb[i]; }
  int n_agg = 0;
  for (int l = 0; l < A.g.n_levels; l++)
    if ((uint64_t)A.g.res[l] * A.g.res[l] * A.g.res[l] <= (uint64_t)(A.g.offset[l + 1] - A.g.offset[l]) && A.g.res[l] <= 64) n_agg = l + 1;
  static bool attr_set = false;
  if (!attr_set) {
    RNB_CUDA(cudaFuncSetAttribute(field_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BwdSmem::bytes), what);
    attr_set = true;
  }
  const int grid = field_grid(S->capacity, 2);
  field_bwd_kernel<<<grid, 128, BwdSmem::bytes, as_stream(stream)>>>(
      A, (const __half*)saved->enc, (const __half*)saved->rgb_in, (const __half*)saved->geo_h, (const __half*)saved->rgb_h1,
      (const __half*)saved->rgb_h2, (const __half*)saved->sraw, (const __half*)rgb4_f16, dL_dsigmas, dL_drgbs, loss_scale,
      grad_table, G, n_agg);
  return check_launch(what);
}

}  // extern "C"
