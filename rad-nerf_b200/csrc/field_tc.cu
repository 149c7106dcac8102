// field_tc.cu -- the fused field on Blackwell's 5th-generation tensor cores (tcgen05 + TMEM).
//
// Same operator as field.cu's field_fwd_kernel (hash-grid encode -> geo MLP -> TruncExp -> SH4 ->
// rgb MLP for all K passes of a packed sample list), re-mapped for sm_100a:
//   * one persistent CTA per SM, kWG warpgroups; a warpgroup owns one 128-sample tile at a time and
//     THREAD = SAMPLE everywhere: the thread gathers its 16x8 table entries, writes its 32 encoded
//     features as four 16-byte chunks straight into the canonical (no-swizzle) K-major operand tile in
//     shared memory, and later reads back ITS row of every layer's accumulator from tensor memory
//     (tcgen05.ld 32x32b: TMEM lane = tile row = thread), applies ReLU / TruncExp / sigmoid, and
//     writes the fp16 row as the next layer's operand.  No mma.sync fragment layouts, no shuffles, no
//     transposition through shared memory; accumulators never occupy registers.
//   * the five dense layers are tcgen05.mma (M = 128, N = 64/32/64/64/16, K = 16 per instruction),
//     issued by one elected thread of the warpgroup, completion signalled through an mbarrier
//     (tcgen05.commit); the 22 KB of weights of the current pass are staged once per pass and CTA.
//   * while one warpgroup waits for its MMA or its gathers, the other warpgroups of the SM are in other
//     phases of their own tiles (gather is L2-latency bound, the layer chain is a dependent sequence).
// Rounding points are those of the operator path (fp16 features / activations / outputs, fp32
// accumulation), so this kernel, field.cu and radnerf_b200.tcnn agree to fp16 resolution.
#include <stdlib.h>

#include "hashgrid.cuh"
#include "tc05.cuh"

namespace rnb {

constexpr int kWG = 5;            // warpgroups per CTA (forward): 640 threads x 96 registers
constexpr int kRows = 128;        // samples per tile

struct TcFieldArgs {
  GridLevels g;
  float xmin[3], xden[3];
  const __half2* table;
  const __half* geo_w[RNB_MAX_K];
  const __half* rgb_w[RNB_MAX_K];
  int K;
  const float* rays_o; const float* rays_d;
  const int32_t* ray_id; const float* ts; const int32_t* seg;
  int64_t capacity;
};

__device__ __forceinline__ uint32_t pack2h(float a, float b) {
  const __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}

// one level of the multiresolution grid -> packed half2 (same arithmetic as hashgrid_fwd_kernel).
// Corners 2p and 2p+1 are x-neighbours.  On dense levels their entries are adjacent; on hashed levels
// idx = x ^ (y*P1) ^ (z*P2), so for even x the neighbour is idx ^ 1.  Either way, half of the time both live in one
// aligned 8-byte pair: the pair is fetched with one load and the second 4-byte gather is issued only by the lanes that
// still need it (fewer L1 tag lookups and sectors -- the unit this kernel is bound by).
__device__ __forceinline__ uint32_t tc_encode_level(const TcFieldArgs& A, int l, float x, float y, float z) {
  const uint32_t off = A.g.offset[l];
  const LevelCorners L = level_corners(x, y, z, A.g.scale[l], A.g.res[l], A.g.offset[l + 1] - off);
  const __half2* t = A.table + off;                      // level offsets are multiples of 8 entries: pairs stay aligned
  uint2 pr[4];
  uint32_t vb[4];
#pragma unroll
  for (int p = 0; p < 4; p++) pr[p] = __ldg(reinterpret_cast<const uint2*>(t) + (L.idx[2 * p] >> 1));
#pragma unroll
  for (int p = 0; p < 4; p++) {
    vb[p] = 0u;
    if ((L.idx[2 * p] ^ L.idx[2 * p + 1]) != 1u) vb[p] = *reinterpret_cast<const uint32_t*>(t + L.idx[2 * p + 1]);
  }
  float a0 = 0.f, a1 = 0.f;
#pragma unroll
  for (int p = 0; p < 4; p++) {
    const bool odd = L.idx[2 * p] & 1u;
    const uint32_t ea = odd ? pr[p].y : pr[p].x;
    const uint32_t eb = ((L.idx[2 * p] ^ L.idx[2 * p + 1]) == 1u) ? (odd ? pr[p].x : pr[p].y) : vb[p];
    const float2 fa = __half22float2(*reinterpret_cast<const __half2*>(&ea));
    const float2 fb = __half22float2(*reinterpret_cast<const __half2*>(&eb));
    a0 = fmaf(L.w[2 * p], fa.x, a0); a1 = fmaf(L.w[2 * p], fa.y, a1);
    a0 = fmaf(L.w[2 * p + 1], fb.x, a0); a1 = fmaf(L.w[2 * p + 1], fb.y, a1);
  }
  return pack2h(a0, a1);
}

// W (n_out, n_in) row-major fp16 in global -> canonical K-major tile [n_out rows][n_in features] in smem
__device__ __forceinline__ void tc_stage_weight(unsigned char* dst, const __half* __restrict__ src, int n_out, int n_in, int tid, int nthreads) {
  const int cpr = n_in / 8;
  for (int i = tid; i < n_out * cpr; i += nthreads) {
    const int n = i / cpr, c = i % cpr;
    *reinterpret_cast<uint4*>(dst + tc::tile_chunk_off(n_out, n, c)) = __ldg(reinterpret_cast<const uint4*>(src + n * n_in) + c);
  }
}

struct TcSmem {
  // weights of the current pass (bytes)
  static constexpr int g1 = 0;                 // 64 x 32
  static constexpr int g2 = g1 + 64 * 32 * 2;  // 32 x 64
  static constexpr int r1 = g2 + 32 * 64 * 2;  // 64 x 32
  static constexpr int r2 = r1 + 64 * 32 * 2;  // 64 x 64
  static constexpr int r3 = r2 + 64 * 64 * 2;  // 16 x 64
  static constexpr int wend = r3 + 16 * 64 * 2;
  // per warpgroup: a 128x32 operand tile (encoding, then rgb input) and a 128x64 tile (the hidden activations in turn).
  // 24 KB per warpgroup, 142 KB per CTA of 5 warpgroups: the rest of the SM's 256 KB stays L1 (~90 KB) for the table
  // gathers, which is worth more than a second 128x64 tile.  Measured forward time on the K=2 training shape:
  // 40 KB/warpgroup: 3 WG 0.158, 4 WG 0.157, 5 WG (227 KB, L1 ~29 KB) 0.215 ms; 24 KB/warpgroup: 4 WG 0.149, 5 WG 0.140,
  // 6 WG (80 registers) 0.173 ms
  static constexpr int a8 = 0;
  static constexpr int b16 = kRows * 32 * 2;
  static constexpr int per_wg = b16 + kRows * 64 * 2;
  static constexpr int wg0 = (wend + 127) / 128 * 128;
  static constexpr int bars = wg0 + kWG * per_wg;       // kWG mbarriers
  static constexpr int tmem_slot = bars + kWG * 8;
  static constexpr int extent = tmem_slot + 16;          // i64 [2]: this pass's base and end, read ONCE per CTA
  static constexpr int bytes = extent + 16;
};

// D[tmem] = A[128 x KDIM] * W[N x KDIM]^T, issued by one thread
template <int N, int KDIM>
__device__ __forceinline__ void tc_layer_mma(uint32_t d_tmem, uint32_t a_smem, uint32_t w_smem) {
  constexpr uint32_t idesc = tc::instr_desc(128, N, 0, 0);
#pragma unroll
  for (int k = 0; k < KDIM / 16; k++)
    tc::mma_f16(d_tmem, tc::smem_desc(a_smem + k * 2 * (kRows * 16), kRows * 16, 128), tc::smem_desc(w_smem + k * 2 * (N * 16), N * 16, 128), idesc, k > 0);
}

// named barrier over the 128 threads of a warpgroup (barrier ids 1..kWG; 0 is __syncthreads)
__device__ __forceinline__ void wg_sync(int wg) { asm volatile("bar.sync %0, 128;" ::"r"(wg + 1) : "memory"); }

// 64 accumulator columns of this thread's row -> ReLU -> fp16 -> 8 operand chunks
__device__ __forceinline__ void tc_relu_row64(uint32_t taddr, unsigned char* tile, int row) {
#pragma unroll
  for (int h = 0; h < 2; h++) {
    uint32_t r[32];
    tc::tmem_ld32_nowait(taddr + h * 32, r);
    tc::tmem_ld_wait();
#pragma unroll
    for (int c = 0; c < 4; c++) {
      uint4 q;
      q.x = pack2h(fmaxf(__uint_as_float(r[8 * c + 0]), 0.f), fmaxf(__uint_as_float(r[8 * c + 1]), 0.f));
      q.y = pack2h(fmaxf(__uint_as_float(r[8 * c + 2]), 0.f), fmaxf(__uint_as_float(r[8 * c + 3]), 0.f));
      q.z = pack2h(fmaxf(__uint_as_float(r[8 * c + 4]), 0.f), fmaxf(__uint_as_float(r[8 * c + 5]), 0.f));
      q.w = pack2h(fmaxf(__uint_as_float(r[8 * c + 6]), 0.f), fmaxf(__uint_as_float(r[8 * c + 7]), 0.f));
      *reinterpret_cast<uint4*>(tile + tc::tile_chunk_off(kRows, row, h * 4 + c)) = q;
    }
  }
}

// One layer step of the forward chain, executed by the warpgroup's elected thread: issue the layer's MMAs and, when
// activations are saved for the backward pass, copy the operand tile the MMAs read -- it already has the layout the
// backward kernel wants -- to global memory with one bulk store.  A tile is overwritten at the earliest in the
// epilogue of the FOLLOWING step; this thread confirms that the previous step's store has finished reading its tile
// before it issues (and commits) this step's MMAs, and everybody waits for those, so no extra signalling is needed
// (one exception: rgb hidden 2 overwrites rgb hidden 1 in the epilogue of the SAME step, see there).
template <bool kSave, int N, int KDIM>
__device__ __forceinline__ void tc_fwd_step(uint64_t* bar, uint32_t tmem_d, uint32_t a_smem, uint32_t w_smem, __half* save_tile, uint64_t l2_stream = 0) {
  if (kSave) tc::bulk_wait_read();
  tc::fence_after_sync();
  tc_layer_mma<N, KDIM>(tmem_d, a_smem, w_smem);
  tc::mma_commit(bar);
  if (kSave) {
    if (l2_stream) tc::bulk_store_hint(save_tile, a_smem, kRows * KDIM * 2, l2_stream);
    else tc::bulk_store(save_tile, a_smem, kRows * KDIM * 2);
    tc::bulk_commit();
  }
}

template <bool kSave>
__global__ void __maxnreg__(96) field_fwd_tc_kernel(TcFieldArgs A, __half* __restrict__ s_enc, __half* __restrict__ s_rgb_in,
                                                                     __half* __restrict__ s_geo_h, __half* __restrict__ s_rgb_h1,
                                                                     __half* __restrict__ s_rgb_h2, __half* __restrict__ s_sraw,
                                                                     float* __restrict__ sigmas, __half* __restrict__ rgb4, uint64_t l2_stream) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, wg = tid >> 7, row = tid & 127, warp = tid >> 5;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TcSmem::bars);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + TcSmem::tmem_slot);
  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid < kWG) tc::mbar_init(bars + tid, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_d = *tmem_slot + wg * 64;                               // this warpgroup's 64 accumulator columns
  const uint32_t tmem_row = tmem_d + ((uint32_t)((warp & 3) * 32) << 16);     // + this warp's 32 lanes
  unsigned char* a8 = smem + TcSmem::wg0 + wg * TcSmem::per_wg + TcSmem::a8;
  unsigned char* b16 = smem + TcSmem::wg0 + wg * TcSmem::per_wg + TcSmem::b16;
  const uint32_t a8_s = tc::smem_u32(a8), b16_s = tc::smem_u32(b16), w_s = tc::smem_u32(smem);
  uint64_t* bar = bars + wg;
  uint32_t phase = 0;
  const int K = A.K;
  int64_t tiles_before = 0;
  if (kSave && l2_stream) l2_stream = tc::l2_evict_first_policy();
  volatile int64_t* extent = reinterpret_cast<volatile int64_t*>(smem + TcSmem::extent);
  for (int pass = 0; pass < K; pass++) {
    // The pass extents are read by ONE thread and shared: every barrier below is taken (or skipped) by the whole CTA even if the
    // sample list is being rewritten under the kernel (a caller error, but it must not hang the GPU).
    __syncthreads();
    if (tid == 0) { extent[0] = __ldg(A.seg + pass); extent[1] = extent[0] + __ldg(A.seg + K + 1 + pass); }
    __syncthreads();
    const int64_t base = extent[0];
    int64_t pass_end = extent[1];
    if (pass_end > A.capacity) pass_end = A.capacity;
    const int64_t n_tiles = pass_end > base ? (pass_end - base + kRows - 1) / kRows : 0;
    if (n_tiles == 0) continue;      // (CTA-uniform) an empty pass -- e.g. every round behind the end of a frame -- costs nothing
    // ---- stage this pass's weights (whole CTA) ----
    {
      const __half* gw = A.geo_w[pass];
      const __half* rw = A.rgb_w[pass];
      tc_stage_weight(smem + TcSmem::g1, gw, 64, 32, tid, kWG * 128);
      tc_stage_weight(smem + TcSmem::g2, gw + 64 * 32, 32, 64, tid, kWG * 128);
      tc_stage_weight(smem + TcSmem::r1, rw, 64, 32, tid, kWG * 128);
      tc_stage_weight(smem + TcSmem::r2, rw + 64 * 32, 64, 64, tid, kWG * 128);
      tc_stage_weight(smem + TcSmem::r3, rw + 64 * 32 + 64 * 64, 16, 64, tid, kWG * 128);
    }
    tc::fence_async_smem();
    __syncthreads();
    // The sample -> ray -> origin/direction lookups are two dependent global loads in front of the first gather: they are
    // issued one tile ahead (while this tile is in its layer chain) and carried in registers.
    // Tiles go round-robin over the grid's warpgroups; each pass starts where the previous one stopped, so the
    // warpgroups that took the odd tile of one pass are not the ones that take it in the next (with a fixed start, the
    // first CTAs carried ceil(tiles / warpgroups) in EVERY pass and set the kernel's duration).
    const int64_t tstride = (int64_t)gridDim.x * kWG;
    // warpgroup ids interleave the CTAs (wg-major): the tiles of a partial last round then spread over all SMs, a few
    // warpgroups each, instead of filling some SMs and leaving others empty
    // (forward 0.140 -> 0.135 ms; the backward kernel, whose tiles are latency- rather than throughput-bound, keeps CTA-major ids)
    int64_t tile0 = (int64_t)wg * gridDim.x + blockIdx.x - tiles_before % tstride;
    if (tile0 < 0) tile0 += tstride;
    tiles_before += n_tiles;
    float nt = 0.f, nox = 0.f, noy = 0.f, noz = 0.f, ndx = 0.f, ndy = 0.f, ndz = 0.f;
    if (tile0 < n_tiles && base + tile0 * kRows + row < pass_end) {
      const int64_t s0 = base + tile0 * kRows + row;
      const int r = __ldg(A.ray_id + s0);
      nt = __ldg(A.ts + s0);
      nox = __ldg(A.rays_o + 3 * r); noy = __ldg(A.rays_o + 3 * r + 1); noz = __ldg(A.rays_o + 3 * r + 2);
      ndx = __ldg(A.rays_d + 3 * r); ndy = __ldg(A.rays_d + 3 * r + 1); ndz = __ldg(A.rays_d + 3 * r + 2);
    }
    for (int64_t tile = tile0; tile < n_tiles; tile += tstride) {
      const int64_t s = base + tile * kRows + row;
      const bool valid = s < pass_end;
      const int64_t s_next = s + tstride * kRows;
      const bool valid_next = tile + tstride < n_tiles && s_next < pass_end;
      int r_next = 0;
      // ---- gather: 16 levels x 8 corners -> 32 fp16 features -> operand tile (4 chunks) ----
      uint4 shq0 = make_uint4(0u, 0u, 0u, 0u), shq1 = shq0;
      if (valid) {
        const float t = nt, ox = nox, oy = noy, oz = noz, dx = ndx, dy = ndy, dz = ndz;
        if (valid_next) { r_next = __ldg(A.ray_id + s_next); nt = __ldg(A.ts + s_next); }
        const float x = fminf(fmaxf(__fdiv_rn(__fmaf_rn(dx, t, ox) - A.xmin[0], A.xden[0]), 0.f), 1.f);
        const float y = fminf(fmaxf(__fdiv_rn(__fmaf_rn(dy, t, oy) - A.xmin[1], A.xden[1]), 0.f), 1.f);
        const float z = fminf(fmaxf(__fdiv_rn(__fmaf_rn(dz, t, oz) - A.xmin[2], A.xden[2]), 0.f), 1.f);
#pragma unroll 1
        for (int c = 0; c < 4; c++) {
          uint4 q;
          q.x = tc_encode_level(A, 4 * c, x, y, z); q.y = tc_encode_level(A, 4 * c + 1, x, y, z);
          q.z = tc_encode_level(A, 4 * c + 2, x, y, z); q.w = tc_encode_level(A, 4 * c + 3, x, y, z);
          *reinterpret_cast<uint4*>(a8 + tc::tile_chunk_off(kRows, row, c)) = q;
        }
        // direction: d/|d| -> (d+1)/2 -> (tcnn) 2u-1 -> SH4   (networks.py:322-323)
        const float nrm = sqrtf(dx * dx + dy * dy + dz * dz);
        const float ux = (__fdiv_rn(dx, nrm) + 1.f) * 0.5f, uy = (__fdiv_rn(dy, nrm) + 1.f) * 0.5f, uz = (__fdiv_rn(dz, nrm) + 1.f) * 0.5f;
        float o[16];
        sh4_eval(ux * 2.f - 1.f, uy * 2.f - 1.f, uz * 2.f - 1.f, o);
        shq0 = make_uint4(pack2h(o[0], o[1]), pack2h(o[2], o[3]), pack2h(o[4], o[5]), pack2h(o[6], o[7]));
        shq1 = make_uint4(pack2h(o[8], o[9]), pack2h(o[10], o[11]), pack2h(o[12], o[13]), pack2h(o[14], o[15]));
      } else {
#pragma unroll
        for (int c = 0; c < 4; c++) *reinterpret_cast<uint4*>(a8 + tc::tile_chunk_off(kRows, row, c)) = make_uint4(0u, 0u, 0u, 0u);
      }
      // ---- geo layer 1: 32 -> 64, ReLU ----
      tc::fence_async_smem();
      tc::fence_before_sync();
      wg_sync(wg);
      if (row == 0) tc_fwd_step<kSave, 64, 32>(bar, tmem_d, a8_s, w_s + TcSmem::g1, s_enc + s * 32, l2_stream);
      if (valid_next) {      // second hop of the next tile's lookups; lands while this tile runs its five layers
        if (!valid) { r_next = __ldg(A.ray_id + s_next); nt = __ldg(A.ts + s_next); }
        nox = __ldg(A.rays_o + 3 * r_next); noy = __ldg(A.rays_o + 3 * r_next + 1); noz = __ldg(A.rays_o + 3 * r_next + 2);
        ndx = __ldg(A.rays_d + 3 * r_next); ndy = __ldg(A.rays_d + 3 * r_next + 1); ndz = __ldg(A.rays_d + 3 * r_next + 2);
      }
      tc::mbar_wait(bar, phase); phase ^= 1;
      tc::fence_after_sync();
      tc_relu_row64(tmem_row, b16, row);
      // ---- geo layer 2: 64 -> 32 (col 0 = raw sigma, cols 1..16 = geometry feature) ----
      tc::fence_async_smem();
      tc::fence_before_sync();
      wg_sync(wg);
      if (row == 0) tc_fwd_step<kSave, 32, 64>(bar, tmem_d, b16_s, w_s + TcSmem::g2, s_geo_h + s * 64, l2_stream);
      tc::mbar_wait(bar, phase); phase ^= 1;
      tc::fence_after_sync();
      {
        uint32_t r[32];
        tc::tmem_ld32_nowait(tmem_row, r);
        tc::tmem_ld_wait();
        const __half raw = __float2half_rn(__uint_as_float(r[0]));
        uint4 f0, f1;
        f0.x = pack2h(__uint_as_float(r[1]), __uint_as_float(r[2])); f0.y = pack2h(__uint_as_float(r[3]), __uint_as_float(r[4]));
        f0.z = pack2h(__uint_as_float(r[5]), __uint_as_float(r[6])); f0.w = pack2h(__uint_as_float(r[7]), __uint_as_float(r[8]));
        f1.x = pack2h(__uint_as_float(r[9]), __uint_as_float(r[10])); f1.y = pack2h(__uint_as_float(r[11]), __uint_as_float(r[12]));
        f1.z = pack2h(__uint_as_float(r[13]), __uint_as_float(r[14])); f1.w = pack2h(__uint_as_float(r[15]), __uint_as_float(r[16]));
        // rgb input = [SH4 | geometry feature]; the 128x32 tile is free again (its MMA completed)
        *reinterpret_cast<uint4*>(a8 + tc::tile_chunk_off(kRows, row, 0)) = shq0;
        *reinterpret_cast<uint4*>(a8 + tc::tile_chunk_off(kRows, row, 1)) = shq1;
        *reinterpret_cast<uint4*>(a8 + tc::tile_chunk_off(kRows, row, 2)) = f0;
        *reinterpret_cast<uint4*>(a8 + tc::tile_chunk_off(kRows, row, 3)) = f1;
        if (valid) {
          sigmas[s] = expf(__half2float(raw));
          if (kSave) s_sraw[s] = raw;
        }
      }
      // ---- rgb layer 1: 32 -> 64, ReLU ----
      tc::fence_async_smem();
      tc::fence_before_sync();
      wg_sync(wg);
      if (row == 0) tc_fwd_step<kSave, 64, 32>(bar, tmem_d, a8_s, w_s + TcSmem::r1, s_rgb_in + s * 32, l2_stream);
      tc::mbar_wait(bar, phase); phase ^= 1;
      tc::fence_after_sync();
      tc_relu_row64(tmem_row, b16, row);
      // ---- rgb layer 2: 64 -> 64, ReLU ----
      tc::fence_async_smem();
      tc::fence_before_sync();
      wg_sync(wg);
      if (row == 0) tc_fwd_step<kSave, 64, 64>(bar, tmem_d, b16_s, w_s + TcSmem::r2, s_rgb_h1 + s * 64, l2_stream);
      tc::mbar_wait(bar, phase); phase ^= 1;
      tc::fence_after_sync();
      if (kSave) {
        // rgb hidden 2 goes into the tile whose previous content (rgb hidden 1) is being copied out by the bulk store
        // issued a moment ago, next to the MMAs that have now completed: confirm the store has read the tile
        if (row == 0) tc::bulk_wait_read();
        wg_sync(wg);
      }
      tc_relu_row64(tmem_row, b16, row);
      // ---- rgb layer 3: 64 -> 16, sigmoid ----
      tc::fence_async_smem();
      tc::fence_before_sync();
      wg_sync(wg);
      if (row == 0) tc_fwd_step<kSave, 16, 64>(bar, tmem_d, b16_s, w_s + TcSmem::r3, s_rgb_h2 + s * 64, l2_stream);
      tc::mbar_wait(bar, phase); phase ^= 1;
      tc::fence_after_sync();
      {
        float v[16];
        tc::tmem_ld16(tmem_row, v);
        if (valid) {
          const uint2 q = make_uint2(pack2h(1.f / (1.f + __expf(-v[0])), 1.f / (1.f + __expf(-v[1]))),
                                     pack2h(1.f / (1.f + __expf(-v[2])), 1.f / (1.f + __expf(-v[3]))));
          *reinterpret_cast<uint2*>(rgb4 + s * 4) = q;
        }
      }
      tc::fence_before_sync();   // the next tile's first MMA overwrites the accumulator columns just read
    }
  }
  if (kSave && row == 0) tc::bulk_wait_all();
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(*tmem_slot, 512);
}


// ================================ density of occupancy-grid cells ====================================
// MNGP.update_density_grid's evaluation step (networks.py:374-399): for every sub-NeRF i and cascade c, a
// jittered point inside each listed cell -> density_i (hash encode -> geo MLP -> TruncExp) -> scratch grid.
// Same tile machinery as the forward kernel, cut after the geo network; thread = cell.
struct GridCellArgs {
  float* tmp[RNB_MAX_K];       // (C, G^3) f32 scratch grid of sub-NeRF i
  const int32_t* cell_idx;     // (K, C, M) morton index, < 0 = skip; NULL: all cells, index = m
  const float* noise;          // (K, C, M, 3) in [0,1); NULL: counter-based generator seeded with `seed`
  float span[8], half[8];      // per cascade: s_c - s_c/G and s_c/G with s_c = min(2^(c-1), scale)  (networks.py:385-390)
  int C, G;
  int64_t M;
  uint64_t seed;
};

__device__ __forceinline__ uint32_t tc_compact_bits(uint32_t x) {
  x = x & 0x49249249u;
  x = (x | (x >> 2)) & 0xc30c30c3u;
  x = (x | (x >> 4)) & 0x0f00f00fu;
  x = (x | (x >> 8)) & 0xff0000ffu;
  x = (x | (x >> 16)) & 0x0000ffffu;
  return x;
}
// uniform [0,1) from a 64-bit counter (splitmix64 finaliser; 24 mantissa bits)
__device__ __forceinline__ float tc_counter_uniform(uint64_t seed, uint64_t ctr) {
  uint64_t z = seed + (ctr + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(uint32_t)(z >> 40) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(kWG * 128, 1) density_cells_tc_kernel(TcFieldArgs A, GridCellArgs Q) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, wg = tid >> 7, row = tid & 127, warp = tid >> 5;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TcSmem::bars);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + TcSmem::tmem_slot);
  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid < kWG) tc::mbar_init(bars + tid, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_d = *tmem_slot + wg * 64;
  const uint32_t tmem_row = tmem_d + ((uint32_t)((warp & 3) * 32) << 16);
  unsigned char* a8 = smem + TcSmem::wg0 + wg * TcSmem::per_wg + TcSmem::a8;
  unsigned char* b16 = smem + TcSmem::wg0 + wg * TcSmem::per_wg + TcSmem::b16;
  const uint32_t a8_s = tc::smem_u32(a8), b16_s = tc::smem_u32(b16), w_s = tc::smem_u32(smem);
  uint64_t* bar = bars + wg;
  uint32_t phase = 0;
  const int64_t G3 = (int64_t)Q.G * Q.G * Q.G;
  const int64_t per_pass = (int64_t)Q.C * Q.M;
  const int64_t n_tiles = (per_pass + kRows - 1) / kRows;
  const float gm1 = (float)(Q.G - 1);
  for (int pass = 0; pass < A.K; pass++) {
    __syncthreads();
    {
      const __half* gw = A.geo_w[pass];
      tc_stage_weight(smem + TcSmem::g1, gw, 64, 32, tid, kWG * 128);
      tc_stage_weight(smem + TcSmem::g2, gw + 64 * 32, 32, 64, tid, kWG * 128);
    }
    tc::fence_async_smem();
    __syncthreads();
    for (int64_t tile = (int64_t)blockIdx.x * kWG + wg; tile < n_tiles; tile += (int64_t)gridDim.x * kWG) {
      const int64_t j = tile * kRows + row;
      bool valid = j < per_pass;
      int c = 0;
      int32_t idx = -1;
      if (valid) {
        c = (int)(j / Q.M);
        const int64_t m = j - (int64_t)c * Q.M;
        idx = Q.cell_idx ? __ldg(Q.cell_idx + (int64_t)pass * per_pass + j) : (int32_t)m;
        valid = idx >= 0;
      }
      if (valid) {
        const int64_t e = (int64_t)pass * per_pass + j;
        float u[3];
#pragma unroll
        for (int k = 0; k < 3; k++) u[k] = Q.noise ? __ldg(Q.noise + 3 * e + k) : tc_counter_uniform(Q.seed, 3 * (uint64_t)e + k);
        float p[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          // (coords / (G-1) * 2 - 1) * (s - half) + (rand * 2 - 1) * half, each operation rounded on its own like the ATen chain
          const float cc = (float)tc_compact_bits((uint32_t)(idx >> k));
          const float centre = __fmul_rn(__fadd_rn(__fmul_rn(__fdiv_rn(cc, gm1), 2.f), -1.f), Q.span[c]);
          const float jit = __fmul_rn(__fadd_rn(__fmul_rn(u[k], 2.f), -1.f), Q.half[c]);
          p[k] = fminf(fmaxf(__fdiv_rn(__fadd_rn(centre, jit) - A.xmin[k], A.xden[k]), 0.f), 1.f);
        }
#pragma unroll 1
        for (int cq = 0; cq < 4; cq++) {
          uint4 q;
          q.x = tc_encode_level(A, 4 * cq, p[0], p[1], p[2]); q.y = tc_encode_level(A, 4 * cq + 1, p[0], p[1], p[2]);
          q.z = tc_encode_level(A, 4 * cq + 2, p[0], p[1], p[2]); q.w = tc_encode_level(A, 4 * cq + 3, p[0], p[1], p[2]);
          *reinterpret_cast<uint4*>(a8 + tc::tile_chunk_off(kRows, row, cq)) = q;
        }
      } else {
#pragma unroll
        for (int cq = 0; cq < 4; cq++) *reinterpret_cast<uint4*>(a8 + tc::tile_chunk_off(kRows, row, cq)) = make_uint4(0u, 0u, 0u, 0u);
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      wg_sync(wg);
      if (row == 0) tc_fwd_step<false, 64, 32>(bar, tmem_d, a8_s, w_s + TcSmem::g1, nullptr);
      tc::mbar_wait(bar, phase); phase ^= 1;
      tc::fence_after_sync();
      tc_relu_row64(tmem_row, b16, row);
      tc::fence_async_smem();
      tc::fence_before_sync();
      wg_sync(wg);
      if (row == 0) tc_fwd_step<false, 32, 64>(bar, tmem_d, b16_s, w_s + TcSmem::g2, nullptr);
      tc::mbar_wait(bar, phase); phase ^= 1;
      tc::fence_after_sync();
      {
        float v[16];
        tc::tmem_ld16(tmem_row, v);
        if (valid) Q.tmp[pass][(int64_t)c * G3 + idx] = expf(__half2float(__float2half_rn(v[0])));
      }
      tc::fence_before_sync();
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(*tmem_slot, 512);
}

// ================================ backward on tcgen05 ================================================
// One persistent CTA per SM: kWGB warpgroups (thread = sample, one 128-sample tile each) plus ONE
// issuer warp whose elected lane issues every tcgen05.mma of the CTA, so that products accumulating
// into shared TMEM columns are strictly ordered.  Per tile and layer l (rgb3, rgb2, rgb1, geo2, geo1):
//   warpgroup:  writes its row of dZ_l into the `dz` operand tile and copies its row of the saved input
//               In_l (global) into the `in` tile, fences, arrives on full[wg];
//   issuer:     dW_l (TMEM, persists across all tiles of the pass, M = 64 out-neurons x N = in-features)
//               += dZ_l^T In_l  -- eight K=16 steps over the tile's 128 samples, BOTH operands read as
//               MN-major views of the very tiles the threads wrote row by row (tc05.cuh);
//               dA (TMEM, per warpgroup) = dZ_l W_l -- W_l staged once per pass in its forward layout
//               and read as an MN-major B operand; tcgen05.commit -> done[wg];
//   warpgroup:  reads ITS row of dA from TMEM, applies relu'(In_l) with its own row of the `in` tile,
//               and that is dZ_{l-1}.  After geo1 the row is d(encoding): the thread scatters it.
// The weight gradients leave TMEM once per pass and CTA (64 x N fp32 each), not once per tile.
constexpr int kWGB = 4;

struct TcBwdSmem {
  static constexpr int wend = TcSmem::wend;                       // same weight tiles as the forward kernel
  static constexpr int wg0 = (wend + 127) / 128 * 128;
  static constexpr int dz = 0;                                    // 128 x 64 fp16
  static constexpr int in = kRows * 64 * 2;                       // 2 x (128 x 64 fp16): saved-input tiles, double-buffered
  static constexpr int in_bytes = kRows * 64 * 2;
  static constexpr int per_wg = 3 * kRows * 64 * 2;
  static constexpr int bars = wg0 + kWGB * per_wg;                // full[kWGB], done[kWGB], loaded[kWGB][2]
  static constexpr int tmem_slot = bars + 4 * kWGB * 8;
  static constexpr int ready = tmem_slot + 16;                    // i32 [kWGB][4]: tiles of this pass whose d(encoding) rows were handed over
  static constexpr int extent = ready + kWGB * 4 * 4;            // i64 [2]: this pass's base and end, read ONCE per CTA
  static constexpr int bytes = extent + 16;
};
// Scatter warps.  The table scatter (16 levels x 8 corners of reductions per sample, plus the run aggregation on the
// coarse levels) used to run in the compute warpgroups behind the five MMA round trips of a tile: a warp was either in its
// MMA chain (mostly waiting) or in its scatter (latency-bound: ~1400 cycles per level and warp), never both (ncu r01: 27 %
// occupancy, issue 36 %, nothing saturated).  It now runs in kScatterWarps dedicated warps of the same CTA: a compute warp
// hands its 32 rows of d(encoding) (fp16) and the samples' positions over through the global-memory region of the tile's
// own saved rgb-hidden-2 activations -- dead by then, L2-resident, 10 of its 16 KB -- publishes a counter in shared memory
// and goes on with its next tile.  One region per tile: no slot reuse, so the compute warpgroups never wait for the
// scatter.  Tried and dropped (profiles/r02_ab_bwd_scatter.txt): a CTA-wide queue of 4-level units that compute warps also
// serve while their MMAs are in flight (0.43-0.48 ms: a stolen unit holds up the warpgroup's next MMA step, and every
// unit pays the L2 latency of its inputs again), and redux.sync / match.any for the run sums (0.76 ms: both are slow with
// partial masks).
constexpr int kHandoffXyz = 16 * kRows * 4;                       // byte offset of float4 (x, y, z, valid)[128] behind half2 d[16][128]
static_assert(kHandoffXyz + kRows * 16 <= kRows * 64 * 2, "the hand-off fits the tile's rgb-hidden-2 region");
constexpr int bwd_threads(int scatter_warps) { return kWGB * 128 + scatter_warps * 32 + 32; }
// TMEM columns: per-warpgroup dA accumulators, then the five persistent weight-gradient accumulators
constexpr int kColD = 0, kColWg1 = 64 * kWGB, kColWg2 = kColWg1 + 32, kColWr1 = kColWg2 + 64, kColWr2 = kColWr1 + 32, kColWr3 = kColWr2 + 64;
static_assert(kColWr3 + 64 <= 512, "TMEM budget");

struct TcGradPtrs { float* geo[RNB_MAX_K]; float* rgb[RNB_MAX_K]; };

__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(tc::smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}

// dW (M = 64 rows of the dz tile's feature axis, N in-features) += dz^T in over the tile's 128 samples
template <int N>
__device__ __forceinline__ void tc_wgrad_mma(uint32_t d_tmem, uint32_t dz_s, uint32_t in_s, bool first) {
  constexpr uint32_t idesc = tc::instr_desc(64, N, 1, 1);
#pragma unroll
  for (int k = 0; k < kRows / 16; k++)
    tc::mma_f16(d_tmem, tc::smem_desc(dz_s + k * 256, 128, kRows * 16), tc::smem_desc(in_s + k * 256, 128, kRows * 16), idesc, !(first && k == 0));
}
// dA (128 x N) = dz (128 x KOUT, K-major) * W (KOUT out-rows x in-features tile, read MN-major from in-feature chunk c0)
template <int N, int KOUT>
__device__ __forceinline__ void tc_dgrad_mma(uint32_t d_tmem, uint32_t dz_s, uint32_t w_s, int c0) {
  constexpr uint32_t idesc = tc::instr_desc(128, N, 0, 1);
#pragma unroll
  for (int k = 0; k < KOUT / 16; k++)
    tc::mma_f16(d_tmem, tc::smem_desc(dz_s + k * 2 * (kRows * 16), kRows * 16, 128),
                tc::smem_desc(w_s + c0 * (KOUT * 16) + k * 256, 128, KOUT * 16), idesc, k > 0);
}

// relu'(h) of this thread's row of a saved 64-wide hidden activation tile, as 64 bits
__device__ __forceinline__ uint2 tc_row_mask64(const unsigned char* in_tile, int row) {
  uint32_t m[2] = {0u, 0u};
#pragma unroll
  for (int c = 0; c < 8; c++) {
    const uint4 hv = *reinterpret_cast<const uint4*>(in_tile + tc::tile_chunk_off(kRows, row, c));
    const uint32_t hw[4] = {hv.x, hv.y, hv.z, hv.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {   // post-ReLU values are >= 0: positive <=> any magnitude bit set
      m[c >> 2] |= ((hw[j] & 0x7FFFu) ? 1u : 0u) << ((c & 3) * 8 + 2 * j);
      m[c >> 2] |= ((hw[j] & 0x7FFF0000u) ? 1u : 0u) << ((c & 3) * 8 + 2 * j + 1);
    }
  }
  return make_uint2(m[0], m[1]);
}

// dZ row = dA row (64 TMEM columns) masked by relu'(h) (bit f of `mask`)
__device__ __forceinline__ void tc_mask_row64(uint32_t taddr, uint2 mask, unsigned char* dz_tile, int row) {
#pragma unroll
  for (int h = 0; h < 2; h++) {
    uint32_t r[32];
    tc::tmem_ld32_nowait(taddr + h * 32, r);
    tc::tmem_ld_wait();
    const uint32_t m = h ? mask.y : mask.x;
#pragma unroll
    for (int c = 0; c < 4; c++) {
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; j++)
        o[j] = pack2h((m >> (8 * c + 2 * j)) & 1u ? __uint_as_float(r[8 * c + 2 * j]) : 0.f,
                      (m >> (8 * c + 2 * j + 1)) & 1u ? __uint_as_float(r[8 * c + 2 * j + 1]) : 0.f);
      *reinterpret_cast<uint4*>(dz_tile + tc::tile_chunk_off(kRows, row, h * 4 + c)) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// gradient seeds enter the fp16 chain saturated; `bad` records that one did not fit (or is nan): the step guard then
// skips the optimizer step, as the reference's GradScaler does for an inf (train_ml.py:304)
__device__ __forceinline__ float tc_sat_h(float v, bool& bad) {
  bad |= !(fabsf(v) <= 65504.f);
  return fminf(fmaxf(v, -65504.f), 65504.f);
}

// weight-gradient accumulator (M = 64: row i lives in TMEM lane 32*(i/16) + i%16) -> global fp32 gradient
template <int N>
__device__ __forceinline__ void tc_flush_wgrad(uint32_t tmem_cols, int n_out, float* __restrict__ gw, float inv_gs, int warp, int lane) {
  const int o = (warp & 3) * 16 + lane;
  const uint32_t taddr = tmem_cols + ((uint32_t)((warp & 3) * 32) << 16);
#pragma unroll
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t r[32];
    tc::tmem_ld32_nowait(taddr + c0, r);      // warp-collective: every lane executes it
    tc::tmem_ld_wait();
    if (lane < 16 && o < n_out) {
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        const float a = __uint_as_float(r[j]) * inv_gs, b = __uint_as_float(r[j + 1]) * inv_gs;
        if (a != 0.f || b != 0.f) asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(gw + o * N + c0 + j), "f"(a), "f"(b) : "memory");
      }
    }
  }
}

template <int kScatterWarps>
__global__ void __launch_bounds__(bwd_threads(kScatterWarps), 1) field_bwd_tc_kernel(
    TcFieldArgs A, const __half* __restrict__ s_enc, const __half* __restrict__ s_rgb_in, const __half* __restrict__ s_geo_h,
    const __half* __restrict__ s_rgb_h1, __half* s_rgb_h2, const __half* __restrict__ s_sraw,
    const __half* __restrict__ rgb4, const float* __restrict__ dsig, const float* __restrict__ drgb, float gs,
    float* __restrict__ grad_table, TcGradPtrs G, int n_aggregate_levels, int agg_long_res, int32_t* __restrict__ flags, uint64_t l2_stream, int agg_half) {
  extern __shared__ __align__(128) unsigned char smem[];
  constexpr int kBwdThreads = bwd_threads(kScatterWarps);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool is_issuer = warp == kWGB * 4 + kScatterWarps;
  const bool is_scatter = warp >= kWGB * 4 && !is_issuer;
  const int wg = tid >> 7, row = tid & 127;
  volatile int32_t* ready = reinterpret_cast<volatile int32_t*>(smem + TcBwdSmem::ready);   // [g][w]: tiles of this pass handed over

  uint64_t* full = reinterpret_cast<uint64_t*>(smem + TcBwdSmem::bars);
  uint64_t* done = full + kWGB;
  uint64_t* loaded = done + kWGB;          // [kWGB][2]: a saved-input tile has landed in buffer b of warpgroup g
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + TcBwdSmem::tmem_slot);
  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid < kWGB) {
    tc::mbar_init(full + tid, 128); tc::mbar_init(done + tid, 1);
    tc::mbar_init(loaded + 2 * tid, 1); tc::mbar_init(loaded + 2 * tid + 1, 1);
  }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t w_s = tc::smem_u32(smem);
  const float inv_gs = 1.0f / gs;
  if (l2_stream) l2_stream = tc::l2_evict_first_policy();
  const int K = A.K;
  // per-role state in two registers (the three roles share one register allocation):
  //   issuer:    ra = parities of full[g] (bit g), rb = saved-input buffer the next step of warpgroup g reads (bit g)
  //   warpgroup: ra = parity of done[wg],          rb = saved-input tiles consumed so far (buffer = q & 1, parity = (q >> 1) & 1)
  uint32_t ra = 0, rb = 0;
#define ph_done ra
#define q_load rb
  int64_t tiles_before = 0;
  volatile int64_t* extent = reinterpret_cast<volatile int64_t*>(smem + TcBwdSmem::extent);
  for (int pass = 0; pass < K; pass++) {
    // ONE read of the pass extents per CTA: the compute warpgroups, the scatter warps and the issuer all derive their tile
    // counts from them and wait for each other on that basis -- they must agree even if the list is rewritten under the kernel
    __syncthreads();
    if (tid == 0) { extent[0] = __ldg(A.seg + pass); extent[1] = extent[0] + __ldg(A.seg + K + 1 + pass); }
    __syncthreads();
    const int64_t base = extent[0];
    int64_t pass_end = extent[1];
    if (pass_end > A.capacity) pass_end = A.capacity;
    const int64_t n_tiles = pass_end > base ? (pass_end - base + kRows - 1) / kRows : 0;
    // ---- stage this pass's weights (whole CTA) ----
    {
      const __half* gw = A.geo_w[pass];
      const __half* rw = A.rgb_w[pass];
      tc_stage_weight(smem + TcSmem::g1, gw, 64, 32, tid, kBwdThreads);
      tc_stage_weight(smem + TcSmem::g2, gw + 64 * 32, 32, 64, tid, kBwdThreads);
      tc_stage_weight(smem + TcSmem::r1, rw, 64, 32, tid, kBwdThreads);
      tc_stage_weight(smem + TcSmem::r2, rw + 64 * 32, 64, 64, tid, kBwdThreads);
      tc_stage_weight(smem + TcSmem::r3, rw + 64 * 32 + 64 * 64, 16, 64, tid, kBwdThreads);
    }
    if (tid < kWGB * 4) ready[tid] = 0;
    tc::fence_async_smem();
    __syncthreads();
    // round-robin over the grid's warpgroups, each pass continuing where the previous one stopped (see the forward kernel)
    const int64_t tile_stride = (int64_t)gridDim.x * kWGB;
    const int64_t rot = tiles_before % tile_stride;
    tiles_before += n_tiles;
    auto wg_first = [&](int g) -> int64_t {
      const int64_t t = (int64_t)blockIdx.x * kWGB + g - rot;
      return t < 0 ? t + tile_stride : t;
    };
    if (is_issuer) {
      // ---------------- issuer warp: every MMA of the CTA, in arrival order ----------------
      if (lane == 0) {
        int64_t left[kWGB]; int step[kWGB];
        int remaining = 0;
#pragma unroll
        for (int g = 0; g < kWGB; g++) {
          const int64_t t0 = wg_first(g);
          left[g] = t0 < n_tiles ? (n_tiles - t0 + tile_stride - 1) / tile_stride : 0;
          step[g] = 0;
          remaining += (int)left[g];
        }
        bool first_w[5] = {true, true, true, true, true};
        while (remaining > 0) {
#pragma unroll
          for (int g = 0; g < kWGB; g++) {
            if (left[g] == 0) continue;
            if (!mbar_test(full + g, (ra >> g) & 1u)) continue;
            ra ^= 1u << g;
            tc::fence_after_sync();
            const uint32_t dz_s = w_s + TcBwdSmem::wg0 + g * TcBwdSmem::per_wg + TcBwdSmem::dz;
            const uint32_t in_s = dz_s + TcBwdSmem::in + ((rb >> g) & 1u) * TcBwdSmem::in_bytes;
            rb ^= 1u << g;
            const uint32_t d_t = tmem + kColD + g * 64;
            switch (step[g]) {
              case 0:   // rgb layer 3: dZ (.,16), In = h2
                tc_wgrad_mma<64>(tmem + kColWr3, dz_s, in_s, first_w[4]); first_w[4] = false;
                tc_dgrad_mma<64, 16>(d_t, dz_s, w_s + TcSmem::r3, 0);
                break;
              case 1:   // rgb layer 2: dZ (.,64), In = h1
                tc_wgrad_mma<64>(tmem + kColWr2, dz_s, in_s, first_w[3]); first_w[3] = false;
                tc_dgrad_mma<64, 64>(d_t, dz_s, w_s + TcSmem::r2, 0);
                break;
              case 2:   // rgb layer 1: dZ (.,64), In = [SH | feature]; only d(feature) is needed (in-chunks 2,3)
                tc_wgrad_mma<32>(tmem + kColWr1, dz_s, in_s, first_w[2]); first_w[2] = false;
                tc_dgrad_mma<16, 64>(d_t, dz_s, w_s + TcSmem::r1, 2);
                break;
              case 3:   // geo layer 2: dZ (.,32), In = geo hidden
                tc_wgrad_mma<64>(tmem + kColWg2, dz_s, in_s, first_w[1]); first_w[1] = false;
                tc_dgrad_mma<64, 32>(d_t, dz_s, w_s + TcSmem::g2, 0);
                break;
              default:  // geo layer 1: dZ (.,64), In = encoding
                tc_wgrad_mma<32>(tmem + kColWg1, dz_s, in_s, first_w[0]); first_w[0] = false;
                tc_dgrad_mma<32, 64>(d_t, dz_s, w_s + TcSmem::g1, 0);
                break;
            }
            tc::mma_commit(done + g);
            if (++step[g] == 5) { step[g] = 0; left[g]--; remaining--; }
          }
        }
      }
      __syncwarp();
    } else if (is_scatter) {
      // ---------------- scatter warps ----------------
      // Unit of work: 32 rows (one compute warp's share) of one tile.  Unit u = n * 16 + g * 4 + w is the w-th quarter of the
      // n-th tile of warpgroup g in this pass; scatter warp j takes the units u = j (mod kScatterWarps) in ascending order
      // and waits for each one's hand-over (the warpgroups advance at about the same pace, so in-order service costs little
      // and needs no bookkeeping).
      for (int u = warp - kWGB * 4; ; u += kScatterWarps) {
        const int n = u >> 4, g = (u >> 2) & 3, w = u & 3;
        if ((int64_t)n * tile_stride >= n_tiles) break;
        const int64_t tile = wg_first(g) + (int64_t)n * tile_stride;
        if (tile >= n_tiles) continue;
        while (ready[g * 4 + w] <= n) __nanosleep(100);
        __threadfence_block();                                   // acquire: the rows written before the counter was bumped
        const char* hand = reinterpret_cast<const char*>(s_rgb_h2 + (base + tile * kRows) * 64);
        const uint32_t* hd = reinterpret_cast<const uint32_t*>(hand) + w * 32 + lane;       // + 128 per level
        // d(encoding) comes back from L2 (~700 cycles, about what a level takes): a ring of four keeps the loads of levels
        // l + 1 .. l + 4 in flight.  The per-level constants are indexed loads from the parameter bank (~100 cycles): one ahead.
        uint32_t d_a = __ldcg(hd), d_b = __ldcg(hd + kRows), d_c = __ldcg(hd + 2 * kRows), d_d = __ldcg(hd + 3 * kRows);
        const float4 q = __ldcg(reinterpret_cast<const float4*>(hand + kHandoffXyz) + w * 32 + lane);
        const bool valid = q.w != 0.f;
        float sc_n = A.g.scale[0];
        uint32_t res_n = A.g.res[0], off_n = A.g.offset[0], end_n = A.g.offset[1];
#pragma unroll 1
        for (int l = 0; l < 16; l++) {
          const uint32_t off = off_n, res_l = res_n, end_l = end_n, dw = d_a;
          const float sc_l = sc_n;
          d_a = d_b; d_b = d_c; d_c = d_d;
          if (l < 12) d_d = __ldcg(hd + (l + 4) * kRows);
          if (l < 15) { sc_n = A.g.scale[l + 1]; res_n = A.g.res[l + 1]; off_n = end_l; end_n = A.g.offset[l + 2]; }
          const LevelCorners L = level_corners(q.x, q.y, q.z, sc_l, res_l, end_l - off);
          const float2 d = __half22float2(*reinterpret_cast<const __half2*>(&dw));
          const float d0 = d.x * inv_gs, d1 = d.y * inv_gs;
          float* gl = grad_table + 2 * (size_t)off;
          if (l < n_aggregate_levels) {   // warp-uniform
            if (agg_half) {
              if (res_l <= (uint32_t)agg_long_res) scatter_level_aggregated_h2<5>(gl, L, *reinterpret_cast<const __half2*>(&dw), 32.f * inv_gs, valid);
              else scatter_level_aggregated_h2<3>(gl, L, *reinterpret_cast<const __half2*>(&dw), 32.f * inv_gs, valid);
            } else if (res_l <= (uint32_t)agg_long_res) scatter_level_aggregated<5>(gl, L, d0, d1, valid);
            else scatter_level_aggregated<3>(gl, L, d0, d1, valid);
          } else if (valid) {
            // x-neighbour corners that share an aligned pair of entries (see tc_encode_level) take ONE 16-byte reduction
#pragma unroll
            for (int p = 0; p < 4; p++) {
              const uint32_t ia = L.idx[2 * p], ib = L.idx[2 * p + 1];
              const float wa = L.w[2 * p], wb = L.w[2 * p + 1];
              if ((ia ^ ib) == 1u) {
                const bool odd = ia & 1u;
                const float lo_w = odd ? wb : wa, hi_w = odd ? wa : wb;
                red_add_f32x4(gl + 2 * (size_t)(ia & ~1u), lo_w * d0, lo_w * d1, hi_w * d0, hi_w * d1);
              } else {
                red_add_f32x2(gl + 2 * (size_t)ia, wa * d0, wa * d1);
                red_add_f32x2(gl + 2 * (size_t)ib, wb * d0, wb * d1);
              }
            }
          }
        }
      }
    } else {
      // ---------------- warpgroups: thread = sample ----------------
      unsigned char* dzb = smem + TcBwdSmem::wg0 + wg * TcBwdSmem::per_wg + TcBwdSmem::dz;
      unsigned char* inb = dzb + TcBwdSmem::in;
      const uint32_t inb_s = tc::smem_u32(inb);
      uint64_t* ld = loaded + 2 * wg;
      const uint32_t tmem_row = tmem + kColD + wg * 64 + ((uint32_t)((warp & 3) * 32) << 16);
      const int64_t t0 = wg_first(wg);
      const int64_t my_tiles = t0 < n_tiles ? (n_tiles - t0 + tile_stride - 1) / tile_stride : 0;
      // Saved-input tiles arrive by bulk copy in the order the layers consume them: per tile h2, h1, rgb_in, geo_h, enc.
      // The forward kernel stored them in the operand-tile layout, so they are MMA operands as they land.  Load number
      // q (counted over the whole kernel) goes to buffer q & 1; it may be issued once the MMAs of load q - 2 are done.
      auto issue_load = [&](int64_t n, int j, uint32_t q) {     // elected thread only
        if (n >= my_tiles) return;
        const int64_t s0 = base + (t0 + n * tile_stride) * kRows;
        const __half* src = j == 0 ? s_rgb_h2 + s0 * 64 : j == 1 ? s_rgb_h1 + s0 * 64 : j == 2 ? s_rgb_in + s0 * 32 : j == 3 ? s_geo_h + s0 * 64
                                                                                                                            : s_enc + s0 * 32;
        const uint32_t bytes = (j == 2 || j == 4) ? kRows * 32 * 2 : kRows * 64 * 2;
        tc::mbar_expect_tx(ld + (q & 1), bytes);
        if (l2_stream) tc::bulk_load_hint(inb_s + (q & 1) * TcBwdSmem::in_bytes, src, bytes, ld + (q & 1), l2_stream);
        else tc::bulk_load(inb_s + (q & 1) * TcBwdSmem::in_bytes, src, bytes, ld + (q & 1));
      };
      auto wait_load = [&]() -> const unsigned char* {          // every thread: the tile of load q_load is in shared memory
        const unsigned char* p = inb + (q_load & 1) * TcBwdSmem::in_bytes;
        tc::mbar_wait(ld + (q_load & 1), (q_load >> 1) & 1);
        q_load++;
        return p;
      };
      auto wait_mma = [&]() { tc::mbar_wait(done + wg, ph_done); ph_done ^= 1; };
      if (row == 0) { issue_load(0, 0, q_load); issue_load(0, 1, q_load + 1); }
      for (int64_t n = 0; n < my_tiles; n++) {
        const int64_t s = base + (t0 + n * tile_stride) * kRows + row;
        const bool valid = s < pass_end;
        bool bad = false;
        // ---- step 0: rgb layer 3 (sigmoid): dZ = dL/drgb * gs * y (1 - y) ----
        {
          uint4 q = make_uint4(0u, 0u, 0u, 0u);
          if (valid) {
            const uint2 yq = __ldg(reinterpret_cast<const uint2*>(rgb4 + s * 4));
            const float2 y01 = __half22float2(*reinterpret_cast<const __half2*>(&yq.x));
            const float2 y23 = __half22float2(*reinterpret_cast<const __half2*>(&yq.y));
            const float d0 = __ldg(drgb + 3 * s), d1 = __ldg(drgb + 3 * s + 1), d2 = __ldg(drgb + 3 * s + 2);
            q.x = pack2h(tc_sat_h(d0 * gs * y01.x * (1.f - y01.x), bad), tc_sat_h(d1 * gs * y01.y * (1.f - y01.y), bad));
            q.y = pack2h(tc_sat_h(d2 * gs * y23.x * (1.f - y23.x), bad), 0.f);
          }
          *reinterpret_cast<uint4*>(dzb + tc::tile_chunk_off(kRows, row, 0)) = q;
          *reinterpret_cast<uint4*>(dzb + tc::tile_chunk_off(kRows, row, 1)) = make_uint4(0u, 0u, 0u, 0u);
        }
        uint2 mask = tc_row_mask64(wait_load(), row);                                     // relu'(h2)
        tc::fence_async_smem(); tc::fence_before_sync(); tc::mbar_arrive(full + wg);
        wait_mma(); tc::fence_after_sync();
        if (row == 0) issue_load(n, 2, q_load + 1);
        // ---- step 1: rgb layer 2 ----
        tc_mask_row64(tmem_row, mask, dzb, row);
        mask = tc_row_mask64(wait_load(), row);                                           // relu'(h1)
        tc::fence_async_smem(); tc::fence_before_sync(); tc::mbar_arrive(full + wg);
        wait_mma(); tc::fence_after_sync();
        if (row == 0) issue_load(n, 3, q_load + 1);
        // ---- step 2: rgb layer 1 ----
        tc_mask_row64(tmem_row, mask, dzb, row);
        wait_load();                                                                      // [SH | feature]: no activation in front
        tc::fence_async_smem(); tc::fence_before_sync(); tc::mbar_arrive(full + wg);
        float dsr = 0.f;
        if (valid) {   // TruncExp backward: g * exp(clamp(x, -15, 15))  (custom_functions.py:170-173)
          const float raw = __half2float(s_sraw[s]);
          dsr = tc_sat_h(__ldg(dsig + s) * gs * expf(fminf(fmaxf(raw, -15.f), 15.f)), bad);
        }
        wait_mma(); tc::fence_after_sync();
        if (row == 0) issue_load(n, 4, q_load + 1);
        // ---- step 3: geo layer 2: dZ = [d raw sigma | d feature (16) | 0 ...] ----
        {
          float v[16];
          tc::tmem_ld16(tmem_row, v);
          uint4 q0, q1, q2;
          q0.x = pack2h(dsr, v[0]); q0.y = pack2h(v[1], v[2]); q0.z = pack2h(v[3], v[4]); q0.w = pack2h(v[5], v[6]);
          q1.x = pack2h(v[7], v[8]); q1.y = pack2h(v[9], v[10]); q1.z = pack2h(v[11], v[12]); q1.w = pack2h(v[13], v[14]);
          q2.x = pack2h(v[15], 0.f); q2.y = 0u; q2.z = 0u; q2.w = 0u;
          *reinterpret_cast<uint4*>(dzb + tc::tile_chunk_off(kRows, row, 0)) = q0;
          *reinterpret_cast<uint4*>(dzb + tc::tile_chunk_off(kRows, row, 1)) = q1;
          *reinterpret_cast<uint4*>(dzb + tc::tile_chunk_off(kRows, row, 2)) = q2;
          *reinterpret_cast<uint4*>(dzb + tc::tile_chunk_off(kRows, row, 3)) = make_uint4(0u, 0u, 0u, 0u);
        }
        mask = tc_row_mask64(wait_load(), row);                                           // relu'(geo hidden)
        tc::fence_async_smem(); tc::fence_before_sync(); tc::mbar_arrive(full + wg);
        wait_mma(); tc::fence_after_sync();
        if (row == 0) issue_load(n + 1, 0, q_load + 1);                                   // next tile's first input
        // ---- step 4: geo layer 1 ----
        tc_mask_row64(tmem_row, mask, dzb, row);
        wait_load();                                                                      // encoding
        tc::fence_async_smem(); tc::fence_before_sync(); tc::mbar_arrive(full + wg);
        float x = 0.f, y = 0.f, z = 0.f;      // the sample's normalised position, for the scatter below
        if (valid) {
          const int ry = __ldg(A.ray_id + s);
          const float t = __ldg(A.ts + s);
          x = fminf(fmaxf(__fdiv_rn(__fmaf_rn(__ldg(A.rays_d + 3 * ry), t, __ldg(A.rays_o + 3 * ry)) - A.xmin[0], A.xden[0]), 0.f), 1.f);
          y = fminf(fmaxf(__fdiv_rn(__fmaf_rn(__ldg(A.rays_d + 3 * ry + 1), t, __ldg(A.rays_o + 3 * ry + 1)) - A.xmin[1], A.xden[1]), 0.f), 1.f);
          z = fminf(fmaxf(__fdiv_rn(__fmaf_rn(__ldg(A.rays_d + 3 * ry + 2), t, __ldg(A.rays_o + 3 * ry + 2)) - A.xmin[2], A.xden[2]), 0.f), 1.f);
        }
        wait_mma(); tc::fence_after_sync();
        if (row == 0) issue_load(n + 1, 1, q_load + 1);
        // ---- step 5: d(encoding) row (fp16, like the operator path) -> the scatter warps ----
        {
          uint32_t r[32];
          tc::tmem_ld32_nowait(tmem_row, r);
          tc::tmem_ld_wait();
          tc::fence_before_sync();
          char* hand = reinterpret_cast<char*>(s_rgb_h2 + (s - row) * 64);      // this tile's rgb-hidden-2 region: consumed in steps 0 and 1
          uint32_t* hd = reinterpret_cast<uint32_t*>(hand) + row;
#pragma unroll
          for (int l = 0; l < 16; l++) {
            const uint32_t w = pack2h(__uint_as_float(r[2 * l]), __uint_as_float(r[2 * l + 1]));
            // an inf / nan in d(encoding) (an overflow anywhere up the fp16 chain ends here) would reach the table gradient
            const uint32_t e = (w & 0x7c007c00u) ^ 0x7c007c00u;
            bad |= valid && (!(e & 0xffffu) || !(e >> 16));
            __stcg(hd + l * kRows, w);                                         // warp: 128 contiguous bytes per level
          }
          __stcg(reinterpret_cast<float4*>(hand + kHandoffXyz) + row, make_float4(x, y, z, valid ? 1.f : 0.f));
          if (flags != nullptr && __any_sync(kFullMask, bad) && lane == 0) atomicOr(flags, RNB_FLAG_NONFINITE);
          __threadfence_block();
          __syncwarp();
          if (lane == 0) ready[wg * 4 + (warp & 3)] = (int32_t)n + 1;
        }
      }
    }
    // ---- end of pass: weight gradients TMEM -> global (once per pass and CTA) ----
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    bool cta_had_tiles = false;
#pragma unroll
    for (int g = 0; g < kWGB; g++) cta_had_tiles |= wg_first(g) < n_tiles;
    if (warp < 4 && cta_had_tiles) {
      float* gg = G.geo[pass]; float* gr = G.rgb[pass];
      tc_flush_wgrad<32>(tmem + kColWg1, 64, gg, inv_gs, warp, lane);
      tc_flush_wgrad<64>(tmem + kColWg2, 32, gg + 64 * 32, inv_gs, warp, lane);
      tc_flush_wgrad<32>(tmem + kColWr1, 64, gr, inv_gs, warp, lane);
      tc_flush_wgrad<64>(tmem + kColWr2, 64, gr + 64 * 32, inv_gs, warp, lane);
      tc_flush_wgrad<64>(tmem + kColWr3, 16, gr + 64 * 32 + 64 * 64, inv_gs, warp, lane);
    }
    tc::fence_before_sync();
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
#undef ph_done
#undef q_load
}

}  // namespace rnb

using namespace rnb;

// L2 eviction policy of the saved-activation tiles.  They are written once by the forward kernel and read once by the backward
// kernel, 260 MB per step -- twice the L2 -- and on their way through they push out what the kernels live off: the fp16
// table (23 MB, gathered 128 times per sample) and the fp32 gradient table (46 MB, the target of every reduction; a
// reduction that misses L2 waits for DRAM).  evict_first on the bulk copies keeps those resident.  RNB_L2_STREAM=0: off.
static uint64_t l2_stream_policy() {
  static int on = -1;
  if (on < 0) { on = 1; if (const char* e = getenv("RNB_L2_STREAM")) on = atoi(e) != 0; }
  return on ? 1ull : 0ull;        // (a flag: the kernels build the policy word themselves with createpolicy)
}

// argument checks shared by the entry points (the kernels are specialised for L = 16 levels, F = 2: 32 features)
static int check_field_args(const rnb_field_params* P, const rnb_samples* S, const char* what) {
  RNB_REQUIRE(P != nullptr && S != nullptr && P->K >= 1 && P->K <= RNB_MAX_K && P->grid.n_levels == 16, what);
  RNB_REQUIRE(P->table_f16 != nullptr, what);
  for (int i = 0; i < P->K; i++) RNB_REQUIRE(P->geo_w_f16[i] && P->rgb_w_f16[i], what);
  RNB_REQUIRE(S->rays_o && S->rays_d && S->ray_id && S->ts && S->seg, what);
  RNB_REQUIRE(S->capacity > 0 && S->capacity % kRows == 0, what);
  return RNB_OK;
}

extern "C" {

int rnb_field_fwd(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved, float* sigmas,
                  void* rgb4_f16, rnb_stream_t stream) {
  const char* what = "rnb_field_fwd";
  TcFieldArgs A;
  if (int rc = check_field_args(P, S, what)) return rc;
  RNB_REQUIRE(sigmas && rgb4_f16, what);
  if (saved != nullptr && saved->enc != nullptr) RNB_REQUIRE(saved->rgb_in && saved->geo_h && saved->rgb_h1 && saved->rgb_h2 && saved->sraw, what);
  if (int rc = to_levels(&P->grid, &A.g, what)) return rc;
  for (int c = 0; c < 3; c++) { A.xmin[c] = P->xyz_min[c]; A.xden[c] = P->xyz_max[c] - P->xyz_min[c]; }
  A.table = (const __half2*)P->table_f16;
  for (int i = 0; i < P->K; i++) { A.geo_w[i] = (const __half*)P->geo_w_f16[i]; A.rgb_w[i] = (const __half*)P->rgb_w_f16[i]; }
  A.K = P->K;
  A.rays_o = S->rays_o; A.rays_d = S->rays_d; A.ray_id = S->ray_id; A.ts = S->ts; A.seg = S->seg; A.capacity = S->capacity;
  const bool save = saved != nullptr && saved->enc != nullptr;
  static bool attr_set = false;
  if (!attr_set) {
    RNB_CUDA(cudaFuncSetAttribute(field_fwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem::bytes), what);
    RNB_CUDA(cudaFuncSetAttribute(field_fwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem::bytes), what);
    attr_set = true;
  }
  const int64_t tiles = S->capacity / kRows;
  int grid = (int)((tiles + kWG - 1) / kWG);
  if (grid > kNumSMs) grid = kNumSMs;
  if (save)
    field_fwd_tc_kernel<true><<<grid, kWG * 128, TcSmem::bytes, as_stream(stream)>>>(A, (__half*)saved->enc, (__half*)saved->rgb_in, (__half*)saved->geo_h,
                                                                                    (__half*)saved->rgb_h1, (__half*)saved->rgb_h2, (__half*)saved->sraw,
                                                                                    sigmas, (__half*)rgb4_f16, l2_stream_policy());
  else
    field_fwd_tc_kernel<false><<<grid, kWG * 128, TcSmem::bytes, as_stream(stream)>>>(A, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, sigmas,
                                                                                     (__half*)rgb4_f16, 0);
  return check_launch(what);
}

int rnb_field_bwd_guarded(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved, const void* rgb4_f16,
                          const float* dL_dsigmas, const float* dL_drgbs, float loss_scale, float* grad_table,
                          float* const* grad_geo, float* const* grad_rgb, int32_t* flags, rnb_stream_t stream) {
  const char* what = "rnb_field_bwd";
  TcFieldArgs A;
  if (int rc = check_field_args(P, S, what)) return rc;
  RNB_REQUIRE(saved && saved->enc && saved->rgb_in && saved->geo_h && saved->rgb_h1 && saved->rgb_h2 && saved->sraw, what);
  RNB_REQUIRE(rgb4_f16 && dL_dsigmas && dL_drgbs && grad_table && grad_geo && grad_rgb && loss_scale > 0.f, what);
  for (int i = 0; i < P->K; i++) RNB_REQUIRE(grad_geo[i] && grad_rgb[i], what);
  if (int rc = to_levels(&P->grid, &A.g, what)) return rc;
  for (int c = 0; c < 3; c++) { A.xmin[c] = P->xyz_min[c]; A.xden[c] = P->xyz_max[c] - P->xyz_min[c]; }
  A.table = (const __half2*)P->table_f16;
  TcGradPtrs G;
  for (int i = 0; i < P->K; i++) {
    A.geo_w[i] = (const __half*)P->geo_w_f16[i]; A.rgb_w[i] = (const __half*)P->rgb_w_f16[i];
    G.geo[i] = grad_geo[i]; G.rgb[i] = grad_rgb[i];
  }
  A.K = P->K;
  A.rays_o = S->rays_o; A.rays_d = S->rays_d; A.ray_id = S->ray_id; A.ts = S->ts; A.seg = S->seg; A.capacity = S->capacity;
  // levels whose cells are coarser than a few marching steps: runs of consecutive samples of a ray share corner
  // entries there, so the warp combines them before issuing reductions.  What counts is the cell size in WORLD units
  // against the step: the resolution thresholds tuned on the unit box (128 / 48: profiles/r01_ab_agg_levels.txt) grow
  // with the extent of the encoded box -- in a scale-16 scene most samples sit near the centre, where a res-921 cell
  // still holds ~6 of them, and without combining them the reductions serialise on a few L2 addresses
  // (field_bwd 1.06 -> 0.70 ms on the 6-cascade workload: profiles/r01_ab_agg_levels.txt)
  const float extent = P->xyz_max[0] - P->xyz_min[0];
  const float grow = fmaxf(1.0f, 1.3f * sqrtf(fmaxf(extent, 0.0f)));
  int n_agg = 0;
  for (int l = 0; l < A.g.n_levels; l++)
    if ((float)A.g.res[l] <= 340.0f * grow) n_agg = l + 1;      // (128 before the scatter moved to its own warps: there a combined level costs issue slots, not reductions)
  if (const char* e = getenv("RNB_AGG_LEVELS")) n_agg = atoi(e);   // tuning knob (profiles/): levels with warp-aggregated reductions
  if (n_agg > A.g.n_levels) n_agg = A.g.n_levels;
  while (n_agg > 0 && A.g.res[n_agg - 1] >= 1024) n_agg--;          // the cell key holds 10-11 bits per axis
  int agg_long_res = (int)(48.0f * grow);                           // levels up to this resolution combine whole-warp runs
  if (const char* e = getenv("RNB_AGG_LONG_RES")) agg_long_res = atoi(e);
  // scatter warps per CTA: 3 (20 warps, 96 registers) or 7 (24 warps, 80 registers) -- warps are allocated in fours, so a
  // 21-warp CTA would be charged 24 (RNB_BWD_SCATTER_WARPS: A/B knob, profiles/)
  static int n_scatter = 0, agg_half = 1;
  if (n_scatter == 0) {
    if (const char* e = getenv("RNB_AGG_HALF")) agg_half = atoi(e);      // run reduction on packed fp16 pairs (0: fp32, A/B)
    n_scatter = 7;
    if (const char* e = getenv("RNB_BWD_SCATTER_WARPS")) n_scatter = atoi(e) == 3 ? 3 : 7;
    RNB_CUDA(cudaFuncSetAttribute(field_bwd_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcBwdSmem::bytes), what);
    RNB_CUDA(cudaFuncSetAttribute(field_bwd_tc_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcBwdSmem::bytes), what);
  }
  const uint64_t l2_stream = l2_stream_policy();
  const int64_t tiles = S->capacity / kRows;
  int grid = (int)((tiles + kWGB - 1) / kWGB);
  if (grid > kNumSMs) grid = kNumSMs;
#define RNB_LAUNCH_BWD(SW)                                                                                                              \
  field_bwd_tc_kernel<SW><<<grid, bwd_threads(SW), TcBwdSmem::bytes, as_stream(stream)>>>(                                                \
      A, (const __half*)saved->enc, (const __half*)saved->rgb_in, (const __half*)saved->geo_h, (const __half*)saved->rgb_h1,              \
      (__half*)saved->rgb_h2, (const __half*)saved->sraw, (const __half*)rgb4_f16, dL_dsigmas, dL_drgbs, loss_scale, grad_table, G, n_agg, \
      agg_long_res, flags, l2_stream, agg_half)
  if (n_scatter == 7) RNB_LAUNCH_BWD(7); else RNB_LAUNCH_BWD(3);
#undef RNB_LAUNCH_BWD
  return check_launch(what);
}

int rnb_field_bwd(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved, const void* rgb4_f16,
                  const float* dL_dsigmas, const float* dL_drgbs, float loss_scale, float* grad_table,
                  float* const* grad_geo, float* const* grad_rgb, rnb_stream_t stream) {
  return rnb_field_bwd_guarded(P, S, saved, rgb4_f16, dL_dsigmas, dL_drgbs, loss_scale, grad_table, grad_geo, grad_rgb, nullptr, stream);
}

}  // extern "C"

// Density of grid cells for rnb_grid_update (grid.cu).  span/half: per-cascade scalars computed by the caller.
int rnb_density_cells_tc_launch(const rnb_field_params* P, float* const* tmp, int C, int G, int64_t M, const float* span, const float* half,
                                const int32_t* cell_idx, const float* noise, uint64_t seed, rnb_stream_t stream) {
  const char* what = "rnb_grid_update(density)";
  TcFieldArgs A;
  RNB_REQUIRE(P != nullptr && P->K >= 1 && P->K <= RNB_MAX_K && P->grid.n_levels == 16 && C >= 1 && C <= 8 && M >= 1, what);
  if (int rc = to_levels(&P->grid, &A.g, what)) return rc;
  for (int c = 0; c < 3; c++) { A.xmin[c] = P->xyz_min[c]; A.xden[c] = P->xyz_max[c] - P->xyz_min[c]; }
  A.table = (const __half2*)P->table_f16;
  GridCellArgs Q;
  for (int i = 0; i < P->K; i++) { A.geo_w[i] = (const __half*)P->geo_w_f16[i]; A.rgb_w[i] = nullptr; Q.tmp[i] = tmp[i]; }
  A.K = P->K;
  A.rays_o = nullptr; A.rays_d = nullptr; A.ray_id = nullptr; A.ts = nullptr; A.seg = nullptr; A.capacity = 0;
  for (int c = 0; c < C; c++) { Q.span[c] = span[c]; Q.half[c] = half[c]; }
  Q.cell_idx = cell_idx; Q.noise = noise; Q.C = C; Q.G = G; Q.M = M; Q.seed = seed;
  static bool attr_set = false;
  if (!attr_set) {
    RNB_CUDA(cudaFuncSetAttribute(density_cells_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem::bytes), what);
    attr_set = true;
  }
  const int64_t tiles = ((int64_t)C * M + kRows - 1) / kRows;
  int grid = (int)((tiles + kWG - 1) / kWG);
  if (grid > kNumSMs) grid = kNumSMs;
  density_cells_tc_kernel<<<grid, kWG * 128, TcSmem::bytes, as_stream(stream)>>>(A, Q);
  return check_launch(what);
}
