// grid.cu -- occupancy-grid maintenance of all K sub-NeRFs without host synchronisation.
//
// Replaces MNGP.update_density_grid (models/networks.py:374-409; NGP :171-201) with its helpers get_all_cells
// (:330-342) and sample_uniform_and_occupied_cells (:344-372).  The reference runs, per sub-NeRF and cascade:
// randint / morton3D / nonzero (host sync) / randint / morton3D_invert / cat, the density network on the points
// (tcnn encode + MLP + ATen glue), an index_put into a zero-filled scratch grid, where/maximum, a masked mean with
// .item() (host sync) and packbits.  Here, for all K x C at once:
//   [steady state only] occ_count / occ_scan / occ_write   ordered compaction of {cell : grid > threshold} (= torch.nonzero)
//                       cells_kernel                        M/2 uniform cells + M/2 picks from the occupied list
//   density_cells_tc_kernel (field_tc.cu)                   jittered point -> hash encode -> geo MLP -> exp -> scratch grid
//   ema_kernel                                              grid = grid < 0 ? grid : max(grid * decay, scratch); partial sums
//   mean_kernel                                             mean over grid > 0, fixed summation order (deterministic)
//   packbits (intersect.cu)                                 threshold min(mean, density_threshold) read from device memory
#include "common.cuh"

namespace rnb {

constexpr int kOccBlock = 1024;   // cells per compaction block

struct GridPtrs { float* g[RNB_MAX_K]; };
struct TmpPtrs { float* t[RNB_MAX_K]; };

__device__ __forceinline__ uint32_t g_expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ __forceinline__ float g_counter_uniform(uint64_t seed, uint64_t ctr) {
  uint64_t z = seed + (ctr + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(uint32_t)(z >> 40) * (1.0f / 16777216.0f);
}

// ---- ordered compaction of occupied cells, per (sub-NeRF, cascade) ------------------------------------------
// block b of plane p = i*C + c counts its kOccBlock cells
__global__ void __launch_bounds__(256) occ_count_kernel(GridPtrs Gp, int C, int64_t G3, float thr, int nb, int32_t* __restrict__ block_counts) {
  const int p = blockIdx.y, b = blockIdx.x;
  const float* g = Gp.g[p / C] + (int64_t)(p % C) * G3 + (int64_t)b * kOccBlock;
  int n = 0;
#pragma unroll
  for (int k = 0; k < kOccBlock / 256; k++) n += g[k * 256 + threadIdx.x] > thr ? 1 : 0;
  n = warp_sum_i(n);
  __shared__ int ws[8];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; w++) t += ws[w];
    block_counts[(int64_t)p * nb + b] = t;
  }
}
// one CTA per plane: exclusive scan of the nb block counts (in place) and the plane total
__global__ void __launch_bounds__(1024) occ_scan_kernel(int nb, int32_t* __restrict__ block_counts, int32_t* __restrict__ occ_count) {
  __shared__ int wsum[32];
  __shared__ int carry;
  const int p = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int32_t* bc = block_counts + (int64_t)p * nb;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < nb ? bc[i] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(kFullMask, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int s = wsum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(kFullMask, s, o); if (lane >= o) s += y; }
      wsum[lane] = s;
    }
    __syncthreads();
    const int before = carry + (warp ? wsum[warp - 1] : 0) + x - v;
    if (i < nb) bc[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) occ_count[p] = carry;
}
// block b writes its occupied cells in ascending order at its scanned offset
__global__ void __launch_bounds__(256) occ_write_kernel(GridPtrs Gp, int C, int64_t G3, float thr, int nb, const int32_t* __restrict__ block_offsets,
                                                        int32_t* __restrict__ occ_list) {
  __shared__ int wcnt[kOccBlock / 32];
  const int p = blockIdx.y, b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* g = Gp.g[p / C] + (int64_t)(p % C) * G3 + (int64_t)b * kOccBlock;
  unsigned bal[kOccBlock / 256];
#pragma unroll
  for (int k = 0; k < kOccBlock / 256; k++) {     // chunk (k*8 + warp) covers cells [(k*8+warp)*32, +32)
    bal[k] = __ballot_sync(kFullMask, g[(k * 8 + warp) * 32 + lane] > thr);
    if (lane == 0) wcnt[k * 8 + warp] = __popc(bal[k]);
  }
  __syncthreads();
  int32_t* out = occ_list + (int64_t)p * G3 + block_offsets[(int64_t)p * nb + b];
#pragma unroll
  for (int k = 0; k < kOccBlock / 256; k++) {
    const int chunk = k * 8 + warp;
    int before = 0;
    for (int q = 0; q < chunk; q++) before += wcnt[q];
    if (bal[k] & (1u << lane)) out[before + __popc(bal[k] & ((1u << lane) - 1u))] = (int32_t)(b * kOccBlock + chunk * 32 + lane);
  }
}

// cells to evaluate in steady state (networks.py:344-372): per plane, M1 uniform cells then M1 picks among the occupied
__global__ void __launch_bounds__(256) cells_kernel(int planes, int64_t M1, int G, int64_t G3, const int32_t* __restrict__ coords1,
                                                    const float* __restrict__ pick_u, uint64_t seed, const int32_t* __restrict__ occ_count,
                                                    const int32_t* __restrict__ occ_list, int32_t* __restrict__ cell_idx) {
  const int64_t total = (int64_t)planes * 2 * M1;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int p = (int)(e / (2 * M1));
    const int64_t m = e - (int64_t)p * 2 * M1;
    int32_t idx;
    if (m < M1) {
      const int64_t q = (int64_t)p * M1 + m;
      uint32_t c[3];
#pragma unroll
      for (int k = 0; k < 3; k++)
        c[k] = coords1 ? (uint32_t)coords1[3 * q + k] : min((uint32_t)(g_counter_uniform(seed ^ 0xC0FFEEull, 3 * (uint64_t)q + k) * (float)G), (uint32_t)(G - 1));
      idx = (int32_t)(g_expand_bits(c[0]) | (g_expand_bits(c[1]) << 1) | (g_expand_bits(c[2]) << 2));
    } else {
      const int64_t q = (int64_t)p * M1 + (m - M1);
      const int n = occ_count[p];
      if (n == 0) {
        idx = -1;                                                   // no occupied cell yet: the reference keeps the uniform half only
      } else {
        const float u = pick_u ? pick_u[q] : g_counter_uniform(seed ^ 0xF00Dull, (uint64_t)q);
        idx = occ_list[(int64_t)p * G3 + min((int)(u * (float)n), n - 1)];
      }
    }
    cell_idx[e] = idx;
  }
}

// ---- EMA-max update + masked-mean partials ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) ema_kernel(GridPtrs Gp, TmpPtrs Tp, int64_t cells, float decay, int zero_tmp, double* __restrict__ part_sum,
                                                  int64_t* __restrict__ part_cnt) {
  const int i = blockIdx.y;
  float* g = Gp.g[i];
  float* t = Tp.t[i];
  double s = 0.0;
  int64_t n = 0;
  for (int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; j < cells; j += (int64_t)gridDim.x * blockDim.x * 4) {
    float4 a = *reinterpret_cast<const float4*>(g + j);
    const float4 b = *reinterpret_cast<const float4*>(t + j);
    a.x = a.x < 0.f ? a.x : fmaxf(a.x * decay, b.x); a.y = a.y < 0.f ? a.y : fmaxf(a.y * decay, b.y);
    a.z = a.z < 0.f ? a.z : fmaxf(a.z * decay, b.z); a.w = a.w < 0.f ? a.w : fmaxf(a.w * decay, b.w);
    *reinterpret_cast<float4*>(g + j) = a;
    if (zero_tmp) *reinterpret_cast<float4*>(t + j) = make_float4(0.f, 0.f, 0.f, 0.f);   // the scratch grid starts the next update at zero
    if (a.x > 0.f) { s += a.x; n++; }
    if (a.y > 0.f) { s += a.y; n++; }
    if (a.z > 0.f) { s += a.z; n++; }
    if (a.w > 0.f) { s += a.w; n++; }
  }
  __shared__ double ss[256];
  __shared__ int64_t sn[256];
  ss[threadIdx.x] = s; sn[threadIdx.x] = n;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { ss[threadIdx.x] += ss[threadIdx.x + o]; sn[threadIdx.x] += sn[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { part_sum[(int64_t)i * gridDim.x + blockIdx.x] = ss[0]; part_cnt[(int64_t)i * gridDim.x + blockIdx.x] = sn[0]; }
}
__global__ void __launch_bounds__(256) mean_kernel(int nparts, const double* __restrict__ part_sum, const int64_t* __restrict__ part_cnt,
                                                   float* __restrict__ mean) {
  const int i = blockIdx.x;
  __shared__ double ss[256];
  __shared__ int64_t sn[256];
  double s = 0.0;
  int64_t n = 0;
  for (int k = threadIdx.x; k < nparts; k += 256) { s += part_sum[(int64_t)i * nparts + k]; n += part_cnt[(int64_t)i * nparts + k]; }
  ss[threadIdx.x] = s; sn[threadIdx.x] = n;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { ss[threadIdx.x] += ss[threadIdx.x + o]; sn[threadIdx.x] += sn[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) mean[i] = sn[0] > 0 ? (float)(ss[0] / (double)sn[0]) : 0.f;
}

}  // namespace rnb

using namespace rnb;

int rnb_density_cells_tc_launch(const rnb_field_params* P, float* const* tmp, int C, int G, int64_t M, const float* span, const float* half,
                                const int32_t* cell_idx, const float* noise, uint64_t seed, rnb_stream_t stream);

extern "C" {

int rnb_grid_update_layout_of(const rnb_engine_config* c, rnb_grid_layout* L) {
  const char* what = "rnb_grid_update_layout_of";
  RNB_REQUIRE(c && L && c->K >= 1 && c->K <= RNB_MAX_K && c->cascades >= 1 && c->cascades <= 8 && c->grid_size >= 16, what);
  const int64_t K = c->K, C = c->cascades, G3 = (int64_t)c->grid_size * c->grid_size * c->grid_size;
  RNB_REQUIRE(G3 % kOccBlock == 0, what);
  int64_t off = 0;
  auto take = [&](int64_t bytes) { const int64_t o = off; off = (off + bytes + 255) / 256 * 256; return o; };
  L->tmp = take(K * C * G3 * 4);
  L->cell_idx = take(K * C * (G3 / 2) * 4);
  L->occ_list = take(K * C * G3 * 4);
  L->block_counts = take(K * C * (G3 / kOccBlock) * 4);
  L->occ_count = take(K * C * 4);
  L->part_sum = take(K * kNumSMs * 8 * 8);
  L->part_cnt = take(K * kNumSMs * 8 * 8);
  L->mean = take(K * 4);
  L->total_bytes = off;
  return RNB_OK;
}

int rnb_grid_update(const rnb_engine_config* c, void* workspace, const void* params_f16, float* const* density_grids,
                    uint8_t* const* bitfields, int warmup, float decay, float density_threshold, const float* noise,
                    const int32_t* coords1, const float* pick_u, uint64_t seed, int workspace_is_fresh, rnb_stream_t stream) {
  const char* what = "rnb_grid_update";
  rnb_grid_layout L;
  if (int rc = rnb_grid_update_layout_of(c, &L)) return rc;
  RNB_REQUIRE(workspace && params_f16 && density_grids && bitfields, what);
  cudaStream_t st = as_stream(stream);
  char* ws = (char*)workspace;
  const int K = c->K, C = c->cascades, G = c->grid_size;
  const int64_t G3 = (int64_t)G * G * G, cells = (int64_t)C * G3;
  GridPtrs Gp; TmpPtrs Tp;
  float* tmp[RNB_MAX_K];
  for (int i = 0; i < K; i++) {
    RNB_REQUIRE(density_grids[i] && bitfields[i], what);
    Gp.g[i] = density_grids[i];
    tmp[i] = (float*)(ws + L.tmp) + (int64_t)i * cells;
    Tp.t[i] = tmp[i];
  }
  float span[8], half[8];
  for (int cc = 0; cc < C; cc++) {                      // networks.py:385-388, evaluated in double like the Python scalars
    const double s = fmin(ldexp(1.0, cc - 1), (double)c->scale);
    const double h = s / G;
    span[cc] = (float)(s - h); half[cc] = (float)h;
  }
  rnb_field_params P;
  P.K = K; P.grid = c->grid;
  for (int k = 0; k < 3; k++) { P.xyz_min[k] = c->xyz_min[k]; P.xyz_max[k] = c->xyz_max[k]; }
  const __half* p16 = (const __half*)params_f16;
  P.table_f16 = p16 + c->table_off;
  for (int i = 0; i < K; i++) { P.geo_w_f16[i] = p16 + c->geo_off[i]; P.rgb_w_f16[i] = p16 + c->rgb_off[i]; }
  const int32_t* cell_idx = nullptr;
  int64_t M = G3;
  if (!warmup) {
    const int64_t M1 = G3 / 4;                                     // networks.py:381: M = G^3 // 4 of each kind
    const int nb = (int)(G3 / kOccBlock);
    int32_t* bc = (int32_t*)(ws + L.block_counts);
    int32_t* oc = (int32_t*)(ws + L.occ_count);
    int32_t* ol = (int32_t*)(ws + L.occ_list);
    int32_t* ci = (int32_t*)(ws + L.cell_idx);
    occ_count_kernel<<<dim3(nb, K * C), 256, 0, st>>>(Gp, C, G3, density_threshold, nb, bc);
    occ_scan_kernel<<<K * C, 1024, 0, st>>>(nb, bc, oc);
    occ_write_kernel<<<dim3(nb, K * C), 256, 0, st>>>(Gp, C, G3, density_threshold, nb, bc, ol);
    cells_kernel<<<kNumSMs * 8, 256, 0, st>>>(K * C, M1, G, G3, coords1, pick_u, seed, oc, ol, ci);
    if (int rc = check_launch(what)) return rc;
    if (workspace_is_fresh) RNB_CUDA(cudaMemsetAsync(ws + L.tmp, 0, (size_t)K * cells * 4, st), what);
    cell_idx = ci;
    M = 2 * M1;
  }
  if (int rc = rnb_density_cells_tc_launch(&P, tmp, C, G, M, span, half, cell_idx, noise, seed, stream)) return rc;
  const int nparts = kNumSMs * 8;
  // in steady state the scratch grid must be zero where no cell was evaluated: the EMA pass re-zeroes it for the next update
  ema_kernel<<<dim3(nparts, K), 256, 0, st>>>(Gp, Tp, cells, decay, 1, (double*)(ws + L.part_sum), (int64_t*)(ws + L.part_cnt));
  mean_kernel<<<K, 256, 0, st>>>(nparts, (const double*)(ws + L.part_sum), (const int64_t*)(ws + L.part_cnt), (float*)(ws + L.mean));
  if (int rc = check_launch(what)) return rc;
  for (int i = 0; i < K; i++)
    if (int rc = rnb_packbits_devthr(density_grids[i], cells / 8, (const float*)(ws + L.mean) + i, density_threshold, bitfields[i], stream)) return rc;
  return RNB_OK;
}

}  // extern "C"
