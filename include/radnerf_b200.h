/*
 * radnerf_b200.h -- C ABI of libradnerf_b200.so (hand-written CUDA for sm_100a).
 *
 * This is the drop-in boundary for Rad-NeRF's sampling -> K sub-NeRFs -> gated compositing
 * hot path.  Every entry point names the reference interface it replaces (paths relative
 * to the thu-nics/Rad-NeRF tree).  Conventions:
 *   - plain pointers + sizes; every data pointer is a DEVICE pointer unless the parameter
 *     name ends in `_host`; buffers are dense row-major with the shapes written beside them;
 *   - nothing is allocated inside the library except through an explicit rnb_*_create();
 *     callers hand in every output buffer (outputs need NOT be zero-filled: kernels write
 *     every element they own, which removes the reference's 256 MiB memsets);
 *   - `stream` is a cudaStream_t (0 = legacy default stream, what the reference uses);
 *   - return value: 0 on success, a negative RNB_ERR_* otherwise; rnb_last_error() gives a
 *     thread-local message.  Launch errors are detected with cudaPeekAtLastError().
 */
#ifndef RADNERF_B200_H_
#define RADNERF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* rnb_stream_t;

#define RNB_OK 0
#define RNB_ERR_BADARG (-1)   /* null pointer / negative size / unsupported configuration */
#define RNB_ERR_LAUNCH (-2)   /* CUDA launch or runtime failure */
#define RNB_ERR_NOMEM (-3)

#define RNB_ABI_VERSION 1

int rnb_abi_version(void);
const char* rnb_last_error(void);
/* compute capability of the current device as major*10+minor (100 on B200); <0 on error */
int rnb_device_arch(void);

/* ------------------------------------------------------------------------------------
 * (1) intersection / morton / packbits  -- replaces vren.ray_aabb_intersect,
 *     vren.ray_sphere_intersect, vren.morton3D, vren.morton3D_invert, vren.packbits
 * ------------------------------------------------------------------------------------ */

/* models/csrc/binding.cpp:4-16 -> intersection.cu:5-100.
 * rays_o, rays_d (n_rays,3) f32; centers, half_sizes (n_voxels,3) f32.
 * out: hit_cnt (n_rays) i32; hits_t (n_rays,max_hits,2) f32, -1 = no hit;
 *      hits_voxel_idx (n_rays,max_hits) i64, -1 = no hit.  Rows are ordered by ascending
 *      hits_t[...,0] exactly as the reference's torch::sort does (unfilled -1 rows first). */
int rnb_ray_aabb_intersect(const float* rays_o, const float* rays_d, const float* centers,
                           const float* half_sizes, int64_t n_rays, int64_t n_voxels, int max_hits,
                           int32_t* hit_cnt, float* hits_t, int64_t* hits_voxel_idx,
                           rnb_stream_t stream);

/* binding.cpp:19-31 -> intersection.cu:103-197.  radii (n_spheres) f32. */
int rnb_ray_sphere_intersect(const float* rays_o, const float* rays_d, const float* centers,
                             const float* radii, int64_t n_rays, int64_t n_spheres, int max_hits,
                             int32_t* hit_cnt, float* hits_t, int64_t* hits_sphere_idx,
                             rnb_stream_t stream);

/* binding.cpp:46-50 -> raymarching.cu:62-88.  coords (n,3) i32 -> indices (n) i32. */
int rnb_morton3D(const int32_t* coords, int64_t n, int32_t* indices, rnb_stream_t stream);
/* binding.cpp:53-57 -> raymarching.cu:90-119.  indices (n) i32 -> coords (n,3) i32. */
int rnb_morton3D_invert(const int32_t* indices, int64_t n, int32_t* coords, rnb_stream_t stream);

/* binding.cpp:34-43 -> raymarching.cu:122-161.  density_grid (8*n_bytes) f32 (or f16 when
 * grid_is_half != 0); bitfield (n_bytes) u8, bit i of byte n = grid[8n+i] > threshold. */
int rnb_packbits(const void* density_grid, int grid_is_half, int64_t n_bytes, float threshold,
                 uint8_t* density_bitfield, rnb_stream_t stream);
/* Same, with the threshold read from device memory: thr = min(*mean_density, threshold_cap)
 * (networks.py:405-406 without the .item() host sync). */
int rnb_packbits_devthr(const float* density_grid, int64_t n_bytes, const float* mean_density,
                        float threshold_cap, uint8_t* density_bitfield, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (2) occupancy-grid ray marching -- replaces vren.raymarching_train / raymarching_test
 * ------------------------------------------------------------------------------------ */

/* binding.cpp:60-81 -> raymarching.cu:166-332, split in two phases so that the caller owns
 * the data-dependent allocation.  Phase 1 marches every ray ONCE (warp per ray) and writes
 *   rays_a (n_rays,3) i64 = [ray_idx, start_idx, N_samples] in ASCENDING RAY ORDER with
 *   start_idx = exclusive prefix sum of N_samples (a deterministic, legal outcome of the
 *   reference's atomicAdd slot reservation, raymarching.cu:237-241),
 *   counter (2) i32 = [total_samples, n_rays]   (the reference's `counter`), and
 *   t_scratch (n_rays*max_samples) f32: t of sample k of ray r at [r*max_samples + k]
 *   (only the first N_samples entries of a row are written; no initialisation needed).
 * hits_t is (n_rays,2) f32 (t1 = -1 for a miss); noise (n_rays) f32 in [0,1). */
int rnb_raymarching_train_count(const float* rays_o, const float* rays_d, const float* hits_t,
                                const uint8_t* density_bitfield, int cascades, float scale,
                                float exp_step_factor, const float* noise, int grid_size,
                                int max_samples, int64_t n_rays,
                                int64_t* rays_a, int32_t* counter, float* t_scratch,
                                rnb_stream_t stream);
/* Phase 2: expands t_scratch into xyzs, dirs (cap,3) f32 and deltas, ts (cap) f32 for the
 * samples whose slot is < cap (cap >= counter[0] gives the reference result); the payload is
 * bit-identical to the reference's second march (x = o + t*d as one FMA, dt = calc_dt(t)). */
int rnb_raymarching_train_write(const float* rays_o, const float* rays_d, float scale,
                                float exp_step_factor, int grid_size, int max_samples,
                                int64_t n_rays, const int64_t* rays_a, const float* t_scratch,
                                int64_t cap, float* xyzs, float* dirs, float* deltas, float* ts,
                                rnb_stream_t stream);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  rnb_raymarching_train_count +
 * rnb_raymarching_train_write in ONE call that allocates nothing: the per-ray sample-t scratch (t_scratch above) is a
 * file-scope device array inside the library of RNB_NOALLOC_MARCH_MAX_FLOATS floats (32 MiB, held by every GPU that loads
 * the library until exit); n_rays * max_samples beyond that is rejected with RNB_ERR_BADARG and a message.  The caller
 * passes inputs and ALL outputs: rays_a (n_rays,3) i64, counter (2) i32, and xyzs / dirs (cap,3), deltas / ts (cap) f32 --
 * samples whose slot is >= cap are dropped (counter[0] holds the true total, so the caller can tell).  Same kernels, so the
 * results are bit-identical to the two-phase pair.  No host synchronisation, no cudaMalloc / cudaFree / allocating
 * library call; the first call on a device resolves the array's device address (cudaGetSymbolAddress, once per device).
 * Every call on a device SHARES the one array: calls that overlap on the GPU (different streams, or concurrent graph
 * launches) race -- serialise them.  The array keeps the last call's contents; nothing in it needs to start at zero
 * (only entries written by this call are read). */
#define RNB_NOALLOC_MARCH_MAX_FLOATS (8 * 1024 * 1024)
int rnb_raymarching_train_noalloc(const float* rays_o, const float* rays_d, const float* hits_t,
                                  const uint8_t* density_bitfield, int cascades, float scale,
                                  float exp_step_factor, const float* noise, int grid_size, int max_samples,
                                  int64_t n_rays, int64_t* rays_a, int32_t* counter, int64_t cap, float* xyzs,
                                  float* dirs, float* deltas, float* ts, rnb_stream_t stream);

/* binding.cpp:84-106 -> raymarching.cu:335-454.  hits_t (n_rays_total,2) f32 is MUTATED
 * ([r][0] <- resume point).  alive_indices (n_alive) i64.  Outputs (n_alive,N_samples[,3])
 * f32 are fully written (unused slots = 0, like the reference's zero-filled slabs) and
 * N_eff_samples (n_alive) i32.  Reproduces hazard Q1 (calc_dt is handed `cascades` where the
 * train kernel hands `scale`, raymarching.cu:370,399). */
int rnb_raymarching_test(const float* rays_o, const float* rays_d, float* hits_t,
                         const int64_t* alive_indices, int64_t n_alive,
                         const uint8_t* density_bitfield, int cascades, float scale,
                         float exp_step_factor, int grid_size, int max_samples, int N_samples,
                         float* xyzs, float* dirs, float* deltas, float* ts,
                         int32_t* N_eff_samples, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (5) volume compositing -- replaces vren.composite_train_fw / _bw / composite_test_fw
 * ------------------------------------------------------------------------------------ */

/* binding.cpp:109-126 -> volumerendering.cu:6-84.  sigmas, deltas, ts (N) f32; rgbs (N,3)
 * f32; rays_a (n_rays,3) i64.  out: total_samples (n_rays) i64 indexed by ray_idx; opacity,
 * depth (n_rays), rgb (n_rays,3) indexed by ray_idx; ws (N) (0 after early termination). */
int rnb_composite_train_fw(const float* sigmas, const float* rgbs, const float* deltas,
                           const float* ts, const int64_t* rays_a, float T_threshold,
                           int64_t n_rays, int64_t N, int64_t* total_samples, float* opacity,
                           float* depth, float* rgb, float* ws, rnb_stream_t stream);

/* binding.cpp:129-163 -> volumerendering.cu:87-203.  out: dL_dsigmas (N), dL_drgbs (N,3). */
int rnb_composite_train_bw(const float* dL_dopacity, const float* dL_ddepth, const float* dL_drgb,
                           const float* dL_dws, const float* sigmas, const float* rgbs,
                           const float* ws, const float* deltas, const float* ts,
                           const int64_t* rays_a, const float* opacity, const float* depth,
                           const float* rgb, float T_threshold, int64_t n_rays, int64_t N,
                           float* dL_dsigmas, float* dL_drgbs, rnb_stream_t stream);

/* binding.cpp:166-194 -> volumerendering.cu:206-286.  sigmas, deltas, ts (n_alive,N_samples);
 * rgbs (n_alive,N_samples,3); alive_indices (n_alive) i64 MUTATED (-1 = finished);
 * opacity, depth (n_rays_total), rgb (n_rays_total,3) accumulated IN PLACE. */
int rnb_composite_test_fw(const float* sigmas, const float* rgbs, const float* deltas,
                          const float* ts, const float* hits_t, int64_t* alive_indices,
                          float T_threshold, const int32_t* N_eff_samples, int64_t n_alive,
                          int N_samples, float* opacity, float* depth, float* rgb,
                          rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (6) distortion loss -- replaces vren.distortion_loss_fw / distortion_loss_bw
 *     (binding.cpp:197-231 -> losses.cu:9-175; caller losses.py:6-37).  ws, deltas, ts (N) f32; rays_a (n_rays,3) i64.
 * ------------------------------------------------------------------------------------ */
/* out: loss (n_rays) indexed by ray_idx (rays with no samples get 0); ws_inclusive_scan, wts_inclusive_scan (N):
 * per-ray running sums of w and w*t.  Samples not covered by any ray are not written (the wrapper zero-fills). */
int rnb_distortion_loss_fw(const float* ws, const float* deltas, const float* ts, const int64_t* rays_a,
                           int64_t n_rays, int64_t N, float* loss, float* ws_inclusive_scan,
                           float* wts_inclusive_scan, rnb_stream_t stream);
/* dL_dloss (n_rays) indexed by ray_idx -> dL_dws (N). */
int rnb_distortion_loss_bw(const float* dL_dloss, const float* ws_inclusive_scan, const float* wts_inclusive_scan,
                           const float* ws, const float* deltas, const float* ts, const int64_t* rays_a,
                           int64_t n_rays, int64_t N, float* dL_dws, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (3) multiresolution hash-grid encoding + SH4 -- replaces tinycudann's
 *     Encoding{Grid/Hash} and Encoding{SphericalHarmonics,4}
 *     (reference call sites: models/networks.py:42-55, 70-77, 234-256; tinycudann itself is
 *      not vendored in the reference -- published algorithm, see oracle/tcnn_cpu.py)
 * ------------------------------------------------------------------------------------ */

#define RNB_MAX_LEVELS 32
typedef struct {
  int32_t n_levels;                    /* L (16)                                           */
  int32_t n_features;                  /* F (2; only 2 is implemented)                     */
  float scale[RNB_MAX_LEVELS];         /* s_l = N_min * b^l - 1                            */
  uint32_t res[RNB_MAX_LEVELS];        /* ceil(s_l) + 1                                    */
  uint32_t offset[RNB_MAX_LEVELS + 1]; /* first ENTRY of level l in the flat table         */
} rnb_grid_meta;

/* x (n,3) f32 in [0,1]; table (n_entries,2) f16 (level-major, entry, feature -- tcnn's
 * `params` layout rounded to fp16); out (n, 2L) f16. */
int rnb_hashgrid_fwd(const float* x, int64_t n, const void* table_f16, const rnb_grid_meta* meta,
                     void* out_f16, rnb_stream_t stream);
/* Scatter-add of dL/dout (n,2L) f16 into grad_table (n_entries,2) f32 (ACCUMULATES: the
 * table is shared by all K sub-NeRFs, the caller zeroes it once per step).  grad_scale
 * multiplies every contribution (1/loss_scale). */
int rnb_hashgrid_bwd(const float* x, int64_t n, const void* dL_dout_f16, const rnb_grid_meta* meta,
                     float grad_scale, float* grad_table, rnb_stream_t stream);
/* u (n,3) f32 in [0,1] (the reference feeds (d/|d|+1)/2, networks.py:322-323) -> (n,16) f16 */
int rnb_sh4_fwd(const float* u, int64_t n, void* out_f16, rnb_stream_t stream);
/* fp32 master parameters -> fp16 compute copy (n elements) */
int rnb_cast_f32_to_f16(const float* src, int64_t n, void* dst_f16, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (4) 64-wide fp16 MLPs on tensor cores -- replaces tinycudann Network{FullyFusedMLP}
 *     (reference call sites: models/networks.py:269-289, 1075-1085)
 * ------------------------------------------------------------------------------------ */
typedef struct {
  int32_t in_width;        /* padded to a multiple of 16 (pad columns hold 1.0, as tcnn) */
  int32_t out_width;       /* padded to a multiple of 16                                 */
  int32_t n_hidden;        /* hidden layers of 64 neurons (ReLU)                         */
  int32_t out_activation;  /* 0 = None, 1 = Sigmoid                                      */
} rnb_mlp_desc;

/* weights: fp16 copy of tcnn's flat `params`: row-major (out,in) matrices
 * [64 x in_width | (n_hidden-1) x 64 x 64 | out_width x 64].  in (n,in_width) f16 ->
 * out (n,out_width) f16.  hidden (n_hidden, n, 64) f16 receives the post-ReLU activations
 * needed by the backward pass (may be NULL for inference). */
int rnb_mlp_fwd(const rnb_mlp_desc* desc, const void* weights_f16, const void* in_f16, int64_t n,
                void* out_f16, void* hidden_f16, rnb_stream_t stream);
/* dL_din (n,in_width) f16 may be NULL (no input gradient wanted); grad_weights (same layout
 * as weights, f32) is ACCUMULATED into, scaled by grad_scale; may be NULL. */
int rnb_mlp_bwd(const rnb_mlp_desc* desc, const void* weights_f16, const void* in_f16,
                const void* hidden_f16, const void* out_f16, const void* dL_dout_f16, int64_t n,
                void* dL_din_f16, float* grad_weights, float grad_scale, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (3+4 fused) the field of all K sub-NeRFs on one packed sample list: hash-grid encode ->
 *     geo MLP -> TruncExp -> SH4 -> rgb MLP in ONE kernel per direction (forward / backward).
 *     Replaces the chain MNGP.forward = xyz_encoder -> geo_net_i -> TruncExp -> dir_encoder ->
 *     cat -> rgb_net_i (models/networks.py:291-328) and its autograd backward (tinycudann
 *     kernel_grid_backward + FullyFusedMLP backward).  The 32-feature row, the hidden
 *     activations and the SH row never leave the SM in the forward pass unless the caller asks
 *     for them to be saved for the backward pass.
 * ------------------------------------------------------------------------------------ */
#define RNB_MAX_K 16

typedef struct {
  int32_t K;                           /* sub-NeRFs sharing the hash table (model_zoo_size)         */
  rnb_grid_meta grid;
  float xyz_min[3], xyz_max[3];        /* MNGP.xyz_min / xyz_max (networks.py:300)                   */
  const void* table_f16;               /* (entries,2) f16                                            */
  const void* geo_w_f16[RNB_MAX_K];    /* geo_net_i: [64x32 | 32x64] f16, tinycudann layout          */
  const void* rgb_w_f16[RNB_MAX_K];    /* rgb_net_i: [64x32 | 64x64 | 16x64] f16                     */
} rnb_field_params;

/* Packed sample list of the K passes.  Pass i owns samples [seg[i], seg[i] + seg[K+1+i]); every
 * seg[i] is a multiple of 128 (a 128-row tile never straddles two passes); seg[K] = end of the
 * last pass rounded up to 128.  `seg` lives in DEVICE memory (written by the marcher's scan), so
 * no host synchronisation is needed between marching and field evaluation. */
typedef struct {
  const float* rays_o;                 /* (n_rays,3) f32                                             */
  const float* rays_d;                 /* (n_rays,3) f32, unnormalised (Q20)                         */
  const int32_t* ray_id;               /* (capacity) i32: sample -> ray                              */
  const float* ts;                     /* (capacity) f32: sample -> t   (x = o + t*d)                */
  const int32_t* seg;                  /* device, (2K+1) i32, see above                              */
  int64_t capacity;                    /* multiple of 128                                            */
} rnb_samples;

/* Activations saved by the forward pass for the backward pass: capacity*W f16 each, OPAQUE between
 * rnb_field_fwd and rnb_field_bwd.  (Default tcgen05 implementation: per 128-sample tile the shared-memory
 * operand image [W/8][128][8], written and read back with one bulk copy per tile and layer; element (s, f)
 * of a tile is at half-index (f/8)*1024 + s*8 + f%8.)  Pass saved = NULL in inference (nothing is saved). */
typedef struct {
  void* enc;      /* (cap,32) hash-grid features                                   */
  void* rgb_in;   /* (cap,32) [SH4(16) | geometry feature(16)]                     */
  void* geo_h;    /* (cap,64) geo hidden (post-ReLU)                               */
  void* rgb_h1;   /* (cap,64)                                                      */
  void* rgb_h2;   /* (cap,64)                                                      */
  void* sraw;     /* (cap)    raw sigma output before TruncExp                     */
} rnb_field_saved;

/* out: sigmas (cap) f32 = exp(raw); rgb4 (cap,4) f16 = sigmoid(rgb_net)[:, :3] (+1 pad). */
int rnb_field_fwd(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved,
                  float* sigmas, void* rgb4_f16, rnb_stream_t stream);
/* dL_dsigmas (cap) f32, dL_drgbs (cap,3) f32 (what composite_train_bw emits).  Gradients are
 * ACCUMULATED (fp32) into grad_table (entries,2), grad_geo[i], grad_rgb[i] (tinycudann layouts).
 * loss_scale: internal fp16 gradient scale (activations gradients are carried in fp16 like
 * tinycudann; parameter gradients are un-scaled before accumulation). */
int rnb_field_bwd(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved,
                  const void* rgb4_f16, const float* dL_dsigmas, const float* dL_drgbs,
                  float loss_scale, float* grad_table, float* const* grad_geo, float* const* grad_rgb,
                  rnb_stream_t stream);
/* The same; additionally ORs RNB_FLAG_NONFINITE into *flags (may be NULL) when a gradient seed saturates fp16 at this
 * loss_scale or an inf / nan reaches the gradient of the encoding (hence of the table). */
int rnb_field_bwd_guarded(const rnb_field_params* P, const rnb_samples* S, const rnb_field_saved* saved,
                          const void* rgb4_f16, const float* dL_dsigmas, const float* dL_drgbs,
                          float loss_scale, float* grad_table, float* const* grad_geo, float* const* grad_rgb,
                          int32_t* flags, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (fused step) One Rad-NeRF training step without host synchronisation.  Replaces the body of
 *     NeRFSystem.training_step (train_ml.py:172-200) below the data loader: Ray_Gate forward
 *     (networks.py:1087-1093), ml_render's K-loop (ml_rendering.py:47-68: AABB, RayMarcher, MNGP.forward,
 *     VolumeRenderer, background blend, gated blend), NeRFLoss (losses.py:52-74), the backward of all
 *     of it, and the Adam update -- about 12 kernel launches, sample counts stay on the device.
 *     The library allocates nothing: rnb_engine_layout() tells the caller how to carve ONE workspace
 *     buffer (offsets in bytes), so the host side can also view every intermediate as a tensor.
 * ------------------------------------------------------------------------------------ */
typedef struct {
  int32_t K;                     /* model_zoo_size; K == 1: single NGP field, no gate                       */
  int32_t n_rays;                /* rays per step on this rank                                               */
  int32_t max_samples;           /* MAX_SAMPLES = 1024 (ml_rendering.py:8)                                   */
  int32_t grid_size;             /* 128                                                                      */
  int32_t cascades;
  float scale;
  float exp_step_factor;
  float T_threshold;             /* 1e-4                                                                     */
  float near_distance;           /* NEAR_DISTANCE = 0.01 (ml_rendering.py:9)                                 */
  int64_t sample_capacity;       /* samples of all K passes the workspace holds (multiple of 128)            */
  rnb_grid_meta grid;
  float xyz_min[3], xyz_max[3], center[3], half_size[3];
  float lambda_opacity, lambda_cv, lambda_dm;
  float loss_scale;              /* fp16 gradient scale inside the MLP chains (e.g. 128*1024)                */
  /* flat parameter vector layout (element offsets): [hash table | geo_i, rgb_i ... | gate] */
  int64_t n_params;
  int64_t table_off;
  int64_t geo_off[RNB_MAX_K];
  int64_t rgb_off[RNB_MAX_K];
  int64_t gate_off;              /* ignored when K == 1                                                      */
  uint64_t noise_seed;           /* seed of the library's own marching noise (used when the caller passes noise = NULL) */
} rnb_engine_config;

/* byte offsets into the caller's workspace (all 256-byte aligned) */
typedef struct {
  int64_t total_bytes;
  int64_t gate_in, gate_hidden, gate_logits, gate_dlogits, gate_prob, dgate;  /* f16 (R,16) | f16 (4,R,64) | f16 (R,16) | f16 (R,16) | f32 (R,K) | f32 (R,K) */
  int64_t importance;            /* f32 (K): sum over rays of the gate (Ray_Gate's importance)                */
  int64_t t_scratch;             /* f32 (K,R,max_samples)                                                     */
  int64_t counts, starts;        /* i32 (K,R): samples per ray / first sample (index into the packed list)    */
  int64_t seg;                   /* i32 (2K+1) pass bases + counts (rnb_samples.seg), then [2K+1] = overflow flag, [2K+2] = total samples */
  int64_t ray_id, ts, deltas;    /* i32 | f32 | f32 (S)                                                       */
  int64_t enc, rgb_in, geo_h, rgb_h1, rgb_h2, sraw;   /* rnb_field_saved                                     */
  int64_t sigmas, rgb4, ws, dsigmas, drgbs;            /* f32 (S) | f16 (S,4) | f32 (S) | f32 (S) | f32 (S,3) */
  int64_t pass_opacity, pass_depth, pass_rgb, pass_total;  /* f32 (K,R) | f32 (K,R) | f32 (K,R,3) (before background) | i32 (K,R) */
  int64_t d_pass_opacity, d_pass_depth, d_pass_rgb;        /* gradient seeds, same shapes                     */
  int64_t rgb, opacity, depth;   /* results: f32 (R,3) | f32 (R) | f32 (R,K)  (ml_rendering.py:72-77)         */
  int64_t loss;                  /* f32 (8): [total, rgb, opacity, cv_importance, depth_mutual, -, -, -]      */
  int64_t march_slot1;           /* second copy of the marcher's outputs [t_scratch .. deltas] (same internal offsets,
                                  * march_slot_bytes long) for marching the NEXT batch while this one trains      */
  int64_t march_slot_bytes;
  int64_t coarse;                /* u64 (K,4096): dilated 64^3-block occupancy (bit x of word y + 64 z), part of the march slot; used
                                  * for the marcher's empty-space test when cascades == 1 and grid_size == 128     */
  int64_t status;                /* i32 (16) step guard (GradScaler equivalent, see rnb_engine_adam_guarded): [0] optimizer steps skipped
                                  * for a non-finite / saturated gradient, [1] steps skipped because the samples exceeded
                                  * sample_capacity, [2] largest (128-aligned) sample count any step needed, [3] optimizer steps
                                  * applied, [4]-[5] internal tickets, [6] != 0: a peer barrier timed out, [7] / [9] marching
                                  * calls so far into slot 0 / 1 (the library's noise streams), [8] learning rate as float bits, read by the optimizer
                                  * kernels when they are called with lr < 0 (a captured CUDA graph can then be replayed under a
                                  * schedule: the host writes the value before the replay).  16 words.  The caller zeroes it once. */
} rnb_engine_layout;

/* Step guard.  The reference trains with Lightning precision=16 (train_ml.py:304): a dynamic GradScaler that skips the
 * optimizer step when a gradient is inf / nan and adapts the loss scale.  Here the gradient buffer carries
 * RNB_GRAD_TAIL extra floats behind the n_params gradients; word [n_params] is this step's FLAG word (i32 bit mask):
 * producers (blend kernel, field backward, rnb_engine_grad_check) OR bits into it, the Adam kernels read it -- in the
 * sharded optimizer every rank reads every peer's word, so all ranks take the same decision -- and, when it is set,
 * clear the gradients WITHOUT touching parameters or moments.  rnb_engine_fwd_bwd[_ex] clears the word at the start of
 * a step with_backward.  The host adapts cfg->loss_scale from the status counters (no synchronisation needed). */
#define RNB_GRAD_TAIL 64
#define RNB_FLAG_NONFINITE 1      /* a gradient seed saturated fp16, or an inf / nan reached a parameter gradient */
#define RNB_FLAG_CAPACITY 2       /* the step's samples did not fit sample_capacity (the tail was dropped)        */

int rnb_engine_layout_of(const rnb_engine_config* cfg, rnb_engine_layout* out);

/* Phase 1 (K > 1): gate forward.  Writes gate_prob and the LOCAL importance sums. params_f16: flat fp16
 * copy of the parameter vector. */
int rnb_engine_gate(const rnb_engine_config* cfg, void* workspace, const void* params_f16,
                    const float* rays_o, const float* rays_d, rnb_stream_t stream);
/* Phase 2: march, field, composite, blend, loss, and the complete backward.  `importance` (K) f32 device:
 * the GLOBAL importance (all-reduced over ranks when data parallel; the workspace's own buffer otherwise).
 * n_rays_global: rays of the whole (multi-rank) batch, the denominator of the loss means.
 * bitfields: HOST array of K device pointers; noise (K,n_rays) f32 device in [0,1), or NULL: the library draws it from its own
 * counter generator (cfg->noise_seed, one stream of draws per marching call: no launch outside the library); bg (K,3) f32 HOST
 * (background colour per pass, Q17).  Gradients are ACCUMULATED into grads_f32 (flat, fp32; n_params + RNB_GRAD_TAIL floats:
 * the flag word of the step guard sits at index n_params and is cleared at the start of a step with_backward). */
int rnb_engine_fwd_bwd(const rnb_engine_config* cfg, void* workspace, const void* params_f16, float* grads_f32,
                       const float* rays_o, const float* rays_d, const float* target_rgb, const float* noise,
                       const uint8_t* const* bitfields, const float* bg_host, int64_t n_rays_global,
                       const float* importance, int with_backward, int march_slot, rnb_stream_t stream);
/* The same with the gate network taken off the critical path.  The gate forward (rnb_engine_gate) is needed only from
 * the blend on, and the gate backward needs only the blend's output, so a caller may run both on a second stream next to
 * the compositing kernels:  gate_ready_event (cudaEvent_t or NULL): the step's stream waits for it right before the
 * blend;  blend_done_event (cudaEvent_t or NULL): recorded on the step's stream right after the blend;  skip_gate_bwd != 0:
 * the caller runs rnb_engine_gate_bwd itself (after blend_done_event) and joins it before the optimizer. */
int rnb_engine_fwd_bwd_ex(const rnb_engine_config* cfg, void* workspace, const void* params_f16, float* grads_f32,
                          const float* rays_o, const float* rays_d, const float* target_rgb, const float* noise,
                          const uint8_t* const* bitfields, const float* bg_host, int64_t n_rays_global,
                          const float* importance, int with_backward, int march_slot, void* gate_ready_event,
                          void* blend_done_event, int skip_gate_bwd, rnb_stream_t stream);
/* Ray_Gate backward for the batch of the last rnb_engine_fwd_bwd[_ex]: accumulates into grads_f32[gate block]. */
int rnb_engine_gate_bwd(const rnb_engine_config* cfg, void* workspace, const void* params_f16, float* grads_f32,
                        rnb_stream_t stream);
/* Marching depends on the rays and the occupancy bitfields only, not on the parameters: the samples of the NEXT
 * batch can be produced on a second stream while the current batch is in its field / backward kernels.
 * rnb_engine_march runs AABB + march + scan + expand for one batch into workspace slot `slot` (0 or 1);
 * rnb_engine_fwd_bwd(march_slot = that slot) then skips its own marching (noise / bitfields may be NULL).
 * march_slot = -1: march inline into slot 0 (the default, single-stream behaviour).  The caller orders the
 * two streams (events) and must not update the bitfields between the two calls. */
int rnb_engine_march(const rnb_engine_config* cfg, void* workspace, const float* rays_o, const float* rays_d,
                     const float* noise, const uint8_t* const* bitfields, int slot, rnb_stream_t stream);
/* Phase 3: Adam (torch.optim.Adam semantics, no weight decay) over the flat vector; writes the fp16 copy
 * and ZEROES the gradient for the next step.  grad_scale multiplies the gradient first. */
int rnb_engine_adam(float* params_f32, void* params_f16, float* grads_f32, float* exp_avg, float* exp_avg_sq,
                    int64_t n, float lr, float beta1, float beta2, float eps, int64_t step, float grad_scale,
                    rnb_stream_t stream);
/* The same behind the step guard: `flags` = the gradient buffer's flag word (grads_f32 + n as i32), `status` = the
 * workspace's status block.  Flag set: gradients are cleared, parameters / moments / fp16 copy stay, status[0] or [1]
 * counts the skip.  Flag clear: the update is applied with the bias corrections of step status[3] + 1 (the count of
 * APPLIED steps lives on the device, like GradScaler's skipped steps do not advance the optimizer), status[3]++. */
int rnb_engine_adam_guarded(float* params_f32, void* params_f16, float* grads_f32, float* exp_avg, float* exp_avg_sq,
                            int64_t n, float lr, float beta1, float beta2, float eps, float grad_scale,
                            const int32_t* flags, int32_t* status, rnb_stream_t stream);
/* ORs RNB_FLAG_NONFINITE into *flags when grads_f32[lo, hi) holds an inf or nan (used on the dense-network block: the
 * table gradient is checked where it is produced, in the field backward kernel). */
int rnb_engine_grad_check(const float* grads_f32, int64_t lo, int64_t hi, int32_t* flags, rnb_stream_t stream);

/* Ray_Gate.forward (networks.py:1087-1093): logits = MLP(cat(rays_o, rays_d) padded with ones to 16), softmax over
 * the K logits in fp32, importance[k] = sum over rays.  gate_hidden_f16 ((4,n_rays,64), saved for the backward
 * pass) may be NULL at inference.  Scratch buffers are the caller's: gate_in (n_rays,16) f16, gate_logits (n_rays,16) f16. */
int rnb_gate_fwd(const void* gate_w_f16, const float* rays_o, const float* rays_d, int64_t n_rays, int K,
                 void* gate_in_f16, void* gate_hidden_f16, void* gate_logits_f16, float* prob, float* importance,
                 rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (test time) Full-frame rendering: replaces ml_render(test_time=True) = K x __render_rays_test
 *     (ml_rendering.py:81-155; vren.raymarching_test binding.cpp:84-106, vren.composite_test_fw
 *     binding.cpp:166-194) + background + gated blend (ml_rendering.py:64-77).  All K sub-NeRFs
 *     advance together; one round = march (warp per alive ray, dense sample packing), fused field,
 *     composite, and one 4K-byte read-back of the alive counts.  cfg: the engine config with
 *     n_rays = rays of this call and sample_capacity >= K*n_rays*8 + 128*K.
 * ------------------------------------------------------------------------------------ */
typedef struct {
  int64_t total_bytes;
  int64_t gate_in, gate_logits, gate_prob, importance;
  int64_t t_cur, t2, acc_opacity, acc_depth, acc_rgb, item_start, item_n, alive0, alive1, counters, seg;
  int64_t ray_id, ts, sigmas, rgb4;
  int64_t coarse;                /* u64 (K,4096): see rnb_engine_layout.coarse */
} rnb_render_layout;
int rnb_render_layout_of(const rnb_engine_config* cfg, rnb_render_layout* out);
/* out: rgb (R,3), depth (R,K) ungated (Q19), opacity (R); pass_rgb (K,R,3) per-sub-NeRF colours after the
 * background blend (may be NULL); stats_host[2] = {samples evaluated, rounds} (may be NULL).  bg_host (K,3).
 * Rounds are planned on the device and launched without read-backs (the reference: two host synchronisations per round and
 * sub-NeRF); the call synchronises the stream ONCE, at its end (statistics).  First call per host thread: 1 KB of pinned
 * memory and two events are created and kept. */
int rnb_render(const rnb_engine_config* cfg, void* workspace, const void* params_f16, const float* rays_o,
               const float* rays_d, const uint8_t* const* bitfields, const float* bg_host, float* rgb, float* depth,
               float* opacity, float* pass_rgb, int64_t* stats_host, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (occupancy) MNGP.update_density_grid (networks.py:374-409; NGP :171-201) for all K sub-NeRFs and all
 *     cascades, without host synchronisation: cell selection (get_all_cells :330-342 in warm-up,
 *     sample_uniform_and_occupied_cells :344-372 afterwards -- torch.nonzero becomes an ordered device-side
 *     compaction), density of one jittered point per cell (fused hash encode + geo MLP + exp on tcgen05),
 *     grid = grid < 0 ? grid : max(grid*decay, new), mean over grid > 0, packbits at min(mean, threshold).
 *     density_grids[i] (C, G^3) f32 and bitfields[i] (C*G^3/8) u8 are updated in place.
 *     Random inputs may be supplied for reproducibility / parity tests, or NULL to use the built-in
 *     counter-based generator seeded with `seed`:
 *       noise   (K, C, M, 3) f32 in [0,1): jitter inside the cell; M = G^3 (warm-up) or G^3/2 (steady state:
 *               first G^3/4 uniform cells, then G^3/4 picks among the occupied cells);
 *       coords1 (K, C, G^3/4, 3) i32 in [0,G): the uniform cells;  pick_u (K, C, G^3/4) f32 in [0,1): the picks
 *               (cell = occupied[floor(u * n_occupied)], the reference draws randint(n_occupied)).
 *     The workspace's `tmp` region must be zero before the first call (pass workspace_is_fresh != 0 to have it
 *     cleared); the library leaves it zero after every call.
 * ------------------------------------------------------------------------------------ */
typedef struct {
  int64_t total_bytes;
  int64_t tmp, cell_idx, occ_list, block_counts, occ_count, part_sum, part_cnt, mean;
} rnb_grid_layout;
int rnb_grid_update_layout_of(const rnb_engine_config* cfg, rnb_grid_layout* out);
int rnb_grid_update(const rnb_engine_config* cfg, void* workspace, const void* params_f16, float* const* density_grids,
                    uint8_t* const* bitfields, int warmup, float decay, float density_threshold, const float* noise,
                    const int32_t* coords1, const float* pick_u, uint64_t seed, int workspace_is_fresh, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (multi-GPU) The gradient exchange and the optimizer as ONE kernel over NVLink peer memory: rank r owns shard r of
 *     the flat parameter vector; its kernel sums that shard of every rank's gradient buffer with peer loads (fixed
 *     order), applies Adam to its fp32 master / moments, and stores the fp16 compute copy of the shard into every
 *     rank's parameter buffer (reduce-scatter + Adam + all-gather; replaces DDP's NCCL all-reduce followed by a
 *     replicated FusedAdam, train_ml.py:143, 309-318).  The caller brackets it with two small collectives (every rank's
 *     backward done before; every rank's kernel done before parameters are read or gradients re-zeroed) and zeroes its
 *     own gradient buffer afterwards.  fp32 master parameters and moments outside a rank's own shard are NOT kept
 *     current (gather them when a checkpoint is written).
 * ------------------------------------------------------------------------------------ */
#define RNB_MAX_PEERS 8
/* cudaMalloc + zero-fill + CUDA IPC handle (64 bytes) for a buffer that other processes of this node map */
int rnb_peer_alloc(int64_t bytes, void** dev_ptr, void* ipc_handle_64);
int rnb_peer_open(const void* ipc_handle_64, void** dev_ptr);      /* map a peer's allocation (enables peer access lazily) */
int rnb_peer_close(void* dev_ptr);
int rnb_peer_free(void* dev_ptr);
int rnb_adam_shard_bounds(int64_t n, int rank, int world, int64_t* lo, int64_t* hi);
/* grads_peers[p], params_f16_peers[p]: rank p's gradient / fp16 parameter buffer as mapped in THIS process
 * ([rank] = the local ones); params_f32, exp_avg, exp_avg_sq: local, full length n (only the shard is touched). */
/* Parameters [0, n_sharded) are sharded over the ranks; [n_sharded, n) -- the gate network, last block of the flat
 * vector -- is updated redundantly by every rank from the same reduced gradient, so that the next step's gate forward and
 * its importance all-reduce (which then doubles as the "every rank's kernel is done" barrier) need no remote parameters.
 * zero_grads (may be NULL): this rank's gradient buffer of the PREVIOUS step (n floats), cleared by the same kernel;
 * with two alternating gradient buffers no separate zero-fill pass and no second barrier are needed. */
int rnb_engine_adam_sharded(int rank, int world, const float* const* grads_peers, void* const* params_f16_peers,
                            float* params_f32, float* exp_avg, float* exp_avg_sq, int64_t n, int64_t n_sharded,
                            float* zero_grads, float lr, float beta1, float beta2, float eps, int64_t step,
                            float grad_scale, rnb_stream_t stream);
/* The same behind the step guard (see RNB_GRAD_TAIL): every peer gradient buffer carries its flag word at float index n;
 * the kernel ORs all of them (peer loads), so every rank skips or applies the same step.  status: this rank's status
 * block (step count of applied steps as in rnb_engine_adam_guarded; [5] internal ticket, [6] != 0: a barrier spin timed out).
 * flag_peers (HOST array of `world` device pointers, or NULL): every rank's RNB_PEER_FLAG_WORDS int32 barrier words, peer-
 * mapped and zero at start.  Given, the ranks are ordered by FLAG BARRIERS over peer memory instead of collectives around the
 * call: a one-warp kernel announces "backward of step barrier_value done" into every rank's words, the optimizer kernel
 * spins (bounded) until every rank has announced before it reads a peer's gradients, its last block announces "shard
 * stored everywhere", and a one-warp kernel holds the stream until every rank has.  barrier_value must grow by one per call
 * (counters are never reset).  NULL: the caller orders the ranks (one barrier in front, one behind). */
#define RNB_PEER_FLAG_WORDS (2 * RNB_MAX_PEERS)
int rnb_engine_adam_sharded_guarded(int rank, int world, const float* const* grads_peers, void* const* params_f16_peers,
                                    float* params_f32, float* exp_avg, float* exp_avg_sq, int64_t n, int64_t n_sharded,
                                    float* zero_grads, float lr, float beta1, float beta2, float eps, float grad_scale,
                                    int32_t* status, void* const* flag_peers, int32_t barrier_value, rnb_stream_t stream);

/* ------------------------------------------------------------------------------------
 * (callers of the path) Training-batch assembly on the device: replaces BaseDataset.__getitem__'s pixel sampling
 *     (datasets/base.py:22-29), the index gathers of train_ml.py:183-185 and get_rays (datasets/ray_utils.py:46-70).
 *     directions (n_pix,3) f32 camera-space directions; poses (n_img,3,4) f32 [R|t]; images (n_img,n_pix,3) f32 or NULL
 *     (then rgb must be NULL); img_idxs / pix_idxs (B) i64 or NULL = draw iid uniform indices from `seed`.
 *     out: rays_o, rays_d (B,3) (unnormalised, as the reference), rgb (B,3), and the indices used (may be NULL).
 * ------------------------------------------------------------------------------------ */
int rnb_batch_rays(const float* directions, const float* poses, const float* images, const int64_t* img_idxs,
                   const int64_t* pix_idxs, int64_t n_img, int64_t n_pix, int64_t B, uint64_t seed, float* rays_o,
                   float* rays_d, float* rgb, int64_t* img_idxs_out, int64_t* pix_idxs_out, rnb_stream_t stream);

/* Optional per-stage device timing of the fused step (CUDA events around every stage; off by default).
 * Stage order: gate, march, scan, expand, field_fwd, composite_fw, blend_loss, composite_bw, field_bwd,
 * gate_bwd, adam, barrier (caller-recorded).  rnb_engine_profile_read synchronises, returns the accumulated
 * milliseconds and call counts since the previous read (arrays of n >= 12 entries) and clears them. */
int rnb_engine_profile(int enable);
int rnb_engine_profile_read(double* ms, int32_t* calls, int n);
/* Caller-recorded spans for work the library does not launch itself (collectives issued by the host side, the
 * peer-memory optimizer): stage 11 = "barrier", 10 = "adam".  begin returns a handle, or -1 when profiling is off. */
int rnb_engine_profile_span_begin(int stage, rnb_stream_t stream);
int rnb_engine_profile_span_end(int handle, rnb_stream_t stream);
/* The same records as a timeline (issue order): stage index, start and end in ms relative to the first record; the
 * events of the step's side streams share the device clock, so overlap and gaps between stages are visible. */
int rnb_engine_profile_timeline(int32_t* stage, double* start_ms, double* end_ms, int n_max, int32_t* n_out);

/* Diagnostics: self-test of the tcgen05 building blocks (csrc/tc05.cuh) on one CTA.
 * X (128,64), W (64,64), Y (128,32) f16 row-major -> out1 (128,64) = X W^T, out2 (128,32) = raw TMEM
 * dump (128 lanes) of the M=64 product X^T Y, out3 (128,64) = X W.  f32 outputs. */
int rnb_tc_probe(const void* X, const void* W, const void* Y, float* out1, float* out2, float* out3,
                 rnb_stream_t stream);

/* Diagnostics: rate of scattered L2 accesses with the field kernels' own instructions at random entries of a table
 * (csrc/microbench.cu; scripts/microbench_l2.py).  kind 0/1/2: red.global.add .f32 / .v2.f32 / .v4.f32 into `entries` x 8
 * bytes; 3/4: ld.global.nc .b32 / .b64 from `entries` x 4 bytes.  Every active lane of blocks x 256 threads issues
 * iters x 8 accesses; *ops_out = lane-level accesses issued. */
int rnb_diag_l2_rate(int kind, void* table, int64_t entries, int iters, int active_lanes, int blocks, int64_t* ops_out,
                     rnb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* RADNERF_B200_H_ */
