// microbench.cu -- rate of scattered L2 accesses, the unit the fused field kernels are bound by (diagnostics; not on
// the product path).  The field kernels gather 4/8-byte entries from the fp16 hash table (forward) and reduce 8/16-byte
// fp32 pairs into the gradient table (backward); both tables live in L2, so neither HBM bandwidth nor tensor throughput
// is their roofline.  These kernels issue the same instructions -- ld.global.nc.b32/.b64 and red.global.add.v2/v4.f32 --
// at pseudo-random entries of a table of the same size from a full machine of otherwise idle warps, which gives the
// sector rate the memory system sustains for that access pattern (scripts/microbench_l2.py -> profiles/).
#include "hashgrid.cuh"

namespace rnb {

__device__ __forceinline__ uint32_t mb_hash(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

// every active lane issues `iters` x 8 reductions of VEC floats at random entries (entry = 2 floats; VEC = 4: aligned pair)
template <int VEC>
__global__ void __launch_bounds__(256) mb_red_kernel(float* __restrict__ table, uint32_t entries, int iters, int active_lanes) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  if ((int)(threadIdx.x & 31) >= active_lanes) return;
  uint32_t h = mb_hash(tid * 2654435761u + 12345u);
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
      h = mb_hash(h + k);
      uint32_t e = h % entries;
      if (VEC == 4) { e &= ~1u; red_add_f32x4(table + 2 * (size_t)e, 1.f, 1.f, 1.f, 1.f); }
      else if (VEC == 2) red_add_f32x2(table + 2 * (size_t)e, 1.f, 1.f);
      else asm volatile("red.global.add.f32 [%0], %1;" ::"l"(table + 2 * (size_t)e), "f"(1.f) : "memory");
    }
  }
}

// every lane issues `iters` x 8 independent loads of BYTES at random entries (entry = 4 bytes) and keeps a checksum
template <int BYTES>
__global__ void __launch_bounds__(256) mb_gather_kernel(const uint32_t* __restrict__ table, uint32_t entries, int iters, uint32_t* __restrict__ sink) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t h = mb_hash(tid * 2654435761u + 777u), acc = 0;
  for (int i = 0; i < iters; i++) {
    uint32_t e[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { h = mb_hash(h + k); e[k] = h % entries; }
#pragma unroll
    for (int k = 0; k < 8; k++) {
      if (BYTES == 8) { const uint2 v = __ldg(reinterpret_cast<const uint2*>(table) + (e[k] >> 1)); acc += v.x ^ v.y; }
      else acc += __ldg(table + e[k]);
    }
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

}  // namespace rnb

using namespace rnb;

extern "C" {

/* kind 0: red.global.add.f32, 1: .v2.f32, 2: .v4.f32 (aligned pairs), 3: ld.global.nc.b32, 4: ld.global.nc.b64.
 * table: `entries` x 8 bytes (reductions) or x 4 bytes (gathers); ops_out = lane-level accesses issued. */
int rnb_diag_l2_rate(int kind, void* table, int64_t entries, int iters, int active_lanes, int blocks, int64_t* ops_out,
                     rnb_stream_t stream) {
  const char* what = "rnb_diag_l2_rate";
  RNB_REQUIRE(table && entries > 1 && entries < (1ll << 31) && iters >= 1 && active_lanes >= 1 && active_lanes <= 32 && blocks >= 1, what);
  cudaStream_t st = as_stream(stream);
  const int64_t lanes = (int64_t)blocks * 256 / 32 * (kind <= 2 ? active_lanes : 32);
  if (ops_out) *ops_out = lanes * iters * 8;
  switch (kind) {
    case 0: mb_red_kernel<1><<<blocks, 256, 0, st>>>((float*)table, (uint32_t)entries, iters, active_lanes); break;
    case 1: mb_red_kernel<2><<<blocks, 256, 0, st>>>((float*)table, (uint32_t)entries, iters, active_lanes); break;
    case 2: mb_red_kernel<4><<<blocks, 256, 0, st>>>((float*)table, (uint32_t)entries, iters, active_lanes); break;
    case 3: mb_gather_kernel<4><<<blocks, 256, 0, st>>>((const uint32_t*)table, (uint32_t)entries, iters, (uint32_t*)table); break;
    case 4: mb_gather_kernel<8><<<blocks, 256, 0, st>>>((const uint32_t*)table, (uint32_t)entries, iters, (uint32_t*)table); break;
    default: RNB_REQUIRE(kind >= 0 && kind <= 4, what);
  }
  return check_launch(what);
}

}  // extern "C"
