// tc_probe.cu -- self-test of the tcgen05 building blocks in tc05.cuh (descriptor encodings, operand
// majorness, accumulator layouts) on one 128-thread CTA.  Used by tests/test_tc05_gpu.py; it pins the
// conventions the fused tensor-core field kernels rely on before they are trusted.
#include "tc05.cuh"

namespace rnb {

// X (128,64) f16 row-major, W (64,64) f16 row-major (out,in), Y (128,32) f16 row-major.
//   out1 (128,64) = X W^T            A K-major (M=128), B K-major          -- forward layer
//   out2 (128,32) = raw TMEM dump of X^T Y  (M=64, N=32), A and B MN-major -- weight gradient
//   out3 (128,64) = X W              A K-major, B MN-major                 -- data gradient
__global__ void __launch_bounds__(128) tc_probe_kernel(const __half* __restrict__ X, const __half* __restrict__ W,
                                                       const __half* __restrict__ Y, float* __restrict__ out1,
                                                       float* __restrict__ out2, float* __restrict__ out3) {
  __shared__ __align__(128) unsigned char sx[128 * 64 * 2];
  __shared__ __align__(128) unsigned char sw[64 * 64 * 2];
  __shared__ __align__(128) unsigned char sy[128 * 32 * 2];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  // stage operands in the chunked layout: thread = row
  for (int c = 0; c < 8; c++) *reinterpret_cast<uint4*>(sx + tc::tile_chunk_off(128, tid, c)) = reinterpret_cast<const uint4*>(X + tid * 64)[c];
  for (int c = 0; c < 4; c++) *reinterpret_cast<uint4*>(sy + tc::tile_chunk_off(128, tid, c)) = reinterpret_cast<const uint4*>(Y + tid * 32)[c];
  if (tid < 64)
    for (int c = 0; c < 8; c++) *reinterpret_cast<uint4*>(sw + tc::tile_chunk_off(64, tid, c)) = reinterpret_cast<const uint4*>(W + tid * 64)[c];
  if (warp == 0) tc::tmem_alloc(&tmem_base_s, 256);
  if (tid == 0) tc::mbar_init(&bar, 1);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_base_s;
  if (tid == 0) {
    const uint32_t ax = tc::smem_u32(sx), aw = tc::smem_u32(sw), ay = tc::smem_u32(sy);
    // 1: D1 = X W^T
    for (int k = 0; k < 4; k++)
      tc::mma_f16(tmem + 0, tc::smem_desc(ax + k * 2 * 2048, 2048, 128), tc::smem_desc(aw + k * 2 * 1024, 1024, 128),
                  tc::instr_desc(128, 64, 0, 0), k > 0);
    // 2: D2 = X^T Y, contraction over the 128 rows, 16 per MMA
    for (int k = 0; k < 8; k++)
      tc::mma_f16(tmem + 64, tc::smem_desc(ax + k * 256, 128, 2048), tc::smem_desc(ay + k * 256, 128, 2048),
                  tc::instr_desc(64, 32, 1, 1), k > 0);
    // 3: D3 = X W, B = W as MN-major (N = in features, K = out rows)
    for (int k = 0; k < 4; k++)
      tc::mma_f16(tmem + 128, tc::smem_desc(ax + k * 2 * 2048, 2048, 128), tc::smem_desc(aw + k * 256, 128, 1024),
                  tc::instr_desc(128, 64, 0, 1), k > 0);
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::fence_after_sync();
  const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
  float v[16];
  for (int c = 0; c < 64; c += 16) {
    tc::tmem_ld16(lane_addr + c, v);
    for (int i = 0; i < 16; i++) out1[tid * 64 + c + i] = v[i];
  }
  for (int c = 0; c < 32; c += 16) {
    tc::tmem_ld16(lane_addr + 64 + c, v);
    for (int i = 0; i < 16; i++) out2[tid * 32 + c + i] = v[i];
  }
  for (int c = 0; c < 64; c += 16) {
    tc::tmem_ld16(lane_addr + 128 + c, v);
    for (int i = 0; i < 16; i++) out3[tid * 64 + c + i] = v[i];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

}  // namespace rnb

using namespace rnb;

extern "C" int rnb_tc_probe(const void* X, const void* W, const void* Y, float* out1, float* out2, float* out3, rnb_stream_t stream) {
  const char* what = "rnb_tc_probe";
  RNB_REQUIRE(X && W && Y && out1 && out2 && out3, what);
  tc_probe_kernel<<<1, 128, 0, as_stream(stream)>>>((const __half*)X, (const __half*)W, (const __half*)Y, out1, out2, out3);
  return check_launch(what);
}
