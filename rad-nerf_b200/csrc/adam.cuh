// adam.cuh -- the step-guard prologue shared by the optimizer kernels (engine.cu, peer.cu); see RNB_GRAD_TAIL in
// include/radnerf_b200.h.  GradScaler equivalent of the reference's Lightning precision=16 (train_ml.py:304): a step whose
// gradients hold an inf / nan (or whose samples were truncated) is skipped on the device, without a host round trip.
#pragma once
#include "common.cuh"

namespace rnb {

// Step guard prologue shared by the Adam kernels: thread 0 of every block reads the flag word(s) and the count of
// applied steps, derives the bias corrections (double, like the host used to), and takes a ticket; the LAST block to
// take one -- every block has read the status by then -- updates the counters.
struct AdamStep { float step_size, inv_sqrt_bc2; int skip; };
__device__ __forceinline__ void adam_prologue(AdamStep* sh, int flag_bits, int32_t* status, int64_t host_step, float lr, float beta1,
                                              float beta2) {
  if (threadIdx.x == 0) {
    const int64_t step = status != nullptr ? (int64_t)*(volatile int32_t*)(status + 3) + 1 : host_step;
    if (lr < 0.f && status != nullptr) lr = __int_as_float(*(volatile int32_t*)(status + 8));   // schedule value written by the host (graph replay)
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    sh->step_size = (float)((double)lr / bc1); sh->inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2)); sh->skip = flag_bits;
    if (status != nullptr) {
      __threadfence();
      if (atomicAdd(status + 4, 1) == (int)gridDim.x - 1) {
        status[4] = 0;
        if (flag_bits & RNB_FLAG_CAPACITY) status[1] += 1;
        else if (flag_bits) status[0] += 1;
        else status[3] += 1;
      }
    }
  }
  __syncthreads();
}

}  // namespace rnb
