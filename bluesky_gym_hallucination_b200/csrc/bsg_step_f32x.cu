// bsg_step_f32x.cu -- fp32 EXT instantiations (wind field, LNAV routes, Merge env layer, observation noise; N <= 64).
#include "bsg_step_kernel.cuh"
#include "bsg_step_launch.h"

namespace bsg {
cudaError_t launch_step_f32_ext(int tpb, bool reso, dim3 grid, dim3 block, size_t smem, cudaStream_t s, int dev, const StepArgs<float> &A) {
    static size_t set_all[kMaxLaunchDev][4] = { { 0 } };      // the shared-memory opt-in is per device
    if (dev < 0 || dev >= kMaxLaunchDev) return cudaErrorInvalidDevice;
    size_t *set = set_all[dev];
#define L(T, RESO, k) return launch_one(step_kernel<float, T, RESO, true, true>, grid, block, smem, s, A, set[k])
    if (tpb == 128) { if (reso) L(128, true, 0); else L(128, false, 1); }
    else if (tpb == 256) { if (reso) L(256, true, 2); else L(256, false, 3); }
#undef L
    return cudaErrorInvalidValue;
}
}  // namespace bsg
