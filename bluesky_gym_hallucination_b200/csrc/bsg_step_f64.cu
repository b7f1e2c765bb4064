// bsg_step_f64.cu -- fp64 parity-mode instantiations (literal operation order of the reference-style numpy code).
#include "bsg_step_kernel.cuh"
#include "bsg_step_launch.h"

namespace bsg {
cudaError_t launch_step_f64(int tpb, bool reso, bool ext, dim3 grid, dim3 block, size_t smem, cudaStream_t s, int dev, const StepArgs<double> &A) {
    static size_t set_all[kMaxLaunchDev][8] = { { 0 } };      // the shared-memory opt-in is per device
    if (dev < 0 || dev >= kMaxLaunchDev) return cudaErrorInvalidDevice;
    size_t *set = set_all[dev];
#define L(T, RESO, EXT, k) return launch_one(step_kernel<double, T, RESO, false, EXT>, grid, block, smem, s, A, set[k])
    if (tpb == 128) {
        if (ext) { if (reso) L(128, true, true, 0); else L(128, false, true, 1); }
        else { if (reso) L(128, true, false, 2); else L(128, false, false, 3); }
    } else if (tpb == 256) {
        if (ext) { if (reso) L(256, true, true, 4); else L(256, false, true, 5); }
        else { if (reso) L(256, true, false, 6); else L(256, false, false, 7); }
    }
#undef L
    return cudaErrorInvalidValue;
}
}  // namespace bsg
