// bsg_step_f32.cu -- fp32 instantiations of the fused step kernel (the headline configs run these).
#include "bsg_step_kernel.cuh"
#include "bsg_step_launch.h"

namespace bsg {
cudaError_t launch_step_f32(int tpb, bool reso, bool sym, dim3 grid, dim3 block, size_t smem, cudaStream_t s, const StepArgs<float> &A) {
    static size_t set[8] = { 0 };
#define L(T, RESO, SYM, k) return launch_one(step_kernel<float, T, RESO, SYM>, grid, block, smem, s, A, set[k])
    if (tpb == 128) {
        if (sym) { if (reso) L(128, true, true, 0); else L(128, false, true, 1); }
        else { if (reso) L(128, true, false, 2); else L(128, false, false, 3); }
    } else if (tpb == 256) {
        if (sym) { if (reso) L(256, true, true, 4); else L(256, false, true, 5); }
        else { if (reso) L(256, true, false, 6); else L(256, false, false, 7); }
    }
#undef L
    return cudaErrorInvalidValue;
}
}  // namespace bsg
