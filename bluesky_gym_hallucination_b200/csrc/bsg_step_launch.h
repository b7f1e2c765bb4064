// bsg_step_launch.h -- host launchers of the step_kernel instantiations.  Each group of instantiations lives in its own
// translation unit (bsg_step_f32.cu, bsg_step_f32x.cu, bsg_step_f64.cu) so that the library builds in parallel; the
// C-ABI (bsg_api.cu) only sees these three functions.
#pragma once
#include <cuda_runtime.h>
#include "bsg_step.cuh"

namespace bsg {
constexpr int kMaxLaunchDev = 64;      // device ordinals the per-device launch caches cover
// `dev` = the device the launch goes to (the handle's): cudaFuncSetAttribute is per device, so is its once-only cache.
// tpb in {128, 256} (640: the wide detection-only symmetric instantiation); sym = symmetric half-ring sweep (fp32, N <= 64); ext = wind / LNAV / Merge / noise instantiations.
// Returns cudaErrorInvalidValue for a combination that is not instantiated.
cudaError_t launch_step_f32(int tpb, bool reso, bool sym, dim3 grid, dim3 block, size_t smem, cudaStream_t s, int dev, const StepArgs<float> &A);
// detection-only symmetric kernel, "wide" instantiation (tpb = 640): envs per CTA / record rows / shared memory of the
// one-wave launch shape that leaves the fullest SM the fewest warps; false when the default shape is at least as good
bool pick_wide_shape(int E, int N, int K, int def_epb, int def_threads, int *epb, int *nbuf, size_t *smem);
cudaError_t launch_step_f32_ext(int tpb, bool reso, dim3 grid, dim3 block, size_t smem, cudaStream_t s, int dev, const StepArgs<float> &A);
cudaError_t launch_step_f64(int tpb, bool reso, bool ext, dim3 grid, dim3 block, size_t smem, cudaStream_t s, int dev, const StepArgs<double> &A);

// one launch of instantiation K; raises the dynamic shared-memory limit of the instantiation once it is needed
template <typename K, typename Arg>
inline cudaError_t launch_one(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t s, const Arg &A, size_t &smem_set) {
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        smem_set = smem;
    }
    kernel<<<grid, block, smem, s>>>(A);
    return cudaGetLastError();
}
}  // namespace bsg
