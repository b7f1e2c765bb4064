// bsg_step_f32.cu -- fp32 instantiations of the fused step kernel (the headline configs run these).
#include "bsg_step_kernel.cuh"
#include "bsg_step_launch.h"

namespace bsg {
cudaError_t launch_step_f32(int tpb, bool reso, bool sym, dim3 grid, dim3 block, size_t smem, cudaStream_t s, int dev, const StepArgs<float> &A) {
    static size_t set_all[kMaxLaunchDev][9] = { { 0 } };      // the shared-memory opt-in is per device
    if (dev < 0 || dev >= kMaxLaunchDev) return cudaErrorInvalidDevice;
    size_t *set = set_all[dev];
#define L(T, RESO, SYM, k) return launch_one(step_kernel<float, T, RESO, SYM>, grid, block, smem, s, A, set[k])
    if (tpb == 128) {
        if (sym) { if (reso) L(128, true, true, 0); else L(128, false, true, 1); }
        else { if (reso) L(128, true, false, 2); else L(128, false, false, 3); }
    } else if (tpb == 256) {
        if (sym) { if (reso) L(256, true, true, 4); else L(256, false, true, 5); }
        else { if (reso) L(256, true, false, 6); else L(256, false, false, 7); }
    } else if (tpb == 640 && sym && !reso) L(640, false, true, 8);
#undef L
    return cudaErrorInvalidValue;
}

// Launch shape of the detection-only symmetric kernel for a batch that fits ONE wave.  The SMs are issue-bound on this
// kernel, so the launch ends when the SM that holds the most WARPS ends: with 128-thread CTAs of whole envs, C2
// (4096 x 20) is 683 CTAs -- 4 or 5 per SM, 16 or 20 warps -- while the batch itself is 2560 full warps = 17.3 per SM.
// Candidates: `epb` envs per CTA (whole warps, <= 640 threads), `c` resident CTAs per SM by registers and shared
// memory; a candidate must hold the batch in one wave (grid <= SMs x c); the one with the fewest warps on the fullest
// SM wins, ties to the smaller CTA -- if it undercuts the default shape by 15 % (Vertical 4096 x 6: 196 CTAs of 4 warps
// = 4 or 8 warps per SM -> 293 CTAs of 3 warps, 6 per SM; C2 would go 20 -> 18 warps, measured a wash, and stays).
bool pick_wide_shape(int E, int N, int K, int def_epb, int def_threads, int *epb_out, int *nbuf_out, size_t *smem_out) {
    int dev = 0, nsm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return false;
    cudaFuncAttributes fa;
    if (cudaFuncGetAttributes(&fa, step_kernel<float, 640, false, true>) != cudaSuccess) return false;
    const int regs = fa.numRegs > 0 ? fa.numRegs : 96;
    const size_t smem_sm = 227 * 1024;
    const int def_grid = (E + def_epb - 1) / def_epb;
    // warps on the fullest SM with the default shape; a wide shape must undercut it by 15 % (measured on C2: 20 -> 18
    // warps is a wash against the 9-warp CTA's longer barriers, Vertical 4096 x 6: 8 -> 6 warps is -8 % per step)
    long best = (long)((def_grid + nsm - 1) / nsm) * (def_threads / 32) * 85 / 100;
    int best_epb = 0, best_nb = 0, best_thr = 0; size_t best_smem = 0;
    for (int epb = 1; epb * N <= 640 && epb <= E; epb++) {
        const int thr = ((epb * N + 31) / 32) * 32, warps = thr / 32;
        const int grid = (E + epb - 1) / epb;
        const int regs_cta = ((regs * 32 + 255) / 256) * 256 * warps;          // allocation granularity: 256 registers per warp
        for (int c = 65536 / regs_cta; c >= 1; c--) {
            if ((long)grid > (long)nsm * c) break;                              // not one wave (fewer CTAs per SM: even less so)
            const size_t budget = smem_sm / c - 1024;                           // 1 KB per CTA is reserved by the system
            int nb = BSG_CH < K ? BSG_CH : (K > 2 ? K : 2);
            while (nb > 2 && step_smem<float>(N, epb, nb, false).total > budget) nb--;
            const size_t sm = step_smem<float>(N, epb, nb, false).total;
            if (sm > budget) continue;
            const long crit = (long)((grid + nsm - 1) / nsm) * warps;
            if (crit < best || (crit == best && best_epb && thr < best_thr) || (crit == best && !best_epb)) {
                best = crit; best_epb = epb; best_nb = nb; best_thr = thr; best_smem = sm;
            }
            break;                                                              // the most resident CTAs this epb allows
        }
    }
    if (!best_epb) return false;
    *epb_out = best_epb; *nbuf_out = best_nb; *smem_out = best_smem;
    return true;
}
#ifdef BSG_TRACE
int trace_read_f32(unsigned long long *out_host, int n_words) {
    return cudaMemcpyFromSymbol(out_host, g_trace, (size_t)n_words * sizeof(unsigned long long)) == cudaSuccess ? 0 : -1;
}
#endif
}  // namespace bsg
