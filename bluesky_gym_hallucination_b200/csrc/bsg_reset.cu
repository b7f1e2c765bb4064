// bsg_reset.cu -- stand-alone device-side reset (bsg_reset): one CTA per selected env, built from the two
// stages of bsg_reset.cuh.  The next-step auto-reset of bsg_step runs the same stages inside step_kernel.
#include "bsg_handle.h"
#include "bsg_reset.cuh"

using namespace bsg;

namespace {

// The stand-alone reset draws episode 0 of the stream keyed by `seed` and leaves the env's episode counter at 1:
// calling it again with the same seed reproduces the same scenario (Gymnasium's reset(seed=s) contract), and the
// auto-reset episodes that follow (fused into bsg_step, keyed by the auto-reset seed) count on from 1.
// CONT (the same-step auto-reset of bsg_step): the env's next episode of the auto-reset stream instead -- the counter is
// read and advanced, as the fused next-step path does, so that successive episodes of an env differ.
template <typename R, bool CONT>
__global__ void reset_kernel(int kind, int fixed_nac, int density_mode, int E, int N, const int32_t *mask, uint64_t seed,
                             uint32_t *episode, double *pos, R *st, unsigned long long *reso, int *nac, double *ev,
                             const ResetConst RC) {
    const int e = blockIdx.x;
    if (mask && !mask[e]) return;                       // uniform per CTA
    __shared__ ResetEnv S;
    __shared__ double wdraw[256], wsort[256];           // bsg_create: max_ac <= 256
    __shared__ uint32_t ep_s;
    if (threadIdx.x == 0) ep_s = CONT ? episode[e] : 0u;   // one reader: the counter is rewritten below
    __syncthreads();
    const uint32_t ep = ep_s;
    double *v = ev + (size_t)e * BSG_NE;
    if (threadIdx.x == 0) {
        reset_env_level(kind, fixed_nac, density_mode, N, e, seed, ep, v, &S, RC);
        nac[e] = S.nac;
        episode[e] = ep + 1u;
    }
    __syncthreads();
    if (kind == BSG_KIND_SECTOR) {
        for (int i = threadIdx.x; i < N; i += blockDim.x) reset_wpt_draw(e, i, seed, ep, &S, wdraw);
        __syncthreads();
        for (int i = threadIdx.x; i < N; i += blockDim.x) reset_wpt_rank(i, &S, wdraw, wsort);
        __syncthreads();
    }
    for (int i = threadIdx.x; i < N; i += blockDim.x) reset_aircraft<R>(kind, E, N, e, i, seed, ep, &S, wsort, v, pos, st, reso, RC);
}

}  // namespace

int bsg_reset_impl(bsg_handle *h, const int32_t *mask_dev, uint64_t seed, void *stream, bool cont) {
    if (!h) return BSG_ERR_ARG;
    const bsg_config &g = h->cfg;
    if (g.env_kind == BSG_KIND_RAW) { h->err = "bsg_reset: RAW handles take their traffic through bsg_set_state"; return BSG_ERR_UNSUPPORTED; }
    CK(h, cudaSetDevice(g.device));
    if (!h->episode) {
        CK(h, cudaMalloc(&h->episode, g.num_envs * sizeof(uint32_t)));
        CK(h, cudaMemset(h->episode, 0, g.num_envs * sizeof(uint32_t)));
    }
    cudaStream_t s = (cudaStream_t)stream;
#define BSG_RESET_LAUNCH(R, CONT)                                                                                       \
    reset_kernel<R, CONT><<<g.num_envs, 64, 0, s>>>(g.env_kind, g.fixed_nac, g.density_mode, g.num_envs, g.max_ac, mask_dev, \
                                                    seed, h->episode, h->pos, (R *)h->st, h->reso, h->nac, h->ev, h->rc)
    if (g.precision == BSG_FP64) { if (cont) BSG_RESET_LAUNCH(double, true); else BSG_RESET_LAUNCH(double, false); }
    else { if (cont) BSG_RESET_LAUNCH(float, true); else BSG_RESET_LAUNCH(float, false); }
#undef BSG_RESET_LAUNCH
    h->launches++;
    CK(h, cudaGetLastError());
    return BSG_OK;
}

extern "C" int bsg_reset(bsg_handle *h, const int32_t *mask_dev, uint64_t seed, void *stream) {
    return bsg_reset_impl(h, mask_dev, seed, stream, false);
}
