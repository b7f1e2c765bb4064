// bsg_handle.h -- the opaque handle behind include/bsgcore.h (shared by the .cu files of the library).
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <cstdint>
#include "../../include/bsgcore.h"
#include "bsg_dev.cuh"

struct bsg_handle {
    bsg_config cfg;
    int obs_dim, act_dim, epb, tpb, nbuf;
    size_t smem;
    int wepb = 0, wnbuf = 0;          // wide launch shape of the detection-only symmetric fp32 kernel (0: default shape)
    size_t wsmem = 0;
    double *pos = nullptr;            // [2][E*N]
    void *st = nullptr;               // R [S_NS][E*N]
    unsigned long long *reso = nullptr;
    int *nac = nullptr;
    double *ev = nullptr;             // [E][BSG_NE]
    double *stage64 = nullptr;        // [E][BSG_NF][N] interchange staging (device)
    // host-buffer step: pinned staging + device mirrors
    void *h_act = nullptr, *h_out = nullptr, *d_act = nullptr, *d_out = nullptr;
    // host-buffer step, pipelined form: two pinned action / output blocks, one event each (bsg_step_host_begin / _wait)
    void *h_acts[2] = { nullptr, nullptr }, *h_outs[2] = { nullptr, nullptr };
    cudaEvent_t hev[2] = { nullptr, nullptr };
    int h_head = 0, h_pending = 0, h_last = 0;       // next block to fill, steps begun and not yet waited for, last waited block
    size_t out_bytes = 0, act_bytes = 0;
    cudaStream_t hs = nullptr;
    uint32_t *episode = nullptr;
    int autoreset = 0, flip = 0;      // next-step auto-reset: done_buf[flip] = envs reset after the last step
    uint64_t ar_seed = 0;
    bsg::ResetConst rc = {};          // host-evaluated ISA constants of the reset path
    bsg::WindArgs wind = {};          // bsg_set_wind
    double noise_sigma = 0.0;         // bsg_set_obs_noise
    uint64_t noise_seed = 0;
    int32_t *done_buf[2] = { nullptr, nullptr };      // [E] reset counter (Philox counter word)
    float *nd_rec = nullptr;          // bsg_ndmap_state: packed straight-line records of the live state [E][6][N]
    int64_t launches = 0;
    int64_t noise_steps = 0;          // step / observe launches only: the Philox counter word of the observation noise
    cudaEvent_t nd_ev = nullptr;      // bsg_ndmap_state on a foreign stream: the host-step stream waits for its read of the state
    // host-buffer step as a replayed CUDA graph (H2D copy, step kernel, D2H copy): one graph per (pinned block, auto-reset
    // flag buffer); `gen` counts the calls that change anything a captured launch holds (wind, noise, auto-reset, outputs)
    cudaGraphExec_t hgraph[2][2] = { { nullptr, nullptr }, { nullptr, nullptr } };
    int64_t hgraph_gen[2][2] = { { -1, -1 }, { -1, -1 } };
    int64_t gen = 0;
    int act_dir_last = -1;
    std::string err;
};

// bsg_reset.cu: bsg_reset (cont = false: episode 0 of `seed`) and the same-step auto-reset of bsg_step (cont = true: the
// env's next episode of the auto-reset stream).  Library-internal.
__attribute__((visibility("hidden"))) int bsg_reset_impl(bsg_handle *h, const int32_t *mask_dev, uint64_t seed, void *stream, bool cont);

#define CK(h, call)                                                                           \
    do {                                                                                      \
        cudaError_t _e = (call);                                                              \
        if (_e != cudaSuccess) {                                                              \
            if (h) (h)->err = std::string(#call) + ": " + cudaGetErrorString(_e);             \
            return BSG_ERR_CUDA;                                                              \
        }                                                                                     \
    } while (0)

