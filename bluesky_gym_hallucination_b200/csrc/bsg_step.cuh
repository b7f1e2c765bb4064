// bsg_step.cuh -- the fused env-step kernel: action intake -> K x (CD sweep [+MVP, resumenav] ->
// kinematics) -> observation / reward / termination / info, one launch for all envs.
// Replaces, per env and per step, the reference's Python:
//   HorizontalCREnv.step  bluesky_gym/envs/horizontal_cr_env.py:103-125 (+150-275)
//   VerticalCREnv.step    bluesky_gym/envs/vertical_cr_env.py:283-305   (+113-256)
//   SectorCREnv.step      bluesky_gym/envs/sector_cr_env.py:117-132     (+134-139, 227-339)
// Mapping: one thread = one (env, aircraft) row; a CTA holds epb whole envs so the env's packed CD
// records live in shared memory (double-buffered, one __syncthreads per substep) and every thread
// sweeps its own row of the N x N pair matrix from broadcast smem reads; row reductions
// (any / count / min+argmin / max) stay in registers.
#pragma once
#include "bsg_kin.cuh"
#include "bsg_pair.cuh"

namespace bsg {

template <typename R> struct StepArgs {
    int kind, E, N, substeps, asas_on, reso_on, reso_off_own, max_steps, epb, observe_only;
    int nbuf;             // record rows per env in shared memory (step_nbuf)
    int reso_mode;        // MVP resolution method: 0 horizontal (default), 1 vertical only, 2 horizontal + vertical (ordered-sweep kernels)
    int force_walk;       // test hook (BSG_FORCE_WALK): resolution kernels drop the pair work list, every owner walks its row
    R dt;
    CdParams P;
    CdParamsF PF;         // P rounded to fp32 (cd_params_fast)
    double *pos; R *st; unsigned long long *reso; int *nac; double *ev;
    const R *actions;
    const int *skip;      // envs reset after the previous step (auto-reset): observe only, reward 0
    int *done_next;       // auto-reset: the library's own copy of `done` (the next launch's skip flags), written beside out.done
    // fused next-step auto-reset: the envs flagged in `skip` draw their new episode at the head of this launch
    // (bsg_reset.cuh stages) instead of in a separate reset launch; `nac` is written by that stage
    int fuse_reset, fixed_nac, density_mode; unsigned long long ar_seed; unsigned int *episode; ResetConst RC;
    bsg_step_out out;
    // EXT kernels only: wind field, observation noise (sigma, Philox key, launch counter)
    WindArgs W;
    R noise_sigma; unsigned long long noise_seed; unsigned int noise_ctr;
};

template <typename R> struct RecOf { using type = RecLit; };
template <> struct RecOf<float> { using type = RecFast; };

// per-env scratch in shared memory
template <typename R> struct EnvShared {
    double own_lat, own_lon;
    R own_hdg, own_tas, own_alt, own_vs, drift, wdist;
    int anyconf[2];
    int nC, nR, ovf;                 // work list of the pair scratch: conflicting / resolving-only items; a slot row overflowed
    int vneed;                       // some aircraft of the env is not level / co-altitude: generic vertical window
    R alt_ref;
    int intr;
    double ev[BSG_EV_POLY];          // the env-level scalar slots, staged once per launch by the agent thread
};

template <typename R>
__device__ __forceinline__ void make_rec(const AcReg<R> &a, double lat0, double lon0, typename RecOf<R>::type &r) {
    if constexpr (sizeof(R) == 8) {
        r.lat = a.lat; r.lon = a.lon; r.alt = a.alt; r.u = a.gse; r.v = a.gsn; r.vs = a.vs; r.hdg = a.hdg; r.pad = 0;
    } else {
        r.x = (float)((a.lon - lon0) * (kD2R * kRe));
        r.y = (float)((a.lat - lat0) * (kD2R * kRe));
        r.alt = a.alt; r.u = a.gse; r.v = a.gsn; r.vs = a.vs;
        r.sa = a.sa; r.ca = a.ca;
    }
}

// distance [NM] and bearing [deg] own -> other for the env layer (kwikqdrdist call sites, e.g.
// horizontal_cr_env.py:169,190,265).  Fast path: differences in fp64, everything else fp32.
template <typename R>
__device__ __forceinline__ void env_qdrdist(double lata, double lona, double latb, double lonb, R &qdr, R &dnm) {
    if constexpr (sizeof(R) == 8) {
        kwikqdrdist(lata, lona, latb, lonb, qdr, dnm);
    } else {
        float dlat = (float)((latb - lata) * kD2R);
        float dlon = (float)((lonb - lona) * kD2R);                  // |dlon| << 180 in every env: no wrap needed
        float cav = cosf((float)((lata + latb) * (kD2R * 0.5)));
        float dxe = dlon * cav;
        dnm = float(kRe / kNm) * sqrtf(fmaf(dlat, dlat, dxe * dxe));
        float q = atan2f(dxe, dlat) * float(kR2D);
        qdr = q < 0.0f ? q + 360.0f : q;
    }
}

template <typename R> __device__ __forceinline__ void rsincos_deg(R deg, R &s, R &c) {
    if constexpr (sizeof(R) == 8) { s = sin(deg * kD2R); c = cos(deg * kD2R); }
    else sincos_deg_fast(deg, s, c);                 // (|deg| <= 720 at every call site)
}

constexpr double kCenterLat = 51.990426702297746, kCenterLon = 4.376124857109851;   // sector_cr_env.py:16

__device__ __forceinline__ bool poly_inside(const double *poly, int nv, double tx, double ty) {
    bool inside = false;
    double x0 = poly[2 * (nv - 1)], y0 = poly[2 * (nv - 1) + 1];
    bool f0 = y0 >= ty;
    for (int k = 0; k < nv; k++) {
        double x1 = poly[2 * k], y1 = poly[2 * k + 1];
        bool f1 = y1 >= ty;
        if (f0 != f1)
            if ((((y1 - ty) * (x0 - x1)) >= ((x1 - tx) * (y0 - y1))) == f1) inside = !inside;
        f0 = f1; x0 = x1; y0 = y1;
    }
    return inside;
}

}  // namespace bsg
