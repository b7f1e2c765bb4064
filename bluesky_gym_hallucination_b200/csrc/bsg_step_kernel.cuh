// bsg_step_kernel.cuh -- body of the fused env-step kernel (see bsg_step.cuh for the mapping).
#pragma once
#include <type_traits>
#include "bsg_step.cuh"
#include "bsg_env.cuh"
#include "bsg_reset.cuh"

#ifndef BSG_PW
#define BSG_PW 5          // half-ring loop: pairs per trip, straight-line so that the dependency chains interleave
#endif
#ifndef BSG_CH
#define BSG_CH 5          // detection-only kernels: substeps integrated ahead of one barrier (record rows per env)
#endif

namespace bsg {

// -DBSG_TRACE: per-CTA %globaltimer stamps at the phase boundaries of step_kernel (tools/gpu_trace.py)
#ifdef BSG_TRACE
__device__ unsigned long long g_trace[8 * 16384];
#define BSG_STAMP(slot) do { if (threadIdx.x == 0 && blockIdx.x < 16384) { unsigned long long t_; \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_trace[blockIdx.x * 8 + (slot)] = t_; } } while (0)
#else
#define BSG_STAMP(slot) do { } while (0)
#endif

// the fp32 CD thresholds, rounded once on the host (launch_step) and carried in StepArgs
inline void cd_params_fast(const CdParams &P, CdParamsF &fast) {
    fast.R2 = (float)(P.rpz * P.rpz); fast.hpz = (float)P.hpz; fast.dtlook = (float)P.dtlook;
    fast.rpz = (float)P.rpz; fast.rpz_m = (float)(P.rpz * P.fach); fast.hpz_m = (float)(P.hpz * P.facv);
    fast.fach2 = (float)(P.fach * P.fach);
}

// CTA size / residency: 128-thread CTAs for the small-N configs (five resident per SM), 256 threads only when one env
// needs them (N > 128).  TPB = 640 is the "wide" detection-only instantiation: the same register cap as 128 x 5
// (65536 / 640 = 102), launched with ANY block size up to 640 threads -- the host picks the number of envs per CTA that
// leaves every SM the same number of warps in ONE wave (pick_wide_shape, bsg_step_f32.cu).
template <typename R, int TPB> struct StepBounds {
    static constexpr int kMinBlocks = TPB == 640 ? 1 : TPB == 128 ? (sizeof(R) == 4 ? 5 : 3) : (sizeof(R) == 4 ? 2 : 1);
};
// Per-env row of CD records in shared memory, `nbuf` rows per env (2 = double-buffered by substep parity; the
// detection-only kernels keep up to BSG_CH rows, one per substep of a chunk).
// fp64 / fp32 with N > 64: N records + 16 B, so that the rows of the envs sharing a warp start in different banks.
// fp32: N + N/2 records as two float4 arrays [N + N/2] (x,y,alt,u | v,vs,ca,sa; conflict-free LDS.128) -- the
// symmetric sweep keeps a copy of the first half ring behind the last live slot, so j = i + d never wraps -- and the
// N headings (resumenav's bouncing test) behind them, in the same substep buffer.
__host__ __device__ constexpr int sym_row_recs(int N) { return N + N / 2; }
template <typename R> __host__ __device__ constexpr size_t rec_row_bytes(int N) {
    return sizeof(R) == 4 ? (size_t)sym_row_recs(N) * sizeof(RecFast) + (((size_t)N * sizeof(float) + 15) & ~(size_t)15) + 16
                          : (size_t)N * sizeof(typename RecOf<R>::type) + 16;
}
__device__ __forceinline__ RecFast rec_from(const float4 &a, const float4 &b) {
    RecFast r; r.x = a.x; r.y = a.y; r.alt = a.z; r.u = a.w; r.v = b.x; r.vs = b.y; r.ca = b.z; r.sa = b.w; return r;
}

// shared-memory carve-up of one CTA (host and device agree through this one function)
// Pair scratch of one env (fp32 symmetric sweep WITH resolution).  The env's unordered conflicting / resolving pairs
// are evaluated ONCE each, by whichever thread of the env is free (work list), and the results are handed to both
// aircraft of the pair:
//   rows  u32 x6 per aircraft: its conflict row, resumenav's keep row, LoS row (absolute bits, two words each), built
//                with shared-memory atomicOr by whoever evaluates the pair
//   items u16    work list, (i' << 5 | d-1) = the pair {i', i'+d} of the forward half ring of thread i': conflicting
//                pairs from the front, resolving-only pairs from the back; at most BSG_PAIR_CAP * N items
//   res   float2 [N][BSG_SLOT_CAP]: aircraft i's MVP displacements, slot = rank of the intruder among its conflicts
//                (ascending index: the owner's sum keeps the order of the reference's loop)
//   tcp   float  [N][BSG_SLOT_CAP]: tCPA of those pairs (last detection only: the CD row outputs)
// An env that overflows the list or a slot row falls back, for that detection, to the walk in which every owner evaluates
// its own row (same bits: the pair arithmetic is exactly antisymmetric).
#ifndef BSG_PAIR_CAP
#define BSG_PAIR_CAP 8
#endif
#ifndef BSG_SLOT_CAP
#define BSG_SLOT_CAP 16
#endif
__host__ __device__ constexpr int pair_cap(int N) { return BSG_PAIR_CAP * N < N * (N / 2) ? BSG_PAIR_CAP * N : N * (N / 2) + 1; }
__host__ __device__ constexpr int slot_cap(int N) { return BSG_SLOT_CAP < N ? BSG_SLOT_CAP : N; }
__host__ __device__ constexpr size_t pair_scratch_bytes(int N) {
    return (((size_t)N * slot_cap(N) * (sizeof(float2) + sizeof(float)) + (size_t)N * 24 + (size_t)pair_cap(N) * sizeof(unsigned short)) + 15) & ~(size_t)15;
}
struct StepSmem { size_t es_off, sd_off, cr_off, px_off, total; };
template <typename R> __host__ __device__ inline StepSmem step_smem(int N, int epb, int nbuf, bool pairs) {
    StepSmem L;
    L.es_off = (size_t)nbuf * epb * rec_row_bytes<R>(N);                   // [nbuf][epb] record rows
    L.sd_off = L.es_off + (size_t)epb * sizeof(EnvShared<R>);              // [epb] env scratch
    L.cr_off = (L.sd_off + (size_t)epb * N * sizeof(R) + 7) & ~(size_t)7;  // [epb*N] hdg / sector distances
    L.px_off = (L.cr_off + 2 * (size_t)epb * N * sizeof(unsigned long long) + 15) & ~(size_t)15;   // [2][epb*N] conflict bit rows (SYM)
    L.total = L.px_off + (pairs ? (size_t)epb * pair_scratch_bytes(N) : 0); // [epb] pair scratch
    return L;
}
// record rows per env: the detection-only fp32 kernels (no resolution feeding back into the kinematics) integrate a
// chunk of substeps ahead of ONE barrier, so they keep one row per substep of the chunk
template <typename R> __host__ inline int step_nbuf(int N, int epb, bool reso) {
    if (sizeof(R) != 4 || N > 64 || reso) return 2;
    int nb = BSG_CH;
    while (nb > 2 && step_smem<R>(N, epb, nb, false).total > 44 * 1024) nb--;   // five 128-thread CTAs per SM
    return nb;
}

// Next-step auto-reset, fused: an env that ended in the previous step draws its new episode at the head of this
// launch (rare: 1-3 % of the envs per step), so a step is ONE launch.  Called by every thread of a CTA that holds
// such an env; kept out of line so the step kernel's register allocation does not see it.  The ResetEnv scratch
// aliases the (not yet used) record rows.
template <typename R>
static __device__ __noinline__ void fused_reset(int kind, int fixed_nac, int density_mode, int E, int N, int epb, unsigned long long seed,
                                                unsigned int *episode, double *ev, int *nac, double *pos, R *st,
                                                unsigned long long *reso, const ResetConst RC, unsigned char *smem_raw,
                                                bool skip, bool valid, int le, int i, int e) {
    // (scalars by value: taking the address of the kernel's parameter struct would copy it to local memory)
    ResetEnv *rs = reinterpret_cast<ResetEnv *>(smem_raw);
    double *wdraw = reinterpret_cast<double *>(rs + epb) + (size_t)le * 2 * N, *wsort = wdraw + N;   // Sector waypoints
    double *evr = ev + (size_t)(valid ? e : 0) * BSG_NE;
    const unsigned int ep = skip ? episode[e] : 0u;
    if (skip && i == 0) {
        reset_env_level(kind, fixed_nac, density_mode, N, e, seed, ep, evr, &rs[le], RC);
        nac[e] = rs[le].nac;
    }
    __syncthreads();
    if (kind == BSG_KIND_SECTOR) {                       // uniform: every thread of the CTA is here
        if (skip) reset_wpt_draw(e, i, seed, ep, &rs[le], wdraw);
        __syncthreads();
        if (skip) reset_wpt_rank(i, &rs[le], wdraw, wsort);
        __syncthreads();
    }
    if (skip) reset_aircraft<R>(kind, E, N, e, i, seed, ep, &rs[le], wsort, evr, pos, st, reso, RC);
    __syncthreads();
    if (skip && i == 0) episode[e] = ep + 1;
}

// RESO (MVP + resumenav) is a compile-time switch: without it the pair loop is a clean, unrollable FP32 body.
// SYM (fp32, N <= 64): every unordered pair is evaluated ONCE -- thread i sweeps the half ring j = i+1 .. i+N/2
// (the fp32 pair form is exactly symmetric under i <-> j, so the ordered matrix it stands for is bit-identical
// to the full sweep).  The flag of pair (i, i+d) is the sign bit of one 3-input max, funnel-shifted into a register
// mask that the thread publishes once per substep (no atomics, no branch); after one more barrier each thread
// completes its own row from the masks of the partners behind it and finishes it in two walks: conflicting pairs
// (row reductions, MVP, resumenav), then pairs that are only still resolving (resumenav).
// EXT: the wrapper / Merge features (wind field, LNAV route, Merge env layer, observation noise; SURVEY f-3, f-4) as a
// compile-time switch, so the instantiations the headline configs run are unchanged by them.
template <typename R, int TPB, bool RESO, bool SYM, bool EXT = false>
__global__ void __launch_bounds__(TPB, (StepBounds<R, TPB>::kMinBlocks)) step_kernel(const StepArgs<R> A) {
    using Rec = typename RecOf<R>::type;
    static_assert(!SYM || sizeof(R) == 4, "the symmetric sweep is the fp32 fast path");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = A.N, epb = A.epb;
    const size_t rowb = rec_row_bytes<R>(N);
    const StepSmem SL = step_smem<R>(N, epb, A.nbuf, RESO && SYM);
    unsigned char *recs = smem_raw;                                        // [nbuf][epb] rows of rowb bytes
    EnvShared<R> *es = reinterpret_cast<EnvShared<R> *>(smem_raw + SL.es_off);   // [epb]
    R *sdist = reinterpret_cast<R *>(smem_raw + SL.sd_off);                // [epb*N] hdg (resumenav) / sector distances
    unsigned long long *crow = reinterpret_cast<unsigned long long *>(smem_raw + SL.cr_off);   // [2][epb*N]

    BSG_STAMP(0);
    const int tid = threadIdx.x;
    const int le = tid / N, i = tid - le * N;
    const int e = blockIdx.x * epb + le;
    const bool valid = le < epb && e < A.E;
    const size_t EN = (size_t)A.E * N, k = (size_t)(valid ? e : 0) * N + i;
    double *ev = A.ev + (size_t)(valid ? e : 0) * BSG_NE;
    constexpr bool reso = RESO;
    // ------------------------------------------------------------------ prologue loads (the inputs are cold: every
    // dependent batch is a DRAM round trip): episode flag, aircraft count, env scalars, action -- and, speculatively in
    // the same batch, the aircraft state.
    AcReg<R> a = {};
    unsigned long long resorow = 0;
    int skipv = 0, nlv = 0;
    double ev_mine = 0.0, ev_hdg = 0.0;                                     // env scalar slot i; the agent's AC_HDG
    R act0 = R(0), act1 = R(0);                                             // the agent's action
    auto load_env = [&]() {
        if (valid) {
            nlv = A.nac[e];
            if (i < BSG_EV_POLY) ev_mine = ev[i];
            if (i == 0) ev_hdg = ev[BSG_EV_AC_HDG];
        }
    };
    if (valid && A.skip) skipv = A.skip[e];
    load_env();
    if (valid && i == 0 && A.actions) {
        const bool two = A.kind == BSG_KIND_SECTOR || A.kind == BSG_KIND_MERGE;
        act0 = A.actions[two ? 2 * e : e];
        if (two) act1 = A.actions[2 * e + 1];
    }
    // the aircraft state rides in the same batch, speculatively: every slot i < N of a valid env exists in memory
    // whether it is live or not, and the rare reset stage reloads it
    auto load_ac = [&]() {
        if (valid) {
            ac_load_raw<R, RESO, EXT>(a, A.pos, A.st, EN, k);
            if (reso) resorow = A.reso[k];
        }
    };
    // (detection-only kernels: their launches are often ONE wave, where nothing hides a second round trip; the
    // resolution kernels run many waves and keep the state load behind the reset stage -- measured, C3 +1 % otherwise)
    constexpr bool SPEC = !RESO;
    if constexpr (SPEC) load_ac();
    const bool skip = valid && skipv;                                      // fresh episode: no action, no integration
    if (A.fuse_reset && __syncthreads_or(skip)) {                          // rare: some env of this CTA starts an episode
        if constexpr (SPEC) { a = AcReg<R>{}; resorow = 0; }               // (nothing of the speculative load lives across the call)
        fused_reset<R>(A.kind, A.fixed_nac, A.density_mode, A.E, A.N, A.epb, A.ar_seed, A.episode, A.ev, A.nac, A.pos, A.st, A.reso,
                       A.RC, smem_raw, skip, valid, le, i, e);
        load_env();
        if constexpr (SPEC) load_ac();
    }
    BSG_STAMP(7);
    const int nl = valid ? nlv : 0;
    const bool live = valid && i < nl;
    if constexpr (SPEC) {
        if (live) ac_derive<R, EXT>(a);
        else { a = AcReg<R>{}; resorow = 0; }                              // dead slots: the zero state
    } else if (live) { load_ac(); ac_derive<R, EXT>(a); }
    // the env-level scalar slots, staged in shared memory by the first lanes of the env
    if (valid) {
        if (i < BSG_EV_POLY) es[le].ev[i] = ev_mine;
        for (int s2 = i + N; s2 < BSG_EV_POLY; s2 += N) es[le].ev[s2] = ev[s2];       // N < 14 only
    }
    // ------------------------------------------------------------------ action intake (agent = aircraft 0)
    if (live && i == 0) {
        es[le].anyconf[0] = 0; es[le].anyconf[1] = 0; es[le].nC = 0; es[le].nR = 0; es[le].ovf = 0; es[le].intr = 0; es[le].vneed = 0; es[le].alt_ref = a.alt;
        es[le].own_lat = a.lat; es[le].own_lon = a.lon;                    // local origin of the fp32 CD records
        if (A.actions && !skip) {
            // AP.selhdgcmd under a wind field: the HDG argument becomes the track it gives (above 50 ft)
            auto hdg_cmd = [&](R hdg) -> R {
                if constexpr (EXT) {
                    if (A.W.n > 0 && a.alt > R(50.0 * kFt)) {
                        R vn, ve, sh, ch;
                        wind_at<R>(A.W, a.lat, a.lon, 0.0, vn, ve);      // (selhdgcmd's scalar altitude: the field at altitude 0)
                        rsincos_deg<R>(hdg, sh, ch);
                        if constexpr (sizeof(R) == 8) return pymod(atan2(a.tas * sh + ve, a.tas * ch + vn) * kR2D, 360.0);
                        else return wrap360(atan2f(fmaf(a.tas, sh, ve), fmaf(a.tas, ch, vn)) * float(kR2D));
                    }
                }
                return hdg;
            };
            if (A.kind == BSG_KIND_HORIZONTAL) {                           // horizontal_cr_env.py:272-275
                a.aptrk = hdg_cmd(R(ev_hdg) + act0 * R(45));
            } else if (A.kind == BSG_KIND_VERTICAL) {                      // vertical_cr_env.py:242-256
                R v = act0 * R(12.5);
                a.selalt = v >= R(0) ? R(1000000) : R(0);
                a.selvs = v;
            } else if (A.kind == BSG_KIND_SECTOR) {                        // sector_cr_env.py:315-322
                R dh = act0 * R(22.5), dv = act1 * R(20.0 / 3);
                a.aptrk = hdg_cmd(bound180<R>(a.hdg + dh));
                R spd_kts = (a.cas + dv) * R(1.94384);
                a.selspd = fabs(spd_kts) > R(1) ? spd_kts * R(kKts) : spd_kts;
            } else if (EXT && A.kind == BSG_KIND_MERGE) {                  // merge_env.py:285-293 (HDG: LNAV off)
                R dh = act0 * R(15), dv = act1 * R(20);
                a.aptrk = hdg_cmd(bound180<R>(a.hdg + dh));
                a.lnav = -1;
                R spd_kts = (a.cas + dv) * R(1.94384);
                a.selspd = fabs(spd_kts) > R(1) ? spd_kts * R(kKts) : spd_kts;
            }
        }
    }
    if constexpr (RESO && SYM) {
        if (valid) {                                       // conflict rows of the pair scratch start empty
            unsigned *cr0 = reinterpret_cast<unsigned *>(smem_raw + SL.px_off + (size_t)le * pair_scratch_bytes(N) + (size_t)N * slot_cap(N) * 12);
            cr0[2 * i] = 0u; cr0[2 * i + 1] = 0u;
        }
    }
    __syncthreads();
    BSG_STAMP(1);
    const double lat0 = valid ? es[le].own_lat : 0.0, lon0 = valid ? es[le].own_lon : 0.0;
    const R alt_ref = valid ? es[le].alt_ref : R(0);
    const CdParams PL = A.P; const CdParamsF PF = A.PF;
    ExtCtx X;
    if constexpr (EXT) {
        const double *sv = es[valid ? le : 0].ev;
        X.W = &A.W; X.merge = A.kind == BSG_KIND_MERGE;
        X.faf_lat = sv[BSG_EV_WPT_LAT]; X.faf_lon = sv[BSG_EV_WPT_LON]; X.rwy_lat = sv[BSG_EV_WPT_X]; X.rwy_lon = sv[BSG_EV_WPT_Y];
    }


    RowAcc<R> acc;
    // record rows of this env: buffer b at row0 + b * bstride (hoisted: 64-bit address arithmetic out of the loops)
    unsigned char *const row0 = recs + (size_t)le * rowb;
    const size_t bstride = (size_t)epb * rowb;
    const int NR = sym_row_recs(N);
    R dve = R(0), dvn = R(0), dvz = R(0), tsolmin = R(1e9);           // MVP sums; vertical term / timesolveV (reso_mode 1, 2)
    bool touched = false, active = false;
    unsigned long long keep = 0;
    auto kin = [&](int s) {
        const bool last = (s == A.substeps - 1);
        if constexpr (sizeof(R) == 8) kin_substep_lit<EXT>(a, A.dt, reso, last, &X);
        else kin_substep_fast<EXT>(a, A.dt, reso, last && i == 0 && (A.kind == BSG_KIND_SECTOR || A.kind == BSG_KIND_MERGE), &X);
    };
    // ---- the symmetric fp32 sweep in three pieces (SYM kernels only) -----------------------------------------
    // my record of this substep into the env's row: two float4 halves + the ring copy behind slot nl
    auto put_rec = [&](unsigned char *row) {
        if constexpr (SYM) {
            RecFast r;
            make_rec<R>(a, lat0, lon0, r);
            float4 *wA = reinterpret_cast<float4 *>(row), *wB = wA + NR;
            const float4 ha = make_float4(r.x, r.y, r.alt, r.u), hb = make_float4(r.v, r.vs, r.ca, r.sa);
            wA[i] = ha; wB[i] = hb;
            if (i < (nl >> 1)) { wA[nl + i] = ha; wB[nl + i] = hb; }
            if (RESO) reinterpret_cast<float *>(wB + NR)[i] = a.hdg;          // resumenav's bouncing test
            // monotone within the launch: once any aircraft climbs, descends or leaves the common level
            if (a.vs != 0.0f || !(fabsf(a.alt - alt_ref) < 0.5f * PF.hpz)) es[le].vneed = 1;
        }
    };
    // phase 1: thread i sweeps ITS half ring j = i+1 .. i+H of the row and returns the relative conflict mask F
    // (bit d-1 <-> partner i+d).  No atomics, no branch: the sign bit of a margin is the flag, one funnel shift per
    // pair pushes it into G (pair d ends at bit H-d), reversed into F after the loop.  BSG_PW pairs per trip.
    auto sweep_row = [&](const unsigned char *row) -> unsigned {
        unsigned F = 0u;
        if constexpr (SYM) {
            const float4 *pA = reinterpret_cast<const float4 *>(row) + i, *pB = pA + NR;
            const RecFast me = rec_from(pA[0], pB[0]);
            const int H = nl >> 1;
            auto sweep = [&](auto vt) {
                constexpr bool V = decltype(vt)::value;
                int d = 1;
                unsigned G = 0u;
                for (; d + (BSG_PW - 1) <= H; d += BSG_PW) {
#pragma unroll
                    for (int q = 0; q < BSG_PW; q++)
                        G = __funnelshift_l(__float_as_uint(pair_fast_sign<V>(me, rec_from(pA[d + q], pB[d + q]), PF)), G, 1);
                }
#pragma unroll 1
                for (; d <= H; d++)
                    G = __funnelshift_l(__float_as_uint(pair_fast_sign<V>(me, rec_from(pA[d], pB[d]), PF)), G, 1);
                F = H > 0 ? __brev(G) >> (32 - H) : 0u;
            };
            // level, co-altitude traffic (Horizontal and Sector envs): the vertical window is (-inf, inf)
            if (es[le].vneed) sweep(std::true_type{}); else sweep(std::false_type{});
        }
        return F;
    };
    // the masks are published once per detection as hr[i] (and the ring copy hr[i + nl], so that the partners behind
    // a thread are read without an index wrap); phase 2 completes the row with bit k-1 of the masks of the nl-1-H
    // partners i-k.  One buffer: the next barrier separates the readers of this detection from the next writers.
    unsigned *const hr = reinterpret_cast<unsigned *>(crow) + (size_t)le * 2 * N;
    // pair scratch of this env (resolution kernels): res | tcp | rows | items
    unsigned char *const px = smem_raw + SL.px_off + (size_t)le * pair_scratch_bytes(N);
    const int CAP = pair_cap(N), SC = slot_cap(N);
    float2 *const px_res = reinterpret_cast<float2 *>(px);                       // [N][SC]
    float *const px_tcp = reinterpret_cast<float *>(px_res + (size_t)N * SC);    // [N][SC]
    unsigned *const px_cr = reinterpret_cast<unsigned *>(px_tcp + (size_t)N * SC);   // [N][2] conflict rows
    unsigned *const px_keep = px_cr + 2 * N;                                     // [N][2] resumenav keep rows
    unsigned *const px_los = px_keep + 2 * N;                                    // [N][2] LoS rows
    unsigned short *const px_items = reinterpret_cast<unsigned short *>(px_los + 2 * N);   // [CAP]
    // phase 2: my row, conflicting / resolving pairs only.  It feeds MVP / resumenav every substep; the CD row
    // outputs report the LAST detection of the step (what bs.traf.cd holds afterwards), so the row reductions run on
    // the last substep only (want_acc).
    auto phase2 = [&](const unsigned char *row, unsigned F, bool want_acc, bool listed) {
        if constexpr (SYM) {
            const float4 *rA = reinterpret_cast<const float4 *>(row), *rB = rA + NR;
            const RecFast me = rec_from(rA[i], rB[i]);
            // A symmetric relation over the ring (conflict flags, resumenav's keep flags) is stored as one word of forward
            // bits per thread (bit d-1 <-> partner i+d), each word twice (w[i] and the ring copy w[i + nl], so that the
            // partners behind a thread are read without an index wrap).  My full row = my forward word + bit k-1 of the
            // words of the nl-1-H partners i-k.  Groups of 8 partners, fully unrolled (LDS with an immediate offset + one
            // LOP3 with an immediate mask per partner); a group may read up to 7 words past partner kmax -- other rows'
            // words, at worst the 31 words below the array, all inside the CTA's shared memory -- and the final mask
            // drops them.
            auto ring_row = [&](const unsigned *w, unsigned Fw) -> unsigned long long {
                unsigned B = 0u;
                const int kmax = nl - 1 - (nl >> 1);                 // <= 31
                const unsigned *hb = w + i + nl;
#pragma unroll
                for (int g8 = 0; g8 < 32; g8 += 8) {
                    if (g8 < kmax) {
#pragma unroll
                        for (int q = 1; q <= 8; q++) B |= hb[-(g8 + q)] & (1u << (g8 + q - 1));
                    }
                }
                B &= (1u << kmax) - 1u;
                // relative -> absolute positions.  Forward: i+d, wrapped ones (i+d >= nl) at i+d-nl.
                unsigned long long fwd = (i < 63 ? (unsigned long long)Fw << (i + 1) : 0ull) & (nl < 64 ? (1ull << nl) - 1ull : ~0ull);
                if (nl - 1 - i < 32) fwd |= (unsigned long long)(Fw >> (nl - 1 - i));
                // Backward: the bit-reversed word, shifted, lands on i-k (and i-k+nl for the wrapped ones)
                const unsigned long long rb64 = (unsigned long long)__brev(B) << 32;     // bit 64-k <-> partner i-k
                unsigned long long back = i > 0 ? rb64 >> (64 - i) : 0ull;               // positions i-k >= 0
                if (i < 32) back |= (rb64 << i) >> (64 - nl);                            // k > i: positions i-k+nl
                return fwd | back;
            };
            // detection only: my conflict row from the ring words; with resolution: the row the sweep built with atomics
            unsigned long long cr;
            if constexpr (RESO) cr = (unsigned long long)px_cr[2 * i] | ((unsigned long long)px_cr[2 * i + 1] << 32);
            else cr = ring_row(hr, F);
            const unsigned long long notme = ~(1ull << i);
            if constexpr (!RESO) {
                // detection only: the row reductions of the step's last detection over my conflicting pairs
                unsigned long long S = cr & notme;
                while (S) {
                    const int j = __ffsll((long long)S) - 1;
                    S &= S - 1;
                    PairCore pc;
                    pair_fast_core(me, rec_from(rA[j], rB[j]), PF, true, pc);
                    acc.add_conflict(j, pc.tcpa, pc.dcpa2, pc.tin, (pc.dist2 < PF.R2) && (fabsf(pc.dalt) < PF.hpz));
                }
            } else if (listed) {
                // with resolution: every pair of my row was evaluated once by the work list (below): resumenav's keep
                // row came back through atomics, my MVP displacements sit in my slot row in ascending intruder index
                keep = (unsigned long long)px_keep[2 * i] | ((unsigned long long)px_keep[2 * i + 1] << 32);
                touched = ((cr | resorow) & notme) != 0ull;
                active = keep != 0ull;
                const int nc = __popcll(cr);
                const float2 *mr = px_res + i * SC;
                if (!want_acc) {
                    for (int q = 0; q < nc; q++) { const float2 m = mr[q]; dve = __fsub_rn(dve, m.x); dvn = __fsub_rn(dvn, m.y); }
                } else {
                    const float *tr = px_tcp + i * SC;
                    unsigned long long S = cr;
                    for (int q = 0; q < nc; q++) {
                        const float2 m = mr[q];
                        dve = __fsub_rn(dve, m.x); dvn = __fsub_rn(dvn, m.y);
                        const int j = __ffsll((long long)S) - 1;
                        S &= S - 1;
                        acc.add_conflict(j, tr[q], 0.0f, 0.0f, false);
                    }
                    acc.nlos = __popc(px_los[2 * i]) + __popc(px_los[2 * i + 1]);
                    if (acc.near_idx >= 0) {                                // dCPA / tLOS of the nearest-in-time conflict
                        PairCore pc;
                        pair_fast_core(me, rec_from(rA[acc.near_idx], rB[acc.near_idx]), PF, true, pc);
                        acc.near_dcpa2 = pc.dcpa2; acc.near_tin = pc.tin;
                    }
                }
                if (A.reso_off_own && i == 0) { dve = 0.0f; dvn = 0.0f; }   // RESOOFF: the agent does not resolve
            } else {
                // list overflow (an env denser than BSG_PAIR_CAP allows): every owner evaluates its own row -- two walks,
                // conflicting pairs first (row reductions, MVP, resumenav), then the pairs that are only still resolving
                const float *hdgrow = reinterpret_cast<const float *>(rB + NR);
                const float R2f = PF.rpz * PF.rpz, R2b = R2f * PF.fach2;
                auto resume = [&](int j, unsigned long long jb, float dist2, float dot) {
                    bool kp = !(dot > 0.0f) || dist2 < R2f;
                    if (!kp && dist2 < R2b) kp = fabsf(a.hdg - hdgrow[j]) < 30.0f;      // bouncing (rare)
                    touched = true;
                    if (kp) { active = true; keep |= jb; }
                };
                unsigned long long S = cr & notme;
                while (S) {
                    const int j = __ffsll((long long)S) - 1;
                    const unsigned long long jb = S & (0ull - S);
                    S ^= jb;
                    PairCore pc;
                    pair_fast_core(me, rec_from(rA[j], rB[j]), PF, true, pc);
                    if (want_acc)
                        acc.add_conflict(j, pc.tcpa, pc.dcpa2, pc.tin, (pc.dist2 < PF.R2) && (fabsf(pc.dalt) < PF.hpz));
                    float d1, d2;
                    mvp_pair_fast(pc.dx, pc.dy, pc.dist2, pc.tcpa, pc.du, pc.dv, PF.rpz_m, d1, d2);
                    dve = __fsub_rn(dve, d1); dvn = __fsub_rn(dvn, d2);
                    resume(j, jb, pc.dist2, pc.dot);
                }
                S = resorow & ~cr & notme;
                while (S) {
                    const int j = __ffsll((long long)S) - 1;
                    const unsigned long long jb = S & (0ull - S);
                    S ^= jb;
                    float dist2, dot;
                    pair_fast_geom(me, rec_from(rA[j], rB[j]), dist2, dot);
                    resume(j, jb, dist2, dot);
                }
                if (A.reso_off_own && i == 0) { dve = 0.0f; dvn = 0.0f; }
            }
        }
    };
    // Work list (resolution kernels).  Thread i owns the unordered pairs of its forward half ring; those that are in
    // conflict (F) or still resolving (my resopairs row: the matrix is symmetric -- conflicts are, and resumenav's
    // decision is -- so my forward bits cover the pair for both sides) become items of the env's list.
    auto list_build = [&](unsigned F) {
        if constexpr (RESO && SYM) {
            const int H = nl >> 1;
            const unsigned long long rot = (i < 63 ? resorow >> (i + 1) : 0ull) | (resorow << (nl - 1 - i));
            const unsigned hm = H >= 32 ? ~0u : (1u << H) - 1u;
            unsigned Wc = F & hm, Wr = (unsigned)rot & hm & ~F;
            // my conflict row: my forward bits at their absolute positions (i+d, wrapped ones at i+d-nl) ...
            unsigned long long fwd = (i < 63 ? (unsigned long long)Wc << (i + 1) : 0ull) & (nl < 64 ? (1ull << nl) - 1ull : ~0ull);
            if (nl - 1 - i < 32) fwd |= (unsigned long long)(Wc >> (nl - 1 - i));
            if ((unsigned)fwd) atomicOr(&px_cr[2 * i], (unsigned)fwd);
            if ((unsigned)(fwd >> 32)) atomicOr(&px_cr[2 * i + 1], (unsigned)(fwd >> 32));
            // the opposite pair of an even ring is a forward pair of both aircraft: the lower index lists it
            if (H > 0 && 2 * H == nl && i >= H) { Wc &= ~(1u << (H - 1)); Wr &= ~(1u << (H - 1)); }
            const int nc = __popc(Wc), nr = __popc(Wr);
            int bc = nc ? atomicAdd(&es[le].nC, nc) : 0, br = nr ? atomicAdd(&es[le].nR, nr) : 0;
            const bool fits = bc + nc + br + nr <= CAP;     // (an overflowing env drops the list for this detection)
            const unsigned mybit = 1u << (i & 31);
            const int myword = i >> 5;
            // ... and my bit in the rows of my forward partners
            while (Wc) {
                const int b = __ffs(Wc) - 1; Wc &= Wc - 1;
                int jw = i + b + 1; jw -= jw >= nl ? nl : 0;
                atomicOr(&px_cr[2 * jw + myword], mybit);
                if (fits) px_items[bc++] = (unsigned short)((i << 5) | b);
            }
            if (fits)
                while (Wr) { const int b = __ffs(Wr) - 1; Wr &= Wr - 1; px_items[CAP - 1 - br++] = (unsigned short)((i << 5) | b); }
        }
    };
    // every thread of the env takes items t = i, i + N, ...: a conflicting pair gets the full canonical pair, MVP and
    // resumenav's decision; a pair that is only still resolving gets the decision alone
    auto list_run = [&](const unsigned char *row, bool want_acc) {
        if constexpr (RESO && SYM) {
            const float4 *rA = reinterpret_cast<const float4 *>(row), *rB = rA + NR;
            const float *hdgrow = reinterpret_cast<const float *>(rB + NR);
            const int nC = es[le].nC, nR = es[le].nR;
            const float R2f = PF.rpz * PF.rpz, R2b = R2f * PF.fach2;
            // resumenav's decision for the pair (symmetric): a kept pair sets its bit in the keep rows of both aircraft
            auto keep_pair = [&](int ia, int jw, float dist2, float dot) {
                bool kp = !(dot > 0.0f) || dist2 < R2f;                     // not past CPA, or in horizontal LoS
                if (!kp && dist2 < R2b) kp = fabsf(hdgrow[ia] - hdgrow[jw]) < 30.0f;       // bouncing (rare)
                if (kp) {
                    atomicOr(&px_keep[2 * ia + (jw >> 5)], 1u << (jw & 31));
                    atomicOr(&px_keep[2 * jw + (ia >> 5)], 1u << (ia & 31));
                }
            };
            for (int t = i; t < nC; t += N) {
                const unsigned it = px_items[t];
                const int ia = it >> 5, ib = ia + (it & 31u) + 1, jw = ib >= nl ? ib - nl : ib;
                PairCore pc;
                pair_fast_core(rec_from(rA[ia], rB[ia]), rec_from(rA[ib], rB[ib]), PF, true, pc);
                float d1, d2;
                mvp_pair_fast(pc.dx, pc.dy, pc.dist2, pc.tcpa, pc.du, pc.dv, PF.rpz_m, d1, d2);
                // slot of the pair in either aircraft's row: the rank of the other among its conflicts
                const unsigned long long ca = (unsigned long long)px_cr[2 * ia] | ((unsigned long long)px_cr[2 * ia + 1] << 32);
                const unsigned long long cb = (unsigned long long)px_cr[2 * jw] | ((unsigned long long)px_cr[2 * jw + 1] << 32);
                const int qa = __popcll(ca & ((1ull << jw) - 1ull)), qb = __popcll(cb & ((1ull << ia) - 1ull));
                if (qa < SC && qb < SC) {
                    px_res[ia * SC + qa] = make_float2(d1, d2);
                    px_res[jw * SC + qb] = make_float2(-d1, -d2);
                    if (want_acc) {
                        px_tcp[ia * SC + qa] = pc.tcpa; px_tcp[jw * SC + qb] = pc.tcpa;
                        if ((pc.dist2 < PF.R2) && (fabsf(pc.dalt) < PF.hpz)) {
                            atomicOr(&px_los[2 * ia + (jw >> 5)], 1u << (jw & 31));
                            atomicOr(&px_los[2 * jw + (ia >> 5)], 1u << (ia & 31));
                        }
                    }
                } else es[le].ovf = 1;
                keep_pair(ia, jw, pc.dist2, pc.dot);
            }
            for (int t = i; t < nR; t += N) {
                const unsigned it = px_items[CAP - 1 - t];
                const int ia = it >> 5, ib = ia + (it & 31u) + 1, jw = ib >= nl ? ib - nl : ib;
                float dist2, dot;
                pair_fast_geom(rec_from(rA[ia], rB[ia]), rec_from(rA[ib], rB[ib]), dist2, dot);
                keep_pair(ia, jw, dist2, dot);
            }
        }
    };

    if constexpr (SYM && !RESO) {
        // ------------------------------------------------------------ K substeps, detection only
        // Without a resolution nothing the detection finds feeds back into the kinematics, so the two are split:
        // a chunk of up to nbuf substeps is integrated back to back (one latency-bound chain per thread, no barrier),
        // every substep leaving its CD record in its own row; after ONE barrier the sweeps of the whole chunk run
        // back to back (independent, FP32-pipe-bound work).  The flags of the step's last detection become the CD row
        // outputs; the earlier detections (BlueSky runs CD every asas_dt) are folded into the env's any-conflict word.
        const int K = A.substeps, nbuf = A.nbuf;
        unsigned Flast = 0u, Fany = 0u;
        int lastrow = 0;
        for (int s0 = 0; s0 < K; s0 += nbuf) {
            const int s1 = s0 + nbuf < K ? s0 + nbuf : K;
            if (A.asas_on && s0 > 0) __syncthreads();                  // the previous chunk's sweeps have read the rows
            if (live && !skip) {
                for (int s = s0; s < s1; s++) {
                    if (A.asas_on) put_rec(row0 + (size_t)(s - s0) * bstride);
                    kin(s);
                }
            }
            if (A.asas_on) {
                __syncthreads();
                if (s0 == 0) BSG_STAMP(5);
                if (live && !skip) {
                    for (int s = s0; s < s1; s++) {
                        const unsigned F = sweep_row(row0 + (size_t)(s - s0) * bstride);
                        if (s == K - 1) Flast = F; else Fany |= F;
                    }
                }
                if (s0 == 0) BSG_STAMP(6);
                lastrow = K - 1 - s0;
            }
        }
        if (A.asas_on && K > 0) {
            if (live && !skip) {
                hr[i] = Flast; hr[i + nl] = Flast;
                if (Fany | Flast) es[le].anyconf[0] = 1;
            }
            __syncthreads();
            if (live && !skip) phase2(row0 + (size_t)lastrow * bstride, Flast, true, false);
        }
    } else {
    // ------------------------------------------------------------------ K substeps
    for (int s = 0; s < A.substeps; s++) {
        const int buf = s & 1;
        if (A.asas_on) {
            unsigned char *const row = row0 + (buf ? bstride : 0);
            Rec *rb = reinterpret_cast<Rec *>(row);
            if (live && !skip) {
                if constexpr (SYM) put_rec(row);
                else {
                    make_rec<R>(a, lat0, lon0, rb[i]);
                    if (reso) sdist[le * N + i] = a.hdg;
                }
            }
            __syncthreads();
            if (RESO && valid && i == 0) { es[le].anyconf[buf ^ 1] = 0; es[le].ovf = 0; }
            acc = RowAcc<R>();
            dve = R(0); dvn = R(0); dvz = R(0); tsolmin = R(1e9);
            touched = false; active = false;
            keep = 0;
            if constexpr (SYM) {
                unsigned F = 0u;
                const bool want_acc = (s == A.substeps - 1);
                const bool need_rows = RESO || want_acc;
                if (live && !skip) {
                    F = sweep_row(row);
                    if constexpr (RESO) {
                        if (F) es[le].anyconf[buf] = 1;
                        px_keep[2 * i] = 0u; px_keep[2 * i + 1] = 0u;
                        if (want_acc) { px_los[2 * i] = 0u; px_los[2 * i + 1] = 0u; }
                        list_build(F);
                    } else { hr[i] = F; hr[i + nl] = F; }
                }
                if (need_rows) __syncthreads();
                bool listed = false;
                if constexpr (RESO) {
                    const bool fits = valid && es[le].nC + es[le].nR <= CAP && !A.force_walk;
                    if (fits && !skip) list_run(row, want_acc);
                    __syncthreads();
                    listed = fits && !es[le].ovf;
                }
                if (need_rows && live && !skip) phase2(row, F, want_acc, listed);
                if constexpr (RESO) {
                    // next detection: my conflict row cleared (its readers are past the barrier above), the list counters
                    // by the agent thread (their readers: `fits` above, before the barrier)
                    if (live) { px_cr[2 * i] = 0u; px_cr[2 * i + 1] = 0u; }
                    if (valid && i == 0) { es[le].nC = 0; es[le].nR = 0; }
                }
            } else {

            if (live && !skip) {
                const Rec me = rb[i];
                // resopairs rows: one 64-bit word per 64 partners.  N <= 64: the word lives in a register for the whole
                // launch; beyond, word w of aircraft k sits at reso[w * E * N + k] and is read / rewritten chunk by chunk
                // (only the owner thread touches its row)
                const bool multi = N > 64;
                for (int j0 = 0; j0 < nl; j0 += 64) {
                const int j1 = j0 + 64 < nl ? j0 + 64 : nl;
                unsigned long long rw = resorow, kw = 0ull;
                if (RESO && multi) rw = A.reso[(size_t)(j0 >> 6) * EN + k];
#pragma unroll 4
                for (int j = j0; j < j1; j++) {
                    const Rec o = rb[j];
                    PairOut<R> po;
                    const bool off = (j != i);
                    if constexpr (sizeof(R) == 8) {
                        pair_lit(me, o, PL, po);
                        po.sw = po.sw && off; po.los = po.los && off;
                        acc.add(j, po);
                    } else {
                        PairCore pc;
                        po.sw = pair_fast_core(me, o, PF, off, pc);
                        po.los = false;
                        po.tcpa = pc.tcpa; po.dcpa2 = pc.dcpa2; po.tin = pc.tin; po.dx = pc.dx; po.dy = pc.dy; po.dist = pc.dist2; po.dalt = pc.dalt;
                        if (__any_sync(__activemask(), po.sw)) {   // conflicting pairs only: row reductions off the hot path
                            if (po.sw) acc.add_conflict(j, pc.tcpa, pc.dcpa2, pc.tin, (pc.dist2 < PF.R2) && (fabsf(pc.dalt) < PF.hpz));
                        }
                    }
                    if constexpr (RESO) {
                        const bool inres = off && (po.sw || ((rw >> (j - j0)) & 1ull));
                        if (inres) {
                            R vrx = o.u - me.u, vry = o.v - me.v;
                            if (po.sw) {
                                R dist, rpz_m, d1, d2;
                                if constexpr (sizeof(R) == 8) { dist = po.dist; rpz_m = PL.rpz * PL.fach; }
                                else { dist = sqrtf(po.dist); rpz_m = PF.rpz_m; }
                                mvp_pair<R>(po.dx, po.dy, dist, po.tcpa, vrx, vry, rpz_m, d1, d2);
                                dve = dve - d1; dvn = dvn - d2;
                                if (A.reso_mode != 0) {                    // vertical term: dv3 halved (cooperative), timesolveV
                                    R d3, tsv;
                                    const R hpz_m = sizeof(R) == 8 ? R(PL.hpz * PL.facv) : R(PF.hpz_m), dtl = sizeof(R) == 8 ? R(PL.dtlook) : R(PF.dtlook);
                                    mvp_vert<R>(po.dalt, o.vs - me.vs, po.tin, hpz_m, dtl, d3, tsv);
                                    tsolmin = tsv < tsolmin ? tsv : tsolmin;
                                    dvz = dvz - R(0.5) * d3;
                                }
                                if (A.reso_off_own && i == 0) { dve = R(0); dvn = R(0); dvz = R(0); }
                            }
                            R ddx, ddy, rpz, fach;
                            if constexpr (sizeof(R) == 8) {
                                ddx = kRe * ((o.lon - me.lon) * kD2R * cos(0.5 * ((o.lat + me.lat) * kD2R)));
                                ddy = kRe * ((o.lat - me.lat) * kD2R);
                                rpz = PL.rpz; fach = PL.fach;
                            } else { ddx = po.dx; ddy = po.dy; rpz = PF.rpz; fach = (float)A.P.fach; }
                            const bool past = (ddx * vrx + ddy * vry) > R(0);
                            const R hdist = sqrt(ddx * ddx + ddy * ddy);
                            const bool bouncing = fabs(a.hdg - sdist[le * N + j]) < R(30) && hdist < rpz * fach;
                            touched = true;
                            if (!past || hdist < rpz || bouncing) { active = true; kw |= 1ull << (j - j0); }
                        }
                    }
                }
                if (RESO && multi) A.reso[(size_t)(j0 >> 6) * EN + k] = kw; else keep = kw;
                }
            }
            }
            if constexpr (RESO) {
                if constexpr (!SYM) {
                    if (acc.nconf > 0) atomicOr(&es[le].anyconf[buf], 1);
                    __syncthreads();
                }
                if (live && !skip) {
                    if (es[le].anyconf[buf]) {                             // MVP.resolve ran for this env
                        R ne = a.gse + dve, nn = a.gsn + dvn;
                        if (A.reso_mode == 1) { ne = a.gse; nn = a.gsn; }      // vertical only: ownship.trk / ownship.gs kept
                        double vmin, vmax;
                        phase_vlimits(a.phase, vmin, vmax);
                        bool wind_on = false;
                        if constexpr (EXT) wind_on = A.W.n > 0;
                        if (A.reso_mode == 1 && !wind_on) a.asas_trk = a.hdg;
                        else if constexpr (sizeof(R) == 8) a.asas_trk = pymod(atan2(ne, nn) * 180 / kPi, 360);
                        else { float t = atan2f(ne, nn) * float(kR2D); a.asas_trk = t < 0.0f ? t + 360.0f : t; }
                        R newgs = (A.reso_mode == 1 && !wind_on) ? a.tas : sqrt(ne * ne + nn * nn);
                        a.asas_tas = fmax(R(vmin), fmin(R(vmax), newgs));
                        const R vscap = fmax(R(kVsMin), fmin(R(kVsMax), A.reso_mode == 0 ? a.vs : a.vs + dvz));
                        a.asas_vs = vscap;
                        if (A.reso_mode == 0) a.asas_alt = a.selalt;
                        else {
                            // resolution altitude (tail of MVP.resolve): the autopilot's level-off altitude when it also
                            // solves the conflict, else the altitude reached after timesolveV at the capped vertical speed
                            auto sgn = [](R x) -> R { return x > R(0) ? R(1) : (x < R(0) ? R(-1) : R(0)); };
                            const R apvs = fabs(a.selvs) > R(0.1) ? a.selvs : R(kVsDef);
                            const R tmp = vscap * tsolmin + a.alt;
                            const R sdv = sgn(vscap - apvs * sgn(a.selalt - a.alt)), sal = sgn(tmp - a.selalt);
                            R ra = (sdv == R(0) || sdv == sal) ? tmp : a.selalt;
                            const R dtl = sizeof(R) == 8 ? R(PL.dtlook) : R(PF.dtlook);
                            if (tsolmin < dtl && fabs(dvz) > R(0)) ra = tmp;
                            a.asas_alt = ra;
                        }
                    }
                    if (touched) a.asas_active = active ? 1 : 0;
                    resorow = keep;
                }
            }
        }

        if (live && !skip) kin(s);
    }
    }
    BSG_STAMP(2);
    if (live && !skip) {
        ac_store<R, RESO, EXT>(a, A.pos, A.st, EN, k);
        if (reso && N <= 64) A.reso[k] = resorow;          // (N > 64: the row's words were rewritten in place, chunk by chunk)
    }
    if (valid && i < N) {      // CD row outputs of the last detection (dead slots: zeros / -1)
        const bsg_step_out &o = A.out;
        if (o.cd_inconf) o.cd_inconf[k] = acc.nconf > 0;
        if (o.cd_nconf) o.cd_nconf[k] = acc.nconf;
        if (o.cd_nlos) o.cd_nlos[k] = acc.nlos;
        if (o.cd_near_idx) o.cd_near_idx[k] = acc.near_idx;
        if (o.cd_near_tcpa) ((R *)o.cd_near_tcpa)[k] = acc.tcpa_out();
        if (o.cd_near_dcpa) ((R *)o.cd_near_dcpa)[k] = sqrt(acc.near_dcpa2);
        if (o.cd_near_tin) ((R *)o.cd_near_tin)[k] = acc.near_tin;
        if (o.cd_tcpamax) ((R *)o.cd_tcpamax)[k] = acc.tcpamax;
    }
    BSG_STAMP(3);
    env_layer<R, EXT>(A, a, es, sdist, le, i, e, nl, valid, live, skip, ev);
    BSG_STAMP(4);
}

}  // namespace bsg
