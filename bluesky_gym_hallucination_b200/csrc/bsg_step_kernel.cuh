// bsg_step_kernel.cuh -- body of the fused env-step kernel (see bsg_step.cuh for the mapping).
#pragma once
#include <type_traits>
#include "bsg_step.cuh"
#include "bsg_env.cuh"
#include "bsg_reset.cuh"

#ifndef BSG_PW
#define BSG_PW 5          // scalar half-ring loop (generic vertical window): pairs per trip (A/B on B200: 5 > 4 > 2)
#endif
#ifndef BSG_PG
#define BSG_PG 3          // packed half-ring loop: f32x2 trips (two partners each) interleaved per iteration
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

template <typename R>
__device__ __forceinline__ void cd_params_for(const CdParams &P, CdParams &lit, CdParamsF &fast) {
    lit = P;
    fast.R2 = (float)(P.rpz * P.rpz); fast.hpz = (float)P.hpz; fast.dtlook = (float)P.dtlook;
    fast.rpz = (float)P.rpz; fast.rpz_m = (float)(P.rpz * P.fach); fast.hpz_m = (float)(P.hpz * P.facv);
}

// CTA size / residency: 128-thread CTAs keep the small-N configs in one balanced wave
// (C2: 683 CTAs over 148 SMs x 5 resident), 256 threads only when one env needs them (N > 128).
template <typename R, int TPB> struct StepBounds {
    static constexpr int kMinBlocks = TPB == 128 ? (sizeof(R) == 4 ? 5 : 3) : (sizeof(R) == 4 ? 2 : 1);
};
// Per-env row of CD records in shared memory (double-buffered by substep parity).
// fp64 / fp32 with N > 64: N records + 16 B, so that the rows of the envs sharing a warp start in different banks.
// fp32 symmetric sweep (N <= 64): structure of arrays, one float array per record field, so that ONE LDS.64 hands the
// packed f32x2 pair form the same field of two neighbouring partners {f[j], f[j+1]}.  LDS.64 wants j even, and thread i
// sweeps j = i+1, i+3, ... : every field is therefore kept TWICE -- copy 0 holds f[k] at index k, copy 1 holds
// f[k+1] at index k -- odd i read copy 0 at i+1, even i read copy 1 at i (both even indices).  Each copy is the ring
// unrolled: the first N/2 + 1 records are repeated behind the last live slot, so j = i + d never wraps.  Copy 1
// starts 18 banks off copy 0 (the even- and the odd-lane halves of a warp's LDS.64 then hit disjoint banks).  Field
// 8 is the heading (resumenav's bouncing test), single copy, in the same double-buffered row.
constexpr int kSoaFields = 8;                    // x, y, u, v, ca, sa, alt, vs
enum { SF_X, SF_Y, SF_U, SF_V, SF_CA, SF_SA, SF_ALT, SF_VS };
__host__ __device__ constexpr int soa_len(int N) { return (N + N / 2 + 3) & ~1; }             // floats per copy
__host__ __device__ constexpr int soa_copy1(int N) { int c = soa_len(N); while ((c & 31) != 18) c += 2; return c; }
__host__ __device__ constexpr int soa_fstride(int N) { return soa_copy1(N) + soa_len(N); }   // floats per field (even)
__host__ __device__ constexpr size_t soa_row_bytes(int N) {
    return (((size_t)kSoaFields * soa_fstride(N) + N + 1) & ~(size_t)1) * sizeof(float) + 16;
}
template <typename R> __host__ __device__ constexpr size_t rec_row_bytes(int N) {
    const size_t aos = (size_t)N * sizeof(typename RecOf<R>::type) + 16;
    if (sizeof(R) == 4 && N <= 64) return soa_row_bytes(N) > aos ? soa_row_bytes(N) : aos;
    return aos;
}

// shared-memory carve-up of one CTA (host and device agree through this one function)
struct StepSmem { size_t es_off, sd_off, cr_off, total; };
template <typename R> __host__ __device__ inline StepSmem step_smem(int N, int epb) {
    StepSmem L;
    L.es_off = 2 * (size_t)epb * rec_row_bytes<R>(N);                      // [2][epb] record rows
    L.sd_off = L.es_off + (size_t)epb * sizeof(EnvShared<R>);              // [epb] env scratch
    L.cr_off = (L.sd_off + (size_t)epb * N * sizeof(R) + 7) & ~(size_t)7;  // [epb*N] hdg / sector distances
    L.total = L.cr_off + 2 * (size_t)epb * N * sizeof(unsigned long long); // [2][epb*N] conflict bit rows (SYM)
    return L;
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
    const StepSmem SL = step_smem<R>(N, epb);
    unsigned char *recs = smem_raw;                                        // [2][epb] rows of rowb bytes
    EnvShared<R> *es = reinterpret_cast<EnvShared<R> *>(smem_raw + SL.es_off);   // [epb]
    R *sdist = reinterpret_cast<R *>(smem_raw + SL.sd_off);                // [epb*N] hdg (resumenav) / sector distances
    unsigned long long *crow = reinterpret_cast<unsigned long long *>(smem_raw + SL.cr_off);   // [2][epb*N]

    BSG_STAMP(0);
    const int tid = threadIdx.x;
    const int le = tid / N, i = tid - le * N;
    const int e = blockIdx.x * epb + le;
    const bool valid = le < epb && e < A.E;
    const bool skip = valid && A.skip && A.skip[e];                       // fresh episode: no action, no integration
    if (A.fuse_reset && __syncthreads_or(skip))                            // rare: some env of this CTA starts an episode
        fused_reset<R>(A.kind, A.fixed_nac, A.density_mode, A.E, A.N, A.epb, A.ar_seed, A.episode, A.ev, A.nac, A.pos, A.st, A.reso,
                       A.RC, smem_raw, skip, valid, le, i, e);
    const int nl = valid ? A.nac[e] : 0;
    const bool live = valid && i < nl;
    const size_t EN = (size_t)A.E * N, k = (size_t)(valid ? e : 0) * N + i;
    double *ev = A.ev + (size_t)(valid ? e : 0) * BSG_NE;
    constexpr bool reso = RESO;

    AcReg<R> a = {};
    unsigned long long resorow = 0;
    if (live) {
        ac_load_raw<R, RESO, EXT>(a, A.pos, A.st, EN, k);
        ac_derive<R, EXT>(a);
        if (reso) resorow = A.reso[k];
    }
    // ------------------------------------------------------------------ action intake (agent = aircraft 0)
    if (live && i == 0) {
        es[le].anyconf[0] = 0; es[le].anyconf[1] = 0; es[le].intr = 0; es[le].vneed = 0; es[le].alt_ref = a.alt;
        es[le].own_lat = a.lat; es[le].own_lon = a.lon;                    // local origin of the fp32 CD records
#pragma unroll
        for (int s2 = 0; s2 < BSG_EV_POLY; s2++) es[le].ev[s2] = ev[s2];   // 14 independent loads, staged in smem
        if (A.actions && !skip) {
            // AP.selhdgcmd under a wind field: the HDG argument becomes the track it gives (above 50 ft)
            auto hdg_cmd = [&](R hdg) -> R {
                if constexpr (EXT) {
                    if (A.W.n > 0 && a.alt > R(50.0 * kFt)) {
                        R vn, ve, sh, ch;
                        wind_at<R>(A.W, a.lat, a.lon, vn, ve);
                        rsincos_deg<R>(hdg, sh, ch);
                        if constexpr (sizeof(R) == 8) return pymod(atan2(a.tas * sh + ve, a.tas * ch + vn) * kR2D, 360.0);
                        else return wrap360(atan2f(fmaf(a.tas, sh, ve), fmaf(a.tas, ch, vn)) * float(kR2D));
                    }
                }
                return hdg;
            };
            if (A.kind == BSG_KIND_HORIZONTAL) {                           // horizontal_cr_env.py:272-275
                a.aptrk = hdg_cmd(R(es[le].ev[BSG_EV_AC_HDG]) + A.actions[e] * R(45));
            } else if (A.kind == BSG_KIND_VERTICAL) {                      // vertical_cr_env.py:242-256
                R v = A.actions[e] * R(12.5);
                a.selalt = v >= R(0) ? R(1000000) : R(0);
                a.selvs = v;
            } else if (A.kind == BSG_KIND_SECTOR) {                        // sector_cr_env.py:315-322
                R dh = A.actions[2 * e] * R(22.5), dv = A.actions[2 * e + 1] * R(20.0 / 3);
                a.aptrk = hdg_cmd(bound180<R>(a.hdg + dh));
                R spd_kts = (a.cas + dv) * R(1.94384);
                a.selspd = fabs(spd_kts) > R(1) ? spd_kts * R(kKts) : spd_kts;
            } else if (EXT && A.kind == BSG_KIND_MERGE) {                  // merge_env.py:285-293 (HDG: LNAV off)
                R dh = A.actions[2 * e] * R(15), dv = A.actions[2 * e + 1] * R(20);
                a.aptrk = hdg_cmd(bound180<R>(a.hdg + dh));
                a.lnav = -1;
                R spd_kts = (a.cas + dv) * R(1.94384);
                a.selspd = fabs(spd_kts) > R(1) ? spd_kts * R(kKts) : spd_kts;
            }
        }
    }
    __syncthreads();
    BSG_STAMP(1);
    const double lat0 = valid ? es[le].own_lat : 0.0, lon0 = valid ? es[le].own_lon : 0.0;
    const R alt_ref = valid ? es[le].alt_ref : R(0);
    CdParams PL; CdParamsF PF;
    cd_params_for<R>(A.P, PL, PF);
    ExtCtx X;
    if constexpr (EXT) {
        const double *sv = es[valid ? le : 0].ev;
        X.W = &A.W; X.merge = A.kind == BSG_KIND_MERGE;
        X.faf_lat = sv[BSG_EV_WPT_LAT]; X.faf_lon = sv[BSG_EV_WPT_LON]; X.rwy_lat = sv[BSG_EV_WPT_X]; X.rwy_lon = sv[BSG_EV_WPT_Y];
    }

    RowAcc<R> acc;
    const int FS = soa_fstride(N), C1 = soa_copy1(N);                      // SYM: floats per field, offset of copy 1
    // record rows of this env in the two substep buffers (hoisted: 64-bit address arithmetic out of the loop)
    unsigned char *const rb_even = recs + (size_t)le * rowb, *const rb_odd = rb_even + (size_t)epb * rowb;
    // ------------------------------------------------------------------ K substeps
    for (int s = 0; s < A.substeps; s++) {
        const int buf = s & 1;
        if (A.asas_on) {
            Rec *rb = reinterpret_cast<Rec *>(buf ? rb_odd : rb_even);
            float *const srow = reinterpret_cast<float *>(buf ? rb_odd : rb_even);     // SYM: the SoA row of this env
            RecFast me;                                                                // SYM: my own record
            if (live && !skip) {
                if constexpr (SYM) {                  // SoA fields, two copies each (+ the ring unrolled behind slot nl)
                    make_rec<R>(a, lat0, lon0, me);
                    const float fv[kSoaFields] = { me.x, me.y, me.u, me.v, me.ca, me.sa, me.alt, me.vs };
                    const bool ring = i <= (nl >> 1);
#pragma unroll
                    for (int f = 0; f < kSoaFields; f++) {
                        float *p0 = srow + f * FS + i, *p1 = p0 + (C1 - 1);            // copy 1 holds f[k + 1] at k
                        p0[0] = fv[f];
                        if (i > 0) p1[0] = fv[f];
                        if (ring) { p0[nl] = fv[f]; p1[nl] = fv[f]; }
                    }
                    if (reso) srow[kSoaFields * FS + i] = a.hdg;                       // resumenav's bouncing test
                    // monotone within the launch: once any aircraft climbs, descends or leaves the common level
                    if (a.vs != 0.0f || !(fabsf(a.alt - alt_ref) < 0.5f * (float)A.P.hpz)) es[le].vneed = 1;
                } else {
                    make_rec<R>(a, lat0, lon0, rb[i]);
                    if (reso) sdist[le * N + i] = a.hdg;
                }
            }
            __syncthreads();
            if (RESO && valid && i == 0) es[le].anyconf[buf ^ 1] = 0;
            acc = RowAcc<R>();
            R dve = R(0), dvn = R(0);
            bool touched = false, active = false;
            unsigned long long keep = 0;
            if constexpr (SYM) {
                // conflict bits without atomics and without a branch: thread i keeps the flags of ITS half ring as the
                // relative mask F (bit d-1 <-> partner i+d) and publishes it once per substep as hr[i] (and the ring copy
                // hr[i + nl], so that the partners behind it are read without an index wrap); phase 2 completes the row
                // with bit k-1 of the masks of the nl-1-H partners i-k.  One buffer: barrier A of the next substep
                // separates the readers of this substep from the next writers.
                unsigned *hr = reinterpret_cast<unsigned *>(crow) + (size_t)le * 2 * N;
                unsigned F = 0u;
                // partner j of the row, any j < nl + H + 1 (copy 0)
                auto rec_at = [&](int j) {
                    RecFast o;
                    const float *p = srow + j;
                    o.x = p[SF_X * FS]; o.y = p[SF_Y * FS]; o.u = p[SF_U * FS]; o.v = p[SF_V * FS];
                    o.ca = p[SF_CA * FS]; o.sa = p[SF_SA * FS]; o.alt = p[SF_ALT * FS]; o.vs = p[SF_VS * FS];
                    return o;
                };
                if (live && !skip) {                            // phase 1: half ring, flags only
                    const int H = nl >> 1;
                    // level, co-altitude traffic (Horizontal and Sector envs): the vertical window is (-inf, inf) and the
                    // sweep runs PACKED -- partners d, d+1 of one trip in the two halves of f32x2 registers (FFMA2 / FADD2
                    // / FMUL2), one LDS.64 per field and trip; BSG_PG trips are interleaved (straight-line code, the
                    // dependency chains overlap: the pair body is latency-bound at 4-5 warps per scheduler).
                    // The sign bit of a margin is the flag: one funnel shift per pair pushes it into G (pair d ends at
                    // bit nd-d); reversed into F (bit d-1) after the loop.
                    if (!es[le].vneed) {
                        const int Q = (H + 1) >> 1;             // trips: partners d = 2q+1, 2q+2 (the last may run one past H)
                        const f32x2 *g = reinterpret_cast<const f32x2 *>(srow + ((i & 1) ? i + 1 : C1 + i));
                        Rec2 m2;
                        m2.x = pk2(me.x, me.x); m2.y = pk2(me.y, me.y); m2.u = pk2(me.u, me.u); m2.v = pk2(me.v, me.v);
                        m2.ca = pk2(me.ca, me.ca); m2.sa = pk2(me.sa, me.sa);
                        Cd2 C;
                        C.R2 = pk2(PF.R2, PF.R2); C.c1em6 = pk2(1e-6f, 1e-6f); C.ndtlook = pk2(-PF.dtlook, -PF.dtlook);
                        const int FH = FS >> 1;
                        auto trip = [&](int q, unsigned &G) {
                            const f32x2 *p = g + q;
                            Rec2 o;
                            o.x = p[SF_X * FH]; o.y = p[SF_Y * FH]; o.u = p[SF_U * FH]; o.v = p[SF_V * FH];
                            o.ca = p[SF_CA * FH]; o.sa = p[SF_SA * FH];
                            float lo, hi;
                            pair_fast_sign2(m2, o, C, lo, hi);
                            G = __funnelshift_l(__float_as_uint(lo), G, 1);
                            G = __funnelshift_l(__float_as_uint(hi), G, 1);
                        };
                        int q = 0;
                        unsigned G = 0u;
                        for (; q + (BSG_PG - 1) < Q; q += BSG_PG) {
#pragma unroll
                            for (int u = 0; u < BSG_PG; u++) trip(q + u, G);
                        }
#pragma unroll 1
                        for (; q < Q; q++) trip(q, G);
                        F = H > 0 ? (__brev(G) >> (32 - 2 * Q)) & ((1u << H) - 1u) : 0u;
                    } else {                                    // generic vertical window: scalar form, BSG_PW pairs per trip
                        int d = 1;
                        unsigned G = 0u;
                        for (; d + (BSG_PW - 1) <= H; d += BSG_PW) {
#pragma unroll
                            for (int u = 0; u < BSG_PW; u++)
                                G = __funnelshift_l(__float_as_uint(pair_fast_sign<true>(me, rec_at(i + d + u), PF)), G, 1);
                        }
#pragma unroll 1
                        for (; d <= H; d++)
                            G = __funnelshift_l(__float_as_uint(pair_fast_sign<true>(me, rec_at(i + d), PF)), G, 1);
                        F = H > 0 ? __brev(G) >> (32 - H) : 0u;
                    }
                    hr[i] = F; hr[i + nl] = F;
                    if (RESO && F) es[le].anyconf[buf] = 1;
                }
                // phase 2 feeds MVP / resumenav every substep; the CD row outputs report the LAST detection of the
                // step (what bs.traf.cd holds afterwards), so the row reductions run on the last substep only
                const bool want_acc = (s == A.substeps - 1);
                const bool need_rows = RESO || want_acc;
                if (need_rows) __syncthreads();
                if (need_rows && live && !skip) {               // phase 2: my row, conflicting / resolving pairs only
                    // B bit (k-1) <-> partner i-k: the pair (i-k, i) is bit k-1 of that partner's mask
                    // Groups of 8 partners, fully unrolled (LDS with an immediate offset + one LOP3 with an immediate mask
                    // per partner); a group may read up to 7 words past partner kmax -- other rows' masks, at worst the
                    // 31 words below this array, all inside the CTA's shared memory -- and the final mask drops them.
                    unsigned B = 0u;
                    {
                        const int kmax = nl - 1 - (nl >> 1);                 // <= 31
                        const unsigned *hb = hr + i + nl;
#pragma unroll
                        for (int g8 = 0; g8 < 32; g8 += 8) {
                            if (g8 < kmax) {
#pragma unroll
                                for (int q = 1; q <= 8; q++) B |= hb[-(g8 + q)] & (1u << (g8 + q - 1));
                            }
                        }
                        B &= (1u << kmax) - 1u;
                    }
                    // relative -> absolute positions.  Forward: i+d, wrapped ones (i+d >= nl) at i+d-nl.
                    unsigned long long fwd = (i < 63 ? (unsigned long long)F << (i + 1) : 0ull) & (nl < 64 ? (1ull << nl) - 1ull : ~0ull);
                    if (nl - 1 - i < 32) fwd |= (unsigned long long)(F >> (nl - 1 - i));
                    // Backward: the bit-reversed word, shifted, lands on i-k (and i-k+nl for the wrapped ones)
                    const unsigned long long rb64 = (unsigned long long)__brev(B) << 32;     // bit 64-k <-> partner i-k
                    unsigned long long back = i > 0 ? rb64 >> (64 - i) : 0ull;               // positions i-k >= 0
                    if (i < 32) back |= (rb64 << i) >> (64 - nl);                            // k > i: positions i-k+nl
                    const unsigned long long cr = fwd | back;
                    const unsigned long long notme = ~(1ull << i);
                    const float R2f = PF.rpz * PF.rpz, R2b = R2f * (float)(A.P.fach * A.P.fach);
                    const float *hdgrow = srow + kSoaFields * FS;
                    // resumenav test of one resolving pair (squared distances; the canonical pair quantities)
                    auto resume = [&](int j, unsigned long long jb, float dist2, float dot) {
                        const bool past = dot > 0.0f;
                        bool kp = !past || dist2 < R2f;
                        if (!kp && dist2 < R2b) kp = fabsf(a.hdg - hdgrow[j]) < 30.0f;      // bouncing (rare)
                        touched = true;
                        if (kp) { active = true; keep |= jb; }
                    };
                    // two walks instead of one with a divergent `if (sw)`: conflicting pairs first (row reductions, MVP,
                    // resumenav), then the pairs that are only still resolving (resumenav alone).  Ascending j in
                    // both, so the MVP sum keeps its order.
                    unsigned long long S = cr & notme;
                    while (S) {
                        const int j = __ffsll((long long)S) - 1;
                        const unsigned long long jb = S & (0ull - S);        // lowest set bit (no 64-bit shift by j)
                        S ^= jb;
                        const RecFast o = rec_at(j);
                        PairCore pc;
                        pair_fast_core(me, o, PF, true, pc);
                        if (want_acc)
                            acc.add_conflict(j, pc.tcpa, pc.dcpa2, pc.tin, (pc.dist2 < PF.R2) && (fabsf(pc.dalt) < PF.hpz));
                        if constexpr (RESO) {
                            float d1, d2;
                            mvp_pair_fast(pc.dx, pc.dy, pc.dist2, pc.tcpa, pc.du, pc.dv, PF.rpz_m, d1, d2);
                            dve = dve - d1; dvn = dvn - d2;
                            resume(j, jb, pc.dist2, pc.dot);
                        }
                    }
                    if constexpr (RESO) {
                        S = resorow & ~cr & notme;
                        while (S) {
                            const int j = __ffsll((long long)S) - 1;
                            const unsigned long long jb = S & (0ull - S);
                            S ^= jb;
                            float dist2, dot;
                            pair_fast_geom(me, rec_at(j), dist2, dot);
                            resume(j, jb, dist2, dot);
                        }
                    }
                    if (RESO && A.reso_off_own && i == 0) { dve = 0.0f; dvn = 0.0f; }   // RESOOFF: the agent does not resolve
                }
            } else {
            if (live && !skip) {
                const Rec me = rb[i];
#pragma unroll 4
                for (int j = 0; j < nl; j++) {
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
                        po.tcpa = pc.tcpa; po.dcpa2 = pc.dcpa2; po.tin = pc.tin; po.dx = pc.dx; po.dy = pc.dy; po.dist = pc.dist2;
                        if (__any_sync(__activemask(), po.sw)) {   // conflicting pairs only: row reductions off the hot path
                            if (po.sw) acc.add_conflict(j, pc.tcpa, pc.dcpa2, pc.tin, (pc.dist2 < PF.R2) && (fabsf(pc.dalt) < PF.hpz));
                        }
                    }
                    if constexpr (RESO) {
                        const bool inres = off && (po.sw || ((resorow >> j) & 1ull));
                        if (inres) {
                            R vrx = o.u - me.u, vry = o.v - me.v;
                            if (po.sw) {
                                R dist, rpz_m, d1, d2;
                                if constexpr (sizeof(R) == 8) { dist = po.dist; rpz_m = PL.rpz * PL.fach; }
                                else { dist = sqrtf(po.dist); rpz_m = PF.rpz_m; }
                                mvp_pair<R>(po.dx, po.dy, dist, po.tcpa, vrx, vry, rpz_m, d1, d2);
                                dve = dve - d1; dvn = dvn - d2;
                                if (A.reso_off_own && i == 0) { dve = R(0); dvn = R(0); }
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
                            if (!past || hdist < rpz || bouncing) { active = true; keep |= 1ull << j; }
                        }
                    }
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
                        double vmin, vmax;
                        phase_vlimits(a.phase, vmin, vmax);
                        if constexpr (sizeof(R) == 8) a.asas_trk = pymod(atan2(ne, nn) * 180 / kPi, 360);
                        else { float t = atan2f(ne, nn) * float(kR2D); a.asas_trk = t < 0.0f ? t + 360.0f : t; }
                        R newgs = sqrt(ne * ne + nn * nn);
                        a.asas_tas = fmax(R(vmin), fmin(R(vmax), newgs));
                        a.asas_vs = fmax(R(kVsMin), fmin(R(kVsMax), a.vs));
                        a.asas_alt = a.selalt;
                    }
                    if (touched) a.asas_active = active ? 1 : 0;
                    resorow = keep;
                }
            }
        }
        if (live && !skip) {
            const bool last = (s == A.substeps - 1);
            if constexpr (sizeof(R) == 8) kin_substep_lit<EXT>(a, A.dt, reso, last, &X);
            else kin_substep_fast<EXT>(a, A.dt, reso, last && i == 0 && (A.kind == BSG_KIND_SECTOR || A.kind == BSG_KIND_MERGE), &X);
        }
    }
    BSG_STAMP(2);
    if (live && !skip) {
        ac_store<R, RESO, EXT>(a, A.pos, A.st, EN, k);
        if (reso) A.reso[k] = resorow;
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
