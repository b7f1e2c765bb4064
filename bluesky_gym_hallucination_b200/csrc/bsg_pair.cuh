// bsg_pair.cuh -- one ordered (ownship i, intruder j) evaluation of BlueSky's state-based conflict
// detection (StateBased.detect) and the MVP displacement for a conflict pair (MVP.MVP,
// SURVEY.md A.7).  Two forms:
//   PairLit  (double): the literal operation order of the reference-style numpy code (qdr via atan2,
//                      dx/dy via sin/cos of qdr, dcpa2 = |dist^2 - tcpa^2 dv2|) -> parity path, <= 1e-12.
//   PairFast (float) : algebraically identical, no transcendental per pair: dx = Re*dlon*cos(mean lat) with
//                      cos((a+b)/2) = ca_i ca_j - sa_i sa_j from per-aircraft half-angle terms, and
//                      dcpa2 = (dx dv - dy du)^2 / dv2 (cross form: 10x better conditioned in fp32 than the
//                      difference of squares).  3 MUFU (rcp, rcp, sqrt) + ~55 FP32 issue slots per pair.
#pragma once
#include "bsg_dev.cuh"

namespace bsg {

struct __align__(16) RecFast { float x, y, alt, u, v, vs, ca, sa; };         // 32 B: two float4
struct __align__(16) RecLit { double lat, lon, alt, u, v, vs, hdg, pad; };   // 64 B

struct CdParams { double rpz, hpz, dtlook, fach, facv; };

template <typename R> struct PairOut {
    R tcpa, dcpa2, tin, dx, dy, dist, dalt;
    bool sw, los;
};

__device__ __forceinline__ void pair_lit(const RecLit &a, const RecLit &b, const CdParams &P, PairOut<double> &o) {
    double qdr, dnm;
    kwikqdrdist(a.lat, a.lon, b.lat, b.lon, qdr, dnm);
    double dist = dnm * kNm;
    double qr = qdr * kD2R;
    double dx = dist * sin(qr), dy = dist * cos(qr);
    double du = b.u - a.u, dv = b.v - a.v;
    // explicit round-to-nearest mul/add: no FMA contraction, the cancelling sums round exactly like numpy's
    double dv2 = __dadd_rn(__dmul_rn(du, du), __dmul_rn(dv, dv));
    if (fabs(dv2) < 1e-6) dv2 = 1e-6;
    double vrel = sqrt(dv2);
    double tcpa = -__dadd_rn(__dmul_rn(du, dx), __dmul_rn(dv, dy)) / dv2;
    double dcpa2 = fabs(__dsub_rn(__dmul_rn(dist, dist), __dmul_rn(__dmul_rn(tcpa, tcpa), dv2)));
    double R2 = P.rpz * P.rpz;
    bool swhor = dcpa2 < R2;
    double dtinhor = sqrt(fmax(0.0, R2 - dcpa2)) / vrel;
    double tinhor = swhor ? tcpa - dtinhor : 1e8;
    double touthor = swhor ? tcpa + dtinhor : -1e8;
    double dalt = b.alt - a.alt;
    double dvs = b.vs - a.vs;
    if (fabs(dvs) < 1e-6) dvs = 1e-6;
    double thi = (dalt + P.hpz) / -dvs, tlo = (dalt - P.hpz) / -dvs;
    double tinver = fmin(thi, tlo), toutver = fmax(thi, tlo);
    double tin = fmax(tinver, tinhor), tout = fmin(toutver, touthor);
    o.sw = swhor && (tin <= tout) && (tout > 0.0) && (tin < P.dtlook);
    o.los = (dist < P.rpz) && (fabs(dalt) < P.hpz);
    o.tcpa = tcpa; o.dcpa2 = dcpa2; o.tin = tin; o.dx = dx; o.dy = dy; o.dist = dist; o.dalt = dalt;
}

struct CdParamsF { float R2, hpz, dtlook, rpz, rpz_m, hpz_m; };

// MUFU approximations (1-2 ulp): the IEEE forms (__frcp_rn, sqrtf) expand to range checks + a slow-path
// call each and break the software pipelining of the pair loop
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sqrt_approx(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// Hot form of the fp32 pair: only what decides the conflict flag.  Instruction budget per pair (B200: the ALU
// pipe -- FSETP/FSEL/FMNMX/integer -- issues at half the rate of the FMA pipe, so comparisons are the scarce
// resource): ~31 FMA-pipe ops, 3 MUFU, 10 ALU ops.  Everything needed only for conflicting pairs (LoS test,
// row min/argmin/max, MVP) is evaluated by the caller under a warp-uniform `any(sw)` branch.
//  * no select of tinhor/touthor: when dcpa2 >= R2 the sqrt argument is negative -> NaN, FMNMX drops it and
//    the flag is gated by swhor anyway;
//  * no clamp of |dvs|: rcp(0) = inf gives tinver/toutver = -/+inf (vertically inside) or both beyond the
//    look-ahead / in the past, the same decisions the reference's 1e-6 clamp produces.
struct PairCore { float tcpa, dcpa2, tin, dx, dy, dist2, dalt; };

// VERT = false: the caller has established that every aircraft of the env is level (vs == 0 exactly) and within
// hpz/2 of a common altitude; the vertical window of the generic form is then (-inf, +inf) identically
// ((dalt +/- hpz) * (-1/0)), so tin / tout are the horizontal ones and the flag is bit-identical.
template <bool VERT = true>
__device__ __forceinline__ bool pair_fast_core(const RecFast &a, const RecFast &b, const CdParamsF &P, bool off, PairCore &o) {
    const float dy = b.y - a.y;
    const float cav = fmaf(a.ca, b.ca, -(a.sa * b.sa));
    const float dx = (b.x - a.x) * cav;
    const float dist2 = fmaf(dx, dx, dy * dy);
    const float du = b.u - a.u, dv = b.v - a.v;
    // |v_rel|^2 with the reference's lower clamp (1e-6) folded in as an additive floor: for |v_rel|^2 > 16 the sum
    // rounds to |v_rel|^2 itself, for equal velocity vectors (common: integer dpsi draws) it is the clamp value
    // 1e-6 exactly; in between (relative speeds below 3 mm/s) the reference's own value is the arbitrary clamp.
    const float dv2 = fmaf(du, du, fmaf(dv, dv, 1e-6f));
    const float idv2 = rcp_approx(dv2);
    const float ndot = -fmaf(du, dx, dv * dy);
    const float tcpa = ndot * idv2;
    const float cr = fmaf(dx, dv, -(dy * du));
    // dist^2 - tcpa^2 dv2 = (cr^2 + dist^2 (dv2 - |v_rel|^2)) / dv2 with dv2 - |v_rel|^2 = 1e-6: the second term is
    // below one ulp of cr^2-scale values unless the velocities are equal, where dCPA must come out as dist
    const float dcpa2 = fmaf(cr, cr, dist2 * 1e-6f) * idv2;
    const bool swhor = (dcpa2 < P.R2) && off;
    const float dtinhor = sqrt_approx((P.R2 - dcpa2) * idv2);
    const float dalt = b.alt - a.alt;
    // explicit FMAs: the same rounding in every kernel this body is inlined into (the compiler would otherwise
    // contract tcpa -/+ dtinhor in some contexts and not in others)
    float tin = fmaf(ndot, idv2, -dtinhor), tout = fmaf(ndot, idv2, dtinhor);
    if constexpr (VERT) {
        const float n = -rcp_approx(b.vs - a.vs);
        const float thi = (dalt + P.hpz) * n, tlo = (dalt - P.hpz) * n;
        tin = fmaxf(fminf(thi, tlo), tin); tout = fminf(fmaxf(thi, tlo), tout);
    }
    o.tcpa = tcpa; o.dcpa2 = dcpa2; o.tin = tin; o.dx = dx; o.dy = dy; o.dist2 = dist2; o.dalt = dalt;
    // VERT = false: tin <= tout holds whenever swhor does (dtinhor >= 0), so the test is dropped (same flag)
    if constexpr (!VERT) return swhor && (tout > 0.0f) && (tin < P.dtlook);
    else return swhor && (tin <= tout) && (tout > 0.0f) && (tin < P.dtlook);
}

// Flag-only form for the half-ring sweep of the fused step kernel: returns a float whose SIGN BIT is the conflict flag,
// so the caller shifts it into its mask with one funnel shift (no FSETP chain, no select: the comparison pipe is the
// scarce one).  The three (four) strict inequalities of pair_fast_core are folded into one 3-input max of the negated
// margins: conflict <=> max(dcpa2 - R2, -tout, tin - dtlook [, tin - tout]) < 0.  Same arithmetic as pair_fast_core for
// every margin; NaN margins (dcpa2 >= R2 -> sqrt of a negative) are dropped by the max and the first margin decides, an
// all-NaN pair (NaN state) gives the canonical NaN = sign clear = no conflict, as the comparisons would.  The only
// deviation is an exact tie tout == +0 (or tin == tout with VERT), inside any epsilon of the thresholds.
__device__ __forceinline__ float fmax3_f(float a, float b, float c) { float r; asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
template <bool VERT = true>
__device__ __forceinline__ float pair_fast_sign(const RecFast &a, const RecFast &b, const CdParamsF &P) {
    const float dy = b.y - a.y;
    const float cav = fmaf(a.ca, b.ca, -(a.sa * b.sa));
    const float dx = (b.x - a.x) * cav;
    const float dist2 = fmaf(dx, dx, dy * dy);
    const float du = b.u - a.u, dv = b.v - a.v;
    const float dv2 = fmaf(du, du, fmaf(dv, dv, 1e-6f));
    const float idv2 = rcp_approx(dv2);
    const float ndot = -fmaf(du, dx, dv * dy);
    const float cr = fmaf(dx, dv, -(dy * du));
    const float dcpa2 = fmaf(cr, cr, dist2 * 1e-6f) * idv2;
    const float dtinhor = sqrt_approx((P.R2 - dcpa2) * idv2);
    float tin = fmaf(ndot, idv2, -dtinhor), tout = fmaf(ndot, idv2, dtinhor);
    if constexpr (VERT) {
        const float dalt = b.alt - a.alt;
        const float n = -rcp_approx(b.vs - a.vs);
        const float thi = (dalt + P.hpz) * n, tlo = (dalt - P.hpz) * n;
        tin = fmaxf(fminf(thi, tlo), tin); tout = fminf(fmaxf(thi, tlo), tout);
        return fmaxf(fmax3_f(dcpa2 - P.R2, -tout, tin - P.dtlook), tin - tout);
    } else return fmax3_f(dcpa2 - P.R2, -tout, tin - P.dtlook);
}

// full form (used where every output is wanted): flag + LoS + the PairOut fields
__device__ __forceinline__ void pair_fast(const RecFast &a, const RecFast &b, const CdParamsF &P, PairOut<float> &o) {
    PairCore c;
    o.sw = pair_fast_core(a, b, P, true, c);
    o.los = (c.dist2 < P.R2) && (fabsf(c.dalt) < P.hpz);
    o.tcpa = c.tcpa; o.dcpa2 = c.dcpa2; o.tin = c.tin; o.dx = c.dx; o.dy = c.dy; o.dist = c.dist2 /* squared */; o.dalt = c.dalt;
}

// MVP displacement of ownship 1 for one conflict pair; drel = (dx, dy, dalt) is intruder relative to ownship.
// Returns the horizontal pair (dv1, dv2); the vertical term is not needed by the default
// horizontal-only resolution (newvs = own vs, asas alt = selalt), so it is not evaluated.
template <typename R>
__device__ __forceinline__ void mvp_pair(R dx, R dy, R dist, R tcpa, R vrelx, R vrely, R rpz_m, R &dv1, R &dv2) {
    R dcx = dx + vrelx * tcpa, dcy = dy + vrely * tcpa;
    R dabsH = sqrt(dcx * dcx + dcy * dcy);
    R iH = rpz_m - dabsH;
    if (dabsH <= R(10)) {
        dabsH = R(10);
        dcx = dy / dist * dabsH;
        dcy = -dx / dist * dabsH;
    }
    R den = fabs(tcpa) * dabsH;
    if (rpz_m < dist && dabsH < dist) {
        R erratum = cos(asin(rpz_m / dist) - asin(dabsH / dist));
        dv1 = ((rpz_m / erratum - dabsH) * dcx) / den;
        dv2 = ((rpz_m / erratum - dabsH) * dcy) / den;
    } else {
        dv1 = (iH * dcx) / den;
        dv2 = (iH * dcy) / den;
    }
}

// fp32 fast form of mvp_pair: same algebra, no inverse trigonometry and no IEEE divide.
//   cos(asin a - asin b) = sqrt((1 - a^2)(1 - b^2)) + a b      (a = rpz_m/dist, b = |dcpa_h|/dist, both in [0, 1))
// with 1 - a^2 = (dist^2 - rpz_m^2)/dist^2 taken from the squared quantities (no cancellation beyond the inputs').
// The branch tests compare squares, so they do not depend on the rounding of the MUFU approximations.
__device__ __forceinline__ float rsqrt_approx(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ void mvp_pair_fast(float dx, float dy, float dist2, float tcpa, float vrelx, float vrely,
                                              float rpz_m, float &dv1, float &dv2) {
    float dcx = fmaf(vrelx, tcpa, dx), dcy = fmaf(vrely, tcpa, dy);
    float dabsH2 = fmaf(dcx, dcx, dcy * dcy);
    float dabsH = sqrt_approx(dabsH2);
    const float iH = rpz_m - dabsH;
    const float idist = rsqrt_approx(dist2);
    if (dabsH2 <= 100.0f) {
        dabsH = 10.0f; dabsH2 = 100.0f;
        dcx = dy * idist * 10.0f;
        dcy = -dx * idist * 10.0f;
    }
    const float iden = rcp_approx(fabsf(tcpa) * dabsH);
    float f = iH;
    const float r2 = rpz_m * rpz_m;
    if (r2 < dist2 && dabsH2 < dist2) {
        const float id2 = idist * idist;
        const float erratum = fmaf(rpz_m * idist, dabsH * idist, sqrt_approx((dist2 - r2) * (dist2 - dabsH2)) * id2);
        f = fmaf(rpz_m, rcp_approx(erratum), -dabsH);
    }
    f *= iden;
    dv1 = f * dcx; dv2 = f * dcy;
}

// per-row running outputs of the sweep (branch-free: the pair loop stays software-pipelined).
// best = tCPA of the soonest conflict so far (+inf while there is none; ties keep the lower j).
template <typename R> struct RowAcc {
    int nconf = 0, nlos = 0, near_idx = -1;
    R near_tcpa = R(INFINITY), near_dcpa2 = R(0), near_tin = R(0), tcpamax = R(0);
    __device__ __forceinline__ void add(int j, const PairOut<R> &o) {
        nlos += o.los ? 1 : 0;
        nconf += o.sw ? 1 : 0;
        tcpamax = fmax(tcpamax, o.sw ? o.tcpa : R(0));
        const bool take = o.sw && (o.tcpa < near_tcpa);
        near_idx = take ? j : near_idx;
        near_tcpa = take ? o.tcpa : near_tcpa;
        near_dcpa2 = take ? o.dcpa2 : near_dcpa2;
        near_tin = take ? o.tin : near_tin;
    }
    // called only for conflicting pairs (under the caller's rare branch); LoS implies conflict off-threshold
    __device__ __forceinline__ void add_conflict(int j, R tcpa, R dcpa2, R tin, bool los) {
        nconf++;
        nlos += los ? 1 : 0;
        tcpamax = fmax(tcpamax, tcpa);
        if (tcpa < near_tcpa) { near_idx = j; near_tcpa = tcpa; near_dcpa2 = dcpa2; near_tin = tin; }
    }
    __device__ __forceinline__ R tcpa_out() const { return near_idx < 0 ? R(0) : near_tcpa; }
};

}  // namespace bsg
