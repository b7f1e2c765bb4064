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

struct CdParamsF { float R2, hpz, dtlook, rpz, rpz_m, hpz_m, fach2; };

// MUFU approximations (1-2 ulp): the IEEE forms (__frcp_rn, sqrtf) expand to range checks + a slow-path
// call each and break the software pipelining of the pair loop
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sqrt_approx(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fmax3_f(float a, float b, float c) { float r; asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }

// ---------------------------------------------------------------------------------------------------------------
// The fp32 pair, canonical form.  ONE sequence of individually rounded operations, written with the explicit
// round-to-nearest intrinsics (never contracted), so that every kernel that evaluates a pair -- the scalar forms
// below, the tiled sweeps -- produces the same bits:
//   dy = yb - ya;  cav = fma(ca_a, ca_b, -(sa_a sa_b));  dx = (xb - xa) cav;  dist2 = fma(dx, dx, dy dy)
//   du = ub - ua;  dv = vb - va;  dv2 = fma(du, du, fma(dv, dv, 1e-6))      (the reference's clamp as an additive floor)
//   n = rcp(-dv2)                                                           (= -1 / |v_rel|^2: no negation anywhere below)
//   dot = fma(du, dx, dv dy);  cr = fma(dx, dv, -(dy du));  t6 = fma(cr, cr, dist2 1e-6)
//   q = fma(t6, n, R2)            = R2 - dcpa2, one rounding (cross form: dcpa2 = (cr^2 + 1e-6 dist^2) / dv2)
//   dt = sqrt(-(q n))             = dtinhor; NaN when q < 0 (no horizontal conflict): the max below drops it
//   tout = fma(dot, n, dt);  tin = fma(dot, n, -dt)                         (tcpa +- dtinhor, tcpa = dot n)
//   [VERT]  nv = rcp(-(vsb - vsa));  thi = (dalt + hpz) nv;  tlo = (dalt - hpz) nv
//           tin = max(min(thi, tlo), tin);  tout = min(max(thi, tlo), tout)
//   margin = max(-q, -tout, tin - dtlook [, tin - tout])
// conflict <=> SIGN BIT of margin (the three / four strict inequalities of StateBased.detect folded into one max of
// the negated margins; NaN margins are dropped by max, an all-NaN pair gives the canonical NaN = sign clear = no
// conflict, as the comparisons would; the only deviation is an exact tie tout == +0, inside any epsilon).
//  * no clamp of |dvs|: rcp(0) = inf gives tinver/toutver = -/+inf (vertically inside) or both beyond the
//    look-ahead / in the past, the same decisions the reference's 1e-6 clamp produces.
// VERT = false: the caller has established that every aircraft of the env is level (vs == 0 exactly) and within
// hpz/2 of a common altitude; the vertical window of the generic form is then (-inf, +inf) identically
// ((dalt +/- hpz) * (-1/0)), so tin / tout are the horizontal ones and the flag is bit-identical.
struct PairCore { float tcpa, dcpa2, tin, dx, dy, dist2, dalt, du, dv, dot; };

template <bool VERT, bool OUT>
__device__ __forceinline__ float pair_fast_margin(const RecFast &a, const RecFast &b, const CdParamsF &P, PairCore *o) {
    const float dy = __fsub_rn(b.y, a.y);
    const float cav = __fmaf_rn(a.ca, b.ca, -__fmul_rn(a.sa, b.sa));
    const float dx = __fmul_rn(__fsub_rn(b.x, a.x), cav);
    const float dist2 = __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
    const float du = __fsub_rn(b.u, a.u), dv = __fsub_rn(b.v, a.v);
    const float dv2 = __fmaf_rn(du, du, __fmaf_rn(dv, dv, 1e-6f));
    const float n = rcp_approx(-dv2);
    const float dot = __fmaf_rn(du, dx, __fmul_rn(dv, dy));
    const float cr = __fmaf_rn(dx, dv, -__fmul_rn(dy, du));
    const float t6 = __fmaf_rn(cr, cr, __fmul_rn(dist2, 1e-6f));
    const float q = __fmaf_rn(t6, n, P.R2);
    const float dt = sqrt_approx(-__fmul_rn(q, n));
    float tout = __fmaf_rn(dot, n, dt), tin = __fmaf_rn(dot, n, -dt);
    float dalt = 0.0f;
    if constexpr (VERT || OUT) dalt = __fsub_rn(b.alt, a.alt);
    float m;
    if constexpr (VERT) {
        const float nv = rcp_approx(-__fsub_rn(b.vs, a.vs));
        const float thi = __fmul_rn(__fadd_rn(dalt, P.hpz), nv), tlo = __fmul_rn(__fadd_rn(dalt, -P.hpz), nv);
        tin = fmaxf(fminf(thi, tlo), tin); tout = fminf(fmaxf(thi, tlo), tout);
        m = fmaxf(fmax3_f(-q, -tout, __fadd_rn(tin, -P.dtlook)), __fsub_rn(tin, tout));
    } else m = fmax3_f(-q, -tout, __fadd_rn(tin, -P.dtlook));
    if constexpr (OUT) {
        o->tcpa = __fmul_rn(dot, n); o->dcpa2 = __fmul_rn(t6, -n); o->tin = tin; o->dx = dx; o->dy = dy; o->dist2 = dist2;
        o->dalt = dalt; o->du = du; o->dv = dv; o->dot = dot;
    }
    return m;
}
// geometry only (resumenav of a pair that is no longer in conflict): squared distance and r . v_rel, canonical roundings
__device__ __forceinline__ void pair_fast_geom(const RecFast &a, const RecFast &b, float &dist2, float &dot) {
    const float dy = __fsub_rn(b.y, a.y);
    const float cav = __fmaf_rn(a.ca, b.ca, -__fmul_rn(a.sa, b.sa));
    const float dx = __fmul_rn(__fsub_rn(b.x, a.x), cav);
    dist2 = __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
    const float du = __fsub_rn(b.u, a.u), dv = __fsub_rn(b.v, a.v);
    dot = __fmaf_rn(du, dx, __fmul_rn(dv, dy));
}

// flag + the quantities the row reductions / MVP need (evaluated by the callers for conflicting pairs only)
template <bool VERT = true>
__device__ __forceinline__ bool pair_fast_core(const RecFast &a, const RecFast &b, const CdParamsF &P, bool off, PairCore &o) {
    return (__float_as_int(pair_fast_margin<VERT, true>(a, b, P, &o)) < 0) && off;
}
// flag-only form: a float whose SIGN BIT is the conflict flag (one funnel shift pushes it into a mask register)
template <bool VERT = true>
__device__ __forceinline__ float pair_fast_sign(const RecFast &a, const RecFast &b, const CdParamsF &P) {
    return pair_fast_margin<VERT, false>(a, b, P, nullptr);
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

// Vertical term of MVP.MVP, evaluated only by the resolution methods that use it (reso_mode 1, 2):
// intrusion iV and time to solve tsolV from the relative altitude / vertical speed; tLOS = tin of the conflict.
// The |vrel_z| > 0 tests are BlueSky's exact-zero switches.
template <typename R>
__device__ __forceinline__ void mvp_vert(R dalt, R vrelz, R tLOS, R hpz_m, R dtlook, R &dv3, R &tsolV) {
    const bool moving = fabs(vrelz) > R(0);
    R iV = moving ? hpz_m : hpz_m - fabs(dalt);
    tsolV = moving ? fabs(dalt / vrelz) : tLOS;
    if (tsolV > dtlook) { tsolV = tLOS; iV = hpz_m; }
    dv3 = moving ? (iV / tsolV) * (-vrelz / fabs(vrelz)) : iV / tsolV;
}

// fp32 fast form of mvp_pair: same algebra, no inverse trigonometry and no IEEE divide.
//   cos(asin a - asin b) = sqrt((1 - a^2)(1 - b^2)) + a b      (a = rpz_m/dist, b = |dcpa_h|/dist, both in [0, 1))
// with 1 - a^2 = (dist^2 - rpz_m^2)/dist^2 taken from the squared quantities (no cancellation beyond the inputs').
// The branch tests compare squares, so they do not depend on the rounding of the MUFU approximations.
__device__ __forceinline__ float rsqrt_approx(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ void mvp_pair_fast(float dx, float dy, float dist2, float tcpa, float vrelx, float vrely,
                                              float rpz_m, float &dv1, float &dv2) {
    // explicit round-to-nearest operations throughout (never contracted with the caller's arithmetic): the result is
    // the same bits wherever the function is inlined, and exactly odd in (dx, dy, vrelx, vrely) -- the displacement of
    // the pair's other ownship is the negative
    float dcx = __fmaf_rn(vrelx, tcpa, dx), dcy = __fmaf_rn(vrely, tcpa, dy);
    float dabsH2 = __fmaf_rn(dcx, dcx, __fmul_rn(dcy, dcy));
    float dabsH = sqrt_approx(dabsH2);
    const float iH = __fsub_rn(rpz_m, dabsH);
    const float idist = rsqrt_approx(dist2);
    if (dabsH2 <= 100.0f) {
        dabsH = 10.0f; dabsH2 = 100.0f;
        dcx = __fmul_rn(__fmul_rn(dy, idist), 10.0f);
        dcy = __fmul_rn(__fmul_rn(-dx, idist), 10.0f);
    }
    const float iden = rcp_approx(__fmul_rn(fabsf(tcpa), dabsH));
    float f = iH;
    const float r2 = __fmul_rn(rpz_m, rpz_m);
    if (r2 < dist2 && dabsH2 < dist2) {
        const float id2 = __fmul_rn(idist, idist);
        const float root = sqrt_approx(__fmul_rn(__fsub_rn(dist2, r2), __fsub_rn(dist2, dabsH2)));
        const float erratum = __fmaf_rn(__fmul_rn(rpz_m, idist), __fmul_rn(dabsH, idist), __fmul_rn(root, id2));
        f = __fmaf_rn(rpz_m, rcp_approx(erratum), -dabsH);
    }
    f = __fmul_rn(f, iden);
    dv1 = __fmul_rn(f, dcx); dv2 = __fmul_rn(f, dcy);
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
