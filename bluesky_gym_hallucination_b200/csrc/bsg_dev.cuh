// bsg_dev.cuh -- device-side constants and scalar models of the batched BlueSky core (sm_100a).
// What is computed follows SURVEY.md Appendix A (BlueSky 1.0.3 behaviour the reference envs rely on:
// bs.sim.step() at bluesky_gym/envs/horizontal_cr_env.py:109, vertical_cr_env.py:289,
// sector_cr_env.py:121); how it is computed is this repo's own: templated on the real type R,
// R = double is the literal "parity" order of operations, R = float the fast path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/bsgcore.h"

namespace bsg {

constexpr double kPi = 3.141592653589793238462643383279502884;
constexpr double kD2R = kPi / 180.0;
constexpr double kR2D = 180.0 / kPi;
constexpr double kRe = 6371000.0;
constexpr double kNm = 1852.0;
constexpr double kFt = 0.3048;
constexpr double kFpm = 0.3048 / 60.0;
constexpr double kKts = 0.514444;
constexpr double kG0 = 9.80665;
constexpr double kRgas = 287.05287;
constexpr double kP0 = 101325.0;
constexpr double kRho0 = 1.225;
constexpr double kGamma = 1.40;
constexpr double kVsDef = 1500.0 * kFpm;
constexpr double kBank = 25.0 * kD2R;
// performance envelope (A320-like, documented in DESIGN.md; same table as the oracle's)
enum Phase { PH_NA, PH_GD, PH_IC, PH_CL, PH_CR, PH_DE, PH_AP };
constexpr double kVminIc = 65.0, kVmaxIc = 82.0, kVminEr = 80.0, kVmaxEr = 180.0, kVminAp = 60.0, kVmaxAp = 100.0;
constexpr double kVsMin = -20.0, kVsMax = 20.0, kHmax = 12131.0, kMmo = 0.82, kAxMax = 2.0;

// device state: positions always fp64 (fp32 degrees cannot hold a 5 NM separation to 1e-5),
// everything else in R.  Layout per env: pos[2][N], st[NS][N]  (field-major => coalesced by aircraft).
enum StField { S_ALT, S_HDG, S_TAS, S_VS, S_CAS, S_SELSPD, S_SELALT, S_SELVS, S_APTRK, S_AX,
               S_ASAS_TRK, S_ASAS_TAS, S_ASAS_VS, S_ASAS_ALT, S_ASAS_ACTIVE, S_PHASE,
               // only read / written by the EXT kernels (wind field, LNAV routes); derived on export otherwise
               S_GSN, S_GSE, S_LNAV,
               // fast-path caches of pure functions (key -> value); unused by the parity path
               S_C_ALT, S_C_SPD, S_C_APTAS, S_C_PH, S_C_SELALT, S_C_LO, S_C_HI, S_NS };

// launch constants of the reset path: the ISA conversions of the fixed creation speeds, evaluated once on the
// host at bsg_create (six pow() per aircraft would otherwise sit on the critical path of a fused reset)
struct ResetConst {
    double tas_h, cas_h, tas_s, coslat_hv, ccl;   // vcas2tas(150,0), vtas2cas(tas_h,0), vcas2tas(150,350), cos(52 deg), cos(sector centre lat)
    double tas_m, faf_lat, faf_lon, rwy_lat, rwy_lon;   // Merge: vcas2tas(100, 10000), FAF and runway (merge_env.py:39-45)
};

// wind field (bsg_set_wind): passed by value in the kernel arguments
// nk > 0: ONE wind point with an altitude profile of nk knots (kalt, vn, ve): WindFieldWrapper(..., alt=[...])
struct WindArgs { int n, obs; double lat[BSG_MAX_WIND], lon[BSG_MAX_WIND], vn[BSG_MAX_WIND], ve[BSG_MAX_WIND]; int nk, pad; double kalt[BSG_MAX_WIND]; };

template <typename R> __device__ __forceinline__ R rsgn(R x) { return R((x > R(0)) - (x < R(0))); }

// numpy/Python float modulo (result takes the sign of the divisor)
__device__ __forceinline__ double pymod(double a, double b) {
    double r = fmod(a, b);
    if (r != 0.0 && ((r < 0.0) != (b < 0.0))) r += b;
    return r;
}
// fast path: a in (-720, 1080) for every call site; two conditional wraps replace fmodf
__device__ __forceinline__ float wrap360(float a) {
    a -= 360.0f * floorf(a * (1.0f / 360.0f));
    a = (a >= 360.0f) ? a - 360.0f : a;
    return (a < 0.0f) ? a + 360.0f : a;
}
// one wrap only: a in (-360, 720) -- differences and sums of two angles already in [0, 360) (same values as wrap360 there)
__device__ __forceinline__ float wrap360_once(float a) {
    a = (a < 0.0f) ? a + 360.0f : a;                 // (a tiny negative angle rounds up to 360.0 here ...
    return (a >= 360.0f) ? a - 360.0f : a;           //  ... and lands on 0, as in wrap360)
}
// (sin, cos) of an angle in degrees, |deg| < 2^20: x = deg / 180, quadrant k = rint(2 x), r = x - k / 2 in [-1/4, 1/4] (exact),
// Taylor polynomials of sin(pi r) to r^9 and cos(pi r) to r^8 (truncation 2e-9 / 3e-8 absolute), quadrant swap / signs by
// bit tests: about 20 instructions against the 30+ of sincospif (which also serves huge arguments, inf and NaN).
__device__ __forceinline__ void sincos_deg_fast(float deg, float &s, float &c) {
    const float x = deg * (1.0f / 180.0f);
    const float kf = rintf(x + x);
    const float r = fmaf(kf, -0.5f, x), t = r * r;
    const float sp = r * fmaf(t, fmaf(t, fmaf(t, fmaf(t, 0.0821458866f, -0.599264529f), 2.55016404f), -5.16771278f), 3.14159265f);
    const float cp = fmaf(t, fmaf(t, fmaf(t, fmaf(t, 0.235330630f, -1.33526277f), 4.05871213f), -4.93480220f), 1.0f);
    const int k = (int)kf;
    const float a = (k & 1) ? cp : sp, b = (k & 1) ? sp : cp;          // sin(pi x) = +-sin / +-cos of pi r by quadrant
    s = (k & 2) ? -a : a;
    c = ((k + 1) & 2) ? -b : b;
}
template <typename R> __device__ __forceinline__ R mod360(R a) {
    if constexpr (sizeof(R) == 8) return pymod(a, 360.0); else return wrap360(a);
}

// ------------------------------------------------------------------ ISA atmosphere (always fp64: pure
// functions evaluated rarely on the fast path thanks to the key/value caches)
__device__ __forceinline__ void vatmos(double h, double &p, double &rho, double &T) {
    T = fmax(288.15 - 0.0065 * h, 216.65);
    double rhotrop = 1.225 * pow(T / 288.15, 4.256848030018761);
    double dhstrat = fmax(0.0, h - 11000.0);
    rho = rhotrop * exp(-dhstrat / 6341.552161);
    p = rho * kRgas * T;
}
__device__ __forceinline__ double vvsound(double h) { return sqrt(kGamma * kRgas * fmax(288.15 - 0.0065 * h, 216.65)); }
static __device__ __noinline__ double vtas2cas(double tas, double h) {
    double p, rho, T;
    vatmos(h, p, rho, T);
    double qdyn = p * (pow(1.0 + rho * tas * tas / (7.0 * p), 3.5) - 1.0);
    double cas = sqrt(7.0 * kP0 / kRho0 * (pow(qdyn / kP0 + 1.0, 2.0 / 7.0) - 1.0));
    return tas < 0 ? -1 * cas : cas;
}
static __device__ __noinline__ double vcas2tas(double cas, double h) {
    double p, rho, T;
    vatmos(h, p, rho, T);
    double qdyn = kP0 * (pow(1.0 + kRho0 * cas * cas / (7.0 * kP0), 3.5) - 1.0);
    double tas = sqrt(7.0 * p / rho * (pow(1.0 + qdyn / p, 2.0 / 7.0) - 1.0));
    return cas < 0 ? -1 * tas : tas;
}
__device__ __forceinline__ double vcasormach2tas(double spd, double h) {
    return fabs(spd) < 1 ? spd * vvsound(h) : vcas2tas(spd, h);
}


// ------------------------------------------------------------------ pow-free ISA conversions (fast path)
// The fp32 fast path rounds these to float, but the agent's commanded speed changes every step, so the memo
// misses once per env-step and three fp64 pow() (~200 instructions each, one live lane) sat on the warp's
// critical path.  Same functions, evaluated without pow (<= 1e-11 relative to the literal forms, i.e. identical
// after rounding to fp32):  (1+x)^3.5 = (1+x)^3 sqrt(1+x);  (1+y)^(2/7) by one Newton step on z^7 = (1+y)^2 from
// an fp32 exp2/log2 start (quadratic: 1e-6 -> 1e-11);  (T/T0)^4.2568 by its binomial series in z = 1 - T/T0
// for z <= 0.1 (h <= 4433 m; 16 terms, remainder < 1e-15), the literal pow above that.
__device__ __forceinline__ double pow35m1(double x) { const double o = 1.0 + x; return o * o * o * sqrt(o) - 1.0; }
__device__ __forceinline__ double pow27m1(double y) {
    const double o = 1.0 + y, c = o * o;
    double z = (double)exp2f((2.0f / 7.0f) * log2f((float)o));
    const double z2 = z * z, z6 = z2 * z2 * z2;
    z -= (double)((float)(z6 * z - c) / (float)(7.0 * z6));
    return z - 1.0;
}
__device__ __forceinline__ void vatmos_fast(double h, double &p, double &rho, double &T) {
    T = fmax(288.15 - 0.0065 * h, 216.65);
    const double z = 1.0 - T / 288.15;
    double th;
    if (z <= 0.1) {
        const double c[16] = { 1.0, 4.256848030018761, 6.9319535603279245, 5.214788578935872, 1.6385491880999719, 0.08417162621046376, -0.010425384972471766, 0.00259614719322528, -0.0008902032859321487, 0.0003702406870467438, -0.00017561078441329702, 9.168722022301611e-05, -5.152173830574456e-05, 3.0687742266075684e-05, -1.9164828160565436e-05, 1.2448388883131001e-05 };
        th = c[15];
#pragma unroll
        for (int k = 14; k >= 0; k--) th = fma(th, -z, c[k]);
    } else th = pow(T / 288.15, 4.256848030018761);
    rho = 1.225 * th;
    if (h > 11000.0) rho *= exp(-(h - 11000.0) / 6341.552161);
    p = rho * kRgas * T;
}
static __device__ __noinline__ double vtas2cas_fast(double tas, double h) {
    double p, rho, T;
    vatmos_fast(h, p, rho, T);
    const double qdyn = p * pow35m1(rho * tas * tas / (7.0 * p));
    const double cas = sqrt(7.0 * kP0 / kRho0 * pow27m1(qdyn / kP0));
    return tas < 0 ? -cas : cas;
}
static __device__ __noinline__ double vcas2tas_fast(double cas, double h) {
    double p, rho, T;
    vatmos_fast(h, p, rho, T);
    const double qdyn = kP0 * pow35m1(kRho0 * cas * cas / (7.0 * kP0));
    const double tas = sqrt(7.0 * p / rho * pow27m1(qdyn / p));
    return cas < 0 ? -tas : tas;
}
__device__ __forceinline__ double vcasormach2tas_fast(double spd, double h) {
    return fabs(spd) < 1 ? spd * vvsound(h) : vcas2tas_fast(spd, h);
}

// ------------------------------------------------------------------ performance envelope
template <typename R> __device__ __forceinline__ int flight_phase(R vs, R alt) {
    R roc, h;
    if constexpr (sizeof(R) == 8) { roc = vs / kFpm; h = alt / kFt; }               // literal order (numpy divides)
    else { roc = vs * float(1.0 / kFpm); h = alt * float(1.0 / kFt); }              // fast path: no IEEE division
    if constexpr (sizeof(R) == 8) {                                 // literal chain of tests
        int ph = PH_NA;
        if (h <= R(75)) ph = PH_GD;
        if (h > R(75) && h <= R(1000) && roc >= R(150)) ph = PH_IC;
        if (h > R(75) && h <= R(1000) && roc <= R(-150)) ph = PH_AP;
        if (h > R(1000) && roc >= R(150)) ph = PH_CL;
        if (h > R(1000) && roc <= R(-150)) ph = PH_DE;
        if (h > R(10000) && roc > R(-150) && roc < R(150)) ph = PH_CR;
        return ph;
    } else {
        // the same table as a decision tree (equal for every input, NaN included: checked exhaustively on the host)
        const bool gd = h <= R(75), air = h > R(75), low = h <= R(1000), up = roc >= R(150), dn = roc <= R(-150);
        const int moving = up ? (low ? PH_IC : PH_CL) : (low ? PH_AP : PH_DE);
        const int level = (h > R(10000) && roc > R(-150) && roc < R(150)) ? PH_CR : PH_NA;
        return gd ? PH_GD : ((air && (up || dn)) ? moving : level);
    }
}
__device__ __forceinline__ void phase_vlimits(int ph, double &vmin, double &vmax) {
    bool enroute = (ph == PH_CL) || (ph == PH_CR) || (ph == PH_DE);
    vmin = 0.0; vmax = kVmaxEr;
    if (ph == PH_IC) vmin = kVminIc;
    if (enroute) vmin = kVminEr;
    if (ph == PH_AP) vmin = kVminAp;
    if (ph == PH_IC || ph == PH_GD) vmax = kVmaxIc;
    if (ph == PH_AP) vmax = kVmaxAp;
}

// ------------------------------------------------------------------ flat-earth geometry (literal form)
// kwikqdrdist of bluesky/tools/geo.py; qdr [deg 0..360), dist [NM]
__device__ __forceinline__ void kwikqdrdist(double lata, double lona, double latb, double lonb, double &qdr, double &dist) {
    double dlat = (latb - lata) * kD2R;
    double dlon = (pymod((lonb - lona) + 180, 360) - 180) * kD2R;
    double cavelat = cos((lata + latb) * kD2R * 0.5);
    double dangle = sqrt(dlat * dlat + dlon * dlon * cavelat * cavelat);
    dist = kRe * dangle / kNm;
    qdr = pymod(atan2(dlon * cavelat, dlat) * kR2D, 360.0);
}
// env-layer angle wrap (bluesky_gym/envs/common/functions.py:17-22): wraps once only
template <typename R> __device__ __forceinline__ R bound180(R a) {
    if (a > R(180)) return -(R(360) - a);
    if (a < R(-180)) return R(360) + a;
    return a;
}

}  // namespace bsg
