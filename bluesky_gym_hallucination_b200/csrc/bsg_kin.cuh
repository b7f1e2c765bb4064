// bsg_kin.cuh -- per-aircraft state in registers and one BlueSky traffic substep
// (Traffic.update: autopilot -> AP-or-ASAS select -> perf.update/limits -> update_airspeed ->
//  update_groundspeed -> update_pos).  R = double is the literal order of the
// numpy code; R = float is the fast path: positions stay fp64, pure atmosphere functions are
// memoised in the state (key -> value) so the steady state costs no pow() at all.
#pragma once
#include "bsg_dev.cuh"
#include "bsg_pair.cuh"

namespace bsg {

template <typename R> struct AcReg {
    double lat, lon;
    R alt, hdg, tas, vs, cas, selspd, selalt, selvs, aptrk, ax;
    R asas_trk, asas_tas, asas_vs, asas_alt;
    int asas_active, phase;
    R c_alt, c_spd, c_aptas, c_ph, c_selalt, c_lo, c_hi;   // fast-path caches
    R gsn, gse;                                            // derived: tas*cos(hdg), tas*sin(hdg); EXT: state (wind)
    R sa, ca;                                              // fast path: sin, cos of lat/2 (CD record + cos(lat))
    int lnav;                                              // EXT: active route waypoint, -1 = LNAV off
};

// EXT kernels (wind field, LNAV route of the Merge env): what one substep needs beyond the aircraft itself
struct ExtCtx {
    const WindArgs *W;
    bool merge;
    double faf_lat, faf_lon, rwy_lat, rwy_lon;             // Merge route: FAF then runway (merge_env.py:42-45,155-156)
};

template <typename R, bool RESO = true, bool EXT = false>
__device__ __forceinline__ void ac_load_raw(AcReg<R> &a, const double *pos, const R *st, size_t EN, size_t k) {
    if constexpr (EXT) { a.gsn = st[S_GSN * EN + k]; a.gse = st[S_GSE * EN + k]; a.lnav = (int)st[S_LNAV * EN + k]; }
    a.lat = pos[k]; a.lon = pos[EN + k];
    a.alt = st[S_ALT * EN + k]; a.hdg = st[S_HDG * EN + k]; a.tas = st[S_TAS * EN + k]; a.vs = st[S_VS * EN + k];
    a.cas = st[S_CAS * EN + k]; a.selspd = st[S_SELSPD * EN + k]; a.selalt = st[S_SELALT * EN + k];
    a.selvs = st[S_SELVS * EN + k]; a.aptrk = st[S_APTRK * EN + k]; a.ax = st[S_AX * EN + k];
    if constexpr (RESO) {         // the ASAS command slots only matter (and only change) with MVP on
        a.asas_trk = st[S_ASAS_TRK * EN + k]; a.asas_tas = st[S_ASAS_TAS * EN + k]; a.asas_vs = st[S_ASAS_VS * EN + k];
        a.asas_alt = st[S_ASAS_ALT * EN + k]; a.asas_active = (int)st[S_ASAS_ACTIVE * EN + k];
    }
    a.phase = (int)st[S_PHASE * EN + k];
    if constexpr (sizeof(R) == 4) {
        a.c_alt = st[S_C_ALT * EN + k]; a.c_spd = st[S_C_SPD * EN + k]; a.c_aptas = st[S_C_APTAS * EN + k];
        a.c_ph = st[S_C_PH * EN + k]; a.c_selalt = st[S_C_SELALT * EN + k]; a.c_lo = st[S_C_LO * EN + k];
        a.c_hi = st[S_C_HI * EN + k];
    }
}
// derived quantities, evaluated once the loads have landed
template <typename R, bool EXT = false> __device__ __forceinline__ void ac_derive(AcReg<R> &a) {
    if constexpr (EXT) {            // the ground-speed vector is state (it holds the wind of the last update)
        if constexpr (sizeof(R) == 4) sincospif((float)(a.lat * (1.0 / 360.0)), &a.sa, &a.ca);
    } else {
        R s, c;
        if constexpr (sizeof(R) == 8) { s = sin(a.hdg * kD2R); c = cos(a.hdg * kD2R); }
        else { sincos_deg_fast(a.hdg, s, c); sincospif((float)(a.lat * (1.0 / 360.0)), &a.sa, &a.ca); }
        a.gsn = a.tas * c; a.gse = a.tas * s;
    }
}

template <typename R, bool RESO = true, bool EXT = false>
__device__ __forceinline__ void ac_store(const AcReg<R> &a, double *pos, R *st, size_t EN, size_t k) {
    if constexpr (EXT) { st[S_GSN * EN + k] = a.gsn; st[S_GSE * EN + k] = a.gse; st[S_LNAV * EN + k] = R(a.lnav); }
    pos[k] = a.lat; pos[EN + k] = a.lon;
    st[S_ALT * EN + k] = a.alt; st[S_HDG * EN + k] = a.hdg; st[S_TAS * EN + k] = a.tas; st[S_VS * EN + k] = a.vs;
    st[S_CAS * EN + k] = a.cas; st[S_SELSPD * EN + k] = a.selspd; st[S_SELALT * EN + k] = a.selalt;
    st[S_SELVS * EN + k] = a.selvs; st[S_APTRK * EN + k] = a.aptrk; st[S_AX * EN + k] = a.ax;
    if constexpr (RESO) {
        st[S_ASAS_TRK * EN + k] = a.asas_trk; st[S_ASAS_TAS * EN + k] = a.asas_tas; st[S_ASAS_VS * EN + k] = a.asas_vs;
        st[S_ASAS_ALT * EN + k] = a.asas_alt; st[S_ASAS_ACTIVE * EN + k] = R(a.asas_active);
    }
    st[S_PHASE * EN + k] = R(a.phase);
    if constexpr (sizeof(R) == 4) {
        st[S_C_ALT * EN + k] = a.c_alt; st[S_C_SPD * EN + k] = a.c_spd; st[S_C_APTAS * EN + k] = a.c_aptas;
        st[S_C_PH * EN + k] = a.c_ph; st[S_C_SELALT * EN + k] = a.c_selalt; st[S_C_LO * EN + k] = a.c_lo;
        st[S_C_HI * EN + k] = a.c_hi;
    }
}

// ------------------------------------------------------------------ EXT: wind field and LNAV (SURVEY f-3, f-4)
// np.interp of the profile knots at altitude x (constant beyond the ends): both wind components from one interval search
// (indexed through the WindArgs reference: a pointer into the kernel parameters would copy them to local memory)
__device__ __forceinline__ void wind_knots(const WindArgs &W, double x, double &vn, double &ve) {
    const int n = W.nk;
    if (x <= W.kalt[0]) { vn = W.vn[0]; ve = W.ve[0]; return; }
    if (x >= W.kalt[n - 1]) { vn = W.vn[n - 1]; ve = W.ve[n - 1]; return; }
    int k = 0;
    while (k < n - 2 && x >= W.kalt[k + 1]) k++;
    const double dx = x - W.kalt[k], w = W.kalt[k + 1] - W.kalt[k];
    vn = W.vn[k] + dx * ((W.vn[k + 1] - W.vn[k]) / w);
    ve = W.ve[k] + dx * ((W.ve[k + 1] - W.ve[k]) / w);
}
// Windfield.getdata: one point = uniform, several = inverse distance-squared weights; one point with an altitude
// profile (W.nk knots) = the knots resampled onto BlueSky's 100 ft altitude axis (0 .. 45 000 ft), linear between the
// two axis levels around `alt` (oracle/refshim/bluesky/traffic/windfield.py).  Callers that BlueSky reaches with a
// numpy scalar altitude (selhdgcmd, the wrapper's wind observation) pass alt = 0, as upstream reads the field there.
template <typename R>
__device__ __forceinline__ void wind_at(const WindArgs &W, double lat, double lon, double alt, R &vn, R &ve) {
    vn = R(0); ve = R(0);
    if (W.n <= 0) return;
    if (W.nk > 0) {
        const double step = 100.0 * kFt, top = 450.0 * step;
        const double idx = fmax(0.0, fmin(top - 1e-20, alt)) / step;
        const double il = floor(idx), f = idx - il, a0 = il * step, a1 = (il + 1.0) * step;
        double n0, e0, n1, e1;
        wind_knots(W, a0, n0, e0); wind_knots(W, a1, n1, e1);
        vn = (R)((1.0 - f) * n0 + f * n1);
        ve = (R)((1.0 - f) * e0 + f * e1);
        return;
    }
    if (W.n == 1) { vn = (R)W.vn[0]; ve = (R)W.ve[0]; return; }
    R w[BSG_MAX_WIND], tot = R(0);
    for (int k = 0; k < W.n; k++) {
        if constexpr (sizeof(R) == 8) {
            const double cavelat = cos(0.5 * (lat + W.lat[k]) * kD2R);
            const double dy = (lat - W.lat[k]) * kD2R;
            const double dx = cavelat * ((pymod((lon - W.lon[k]) + 180.0, 360.0) - 180.0) * kD2R);
            w[k] = 1.0 / (1e-20 + dx * dx + dy * dy);
        } else {                            // differences in fp64, the rest fp32 (|dlon| << 180: no wrap)
            const float cavelat = cosf((float)(0.5 * (lat + W.lat[k]) * kD2R));
            const float dy = (float)((lat - W.lat[k]) * kD2R), dx = cavelat * (float)((lon - W.lon[k]) * kD2R);
            w[k] = 1.0f / (1e-20f + fmaf(dx, dx, dy * dy));
        }
        tot += w[k];
    }
    for (int k = 0; k < W.n; k++) { vn += w[k] / tot * (R)W.vn[k]; ve += w[k] / tot * (R)W.ve[k]; }
}

__device__ __forceinline__ double rwgs84(double latd) {
    const double lat = latd * kD2R, a = 6378137.0, b = 6356752.314245;
    const double coslat = cos(lat), sinlat = sin(lat);
    const double an = a * a * coslat, bn = b * b * sinlat, ad = a * coslat, bd = b * sinlat;
    return sqrt((an * an + bn * bn) / (ad * ad + bd * bd));
}
// geo.qdrdist: great-circle bearing [deg, -180..180] and distance [m] on the WGS-84 radius of the mean latitude
// (what BlueSky's LNAV steers by).  fp64: the literal arccos form.  fp32: the same quantities in the
// cancellation-free forms (haversine; cos1 sin2 - sin1 cos2 cos(dlon) = sin(dlat) + 2 sin1 cos2 sin^2(dlon/2))
// on coordinate differences taken in fp64.
template <typename R>
__device__ __forceinline__ void lnav_qdrdist(double latd1, double lond1, double latd2, double lond2, R &qdr, R &dist) {
    if constexpr (sizeof(R) == 8) {
        const double res1 = rwgs84(0.5 * (latd1 + latd2));
        const double a = 6378137.0, r1 = rwgs84(latd1), r2 = rwgs84(latd2);
        const double res2 = 0.5 * (fabs(latd1) * (r1 + a) + fabs(latd2) * (r2 + a)) / fmax(0.000001, fabs(latd1) + fabs(latd2));
        const double r = (latd1 * latd2 >= 0.0) ? res1 : res2;
        const double lat1 = latd1 * kD2R, lon1 = lond1 * kD2R, lat2 = latd2 * kD2R, lon2 = lond2 * kD2R;
        const double coslat1 = cos(lat1), coslat2 = cos(lat2);
        double c = coslat1 * coslat2 * cos(lon2 - lon1) + sin(lat1) * sin(lat2);
        c = fmin(1.0, fmax(-1.0, c));
        dist = r * acos(c);
        qdr = atan2(sin(lon2 - lon1) * coslat2, coslat1 * sin(lat2) - sin(lat1) * coslat2 * cos(lon2 - lon1)) * kR2D;
    } else {
        const float dlat = (float)((latd2 - latd1) * kD2R), dlon = (float)((lond2 - lond1) * kD2R);
        float s1, c1, s2, c2, sm, cm, shl, chl, sho, cho;
        sincosf((float)(latd1 * kD2R), &s1, &c1);
        sincosf((float)(latd2 * kD2R), &s2, &c2);
        sincosf((float)(0.5 * (latd1 + latd2) * kD2R), &sm, &cm);
        sincosf(0.5f * dlat, &shl, &chl);
        sincosf(0.5f * dlon, &sho, &cho);
        const float a = 6378137.0f, b = 6356752.314245f;             // same hemisphere in every env (lat ~ 52..56)
        const float an = a * a * cm, bn = b * b * sm, ad = a * cm, bd = b * sm;
        const float r = sqrtf(fmaf(an, an, bn * bn) / fmaf(ad, ad, bd * bd));
        const float hav = fmaf(c1 * c2, sho * sho, shl * shl);
        dist = 2.0f * r * asinf(fminf(1.0f, sqrtf(hav)));
        const float y = 2.0f * sho * cho * c2, x = fmaf(2.0f * s1 * c2, sho * sho, 2.0f * shl * chl);
        qdr = atan2f(y, x) * float(kR2D);
    }
}
template <typename R> __device__ __forceinline__ R degto180(R a) {
    if constexpr (sizeof(R) == 8) return pymod(a + 180.0, 360.0) - 180.0; else return wrap360(a + 180.0f) - 180.0f;
}
// Autopilot LNAV on the two-waypoint Merge route (oracle/refshim/bluesky/traffic/traffic.py _update_fms): each leg is
// activated with no known next leg, so the turn-anticipation distance is zero and a waypoint counts as passed when
// the aircraft flies away from it within 1.02 turn radii; past the runway LNAV switches off (track kept).
template <typename R>
__device__ __forceinline__ void lnav_update(AcReg<R> &a, const ExtCtx &X) {
    const bool first = a.lnav == 0;
    R qdr, dist;
    lnav_qdrdist<R>(a.lat, a.lon, first ? X.faf_lat : X.rwy_lat, first ? X.faf_lon : X.rwy_lon, qdr, dist);
    R turnrad;
    if constexpr (sizeof(R) == 8) turnrad = a.tas * a.tas / (fmax(0.01, tan(kBank)) * kG0);
    else turnrad = a.tas * a.tas * float(1.0 / (0.46630765815499864 * kG0));            // tan(25 deg)
    R trk = a.hdg;
    if (X.W->n > 0) {
        if constexpr (sizeof(R) == 8) trk = pymod(atan2(a.gse, a.gsn) * kR2D, 360.0);
        else trk = wrap360(atan2f(a.gse, a.gsn) * float(kR2D));
    }
    const bool away = fabs(degto180<R>(mod360<R>(trk) - mod360<R>(qdr))) > R(90);
    if (away && dist < turnrad * R(1.02)) {
        if (!first) a.lnav = -1;
        else { a.lnav = 1; lnav_qdrdist<R>(a.lat, a.lon, X.rwy_lat, X.rwy_lon, qdr, dist); }
    }
    if (a.lnav >= 0) a.aptrk = mod360<R>(qdr);
}

// literal substep (fp64), same operation order as the oracle / numpy code
template <bool EXT = false>
__device__ __forceinline__ void kin_substep_lit(AcReg<double> &a, double dt, bool reso_on, bool last, const ExtCtx *X = nullptr) {
    if constexpr (EXT) { if (X->merge && a.lnav >= 0) lnav_update<double>(a, *X); }
    double ap_tas = vcasormach2tas(a.selspd, a.alt);
    double ap_vs = fabs(a.selvs) > 0.1 ? a.selvs : kVsDef;
    bool act = reso_on && a.asas_active;
    double trk = act ? a.asas_trk : a.aptrk;
    double t_tas = act ? a.asas_tas : ap_tas;
    double t_alt = act ? a.asas_alt : a.selalt;
    double t_vs = fabs(act ? a.asas_vs : ap_vs);
    double t_hdg = pymod(trk, 360.0);
    double vwn = 0.0, vwe = 0.0;
    bool wind = false;
    if constexpr (EXT) wind = X->W->n > 0;
    if (wind) {          // APorASAS.update with a wind field: ASAS ground speed -> TAS, wind-correction angle
        wind_at<double>(*X->W, a.lat, a.lon, a.alt, vwn, vwe);
        if (act) {
            const double tn = a.asas_tas * cos(a.asas_trk * kD2R) - vwn, te = a.asas_tas * sin(a.asas_trk * kD2R) - vwe;
            t_tas = sqrt(tn * tn + te * te);
        }
        const double Vw = sqrt(vwn * vwn + vwe * vwe), winddir = atan2(vwe, vwn);
        const double drift = trk * kD2R - winddir;
        const double steer = asin(fmin(1.0, fmax(-1.0, Vw * sin(drift) / fmax(0.001, a.tas))));
        t_hdg = pymod(trk + steer * kR2D, 360.0);
    }
    double vmin, vmax;
    int ph = flight_phase<double>(a.vs, a.alt);
    a.phase = ph;
    phase_vlimits(ph, vmin, vmax);
    double allow_h = t_alt > kHmax ? kHmax : t_alt;
    double icas = vtas2cas(t_tas, allow_h);
    double acas = icas < vmin ? vmin : icas;
    acas = icas > vmax ? vmax : acas;
    double atas = vcas2tas(acas, allow_h);
    if (atas / vvsound(allow_h) > kMmo) atas = kMmo * vvsound(allow_h);
    double vs_acc = (1 - a.ax / kAxMax) * kVsMax;
    double avs = (t_vs > 0 && t_vs > kVsMax) ? vs_acc : t_vs;
    avs = (t_vs < 0 && t_vs < kVsMin) ? vs_acc : avs;
    double dspd = atas - a.tas;
    bool need_ax = fabs(dspd) > fabs(dt * kAxMax);
    double ax = need_ax * rsgn(dspd) * kAxMax;
    double tas = need_ax ? a.tas + ax * dt : atas;
    a.ax = ax;
    double turnrate = kG0 * tan(kBank) / fmax(tas, 1e-8) * kR2D;
    double delhdg = pymod(t_hdg - a.hdg + 180, 360) - 180;
    bool swhdg = fabs(delhdg) > fabs(dt * turnrate);
    double hdg = pymod(swhdg ? a.hdg + dt * turnrate * rsgn(delhdg) : t_hdg, 360.0);
    double dalt = allow_h - a.alt;
    bool swalt = fabs(dalt) > 1.05 * fmax(fabs(dt * avs), fabs(dt * a.vs));
    double tvs = swalt * rsgn(dalt) * fabs(avs);
    double dvs = tvs - a.vs;
    bool need_az = fabs(dvs) > 300 * kFpm;
    double az = need_az * rsgn(dvs) * (300 * kFpm);
    double vs = need_az ? a.vs + az * dt : tvs;
    if (!isfinite(vs)) vs = 0;
    double gsn = tas * cos(hdg * kD2R), gse = tas * sin(hdg * kD2R);
    if (wind && a.alt > 50.0 * kFt) { gsn = gsn + vwn; gse = gse + vwe; }     // wind acts above 50 ft
    double alt = swalt ? rint((a.alt + vs * dt) * 1e6) / 1e6 : allow_h;
    double lat = a.lat + (dt * gsn / kRe) * kR2D;
    double coslat = cos(lat * kD2R);
    double lon = a.lon + (dt * gse / coslat / kRe) * kR2D;
    if (last) a.cas = vtas2cas(tas, a.alt);      // update_airspeed: cas from the altitude before update_pos
    a.tas = tas; a.hdg = hdg; a.vs = vs; a.alt = alt; a.lat = lat; a.lon = lon; a.gsn = gsn; a.gse = gse;
}

// fast substep (fp32 state, fp64 positions, memoised atmosphere)
template <bool EXT = false>
__device__ __forceinline__ void kin_substep_fast(AcReg<float> &a, float dt, bool reso_on, bool want_cas, const ExtCtx *X = nullptr) {
    if constexpr (EXT) { if (X->merge && a.lnav >= 0) lnav_update<float>(a, *X); }
    if (a.alt != a.c_alt || a.selspd != a.c_spd) {           // autopilot target TAS: pure f(selspd, alt)
        a.c_aptas = (float)vcasormach2tas_fast((double)a.selspd, (double)a.alt);
        a.c_alt = a.alt; a.c_spd = a.selspd;
    }
    float ap_vs = fabsf(a.selvs) > 0.1f ? a.selvs : float(kVsDef);
    bool act = reso_on && a.asas_active;
    float trk = act ? a.asas_trk : a.aptrk;
    float t_tas = act ? a.asas_tas : a.c_aptas;
    float t_vs = fabsf(act ? a.asas_vs : ap_vs);
    float t_hdg = wrap360(trk);
    float vwn = 0.0f, vwe = 0.0f;
    bool wind = false;
    if constexpr (EXT) wind = X->W->n > 0;
    if (wind) {
        wind_at<float>(*X->W, a.lat, a.lon, (double)a.alt, vwn, vwe);
        float st, ct;
        if (act) {
            sincospif(a.asas_trk * (1.0f / 180.0f), &st, &ct);
            const float tn = fmaf(a.asas_tas, ct, -vwn), te = fmaf(a.asas_tas, st, -vwe);
            t_tas = sqrtf(fmaf(tn, tn, te * te));
        }
        // Vw sin(trk - winddir) = sin(trk) vwn - cos(trk) vwe: no inverse tangent of the wind direction
        sincospif(t_hdg * (1.0f / 180.0f), &st, &ct);
        const float cross = fmaf(st, vwn, -(ct * vwe));
        const float steer = asinf(fminf(1.0f, fmaxf(-1.0f, cross / fmaxf(0.001f, a.tas))));
        t_hdg = wrap360(fmaf(steer, float(kR2D), trk));
    }
    int ph = flight_phase<float>(a.vs, a.alt);
    a.phase = ph;
    const float t_alt = act ? a.asas_alt : a.selalt;         // (== selalt under the default, horizontal resolution method)
    float allow_h = fminf(t_alt, float(kHmax));
    if (float(ph) != a.c_ph || t_alt != a.c_selalt) {        // TAS-domain clamp limits: pure f(phase, allow_h)
        double vmin, vmax;
        phase_vlimits(ph, vmin, vmax);
        double lo = vmin > 0.0 ? vcas2tas_fast(vmin, (double)allow_h) : 0.0;
        double hi = fmin(vcas2tas_fast(vmax, (double)allow_h), kMmo * vvsound((double)allow_h));
        a.c_lo = (float)lo; a.c_hi = (float)hi; a.c_ph = float(ph); a.c_selalt = t_alt;
    }
    float atas = fminf(fmaxf(t_tas, a.c_lo), a.c_hi);
    float avs = (t_vs > float(kVsMax)) ? (1.0f - a.ax / float(kAxMax)) * float(kVsMax) : t_vs;
    float dspd = atas - a.tas;
    bool need_ax = fabsf(dspd) > dt * float(kAxMax);
    float ax = need_ax ? copysignf(float(kAxMax), dspd) : 0.0f;
    float tas = need_ax ? fmaf(ax, dt, a.tas) : atas;
    a.ax = ax;
    const float kTurn = float(9.80665 * 0.46630765815499864 * kR2D);   // g0 * tan(25 deg) in deg/s * m/s
    float turn = dt * kTurn * rcp_approx(fmaxf(tas, 1e-8f));
    float delhdg = wrap360_once(t_hdg - a.hdg + 180.0f) - 180.0f;      // both in [0, 360): one wrap
    bool swhdg = fabsf(delhdg) > turn;
    float hdg = wrap360_once(swhdg ? a.hdg + copysignf(turn, delhdg) : t_hdg);   // |turn| < 360: one wrap
    float dalt = allow_h - a.alt;
    bool swalt = fabsf(dalt) > 1.05f * dt * fmaxf(avs, fabsf(a.vs));
    float tvs = swalt ? copysignf(avs, dalt) : 0.0f;
    float dvs = tvs - a.vs;
    bool need_az = fabsf(dvs) > float(300 * kFpm);
    float vs = need_az ? a.vs + copysignf(float(300 * kFpm), dvs) * dt : tvs;
    float s, c;
    sincos_deg_fast(hdg, s, c);                              // hdg in [0, 360): exact quadrant reduction
    float gsn = tas * c, gse = tas * s;
    if (wind && a.alt > float(50.0 * kFt)) { gsn += vwn; gse += vwe; }
    float alt_old = a.alt;
    a.alt = swalt ? fmaf(vs, dt, a.alt) : allow_h;
    a.lat = a.lat + (double)(dt * gsn) * (kR2D / kRe);
    // half-angle pair (sin, cos)(lat/2) for the next CD record and cos(lat) here: exact at the head of every launch
    // (ac_derive), advanced within the step by the first-order rotation (half step 1e-4 rad: the dropped second-order
    // term is 5e-9 relative, below fp32 resolution)
    const float hr = dt * gsn * float(0.5 / kRe);
    const float ca0 = a.ca;
    a.ca = fmaf(-a.sa, hr, ca0); a.sa = fmaf(ca0, hr, a.sa);
    float coslat = fmaf(a.ca, a.ca, -(a.sa * a.sa));
    a.lon = a.lon + (double)(dt * gse * rcp_approx(coslat)) * (kR2D / kRe);
    if (want_cas) a.cas = (float)vtas2cas_fast((double)tas, (double)alt_old);
    a.tas = tas; a.hdg = hdg; a.vs = vs; a.gsn = gsn; a.gse = gse;
}

}  // namespace bsg
