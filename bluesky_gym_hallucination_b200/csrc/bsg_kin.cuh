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
    R gsn, gse;                                            // derived: tas*cos(hdg), tas*sin(hdg)
    R sa, ca;                                              // fast path: sin, cos of lat/2 (CD record + cos(lat))
};

template <typename R, bool RESO = true>
__device__ __forceinline__ void ac_load_raw(AcReg<R> &a, const double *pos, const R *st, size_t EN, size_t k) {
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
template <typename R> __device__ __forceinline__ void ac_derive(AcReg<R> &a) {
    R s, c;
    if constexpr (sizeof(R) == 8) { s = sin(a.hdg * kD2R); c = cos(a.hdg * kD2R); }
    else { sincospif(a.hdg * (1.0f / 180.0f), &s, &c); sincospif((float)(a.lat * (1.0 / 360.0)), &a.sa, &a.ca); }
    a.gsn = a.tas * c; a.gse = a.tas * s;
}

template <typename R, bool RESO = true>
__device__ __forceinline__ void ac_store(const AcReg<R> &a, double *pos, R *st, size_t EN, size_t k) {
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

// literal substep (fp64), same operation order as the oracle / numpy code
__device__ __forceinline__ void kin_substep_lit(AcReg<double> &a, double dt, bool reso_on, bool last) {
    double ap_tas = vcasormach2tas(a.selspd, a.alt);
    double ap_vs = fabs(a.selvs) > 0.1 ? a.selvs : kVsDef;
    bool act = reso_on && a.asas_active;
    double trk = act ? a.asas_trk : a.aptrk;
    double t_tas = act ? a.asas_tas : ap_tas;
    double t_alt = act ? a.asas_alt : a.selalt;
    double t_vs = fabs(act ? a.asas_vs : ap_vs);
    double t_hdg = pymod(trk, 360.0);
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
    double alt = swalt ? rint((a.alt + vs * dt) * 1e6) / 1e6 : allow_h;
    double lat = a.lat + (dt * gsn / kRe) * kR2D;
    double coslat = cos(lat * kD2R);
    double lon = a.lon + (dt * gse / coslat / kRe) * kR2D;
    if (last) a.cas = vtas2cas(tas, a.alt);      // update_airspeed: cas from the altitude before update_pos
    a.tas = tas; a.hdg = hdg; a.vs = vs; a.alt = alt; a.lat = lat; a.lon = lon; a.gsn = gsn; a.gse = gse;
}

// fast substep (fp32 state, fp64 positions, memoised atmosphere)
__device__ __forceinline__ void kin_substep_fast(AcReg<float> &a, float dt, bool reso_on, bool want_cas) {
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
    int ph = flight_phase<float>(a.vs, a.alt);
    a.phase = ph;
    float allow_h = fminf(a.selalt, float(kHmax));           // asas alt == selalt for horizontal resolutions
    if (float(ph) != a.c_ph || a.selalt != a.c_selalt) {     // TAS-domain clamp limits: pure f(phase, allow_h)
        double vmin, vmax;
        phase_vlimits(ph, vmin, vmax);
        double lo = vmin > 0.0 ? vcas2tas_fast(vmin, (double)allow_h) : 0.0;
        double hi = fmin(vcas2tas_fast(vmax, (double)allow_h), kMmo * vvsound((double)allow_h));
        a.c_lo = (float)lo; a.c_hi = (float)hi; a.c_ph = float(ph); a.c_selalt = a.selalt;
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
    float delhdg = wrap360(t_hdg - a.hdg + 180.0f) - 180.0f;
    bool swhdg = fabsf(delhdg) > turn;
    float hdg = wrap360(swhdg ? a.hdg + copysignf(turn, delhdg) : t_hdg);
    float dalt = allow_h - a.alt;
    bool swalt = fabsf(dalt) > 1.05f * dt * fmaxf(avs, fabsf(a.vs));
    float tvs = swalt ? copysignf(avs, dalt) : 0.0f;
    float dvs = tvs - a.vs;
    bool need_az = fabsf(dvs) > float(300 * kFpm);
    float vs = need_az ? a.vs + copysignf(float(300 * kFpm), dvs) * dt : tvs;
    float s, c;
    sincospif(hdg * (1.0f / 180.0f), &s, &c);                // hdg in [0, 360): exact range reduction in sincospi
    float gsn = tas * c, gse = tas * s;
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
