/* ORACLE (test infrastructure, not product code).
 * fp64 restatement of BlueSky's atmosphere, flat-earth geometry, performance
 * envelope and one traffic substep, in the operation order of oracle/refshim
 * (SURVEY.md A.1-A.3, A.6).  parity unpinned vs upstream bluesky-simulator==1.0.3. */
#include <math.h>
#include <string.h>
#include "bsgo_int.h"

double bsgo_pymod(double a, double b) {
    double r = fmod(a, b);
    if (r != 0.0 && ((r < 0.0) != (b < 0.0))) r += b;
    return r;
}
static double sgn(double x) { return (x > 0.0) - (x < 0.0); }

/* ------------------------------------------------------------------ aero */
static void vatmos(double h, double *p, double *rho, double *T) {
    *T = fmax(288.15 - 0.0065 * h, 216.65);
    double rhotrop = 1.225 * pow(*T / 288.15, 4.256848030018761);
    double dhstrat = fmax(0.0, h - 11000.0);
    *rho = rhotrop * exp(-dhstrat / 6341.552161);
    *p = *rho * C_R * *T;
}
static double vvsound(double h) { return sqrt(C_GAMMA * C_R * fmax(288.15 - 0.0065 * h, 216.65)); }
double bsgo_vtas2cas(double tas, double h) {
    double p, rho, T;
    vatmos(h, &p, &rho, &T);
    double qdyn = p * (pow(1.0 + rho * tas * tas / (7.0 * p), 3.5) - 1.0);
    double cas = sqrt(7.0 * C_P0 / C_RHO0 * (pow(qdyn / C_P0 + 1.0, 2.0 / 7.0) - 1.0));
    return tas < 0 ? -1 * cas : cas;
}
double bsgo_vcas2tas(double cas, double h) {
    double p, rho, T;
    vatmos(h, &p, &rho, &T);
    double qdyn = C_P0 * (pow(1.0 + C_RHO0 * cas * cas / (7.0 * C_P0), 3.5) - 1.0);
    double tas = sqrt(7.0 * p / rho * (pow(1.0 + qdyn / p, 2.0 / 7.0) - 1.0));
    return cas < 0 ? -1 * tas : tas;
}
double bsgo_vcasormach2tas(double spd, double h) {
    return fabs(spd) < 1 ? spd * vvsound(h) : bsgo_vcas2tas(spd, h);
}

/* ------------------------------------------------------------------- geo */
void bsgo_kwikqdrdist(double lata, double lona, double latb, double lonb, double *qdr, double *dist_nm) {
    double dlat = (latb - lata) * C_D2R;
    double dlon = (bsgo_pymod((lonb - lona) + 180, 360) - 180) * C_D2R;
    double cavelat = cos((lata + latb) * C_D2R * 0.5);
    double dangle = sqrt(dlat * dlat + dlon * dlon * cavelat * cavelat);
    *dist_nm = C_RE * dangle / C_NM;
    *qdr = bsgo_pymod(atan2(dlon * cavelat, dlat) * C_R2D, 360.0);
}

static double rwgs84(double latd) {
    double lat = latd * C_D2R, a = 6378137.0, b = 6356752.314245;
    double coslat = cos(lat), sinlat = sin(lat);
    double an = a * a * coslat, bn = b * b * sinlat, ad = a * coslat, bd = b * sinlat;
    return sqrt((an * an + bn * bn) / (ad * ad + bd * bd));
}
/* geo.qdrdist (oracle/refshim/bluesky/tools/geo.py): great circle, WGS-84 radius of the mean latitude */
void bsgo_qdrdist(double latd1, double lond1, double latd2, double lond2, double *qdr, double *dist_nm) {
    double res1 = rwgs84(0.5 * (latd1 + latd2));
    double a = 6378137.0, r1 = rwgs84(latd1), r2 = rwgs84(latd2);
    double res2 = 0.5 * (fabs(latd1) * (r1 + a) + fabs(latd2) * (r2 + a)) / fmax(0.000001, fabs(latd1) + fabs(latd2));
    double r = (latd1 * latd2 >= 0.0) ? res1 : res2;
    double lat1 = latd1 * C_D2R, lon1 = lond1 * C_D2R, lat2 = latd2 * C_D2R, lon2 = lond2 * C_D2R;
    double coslat1 = cos(lat1), coslat2 = cos(lat2);
    double c = coslat1 * coslat2 * cos(lon2 - lon1) + sin(lat1) * sin(lat2);
    c = fmin(1.0, fmax(-1.0, c));
    *dist_nm = r * acos(c) / C_NM;
    *qdr = atan2(sin(lon2 - lon1) * coslat2, coslat1 * sin(lat2) - sin(lat1) * coslat2 * cos(lon2 - lon1)) * C_R2D;
}
/* np.interp(x, xp, fp): piecewise linear, constant beyond the ends */
static double interp1(double x, const double *xp, const double *fp, int n) {
    if (x <= xp[0]) return fp[0];
    if (x >= xp[n - 1]) return fp[n - 1];
    int k = 0;
    while (k < n - 2 && x >= xp[k + 1]) k++;
    return fp[k] + (x - xp[k]) * ((fp[k + 1] - fp[k]) / (xp[k + 1] - xp[k]));
}
/* Windfield.getdata (oracle/refshim/bluesky/traffic/windfield.py).  With an altitude profile (nknot > 0: one point):
 * the knots resampled on the 100 ft axis, linear between the two axis levels around alt. */
void bsgo_wind(const bsgo_cfg *cfg, double lat, double lon, double alt, double *vn, double *ve) {
    *vn = 0.0; *ve = 0.0;
    if (cfg->nwind <= 0) return;
    if (cfg->nknot > 0) {
        const double step = 100.0 * C_FT, top = 450.0 * step;                  /* altaxis = 0 .. 45 000 ft */
        double idx = fmax(0.0, fmin(top - 1e-20, alt)) / step;
        double il = floor(idx), f = idx - il;
        double a0 = il * step, a1 = (il + 1.0) * step;
        *vn = (1.0 - f) * interp1(a0, cfg->walt, cfg->wvn, cfg->nknot) + f * interp1(a1, cfg->walt, cfg->wvn, cfg->nknot);
        *ve = (1.0 - f) * interp1(a0, cfg->walt, cfg->wve, cfg->nknot) + f * interp1(a1, cfg->walt, cfg->wve, cfg->nknot);
        return;
    }
    if (cfg->nwind == 1) { *vn = cfg->wvn[0]; *ve = cfg->wve[0]; return; }
    double w[BSGO_MAXWIND], tot = 0.0;
    for (int k = 0; k < cfg->nwind; k++) {
        double cavelat = cos(0.5 * (lat + cfg->wlat[k]) * C_D2R);
        double dy = (lat - cfg->wlat[k]) * C_D2R;
        double dx = cavelat * ((bsgo_pymod((lon - cfg->wlon[k]) + 180.0, 360.0) - 180.0) * C_D2R);
        w[k] = 1.0 / (1e-20 + dx * dx + dy * dy);
        tot += w[k];
    }
    for (int k = 0; k < cfg->nwind; k++) { *vn += w[k] / tot * cfg->wvn[k]; *ve += w[k] / tot * cfg->wve[k]; }
}
static double degto180(double a) { return bsgo_pymod(a + 180.0, 360.0) - 180.0; }

/* ------------------------------------------------------------------ perf */
int bsgo_phase(double tas, double vs, double alt) {
    double roc = vs / C_FPM, h = alt / C_FT;
    int ph = PH_NA;
    (void)tas;
    if (h <= 75.0) ph = PH_GD;
    if (h > 75.0 && h <= 1000.0 && roc >= 150.0) ph = PH_IC;
    if (h > 75.0 && h <= 1000.0 && roc <= -150.0) ph = PH_AP;
    if (h > 1000.0 && roc >= 150.0) ph = PH_CL;
    if (h > 1000.0 && roc <= -150.0) ph = PH_DE;
    if (h > 10000.0 && roc > -150.0 && roc < 150.0) ph = PH_CR;
    return ph;
}
void bsgo_vlimits(int ph, double *vmin, double *vmax) {
    int enroute = (ph == PH_CL) || (ph == PH_CR) || (ph == PH_DE);
    double lo = 0.0, hi = ENV_VMAXER;
    if (ph == PH_IC) lo = ENV_VMINIC;
    if (enroute) lo = ENV_VMINER;
    if (ph == PH_AP) lo = ENV_VMINAP;
    if (ph == PH_IC || ph == PH_GD) hi = ENV_VMAXIC;
    if (ph == PH_AP) hi = ENV_VMAXAP;
    *vmin = lo;
    *vmax = hi;
}

/* One substep for aircraft i of one env, AFTER asas outputs (F_ASAS_*) are final.
 * a = field-major block [BSGO_NF][n]. */
void bsgo_kin_substep(const bsgo_cfg *cfg, double *a, const double *ev, int n, int i, double dt, int reso_on) {
#define A(f) a[(f) * n + i]
    const int wind = cfg->nwind > 0;
    /* LNAV (Autopilot._update_fms of the shim): the two-waypoint Merge route FAF -> runway.  Each leg is activated
     * with no known next leg (next_qdr = -999), so the turn-anticipation distance is zero and a waypoint counts
     * as passed when the aircraft flies away from it within 1.02 turn radii. */
    if (cfg->kind == KIND_M && A(F_LNAV) >= 0.0) {
        int w = (int)A(F_LNAV);
        double wlat = w == 0 ? ev[EV_WPT_LAT] : ev[EV_WPT_X], wlon = w == 0 ? ev[EV_WPT_LON] : ev[EV_WPT_Y];
        double qdr, dnm;
        bsgo_qdrdist(A(F_LAT), A(F_LON), wlat, wlon, &qdr, &dnm);
        double turnrad = A(F_TAS) * A(F_TAS) / (fmax(0.01, tan(C_BANK)) * C_G0);
        double trk = wind ? bsgo_pymod(atan2(A(F_GSE), A(F_GSN)) * C_R2D, 360.0) : A(F_HDG);
        int away = fabs(degto180(bsgo_pymod(trk, 360.0) - bsgo_pymod(qdr, 360.0))) > 90.0;
        if (away && dnm * C_NM < turnrad * 1.02) {
            if (w >= 1) A(F_LNAV) = -1.0;
            else {
                A(F_LNAV) = 1.0;
                bsgo_qdrdist(A(F_LAT), A(F_LON), ev[EV_WPT_X], ev[EV_WPT_Y], &qdr, &dnm);
            }
        }
        if (A(F_LNAV) >= 0.0) A(F_APTRK) = bsgo_pymod(qdr, 360.0);
    }
    /* autopilot targets (VNAV off) */
    double ap_tas = bsgo_vcasormach2tas(A(F_SELSPD), A(F_ALT));
    double ap_alt = A(F_SELALT);
    double ap_vs = fabs(A(F_SELVS)) > 0.1 ? A(F_SELVS) : C_VSDEF;
    double ap_trk = A(F_APTRK);
    /* AP or ASAS */
    int act = reso_on && A(F_ASAS_ACTIVE) != 0.0;
    double trk = act ? A(F_ASAS_TRK) : ap_trk;
    double t_tas = act ? A(F_ASAS_TAS) : ap_tas;
    double t_alt = act ? A(F_ASAS_ALT) : ap_alt;
    double t_vs = fabs(act ? A(F_ASAS_VS) : ap_vs);
    double t_hdg = bsgo_pymod(trk, 360.0);
    double vwn = 0.0, vwe = 0.0;
    if (wind) {          /* APorASAS.update with a wind field: ASAS ground speed -> TAS, wind-correction angle */
        bsgo_wind(cfg, A(F_LAT), A(F_LON), A(F_ALT), &vwn, &vwe);
        if (act) {
            double tn = A(F_ASAS_TAS) * cos(A(F_ASAS_TRK) * C_D2R) - vwn, te = A(F_ASAS_TAS) * sin(A(F_ASAS_TRK) * C_D2R) - vwe;
            t_tas = sqrt(tn * tn + te * te);
        }
        double Vw = sqrt(vwn * vwn + vwe * vwe), winddir = atan2(vwe, vwn);
        double drift = trk * C_D2R - winddir;
        double steer = asin(fmin(1.0, fmax(-1.0, Vw * sin(drift) / fmax(0.001, A(F_TAS)))));
        t_hdg = bsgo_pymod(trk + steer * C_R2D, 360.0);
    }
    /* perf.update + limits */
    double vmin, vmax;
    int ph = bsgo_phase(A(F_TAS), A(F_VS), A(F_ALT));
    A(F_PHASE) = (double)ph;
    bsgo_vlimits(ph, &vmin, &vmax);
    double allow_h = t_alt > ENV_HMAX ? ENV_HMAX : t_alt;
    double icas = bsgo_vtas2cas(t_tas, allow_h);
    double acas = icas < vmin ? vmin : icas;
    acas = icas > vmax ? vmax : acas;
    double atas = bsgo_vcas2tas(acas, allow_h);
    if (atas / vvsound(allow_h) > ENV_MMO) atas = ENV_MMO * vvsound(allow_h);
    double vs_acc = (1 - A(F_AX) / ENV_AXMAX) * ENV_VSMAX;
    double avs = (t_vs > 0 && t_vs > ENV_VSMAX) ? vs_acc : t_vs;
    avs = (t_vs < 0 && t_vs < ENV_VSMIN) ? vs_acc : avs;
    /* update_airspeed */
    double dspd = atas - A(F_TAS);
    int need_ax = fabs(dspd) > fabs(dt * ENV_AXMAX);
    double ax = need_ax * sgn(dspd) * ENV_AXMAX;
    double tas = need_ax ? A(F_TAS) + ax * dt : atas;
    A(F_AX) = ax;
    double turnrate = C_G0 * tan(C_BANK) / fmax(tas, 1e-8) * C_R2D;
    double delhdg = bsgo_pymod(t_hdg - A(F_HDG) + 180, 360) - 180;
    int swhdg = fabs(delhdg) > fabs(dt * turnrate);
    double hdg = bsgo_pymod(swhdg ? A(F_HDG) + dt * turnrate * sgn(delhdg) : t_hdg, 360.0);
    double dalt = allow_h - A(F_ALT);
    int swalt = fabs(dalt) > 1.05 * fmax(fabs(dt * avs), fabs(dt * A(F_VS)));
    double tvs = swalt * sgn(dalt) * fabs(avs);
    double dvs = tvs - A(F_VS);
    int need_az = fabs(dvs) > 300 * C_FPM;
    double az = need_az * sgn(dvs) * (300 * C_FPM);
    double vs = need_az ? A(F_VS) + az * dt : tvs;
    if (!isfinite(vs)) vs = 0;
    /* update_groundspeed + update_pos */
    double gsn = tas * cos(hdg * C_D2R), gse = tas * sin(hdg * C_D2R);
    if (wind && A(F_ALT) > 50.0 * C_FT) { gsn = gsn + vwn; gse = gse + vwe; }    /* wind acts above 50 ft */
    double alt = swalt ? rint((A(F_ALT) + vs * dt) * 1e6) / 1e6 : allow_h;
    double lat = A(F_LAT) + (dt * gsn / C_RE) * C_R2D;
    double coslat = cos(lat * C_D2R);
    double lon = A(F_LON) + (dt * gse / coslat / C_RE) * C_R2D;
    A(F_CAS) = bsgo_vtas2cas(tas, A(F_ALT)); /* update_airspeed: cas from the altitude before update_pos */
    A(F_TAS) = tas; A(F_HDG) = hdg; A(F_VS) = vs; A(F_ALT) = alt; A(F_LAT) = lat; A(F_LON) = lon;
    A(F_GSN) = gsn; A(F_GSE) = gse;
#undef A
}
