/* ORACLE (test infrastructure, not product code).
 * Per-env ASAS pass = StateBased.detect -> MVP.resolve (if any conflict) -> resumenav,
 * restating oracle/refshim/bluesky/traffic/{asas,mvp}.py.
 * parity unpinned vs upstream bluesky-simulator==1.0.3. */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "bsgo_int.h"

static double bsgo_sign(double x) { return x > 0.0 ? 1.0 : (x < 0.0 ? -1.0 : 0.0); }
static double relm(double a, double b) { double s = fmax(fabs(a), fabs(b)); return s > 0 ? fabs(a - b) / s : 0.0; }

static void mvp_pair(double qdr, double dist, double tcpa, double tLOS,
                     double alt1, double alt2, const double v1[3], const double v2[3],
                     double rpz, double hpz, double dtlook, double fach, double facv,
                     double dv[3], double *tsolV_o, bsgo_trace *tr, int vert) {
    double rpz_m = rpz * fach, hpz_m = hpz * facv;
    double q = qdr * C_D2R;
    double drel[3] = { sin(q) * dist, cos(q) * dist, alt2 - alt1 };
    double vrel[3] = { v2[0] - v1[0], v2[1] - v1[1], v2[2] - v1[2] };
    double dcpa[3] = { drel[0] + vrel[0] * tcpa, drel[1] + vrel[1] * tcpa, drel[2] + vrel[2] * tcpa };
    double dabsH = sqrt(dcpa[0] * dcpa[0] + dcpa[1] * dcpa[1]);
    double iH = rpz_m - dabsH;
    if (tr) {
        const int i = tr->row;
        double vr2 = vrel[0] * vrel[0] + vrel[1] * vrel[1], vr = sqrt(vr2);
        /* Conditioning of this pair's displacement w.r.t. the rounding of its INPUTS.  A precision with unit roundoff
           eps holds positions to eps * Rabs (Rabs: distance of the pair from the env's coordinate origin, the agent) and
           velocities to eps * V (V: the larger ground speed) -- absolute errors that do not shrink with the pair's
           range or relative speed.  First order, with  tcpa = -(r.w)/|w|^2,  dcpa = r + w tcpa,
           dv = (rpz_m / erratum - |dcpa|) dcpa / (|tcpa| |dcpa|):
             d_tcpa = Rabs / vr + V (dist / vr^2 + |tcpa| / vr)                                    [eps s]
             d_dcpa = Rabs + V |tcpa| + vr d_tcpa                                                  [eps m]
             kappa  = (rpz_m / |tcpa|) (d_tcpa / |tcpa| + d_dcpa / |dcpa|)                         [eps m/s]
           Near-parallel pairs inside or near the zone (vr -> 0) and pairs at their CPA (tcpa -> 0) are ill-conditioned
           in any precision; |dv error| <~ eps * kappa. */
        double V = fmax(sqrt(v1[0] * v1[0] + v1[1] * v1[1]), sqrt(v2[0] * v2[0] + v2[1] * v2[1]));
        double Rabs = fmax(tr->rabs, dist), at = fmax(fabs(tcpa), 1e-3), vrc = fmax(vr, 1e-3);
        double d_t = Rabs / vrc + V * (dist / (vrc * vrc) + at / vrc);
        double d_c = Rabs + V * at + vrc * d_t;
        double k = (rpz_m / at) * (d_t / at + d_c / (dabsH <= 10.0 ? dist : dabsH));
        if (tr->kappa && !(k <= tr->kappa[i])) tr->kappa[i] = k;            /* NaN sticks */
        /* thresholds of this function: |dCPA_h| <= 10 m (an ABSOLUTE distance whose fp32 rounding error scales with the
           pair's range, so its margin is taken relative to the zone radius like the detection's own dCPA threshold),
           rpz_m < dist and |dCPA_h| < dist (the erratum branch) */
        double m = fmin(fabs(dabsH - 10.0) / rpz_m, fmin(relm(rpz_m, dist), relm(dabsH, dist)));
        if (vr2 < 1e-4) m = 0.0;                                            /* degenerate pair */
        if (tr->res_margin && m < tr->res_margin[i]) tr->res_margin[i] = m;
    }
    if (dabsH <= 10.0) {
        dabsH = 10.0;
        dcpa[0] = drel[1] / dist * dabsH;
        dcpa[1] = -drel[0] / dist * dabsH;
    }
    if (rpz_m < dist && dabsH < dist) {
        double erratum = cos(asin(rpz_m / dist) - asin(dabsH / dist));
        dv[0] = ((rpz_m / erratum - dabsH) * dcpa[0]) / (fabs(tcpa) * dabsH);
        dv[1] = ((rpz_m / erratum - dabsH) * dcpa[1]) / (fabs(tcpa) * dabsH);
    } else {
        dv[0] = (iH * dcpa[0]) / (fabs(tcpa) * dabsH);
        dv[1] = (iH * dcpa[1]) / (fabs(tcpa) * dabsH);
    }
    double iV, tsolV;
    if (fabs(vrel[2]) > 0.0) { iV = hpz_m; tsolV = fabs(drel[2] / vrel[2]); }
    else { iV = hpz_m - fabs(drel[2]); tsolV = tLOS; }
    if (tr && vert && tr->res_margin) {
        /* thresholds of the vertical term (they steer a resolution only when the method uses it): the |vrel_z| > 0
           switch -- an EXACT-zero test: two aircraft that resolve against each other symmetrically hold vertical speeds
           that agree to the last bit or not depending on the rounding of tLOS(i,j) vs tLOS(j,i), in any precision --
           and tsolV > dtlookahead */
        double m = fabs(vrel[2]) / fmax(fmax(fabs(v1[2]), fabs(v2[2])), 1.0);
        if (fabs(v1[2]) == 0.0 && fabs(v2[2]) == 0.0) m = 1e30;           /* both exactly level: a stable zero */
        m = fmin(m, relm(tsolV, dtlook));
        if (m < tr->res_margin[tr->row]) tr->res_margin[tr->row] = m;
    }
    if (tsolV > dtlook) { tsolV = tLOS; iV = hpz_m; }
    if (fabs(vrel[2]) > 0.0) dv[2] = (iV / tsolV) * (-vrel[2] / fabs(vrel[2]));
    else dv[2] = iV / tsolV;
    *tsolV_o = tsolV;
}

void bsgo_asas_env(const bsgo_cfg *cfg, double *a, uint64_t *reso, size_t wstride, int nlive, bsgo_row *rows) {
    bsgo_asas_env_trace(cfg, a, reso, wstride, nlive, rows, 0);
}

/* resopairs rows: bit j of aircraft i = word (j >> 6) of its row, words `wstride` apart (reso[W][E][N], W = ceil(N/64)) */
void bsgo_asas_env_trace(const bsgo_cfg *cfg, double *a, uint64_t *reso, size_t wstride, int nlive, bsgo_row *rows, bsgo_trace *tr) {
    int n = cfg->n;
    const int W = (n + 63) / 64;
    if (tr) {
        if (tr->near_conf) memset(tr->near_conf, 0, n);
        if (tr->near_los) memset(tr->near_los, 0, n);
    }
#define A(f, i) a[(f) * n + (i)]
    double *u = malloc(sizeof(double) * nlive * 6);
    double *v = u + nlive, *dve = v + nlive, *dvn = dve + nlive, *dvz = dvn + nlive, *tsolmin = dvz + nlive;
    uint64_t *confrow = calloc((size_t)(nlive > 0 ? nlive : 1) * W, sizeof(uint64_t));      /* [nlive][W] */
    int anyconf = 0;
    const int wind = cfg->nwind > 0;
    for (int i = 0; i < nlive; i++) {
        /* gs = tas, trk = hdg (no wind); with a wind field the ground-speed vector of the last update */
        u[i] = wind ? A(F_GSE, i) : A(F_TAS, i) * sin(A(F_HDG, i) * C_D2R);
        v[i] = wind ? A(F_GSN, i) : A(F_TAS, i) * cos(A(F_HDG, i) * C_D2R);
        dve[i] = dvn[i] = dvz[i] = 0.0;
        tsolmin[i] = 1e9;                 /* timesolveV */
    }
    for (int i = 0; i < nlive; i++) {
        bsgo_row *r = &rows[i];
        r->inconf = 0; r->nconf = 0; r->nlos = 0; r->near_idx = -1;
        r->near_tcpa = r->near_dcpa = r->near_tin = r->tcpamax = 0.0;
        /* velocity vectors as MVP sees them: gseast/gsnorth of the last update_groundspeed */
        double v1[3] = { u[i], v[i], A(F_VS, i) };
        for (int j = 0; j < nlive; j++) {
            if (j == i) continue;
            double qdr, dist, tcpa, dcpa2, tin, mc = 1e30, ml = 1e30;
            int sw, los;
            bsgo_pair(A(F_LAT, i), A(F_LON, i), A(F_ALT, i), u[i], v[i], A(F_VS, i),
                      A(F_LAT, j), A(F_LON, j), A(F_ALT, j), u[j], v[j], A(F_VS, j),
                      cfg->rpz, cfg->hpz, cfg->dtlook, &qdr, &dist, &tcpa, &dcpa2, &tin, 0, &sw, &los, tr ? &mc : 0, tr ? &ml : 0);
            if (tr) {
                if (mc < tr->eps && tr->near_conf) tr->near_conf[i] = 1;
                if (ml < tr->eps && tr->near_los) tr->near_los[i] = 1;
                if (cfg->reso_on && tr->res_margin && mc < tr->res_margin[i]) tr->res_margin[i] = mc;
            }
            if (los) r->nlos++;
            if (!sw) continue;
            anyconf = 1;
            r->inconf = 1; r->nconf++;
            if (tcpa > r->tcpamax) r->tcpamax = tcpa;
            if (r->near_idx < 0 || tcpa < r->near_tcpa) {
                r->near_idx = j; r->near_tcpa = tcpa; r->near_dcpa = sqrt(dcpa2); r->near_tin = tin;
            }
            confrow[(size_t)i * W + (j >> 6)] |= (uint64_t)1 << (j & 63);
            if (cfg->reso_on) {
                double v2[3] = { u[j], v[j], A(F_VS, j) };
                double dv[3], tsolV;
                if (tr) {
                    /* range of the pair's far end from the env's coordinate origin (aircraft 0: the fp32 records are relative to it) */
                    double x0 = C_RE * C_D2R * (A(F_LON, i) - A(F_LON, 0)) * cos(C_D2R * A(F_LAT, 0)), y0 = C_RE * C_D2R * (A(F_LAT, i) - A(F_LAT, 0));
                    tr->row = i; tr->rabs = sqrt(x0 * x0 + y0 * y0) + dist;
                }
                mvp_pair(qdr, dist, tcpa, tin, A(F_ALT, i), A(F_ALT, j), v1, v2,
                         cfg->rpz, cfg->hpz, cfg->dtlook, cfg->resofach, cfg->resofacv, dv, &tsolV, tr, cfg->reso_mode != 0);
                if (tsolV < tsolmin[i]) tsolmin[i] = tsolV;
                dv[2] = 0.5 * dv[2];
                dve[i] = dve[i] - dv[0]; dvn[i] = dvn[i] - dv[1]; dvz[i] = dvz[i] - dv[2];
                if (cfg->reso_off_own && i == 0) { dve[i] = 0.0; dvn[i] = 0.0; dvz[i] = 0.0; }
            }
        }
    }
    if (cfg->reso_on) {
        if (anyconf) {
            for (int i = 0; i < nlive; i++) {
                double ge = u[i], gn = v[i];
                double ne = ge + dve[i], nn = gn + dvn[i];
                double newtrack, newgs, newvs;
                if (cfg->reso_mode == 1) {        /* vertical only: ownship.trk / ownship.gs kept */
                    newtrack = wind ? bsgo_pymod(atan2(ge, gn) * 180 / C_PI, 360) : A(F_HDG, i);
                    newgs = wind ? sqrt(ge * ge + gn * gn) : A(F_TAS, i);
                } else {
                    newtrack = bsgo_pymod(atan2(ne, nn) * 180 / C_PI, 360);
                    newgs = sqrt(ne * ne + nn * nn);
                }
                newvs = cfg->reso_mode == 0 ? A(F_VS, i) : A(F_VS, i) + dvz[i];
                double vmin, vmax;
                bsgo_vlimits((int)A(F_PHASE, i), &vmin, &vmax);
                const double vscapped = fmax(ENV_VSMIN, fmin(ENV_VSMAX, newvs));
                A(F_ASAS_TRK, i) = newtrack;
                A(F_ASAS_TAS, i) = fmax(vmin, fmin(vmax, newgs));
                A(F_ASAS_VS, i) = vscapped;
                if (cfg->reso_mode == 0) A(F_ASAS_ALT, i) = A(F_SELALT, i);
                else {
                    /* resolution altitude (tail of MVP.resolve): the autopilot's level-off altitude when it also solves
                       the conflict, else the altitude reached after timesolveV at the capped vertical speed */
                    const double apvs = fabs(A(F_SELVS, i)) > 0.1 ? A(F_SELVS, i) : C_VSDEF;
                    const double sel = A(F_SELALT, i), alt = A(F_ALT, i);
                    const double tmp = vscapped * tsolmin[i] + alt;
                    const double sdv = bsgo_sign(vscapped - apvs * bsgo_sign(sel - alt)), sal = bsgo_sign(tmp - sel);
                    double ra = (sdv == 0.0 || sdv == sal) ? tmp : sel;
                    if (tsolmin[i] < cfg->dtlook && fabs(dvz[i]) > 0.0) ra = tmp;
                    if (tr && tr->res_margin) {       /* the sign tests and the look-ahead test of the altitude choice */
                        double m = fmin(relm(vscapped, apvs * bsgo_sign(sel - alt)), fmin(relm(tmp, sel), relm(tsolmin[i], cfg->dtlook)));
                        if (fabs(sel - alt) < 1e-6 * fmax(fabs(sel), 1.0)) m = 0.0;
                        if (fabs(dvz[i]) > 0.0 && m < tr->res_margin[i]) tr->res_margin[i] = m;
                    }
                    A(F_ASAS_ALT, i) = ra;
                }
            }
        }
        /* resumenav: resopairs |= confpairs; keep a pair while not past CPA, or in horizontal LoS, or bouncing */
        for (int i = 0; i < nlive; i++) {
            uint64_t keep[4] = { 0, 0, 0, 0 };                        /* n <= 256 */
            int touched = 0, active = 0;
            for (int j = 0; j < nlive; j++) {
                const uint64_t row = reso[(size_t)(j >> 6) * wstride + i] | confrow[(size_t)i * W + (j >> 6)];
                if (!((row >> (j & 63)) & 1)) continue;
                touched = 1;
                double dx = C_RE * ((A(F_LON, j) - A(F_LON, i)) * C_D2R * cos(0.5 * ((A(F_LAT, j) + A(F_LAT, i)) * C_D2R)));
                double dy = C_RE * ((A(F_LAT, j) - A(F_LAT, i)) * C_D2R);
                double vx = u[j] - u[i], vy = v[j] - v[i];
                int past_cpa = (dx * vx + dy * vy) > 0.0;
                double hdist = sqrt(dx * dx + dy * dy);
                int hor_los = hdist < cfg->rpz;
                int bouncing = fabs(A(F_HDG, i) - A(F_HDG, j)) < 30.0 && hdist < cfg->rpz * cfg->resofach;
                if (tr) {
                    double vv = sqrt(vx * vx + vy * vy);
                    /* |cos| of (r, v_rel): the past-CPA sign; identical velocity vectors give an exact 0 in any precision */
                    double m = vv > 0.0 ? fabs(dx * vx + dy * vy) / fmax(hdist * vv, 1e-300) : 1e30;
                    m = fmin(m, relm(hdist, cfg->rpz));
                    if (past_cpa && !hor_los) {
                        m = fmin(m, relm(hdist, cfg->rpz * cfg->resofach));
                        if (hdist < cfg->rpz * cfg->resofach * (1 + tr->eps)) m = fmin(m, relm(fabs(A(F_HDG, i) - A(F_HDG, j)), 30.0));
                    }
                    if (tr->res_margin && m < tr->res_margin[i]) tr->res_margin[i] = m;
                }
                if (!past_cpa || hor_los || bouncing) { active = 1; keep[j >> 6] |= (uint64_t)1 << (j & 63); }
            }
            if (touched) A(F_ASAS_ACTIVE, i) = active ? 1.0 : 0.0;
            for (int w = 0; w < W; w++) reso[(size_t)w * wstride + i] = keep[w];
        }
    }
    free(u);
    free(confrow);
#undef A
}
