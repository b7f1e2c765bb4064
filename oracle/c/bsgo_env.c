/* ORACLE (test infrastructure, not product code).
 * fp64 restatement of the reference env layer, generalised from 5 intruders to n-1:
 *   Horizontal: /root/reference/bluesky_gym/envs/horizontal_cr_env.py:150-275
 *   Vertical  : /root/reference/bluesky_gym/envs/vertical_cr_env.py:113-256
 *   Sector    : /root/reference/bluesky_gym/envs/sector_cr_env.py:134-139,227-339
 *   Merge     : /root/reference/bluesky_gym/envs/merge_env.py:160-293 (19 intruders, the 5 nearest observed)
 *   wind obs  : /root/reference/bluesky_gym/wrappers/wind.py:55-64
 *   helpers   : /root/reference/bluesky_gym/envs/common/functions.py:4-22,98-132
 * Pinned by the tests/golden npz fixtures (unmodified reference classes run on oracle/refshim). */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "bsgo_int.h"

#define NM2KM 1.852
#define AC_SPD 150.0
#define CENTER_LAT 51.990426702297746
#define CENTER_LON 4.376124857109851

int bsgo_obs_dim(int kind, int n) {
    int k = n - 1;
    if (kind == KIND_H) return 5 * k + 3;
    if (kind == KIND_V) return 4 + 7 * k;
    if (kind == KIND_S) return 3 + 7 * 4;
    if (kind == KIND_M) return 5 + 7 * 5;
    return 0;
}
int bsgo_obs_dim_cfg(const bsgo_cfg *c) { return bsgo_obs_dim(c->kind, c->n) + (c->wind_obs ? 2 : 0); }
/* AP.selhdgcmd: the HDG argument becomes the track it gives in the current wind (above 50 ft) */
static double hdg_cmd(const bsgo_cfg *c, const double *a, double hdg) {
    int n = c->n;
    if (c->nwind > 0 && a[F_ALT * n] > 50.0 * C_FT) {
        double vn, ve;
        bsgo_wind(c, a[F_LAT * n], a[F_LON * n], 0.0, &vn, &ve);     /* scalar numpy altitude: the field at altitude 0 (windfield.py) */
        double gn = a[F_TAS * n] * cos(hdg * C_D2R) + vn, ge = a[F_TAS * n] * sin(hdg * C_D2R) + ve;
        return bsgo_pymod(atan2(ge, gn) * C_R2D, 360.0);
    }
    return hdg;
}
static double bound180(double a) {  /* functions.py:17-22: wraps once only */
    if (a > 180) return -(360 - a);
    if (a < -180) return 360 + a;
    return a;
}
static void ll2nm(double lat, double lon, double *x, double *y) {  /* functions.py:130-131 */
    *x = (lat - CENTER_LAT) * 60;
    *y = (lon - CENTER_LON) * 60 * cos(CENTER_LAT * C_D2R);
}
static int poly_inside(const double *poly, int nv, double tx, double ty) {
    int inside = 0;
    double x0 = poly[2 * (nv - 1)], y0 = poly[2 * (nv - 1) + 1];
    int f0 = y0 >= ty;
    for (int k = 0; k < nv; k++) {
        double x1 = poly[2 * k], y1 = poly[2 * k + 1];
        int f1 = y1 >= ty;
        if (f0 != f1)
            if ((((y1 - ty) * (x0 - x1)) >= ((x1 - tx) * (y0 - y1))) == f1) inside = !inside;
        f0 = f1; x0 = x1; y0 = y1;
    }
    return inside;
}

#define A(f, i) a[(f) * n + (i)]
/* ------------------------------------------------------------- actions */
static void apply_action(const bsgo_cfg *c, double *a, double *ev, const double *act) {
    int n = c->n;
    if (c->kind == KIND_H) {            /* horizontal_cr_env.py:272-275: HDG from the cached ac_hdg */
        A(F_APTRK, 0) = hdg_cmd(c, a, ev[EV_AC_HDG] + act[0] * 45);
    } else if (c->kind == KIND_V) {     /* vertical_cr_env.py:242-256 */
        double v = act[0] * 12.5;
        A(F_SELALT, 0) = v >= 0 ? 1000000 : 0;
        A(F_SELVS, 0) = v;
    } else if (c->kind == KIND_S) {     /* sector_cr_env.py:315-322 (SPD arg in kts, parsed back with kts) */
        double dh = act[0] * 22.5, dv = act[1] * (20.0 / 3);
        A(F_APTRK, 0) = hdg_cmd(c, a, bound180(A(F_HDG, 0) + dh));
        double spd_kts = (A(F_CAS, 0) + dv) * 1.94384;
        A(F_SELSPD, 0) = fabs(spd_kts) > 1.0 ? spd_kts * C_KTS : spd_kts;
    } else if (c->kind == KIND_M) {     /* merge_env.py:285-293; the HDG command switches LNAV off */
        double dh = act[0] * 15, dv = act[1] * 20;
        A(F_APTRK, 0) = hdg_cmd(c, a, bound180(A(F_HDG, 0) + dh));
        A(F_LNAV, 0) = -1.0;
        double spd_kts = (A(F_CAS, 0) + dv) * 1.94384;
        A(F_SELSPD, 0) = fabs(spd_kts) > 1.0 ? spd_kts * C_KTS : spd_kts;
    }
}

/* ------------------------------------------------------ obs (+ drift cache) */
static double observe_env(const bsgo_cfg *c, const double *a, int nlive, double *ev, double *obs) {
    int n = c->n, K = n - 1;
    double drift = 0.0;
    if (c->kind == KIND_H || c->kind == KIND_V) {
        double *o_dist, *o_cos, *o_sin, *o_dalt = 0, *o_x, *o_y, *o_z = 0;
        double hdg0 = A(F_HDG, 0);
        if (c->kind == KIND_H) {
            o_dist = obs; o_cos = obs + K; o_sin = obs + 2 * K; o_x = obs + 3 * K; o_y = obs + 4 * K;
            ev[EV_AC_HDG] = hdg0;
        } else {
            double *b = obs + 4;
            o_dist = b; o_cos = b + K; o_sin = b + 2 * K; o_dalt = b + 3 * K; o_x = b + 4 * K; o_y = b + 5 * K; o_z = b + 6 * K;
        }
        double norm = c->kind == KIND_H ? 150.0 : 200.0;
        for (int k = 0; k < K; k++) {
            int j = k + 1;
            if (j >= nlive) { o_dist[k] = o_cos[k] = o_sin[k] = o_x[k] = o_y[k] = 0; if (o_dalt) o_dalt[k] = o_z[k] = 0; continue; }
            double qdr, d;
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), A(F_LAT, j), A(F_LON, j), &qdr, &d);
            o_dist[k] = d * NM2KM / norm;
            double brg = bound180(hdg0 - qdr);
            o_cos[k] = cos(brg * C_D2R);
            o_sin[k] = sin(brg * C_D2R);
            double hd = hdg0 - A(F_HDG, j);
            double gsj = A(F_TAS, j), gs0 = A(F_TAS, 0);                  /* bs.traf.gs: the ground speed under wind */
            if (c->nwind > 0) {
                gsj = sqrt(A(F_GSN, j) * A(F_GSN, j) + A(F_GSE, j) * A(F_GSE, j));
                gs0 = sqrt(A(F_GSN, 0) * A(F_GSN, 0) + A(F_GSE, 0) * A(F_GSE, 0));
            }
            o_x[k] = (-cos(hd * C_D2R) * gsj) / AC_SPD;
            o_y[k] = (gs0 - sin(hd * C_D2R) * gsj) / AC_SPD;
            if (o_dalt) { o_dalt[k] = (A(F_ALT, j) - A(F_ALT, 0)) / 3000.0; o_z[k] = A(F_VS, j) - A(F_VS, 0); }
        }
        if (c->kind == KIND_H) {
            double qdr, d;
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), ev[EV_WPT_LAT], ev[EV_WPT_LON], &qdr, &d);
            obs[5 * K] = d * NM2KM / 150.0;
            drift = bound180(hdg0 - qdr);
            obs[5 * K + 1] = cos(drift * C_D2R);
            obs[5 * K + 2] = sin(drift * C_D2R);
        } else {
            double qdr, d;
            bsgo_kwikqdrdist(52, 4, A(F_LAT, 0), A(F_LON, 0), &qdr, &d);
            double rwy = 200 - d * NM2KM;
            obs[0] = (A(F_ALT, 0) - 1500) / 3000.0;
            obs[1] = (A(F_VS, 0) - 0) / 5.0;
            obs[2] = (ev[EV_TARGET_ALT] - 1500) / 3000.0;
            obs[3] = (rwy - 100) / 200.0;
        }
    } else if (c->kind == KIND_S) {
        /* own waypoint: nm_to_latlong(CENTER, wpts[0]) (functions.py:112-113) */
        double wlat = CENTER_LAT + ev[EV_WPT_X] / 60;
        double wlon = CENTER_LON + ev[EV_WPT_Y] / (60 * cos(CENTER_LAT * C_D2R));
        double qdr, d, hdg0 = A(F_HDG, 0);
        bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), wlat, wlon, &qdr, &d);
        drift = bound180(hdg0 - qdr);
        obs[0] = cos(drift * C_D2R);
        obs[1] = sin(drift * C_D2R);
        obs[2] = (A(F_TAS, 0) - 150) / 6;
        double vx = cos(hdg0 * C_D2R) * A(F_TAS, 0), vy = sin(hdg0 * C_D2R) * A(F_TAS, 0);
        double x0, y0;
        ll2nm(A(F_LAT, 0), A(F_LON, 0), &x0, &y0);
        x0 = x0 * NM2KM * 1000; y0 = y0 * NM2KM * 1000;
        double *dist = malloc(sizeof(double) * (nlive > 1 ? nlive : 1));
        for (int j = 1; j < nlive; j++) {
            double x, y;
            ll2nm(A(F_LAT, j), A(F_LON, j), &x, &y);
            x = x * NM2KM * 1000; y = y * NM2KM * 1000;
            double ddx = x - x0, ddy = y - y0;
            dist[j] = sqrt(ddx * ddx + ddy * ddy);   /* functions.py:148: sqrt(sum((p2-p1)**2)) */
        }
        for (int s = 0; s < 4; s++) {                /* rank by (distance, index): the 4 nearest */
            int best = -1;
            for (int j = 1; j < nlive; j++) {
                int rank = 0;
                for (int m = 1; m < nlive; m++)
                    if (dist[m] < dist[j] || (dist[m] == dist[j] && m < j)) rank++;
                if (rank == s) { best = j; break; }
            }
            double *o = obs + 3;
            if (best < 0) { for (int q = 0; q < 7; q++) o[q * 4 + s] = 0; continue; }
            double x, y;
            ll2nm(A(F_LAT, best), A(F_LON, best), &x, &y);
            x = x * NM2KM * 1000; y = y * NM2KM * 1000;
            double vxi = cos(A(F_HDG, best) * C_D2R) * A(F_TAS, best), vyi = sin(A(F_HDG, best) * C_D2R) * A(F_TAS, best);
            double trk = atan2(vyi - vy, vxi - vx);
            o[0 * 4 + s] = (x - x0) / 13000;
            o[1 * 4 + s] = (y - y0) / 13000;
            o[2 * 4 + s] = (vxi - vx) / 32;
            o[3 * 4 + s] = (vyi - vy) / 66;
            o[4 * 4 + s] = cos(trk);
            o[5 * 4 + s] = sin(trk);
            o[6 * 4 + s] = (dist[best] - 50000.) / 15000.;
        }
        free(dist);
    } else if (c->kind == KIND_M) {     /* merge_env.py:160-236 */
        int w = ev[EV_WPT_REACH] == 0.0 ? 0 : 1;
        double wlat = w == 0 ? ev[EV_WPT_LAT] : ev[EV_WPT_X], wlon = w == 0 ? ev[EV_WPT_LON] : ev[EV_WPT_Y];
        double qdr, d, hdg0 = A(F_HDG, 0);
        bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), wlat, wlon, &qdr, &d);
        drift = bound180(hdg0 - qdr);
        obs[0] = cos(drift * C_D2R);
        obs[1] = sin(drift * C_D2R);
        obs[2] = A(F_TAS, 0);
        obs[3] = d / 250;
        obs[4] = ev[EV_WPT_REACH];
        double vx = cos(hdg0 * C_D2R) * A(F_TAS, 0), vy = sin(hdg0 * C_D2R) * A(F_TAS, 0);
        double *dist = malloc(sizeof(double) * (nlive > 1 ? nlive : 1));
        for (int j = 1; j < nlive; j++) {
            double q2;
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), A(F_LAT, j), A(F_LON, j), &q2, &dist[j]);
        }
        for (int s = 0; s < 5; s++) {                /* stable argsort: rank by (distance, index) */
            int best = -1;
            for (int j = 1; j < nlive; j++) {
                int rank = 0;
                for (int m = 1; m < nlive; m++)
                    if (dist[m] < dist[j] || (dist[m] == dist[j] && m < j)) rank++;
                if (rank == s) { best = j; break; }
            }
            double *o = obs + 5;
            if (best < 0) { for (int q = 0; q < 7; q++) o[q * 5 + s] = 0; continue; }
            double brg, dd;
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), A(F_LAT, best), A(F_LON, best), &brg, &dd);
            double vxi = cos(A(F_HDG, best) * C_D2R) * A(F_TAS, best), vyi = sin(A(F_HDG, best) * C_D2R) * A(F_TAS, best);
            double trk = atan2(vyi - vy, vxi - vx);
            o[0 * 5 + s] = (dd * NM2KM * 1000) * cos(brg * C_D2R) / 1000000;
            o[1 * 5 + s] = (dd * NM2KM * 1000) * sin(brg * C_D2R) / 1000000;
            o[2 * 5 + s] = (vxi - vx) / 150;
            o[3 * 5 + s] = (vyi - vy) / 150;
            o[4 * 5 + s] = cos(trk);
            o[5 * 5 + s] = sin(trk);
            o[6 * 5 + s] = dist[best] / 250;
        }
        free(dist);
    }
    if (c->wind_obs) {                  /* wrappers/wind.py:55-64: wind in the agent's body axes / MAX_WIND */
        double *o = obs + bsgo_obs_dim(c->kind, c->n), wn, we, h = A(F_HDG, 0) * C_D2R;
        bsgo_wind(c, A(F_LAT, 0), A(F_LON, 0), 0.0, &wn, &we);         /* (scalar numpy altitude: altitude 0, as in hdg_cmd) */
        o[0] = (wn * cos(h) + we * sin(h)) / 50;
        o[1] = (-wn * sin(h) + we * cos(h)) / 50;
    }
    return drift;
}

/* ------------------------------------------------ reward / termination */
static void reward_env(const bsgo_cfg *c, const double *a, int nlive, double *ev, const double *obs,
                       double drift, double *reward, int32_t *term, int32_t *trunc) {
    int n = c->n, K = n - 1;
    *term = 0; *trunc = 0;
    if (c->kind == KIND_H) {            /* horizontal_cr_env.py:225-270 */
        double reach = 0;
        double wdist_km = obs[5 * K] * 150.0;    /* not used for the test: recompute exactly below */
        double qdr, d;
        bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), ev[EV_WPT_LAT], ev[EV_WPT_LON], &qdr, &d);
        wdist_km = d * NM2KM;
        if (wdist_km < 5 && ev[EV_WPT_REACH] != 1) { ev[EV_WPT_REACH] = 1; reach += 1; }
        double dr = fabs(drift * C_D2R);
        ev[EV_DRIFT_SUM] += dr; ev[EV_DRIFT_CNT] += 1;
        double intr = 0;
        for (int j = 1; j < nlive; j++) {
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), A(F_LAT, j), A(F_LON, j), &qdr, &d);
            if (d < 5) { ev[EV_TOTAL_INTR] += 1; intr += -1; }
        }
        *reward = reach + dr * -0.1 + intr;
        ev[EV_TOTAL_REWARD] += *reward;
        *term = ev[EV_WPT_REACH] == 1 ? 1 : 0;
    } else if (c->kind == KIND_V) {     /* vertical_cr_env.py:213-240 */
        double qdr, d, intr = 0;
        for (int j = 1; j < nlive; j++) {
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), A(F_LAT, j), A(F_LON, j), &qdr, &d);
            double vert = A(F_ALT, 0) - A(F_ALT, j);
            if (d < 5 && fabs(vert) < 1000 * 0.3048) { ev[EV_TOTAL_INTR] += 1; intr += -50; }
        }
        bsgo_kwikqdrdist(52, 4, A(F_LAT, 0), A(F_LON, 0), &qdr, &d);
        double rwy = 200 - d * NM2KM, alt = A(F_ALT, 0), pen = 0;
        if (rwy > 0 && alt > 0) pen = fabs(ev[EV_TARGET_ALT] - alt) * (-5.0 / 3000);
        else if (alt <= 0) { pen = -100; ev[EV_FINAL_ALT] = -100; *term = 1; }
        else if (rwy <= 0) { pen = alt * (-50.0 / 3000); ev[EV_FINAL_ALT] = alt; *term = 1; }
        *reward = pen + intr;
        ev[EV_TOTAL_REWARD] += *reward;
    } else if (c->kind == KIND_S) {     /* sector_cr_env.py:227-235,324-339,134-139 */
        double dr = fabs(drift * C_D2R);
        ev[EV_DRIFT_SUM] += dr; ev[EV_DRIFT_CNT] += 1;
        double qdr, d, intr = 0;
        for (int j = 1; j < nlive; j++) {
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), A(F_LAT, j), A(F_LON, j), &qdr, &d);
            if (d < 5) { ev[EV_TOTAL_INTR] += 1; intr += -1; }
        }
        *reward = dr * -0.1 + intr;
        ev[EV_TOTAL_REWARD] += *reward;
        *trunc = poly_inside(ev + EV_POLY, (int)ev[EV_NVERT], A(F_LAT, 0), A(F_LON, 0)) ? 0 : 1;
    } else if (c->kind == KIND_M) {     /* merge_env.py:238-283 */
        double reach = 0, wdist = obs[3] * 250;
        double qdr, d;
        {   /* self.waypoint_dist of the _get_obs that just ran: recompute exactly */
            int w = ev[EV_WPT_REACH] == 0.0 ? 0 : 1;
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), w == 0 ? ev[EV_WPT_LAT] : ev[EV_WPT_X],
                             w == 0 ? ev[EV_WPT_LON] : ev[EV_WPT_Y], &qdr, &wdist);
        }
        if (wdist < 10 && ev[EV_WPT_REACH] != 1) { ev[EV_WPT_REACH] = 1; ev[EV_FINAL_ALT] = 1; reach += 1; }
        else if (wdist < 20 && ev[EV_WPT_REACH] == 1) { ev[EV_FINAL_ALT] = 2; *term = 1; }
        double dr = fabs(drift * C_D2R);
        ev[EV_DRIFT_SUM] += dr; ev[EV_DRIFT_CNT] += 1;
        double intr = 0;
        for (int j = 1; j < nlive; j++) {
            bsgo_kwikqdrdist(A(F_LAT, 0), A(F_LON, 0), A(F_LAT, j), A(F_LON, j), &qdr, &d);
            if (d < 4) { ev[EV_TOTAL_INTR] += 1; intr += -1; }
        }
        *reward = reach + dr * -0.1 + intr;
        ev[EV_TOTAL_REWARD] += *reward;
    } else {
        *reward = 0;
    }
}
#undef A

void bsgo_observe(const bsgo_cfg *cfg, int E, const double *ac, const int32_t *nac, double *ev, double *obs) {
    int od = bsgo_obs_dim_cfg(cfg);
    for (int e = 0; e < E; e++)
        observe_env(cfg, ac + (size_t)e * BSGO_NF * cfg->n, nac[e], ev + (size_t)e * BSGO_NE, obs + (size_t)e * od);
}

void bsgo_step(const bsgo_cfg *cfg, int E, double *ac, uint64_t *reso, const int32_t *nac, double *ev,
               const double *actions, double *obs, double *reward, int32_t *term, int32_t *trunc,
               bsgo_cdout cd, int nthreads) {
    bsgo_step_trace(cfg, E, ac, reso, nac, ev, actions, obs, reward, term, trunc, cd, 0.0, 0, 0, 0, 0, nthreads);
}

void bsgo_step_trace(const bsgo_cfg *cfg, int E, double *ac, uint64_t *reso, const int32_t *nac, double *ev,
                     const double *actions, double *obs, double *reward, int32_t *term, int32_t *trunc,
                     bsgo_cdout cd, double eps, uint8_t *near_conf, uint8_t *near_los, double *res_margin, double *kappa,
                     int nthreads) {
    int n = cfg->n, od = bsgo_obs_dim_cfg(cfg);
    const int want_tr = near_conf || near_los || res_margin || kappa;
    int na = (cfg->kind == KIND_S || cfg->kind == KIND_M) ? 2 : 1;
    if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 8)
    for (int e = 0; e < E; e++) {
        double *a = ac + (size_t)e * BSGO_NF * n;
        double *v = ev + (size_t)e * BSGO_NE;
        uint64_t *rs = reso ? reso + (size_t)e * n : 0;
        int nl = nac[e];
        bsgo_row *rows = malloc(sizeof(bsgo_row) * (n > 0 ? n : 1));
        memset(rows, 0, sizeof(bsgo_row) * n);
        for (int i = 0; i < n; i++) rows[i].near_idx = -1;
        if (cfg->kind != KIND_RAW && actions) apply_action(cfg, a, v, actions + (size_t)e * na);
        bsgo_trace trc = { eps, near_conf ? near_conf + (size_t)e * n : 0, near_los ? near_los + (size_t)e * n : 0,
                           res_margin ? res_margin + (size_t)e * n : 0, kappa ? kappa + (size_t)e * n : 0, 0 };
        for (int i = 0; i < n; i++) { if (trc.res_margin) trc.res_margin[i] = 1e30; if (trc.kappa) trc.kappa[i] = 0.0; }
        for (int s = 0; s < cfg->substeps; s++) {
            if (cfg->asas_on) bsgo_asas_env_trace(cfg, a, rs, (size_t)E * n, nl, rows, want_tr ? &trc : 0);
            for (int i = 0; i < nl; i++) bsgo_kin_substep(cfg, a, v, n, i, cfg->simdt, cfg->reso_on && cfg->asas_on);
        }
        double rew = 0;
        int32_t te = 0, tr = 0;
        if (cfg->kind != KIND_RAW) {
            double drift = observe_env(cfg, a, nl, v, obs + (size_t)e * od);
            reward_env(cfg, a, nl, v, obs + (size_t)e * od, drift, &rew, &te, &tr);
        }
        v[EV_STEPS] += 1;
        if (reward) reward[e] = rew;
        if (term) term[e] = te;
        if (trunc) trunc[e] = tr;
        for (int i = 0; i < n; i++) {
            size_t k = (size_t)e * n + i;
            if (cd.inconf) cd.inconf[k] = rows[i].inconf;
            if (cd.nconf) cd.nconf[k] = rows[i].nconf;
            if (cd.nlos) cd.nlos[k] = rows[i].nlos;
            if (cd.near_idx) cd.near_idx[k] = rows[i].near_idx;
            if (cd.near_tcpa) cd.near_tcpa[k] = rows[i].near_tcpa;
            if (cd.near_dcpa) cd.near_dcpa[k] = rows[i].near_dcpa;
            if (cd.near_tin) cd.near_tin[k] = rows[i].near_tin;
            if (cd.tcpamax) cd.tcpamax[k] = rows[i].tcpamax;
        }
        free(rows);
    }
}
