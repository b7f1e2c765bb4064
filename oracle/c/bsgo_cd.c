/* ORACLE (test infrastructure, not product code).
 * fp64 restatement of StateBased.detect, MVP.resolve/MVP.MVP and
 * ConflictResolution.resumenav (A.7), pair by pair, in the operation order of
 * oracle/refshim/bluesky/traffic/{asas,mvp}.py.  parity unpinned vs upstream. */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "bsgo_int.h"

static double relgap(double a, double b, double floor_) {
    double s = fmax(fmax(fabs(a), fabs(b)), floor_);
    return fabs(a - b) / s;
}

/* one ordered pair (i -> j), i != j */
void bsgo_pair(double lati, double loni, double alti, double ui, double vi, double vsi,
               double latj, double lonj, double altj, double uj, double vj, double vsj,
               double rpz, double hpz, double dtlook,
               double *qdr_o, double *dist_o, double *tcpa_o, double *dcpa2_o, double *tin_o, double *tout_o,
               int *swconfl_o, int *swlos_o, double *mconf_o, double *mlos_o) {
    double qdr, dnm;
    bsgo_kwikqdrdist(lati, loni, latj, lonj, &qdr, &dnm);
    double dist = dnm * C_NM;
    double qr = qdr * C_D2R;
    double dx = dist * sin(qr), dy = dist * cos(qr);
    double du = uj - ui, dv = vj - vi;
    double dv2 = du * du + dv * dv;
    if (fabs(dv2) < 1e-6) dv2 = 1e-6;
    double vrel = sqrt(dv2);
    double tcpa = -(du * dx + dv * dy) / dv2;
    double dcpa2 = fabs(dist * dist - tcpa * tcpa * dv2);
    double R2 = rpz * rpz;
    int swhor = dcpa2 < R2;
    double dtinhor = sqrt(fmax(0.0, R2 - dcpa2)) / vrel;
    double tinhor = swhor ? tcpa - dtinhor : 1e8;
    double touthor = swhor ? tcpa + dtinhor : -1e8;
    double dalt = altj - alti;
    double dvs = vsj - vsi;
    if (fabs(dvs) < 1e-6) dvs = 1e-6;
    double thi = (dalt + hpz) / -dvs, tlo = (dalt - hpz) / -dvs;
    double tinver = fmin(thi, tlo), toutver = fmax(thi, tlo);
    double tin = fmax(tinver, tinhor), tout = fmin(toutver, touthor);
    int sw = swhor && (tin <= tout) && (tout > 0.0) && (tin < dtlook);
    int los = (dist < rpz) && (fabs(dalt) < hpz);
    if (qdr_o) *qdr_o = qdr;
    if (dist_o) *dist_o = dist;
    if (tcpa_o) *tcpa_o = tcpa;
    if (dcpa2_o) *dcpa2_o = dcpa2;
    if (tin_o) *tin_o = tin;
    if (tout_o) *tout_o = tout;
    if (swconfl_o) *swconfl_o = sw;
    if (swlos_o) *swlos_o = los;
    if (mconf_o) {
        double m = relgap(dcpa2, R2, 0.0);
        if (swhor) {
            m = fmin(m, relgap(tin, tout, 1.0));
            m = fmin(m, relgap(tout, 0.0, 1.0));
            m = fmin(m, relgap(tin, dtlook, 1.0));
        }
        *mconf_o = m;
    }
    if (mlos_o) *mlos_o = fmin(relgap(dist, rpz, 0.0), relgap(fabs(dalt), hpz, 0.0));
}

void bsgo_detect_dense(int n, const double *lat, const double *lon, const double *alt,
                       const double *trk, const double *gs, const double *vs,
                       double rpz, double hpz, double dtlook,
                       double *tcpa, double *dcpa2, double *tinconf, double *toutconf,
                       uint8_t *swconfl, uint8_t *swlos, double *mconf, double *mlos) {
    double *u = malloc(sizeof(double) * n), *v = malloc(sizeof(double) * n);
    for (int i = 0; i < n; i++) {
        u[i] = gs[i] * sin(trk[i] * C_D2R);
        v[i] = gs[i] * cos(trk[i] * C_D2R);
    }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            size_t k = (size_t)i * n + j;
            double t = 0, d2 = 0, ti = 0, to = 0, mc = 1e30, ml = 1e30;
            int sw = 0, lo = 0;
            if (i != j)
                bsgo_pair(lat[i], lon[i], alt[i], u[i], v[i], vs[i], lat[j], lon[j], alt[j], u[j], v[j], vs[j],
                          rpz, hpz, dtlook, 0, 0, &t, &d2, &ti, &to, &sw, &lo, &mc, &ml);
            if (tcpa) tcpa[k] = t;
            if (dcpa2) dcpa2[k] = d2;
            if (tinconf) tinconf[k] = ti;
            if (toutconf) toutconf[k] = to;
            if (swconfl) swconfl[k] = (uint8_t)sw;
            if (swlos) swlos[k] = (uint8_t)lo;
            if (mconf) mconf[k] = mc;
            if (mlos) mlos[k] = ml;
        }
    free(u);
    free(v);
}

static void row_init(bsgo_row *r) {
    r->inconf = 0; r->nconf = 0; r->nlos = 0; r->near_idx = -1;
    r->near_tcpa = 0; r->near_dcpa = 0; r->near_tin = 0; r->tcpamax = 0;
}
static void row_add(bsgo_row *r, int j, int sw, int los, double tcpa, double dcpa2, double tin) {
    if (los) r->nlos++;
    if (!sw) return;
    r->inconf = 1;
    r->nconf++;
    if (tcpa > r->tcpamax) r->tcpamax = tcpa;       /* max(tcpa*swconfl) over the row, floor 0 */
    if (r->near_idx < 0 || tcpa < r->near_tcpa) {
        r->near_idx = j; r->near_tcpa = tcpa; r->near_dcpa = sqrt(dcpa2); r->near_tin = tin;
    }
}

/* near_conf / near_los (may be NULL): [E][n] flags, row i holds a pair whose margin to a conflict / LoS threshold is
 * below eps -- the enumeration of near-threshold pairs without the dense n x n matrices (giant airspaces) */
void bsgo_cd_rows_near(int E, int n, const double *lat, const double *lon, const double *alt,
                       const double *trk, const double *gs, const double *vs,
                       double rpz, double hpz, double dtlook, int row0, int row1,
                       bsgo_cdout out, double eps, uint8_t *near_conf, uint8_t *near_los, int nthreads);

void bsgo_cd_rows(int E, int n, const double *lat, const double *lon, const double *alt,
                  const double *trk, const double *gs, const double *vs,
                  double rpz, double hpz, double dtlook, int row0, int row1,
                  bsgo_cdout out, int nthreads) {
    bsgo_cd_rows_near(E, n, lat, lon, alt, trk, gs, vs, rpz, hpz, dtlook, row0, row1, out, 0.0, 0, 0, nthreads);
}

void bsgo_cd_rows_near(int E, int n, const double *lat, const double *lon, const double *alt,
                       const double *trk, const double *gs, const double *vs,
                       double rpz, double hpz, double dtlook, int row0, int row1,
                       bsgo_cdout out, double eps, uint8_t *near_conf, uint8_t *near_los, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    const int want_near = near_conf || near_los;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 16) collapse(2)
    for (int e = 0; e < E; e++)
        for (int i = row0; i < row1; i++) {
            size_t b = (size_t)e * n;
            bsgo_row r;
            row_init(&r);
            int nc = 0, nl = 0;
            double ui = gs[b + i] * sin(trk[b + i] * C_D2R), vi = gs[b + i] * cos(trk[b + i] * C_D2R);
            for (int j = 0; j < n; j++) {
                if (j == i) continue;
                double uj = gs[b + j] * sin(trk[b + j] * C_D2R), vj = gs[b + j] * cos(trk[b + j] * C_D2R);
                double t, d2, ti, mc = 1e30, ml = 1e30;
                int sw, lo;
                bsgo_pair(lat[b + i], lon[b + i], alt[b + i], ui, vi, vs[b + i],
                          lat[b + j], lon[b + j], alt[b + j], uj, vj, vs[b + j],
                          rpz, hpz, dtlook, 0, 0, &t, &d2, &ti, 0, &sw, &lo, want_near ? &mc : 0, want_near ? &ml : 0);
                row_add(&r, j, sw, lo, t, d2, ti);
                if (mc < eps) nc = 1;
                if (ml < eps) nl = 1;
            }
            size_t k = b + i;
            if (near_conf) near_conf[k] = (uint8_t)nc;
            if (near_los) near_los[k] = (uint8_t)nl;
            if (out.inconf) out.inconf[k] = r.inconf;
            if (out.nconf) out.nconf[k] = r.nconf;
            if (out.nlos) out.nlos[k] = r.nlos;
            if (out.near_idx) out.near_idx[k] = r.near_idx;
            if (out.near_tcpa) out.near_tcpa[k] = r.near_tcpa;
            if (out.near_dcpa) out.near_dcpa[k] = r.near_dcpa;
            if (out.near_tin) out.near_tin[k] = r.near_tin;
            if (out.tcpamax) out.tcpamax[k] = r.tcpamax;
        }
}
