/* ORACLE internal constants (test infrastructure). Values restate
 * bluesky/tools/aero.py and the envelope table of oracle/refshim/bluesky/traffic/perf.py. */
#ifndef BSGO_INT_H
#define BSGO_INT_H
#include "bsgo.h"
#define C_PI 3.141592653589793238462643383279502884
#define C_D2R (C_PI / 180.0)
#define C_R2D (180.0 / C_PI)
#define C_RE 6371000.0
#define C_NM 1852.0
#define C_FT 0.3048
#define C_FPM (0.3048 / 60.0)
#define C_KTS 0.514444
#define C_G0 9.80665
#define C_R 287.05287
#define C_P0 101325.0
#define C_RHO0 1.225
#define C_GAMMA 1.40
#define C_VSDEF (1500.0 * C_FPM)
#define C_BANK (25.0 * C_D2R)
enum { PH_NA, PH_GD, PH_IC, PH_CL, PH_CR, PH_DE, PH_AP };
#define ENV_VMINIC 65.0
#define ENV_VMAXIC 82.0
#define ENV_VMINER 80.0
#define ENV_VMAXER 180.0
#define ENV_VMINAP 60.0
#define ENV_VMAXAP 100.0
#define ENV_VSMIN (-20.0)
#define ENV_VSMAX 20.0
#define ENV_HMAX 12131.0
#define ENV_MMO 0.82
#define ENV_AXMAX 2.0
double bsgo_pymod(double a, double b);
int bsgo_phase(double tas, double vs, double alt);
void bsgo_vlimits(int ph, double *vmin, double *vmax);
void bsgo_kin_substep(const bsgo_cfg *cfg, double *a, const double *ev, int n, int i, double dt, int reso_on);
/* CD + MVP + resumenav for one env; updates F_ASAS_* and reso rows; fills row outputs at [0..n) */
typedef struct {
    uint8_t inconf; int32_t nconf, nlos, near_idx; double near_tcpa, near_dcpa, near_tin, tcpamax;
} bsgo_row;
void bsgo_asas_env(const bsgo_cfg *cfg, double *a, uint64_t *reso, size_t wstride, int nlive, bsgo_row *rows);
/* Trace of one env over the substeps of a step (parity tests: what an fp32 path may legitimately decide differently).
 * near_conf / near_los [n]: rows holding a pair within eps of a conflict / LoS threshold at the LAST detection;
 * res_margin [n]: per ownship row, the smallest relative margin, over all substeps and partners, of the decisions that
 *   steer its resolution -- conflict flags (every substep feeds MVP), MVP's |dCPA_h| <= 10 m switch and its erratum
 *   branch test, resumenav's past-CPA sign, horizontal-LoS and bouncing tests; 0 for a row with a DEGENERATE conflict
 *   pair (|v_rel|^2 under 1e-4 m2/s2: tCPA is the quotient of two rounding residues over the reference's arbitrary
 *   1e-6 clamp, in any precision);
 * kappa [n]: per ownship row, the largest amplification MVP applies to a relative input rounding over all substeps
 *   and conflict pairs [m/s]: (dist + |v_rel||tCPA|) rpz_m / (|tCPA| |dCPA_h|), or with dist in place of |dCPA_h| in
 *   the perpendicular-escape branch (|dCPA_h| <= 10 m), whose direction comes from the relative position alone. */
typedef struct { double eps; uint8_t *near_conf, *near_los; double *res_margin, *kappa; int row; double rabs; } bsgo_trace;
void bsgo_asas_env_trace(const bsgo_cfg *cfg, double *a, uint64_t *reso, size_t wstride, int nlive, bsgo_row *rows, bsgo_trace *tr);
void bsgo_pair(double lati, double loni, double alti, double ui, double vi, double vsi,
               double latj, double lonj, double altj, double uj, double vj, double vsj,
               double rpz, double hpz, double dtlook,
               double *qdr, double *dist, double *tcpa, double *dcpa2, double *tinconf, double *toutconf,
               int *swconfl, int *swlos, double *margin_conf, double *margin_los);
#endif
