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
void bsgo_asas_env(const bsgo_cfg *cfg, double *a, uint64_t *reso, int nlive, bsgo_row *rows);
void bsgo_pair(double lati, double loni, double alti, double ui, double vi, double vsi,
               double latj, double lonj, double altj, double uj, double vj, double vsj,
               double rpz, double hpz, double dtlook,
               double *qdr, double *dist, double *tcpa, double *dcpa2, double *tinconf, double *toutconf,
               int *swconfl, int *swlos, double *margin_conf, double *margin_los);
#endif
