/* ORACLE (test infrastructure, not product code; only tests/, smoke() and
 * bench.py's cpu_baseline/reference legs may load it).
 *
 * Plain-C, fp64 restatement of the hot path: BlueSky traffic substep
 * (SURVEY.md A.3/A.6), state-based CD (A.5), MVP + resumenav (A.7), and the
 * reference env layer's action/obs/reward/termination
 * (/root/reference/bluesky_gym/envs/{horizontal,vertical,sector}_cr_env.py).
 * Tier A (env layer) is pinned by goldens generated from the unmodified
 * reference classes run on oracle/refshim; Tier B (BlueSky) is a restatement:
 * parity unpinned vs upstream bluesky-simulator==1.0.3.
 *
 * Data model (shared with the CUDA core's set/get-state interface):
 *   ac   : double [E][BSGO_NF][N]   per-aircraft fields, field-major inside an env
 *   reso : uint64 [W][E][N]         resopairs bit rows, W = ceil(N/64) words per aircraft, word-major (bit j of row i =
 *                                   pair (i,j) kept: word j >> 6, bit j & 63); N <= 64: plain [E][N]; N <= 256
 *   nac  : int32  [E]               live aircraft per env (<= N); aircraft 0 is the agent
 *   ev   : double [E][BSGO_NE]      env-level slots
 */
#ifndef BSGO_H
#define BSGO_H
#include <stdint.h>

enum { F_LAT, F_LON, F_ALT, F_HDG, F_TAS, F_VS, F_CAS, F_SELSPD, F_SELALT, F_SELVS, F_APTRK, F_AX,
       F_ASAS_TRK, F_ASAS_TAS, F_ASAS_VS, F_ASAS_ALT, F_ASAS_ACTIVE,
       F_PHASE /* flight phase of the last perf.update (MVP caps use it one substep later) */,
       F_GSN, F_GSE /* ground-speed vector of the last update_groundspeed (= tas*(cos,sin)(hdg) without wind) */,
       F_LNAV /* LNAV: index of the active route waypoint, -1 = LNAV off (Merge intruders: 0 = FAF, 1 = runway) */,
       BSGO_NF };

#define BSGO_MAXV 32
/* Merge reuses: EV_WPT_LAT/LON = FAF, EV_WPT_X/Y = runway lat/lon, EV_WPT_REACH = wpt_reach, EV_FINAL_ALT = faf_reached */
enum { EV_WPT_LAT, EV_WPT_LON, EV_WPT_REACH, EV_AC_HDG,      /* Horizontal */
       EV_TARGET_ALT,                                         /* Vertical */
       EV_WPT_X, EV_WPT_Y, EV_NVERT,                          /* Sector: own waypoint [NM], polygon size */
       EV_TOTAL_REWARD, EV_TOTAL_INTR, EV_DRIFT_SUM, EV_DRIFT_CNT, EV_FINAL_ALT, EV_STEPS,
       EV_POLY,                                               /* 2*BSGO_MAXV slots: lat0,lon0,lat1,lon1,... */
       BSGO_NE = EV_POLY + 2 * BSGO_MAXV };

enum { KIND_H = 0, KIND_V = 1, KIND_S = 2, KIND_RAW = 3, KIND_M = 4 /* MergeEnv (merge_env.py) */ };

#define BSGO_MAXWIND 8
typedef struct {
    int32_t kind, n, substeps, asas_on, reso_on, reso_off_own;
    double simdt, rpz, hpz, dtlook, resofach, resofacv;
    /* wind field of the WindFieldWrapper (bluesky_gym/wrappers/wind.py:28): nwind points, no altitude profile;
     * wind_obs = augment_obs (two more observation slots: wind_u, wind_v) */
    int32_t nwind, wind_obs;
    double wlat[BSGO_MAXWIND], wlon[BSGO_MAXWIND], wvn[BSGO_MAXWIND], wve[BSGO_MAXWIND];
    /* MVP resolution method (stack commands RMETHH / RMETHV): 0 = horizontal, heading + speed (BlueSky's default),
     * 1 = vertical speed only (RMETHV V/S), 2 = horizontal + vertical (RMETHH OFF, RMETHV OFF) */
    int32_t reso_mode;
    /* altitude profile of the (single) wind point, WindFieldWrapper(..., alt=[...]): nknot > 0 knots (walt, wvn, wve);
     * resampled onto BlueSky's 100 ft altitude axis and interpolated linearly between axis levels (refshim windfield.py) */
    int32_t nknot;
    double walt[BSGO_MAXWIND];
} bsgo_cfg;

/* per-aircraft CD outputs of the last detection of a step */
typedef struct {
    uint8_t *inconf;      /* [E][N] */
    int32_t *nconf;       /* [E][N] conflict pairs in row i */
    int32_t *nlos;        /* [E][N] LoS pairs in row i */
    int32_t *near_idx;    /* [E][N] argmin_j tcpa over conflict pairs, -1 if none (ties: lowest j) */
    double *near_tcpa, *near_dcpa, *near_tin;   /* [E][N] */
    double *tcpamax;      /* [E][N] */
} bsgo_cdout;

int bsgo_obs_dim(int kind, int n);   /* without the wind_obs slots */
int bsgo_obs_dim_cfg(const bsgo_cfg *cfg);

/* aero */
double bsgo_vcas2tas(double cas, double h);
double bsgo_vtas2cas(double tas, double h);
double bsgo_vcasormach2tas(double spd, double h);
/* geo */
void bsgo_kwikqdrdist(double lata, double lona, double latb, double lonb, double *qdr, double *dist_nm);
void bsgo_qdrdist(double lat1, double lon1, double lat2, double lon2, double *qdr, double *dist_nm);   /* WGS-84 great circle */
void bsgo_wind(const bsgo_cfg *cfg, double lat, double lon, double alt, double *vn, double *ve);

/* one full CD matrix for one env (dense outputs, row-major n x n); any pointer may be NULL */
void bsgo_detect_dense(int n, const double *lat, const double *lon, const double *alt,
                       const double *trk, const double *gs, const double *vs,
                       double rpz, double hpz, double dtlook,
                       double *tcpa, double *dcpa2, double *tinconf, double *toutconf,
                       uint8_t *swconfl, uint8_t *swlos, double *margin_conf, double *margin_los);

/* batched CD row outputs from raw state (KIND_RAW sweep; rows [row0,row1) of each env) */
void bsgo_cd_rows(int E, int n, const double *lat, const double *lon, const double *alt,
                  const double *trk, const double *gs, const double *vs,
                  double rpz, double hpz, double dtlook, int row0, int row1,
                  bsgo_cdout out, int nthreads);

/* same, plus the near-threshold enumeration without dense matrices: near_conf / near_los [E][n] (may be NULL) flag the
 * rows that hold a pair whose relative margin to a conflict / LoS threshold is below eps */
void bsgo_cd_rows_near(int E, int n, const double *lat, const double *lon, const double *alt,
                       const double *trk, const double *gs, const double *vs,
                       double rpz, double hpz, double dtlook, int row0, int row1,
                       bsgo_cdout out, double eps, uint8_t *near_conf, uint8_t *near_los, int nthreads);

/* K traffic substeps (+CD/MVP when enabled) then the env layer, for E envs */
void bsgo_step(const bsgo_cfg *cfg, int E, double *ac, uint64_t *reso, const int32_t *nac, double *ev,
               const double *actions, double *obs, double *reward, int32_t *term, int32_t *trunc,
               bsgo_cdout cd, int nthreads);
/* same step with a trace for the parity tests (any trace pointer may be NULL): near_conf / near_los [E][N] = rows holding a
 * pair within eps of a conflict / LoS threshold at the LAST detection of the step; res_margin [E][N] = per ownship, the
 * smallest relative margin over all substeps of the decisions that steer its resolution (conflict flags, MVP's 10 m
 * switch and erratum branch, resumenav's tests; 0 = a degenerate equal-velocity conflict pair); kappa [E][N] = per
 * ownship, MVP's largest amplification of a relative input rounding [m/s] (bsgo_int.h: bsgo_trace) */
void bsgo_step_trace(const bsgo_cfg *cfg, int E, double *ac, uint64_t *reso, const int32_t *nac, double *ev,
                     const double *actions, double *obs, double *reward, int32_t *term, int32_t *trunc,
                     bsgo_cdout cd, double eps, uint8_t *near_conf, uint8_t *near_los, double *res_margin, double *kappa,
                     int nthreads);
/* obs/info only (what reset() returns) */
void bsgo_observe(const bsgo_cfg *cfg, int E, const double *ac, const int32_t *nac, double *ev, double *obs);

/* ND-map (builder-owned spec): fp32 sample arithmetic, integer grid */
typedef struct {
    int32_t nx, ny, nz, nt;
    float x0, y0, z0, inv_cx, inv_cy, inv_cz, dt;
} bsgo_ndcfg;
void bsgo_ndmap(const bsgo_ndcfg *g, int E, int n, const float *rec /* [E][6][n]: x,y,z,u,v,w */,
                uint8_t *occ /* [E][nt][nz][ny][nx] */, uint8_t *conf /* same shape, 0/1 */,
                int32_t *ncell /* [E][nt] conflict cells per time slab */, int nthreads);

/* Philox4x32-10 scenario generator (same stream on host and device) */
void bsgo_philox(uint32_t key0, uint32_t key1, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t out[4]);
#endif
