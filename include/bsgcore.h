/* bsgcore.h -- C-ABI of libbsgcore.so, the B200-native batched BlueSky-Gym simulation core.
 *
 * The reference (Somnathab3/bluesky-gym-hallucination) is pure Python: it has no FFI.  The
 * calls this library replaces are the ones the reference's env layer makes into the
 * process-global BlueSky singleton, plus the env layer's own
 * per-step Python:
 *   bs.traf.reset/cre/creconfs            horizontal_cr_env.py:85,91,133; vertical_cr_env.py:198,270;
 *                                         sector_cr_env.py:88,211,217           -> bsg_reset / bsg_set_state
 *   bs.stack.stack("HDG/SPD ..."), selalt/selvs writes
 *                                         horizontal_cr_env.py:272-275; vertical_cr_env.py:242-256;
 *                                         sector_cr_env.py:315-322              -> actions of bsg_step
 *   bs.sim.step() x ACTION_FREQUENCY      horizontal_cr_env.py:107-109; vertical_cr_env.py:287-289;
 *                                         sector_cr_env.py:119-121              -> bsg_step (fused)
 *   _get_obs/_get_reward/_get_info        horizontal_cr_env.py:150-270; vertical_cr_env.py:113-240;
 *                                         sector_cr_env.py:227-339              -> outputs of bsg_step
 *   ASAS StateBased.detect / MVP.resolve  (BlueSky, "ASAS ON"/"RESO MVP"; never enabled by the reference)
 *                                                                               -> asas_on / reso_on, bsg_cd_sweep
 *   ND-map 4D conflict grid               (absent from the reference tree; builder-owned spec) -> bsg_ndmap
 * INTEGRATION.md shows the ctypes binding a reference maintainer would add.
 *
 * Conventions: every entry returns 0 (BSG_OK) or a negative bsg_status and never throws or
 * aborts; CUDA errors are captured into bsg_last_error().  Pointers named *_dev are device
 * pointers on the handle's device, *_host are host pointers; the caller owns all I/O buffers.
 * `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); calls taking a
 * stream are asynchronous on it.  A handle is bound to one device and is not thread-safe.
 * Element type of "real" device/host buffers follows the handle's precision:
 * float (BSG_FP32, the fast path) or double (BSG_FP64, the parity path).
 */
#ifndef BSGCORE_H
#define BSGCORE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define BSG_ABI_VERSION 4

typedef enum { BSG_OK = 0, BSG_ERR_ARG = -1, BSG_ERR_CUDA = -2, BSG_ERR_NOMEM = -3, BSG_ERR_UNSUPPORTED = -4,
               BSG_ERR_NODEVICE = -5,
               BSG_ERR_TOOLARGE = -6 /* bsg_cd_sym_static only: more rows than the library's built-in scratch holds */ } bsg_status;
enum { BSG_KIND_HORIZONTAL = 0, BSG_KIND_VERTICAL = 1, BSG_KIND_SECTOR = 2, BSG_KIND_RAW = 3,
       BSG_KIND_MERGE = 4 /* MergeEnv-v0: bluesky_gym/envs/merge_env.py (20 aircraft, intruders on an LNAV route) */ };
enum { BSG_FP32 = 0, BSG_FP64 = 1 };

/* per-aircraft state fields of the fp64 host interchange layout  ac[E][BSG_NF][N] */
enum { BSG_F_LAT, BSG_F_LON, BSG_F_ALT, BSG_F_HDG, BSG_F_TAS, BSG_F_VS, BSG_F_CAS, BSG_F_SELSPD, BSG_F_SELALT,
       BSG_F_SELVS, BSG_F_APTRK, BSG_F_AX, BSG_F_ASAS_TRK, BSG_F_ASAS_TAS, BSG_F_ASAS_VS, BSG_F_ASAS_ALT,
       BSG_F_ASAS_ACTIVE, BSG_F_PHASE,
       BSG_F_GSN, BSG_F_GSE,   /* ground-speed vector of the last update (= tas * (cos, sin)(hdg) without wind) */
       BSG_F_LNAV,             /* LNAV: index of the active route waypoint, -1 = off (Merge: 0 = FAF, 1 = runway) */
       BSG_NF };
/* env-level slots of ev[E][BSG_NE].  Merge reuses: WPT_LAT/LON = FAF, WPT_X/Y = runway lat/lon [deg],
 * WPT_REACH = wpt_reach, FINAL_ALT = faf_reached (info['faf_reach']) */
#define BSG_MAXV 32
enum { BSG_EV_WPT_LAT, BSG_EV_WPT_LON, BSG_EV_WPT_REACH, BSG_EV_AC_HDG, BSG_EV_TARGET_ALT, BSG_EV_WPT_X,
       BSG_EV_WPT_Y, BSG_EV_NVERT, BSG_EV_TOTAL_REWARD, BSG_EV_TOTAL_INTR, BSG_EV_DRIFT_SUM, BSG_EV_DRIFT_CNT,
       BSG_EV_FINAL_ALT, BSG_EV_STEPS, BSG_EV_POLY, BSG_NE = BSG_EV_POLY + 2 * BSG_MAXV };

typedef struct bsg_handle bsg_handle;

typedef struct bsg_config {
    int32_t struct_size;      /* sizeof(bsg_config), for ABI evolution */
    int32_t env_kind;         /* BSG_KIND_* */
    int32_t num_envs;         /* E */
    int32_t max_ac;           /* N: aircraft slots per env (aircraft 0 = agent) */
    int32_t precision;        /* BSG_FP32 | BSG_FP64 */
    int32_t substeps;         /* ACTION_FREQUENCY; 0 = the reference's value for env_kind (10/30/5) */
    int32_t asas_on;          /* state-based CD every substep (BlueSky "ASAS ON") */
    int32_t reso_on;          /* MVP resolution + resumenav (BlueSky "RESO MVP"); needs asas_on */
    int32_t reso_off_own;     /* agent does not resolve (BlueSky "RESOOFF KL001") */
    int32_t max_episode_steps;/* TimeLimit: truncate when reached; 0 = never */
    int32_t device;           /* CUDA device ordinal */
    int32_t fixed_nac;        /* Sector reset: 1 = every env gets exactly max_ac aircraft (the synthetic scale-up
                                 configs, SURVEY 8d C3 "N fixed"); 0 = the reference's density law (sector_cr_env.py:98-100) */
    int32_t density_mode;     /* Sector reset, SectorCREnv(ac_density_mode=...) (sector_cr_env.py:45,98-103): 0 = "normal"
                                 N(0.005, 0.001) aircraft per NM^2, 1 = "uniform" U(0.003, 0.007) */
    int32_t reso_mode;        /* MVP resolution method (BlueSky stack RMETHH / RMETHV; same offset as ABI 3's reserved0):
                                 0 = horizontal, heading + speed (BlueSky's default), 1 = vertical speed only (RMETHV V/S),
                                 2 = horizontal + vertical (RMETHH OFF, RMETHV OFF).  1 and 2 evaluate MVP's vertical
                                 term, timesolveV and the resolution altitude; they run on the ordered-sweep
                                 kernels (any precision; no wind / Merge in fp32) */
    double simdt;             /* 0 = the reference's DT for env_kind (5/1/1 s) */
    double rpz_m, hpz_m;      /* protected zone; 0 = 5 NM / 1000 ft */
    double dtlookahead;       /* 0 = 300 s */
    double resofach, resofacv;/* 0 = 1.05 */
} bsg_config;

/* device output pointers of bsg_step; any pointer may be NULL (not written) */
typedef struct bsg_step_out {
    void *obs;                /* real [E][obs_dim] */
    void *reward;             /* real [E] */
    int32_t *terminated;      /* [E] 0/1 */
    int32_t *truncated;       /* [E] 0/1 */
    void *info;               /* real [E][4]: total_reward, total_intrusions, average_drift|final_altitude,
                                 steps (Merge: faf_reach) */
    uint8_t *cd_inconf;       /* [E][N] row has a conflict (last CD call of the step) */
    int32_t *cd_nconf;        /* [E][N] conflict pairs in the row */
    int32_t *cd_nlos;         /* [E][N] loss-of-separation pairs in the row */
    int32_t *cd_near_idx;     /* [E][N] argmin_j tCPA over the row's conflict pairs, -1 if none */
    void *cd_near_tcpa, *cd_near_dcpa, *cd_near_tin;   /* real [E][N] */
    void *cd_tcpamax;         /* real [E][N] */
    int32_t *done;            /* [E] terminated | truncated: the mask bsg_reset takes for auto-reset */
} bsg_step_out;

/* stand-alone all-pairs CD sweep (giant-airspace config; pair matrix rows [row0,row1) of every env) */
typedef struct bsg_cd_params {
    int32_t num_envs, num_ac, row0, row1;
    float rpz_m, hpz_m, dtlookahead;
    int32_t reserved;
} bsg_cd_params;

typedef struct bsg_ndmap_params {
    int32_t num_envs, num_ac, nx, ny, nz, nt;
    float x0, y0, z0, cell_x, cell_y, cell_z, dt;
    int32_t reserved;
} bsg_ndmap_params;

int bsg_abi_version(void);
const char *bsg_status_string(int status);
int bsg_device_count(int *count);

int bsg_create(const bsg_config *cfg, bsg_handle **out);
int bsg_destroy(bsg_handle *h);
const char *bsg_last_error(const bsg_handle *h);
int bsg_obs_dim(const bsg_handle *h);
int bsg_act_dim(const bsg_handle *h);
int bsg_launch_count(const bsg_handle *h, int64_t *kernels_launched);

/* Wind field of the reference's WindFieldWrapper (bluesky_gym/wrappers/wind.py:26-29 -> bs.traf.wind.addpointvne):
 * npoints wind vectors [m/s] at (lat, lon) [deg], no altitude profile.  One point = uniform wind, several = inverse
 * distance-squared weighting; npoints = 0 switches the wind off.  The wind acts above 50 ft: ground speed = TAS + wind,
 * the autopilot holds the commanded TRACK (wind-correction angle), HDG actions are converted to the track they give.
 * augment_obs != 0 appends wind_u, wind_v (wind in the agent's body axes / 50, wind.py:55-64) to every observation:
 * bsg_obs_dim() grows by 2.  Call before the first step of an episode (the wrapper adds the wind right after reset). */
#define BSG_MAX_WIND 8
int bsg_set_wind(bsg_handle *h, int32_t npoints, const double *lat_host, const double *lon_host,
                 const double *vnorth_host, const double *veast_host, int32_t augment_obs);
/* The same wrapper with its alt= argument (wind.py:9,28: addpointvne(lat, lon, vnorth, veast, alt)): ONE wind point with an
 * altitude profile of nknots <= BSG_MAX_WIND knots (alt_m increasing [m], vnorth / veast [m/s] at the knots).  As in
 * BlueSky's Windfield the knots are resampled onto a 100 ft altitude axis (0 .. 45 000 ft; constant beyond the profile's
 * ends) and the field is linear between the two axis levels around an aircraft's altitude; the wrapper's wind
 * observation and the HDG -> track conversion read the field at altitude 0 (upstream takes the altitude from array
 * arguments only).  Replaces any wind set before. */
int bsg_set_wind_profile(bsg_handle *h, double lat, double lon, int32_t nknots, const double *alt_m_host,
                         const double *vnorth_host, const double *veast_host, int32_t augment_obs);
/* Gaussian observation noise of the reference's NoisyObservationWrapper (bluesky_gym/wrappers/uncertainty.py:19-32):
 * every observation float gets + N(0, sigma) from a Philox4x32-10 stream keyed by (seed; env, step launch, slot).
 * sigma = 0 switches it off. */
int bsg_set_obs_noise(bsg_handle *h, double sigma, uint64_t seed);

/* fp64 interchange with the host (parity tests, checkpoint/resume). NULL pointers are skipped. Synchronous.
 * reso_host: resopairs bit rows (bit j of aircraft i's row = pair (i, j) is being resolved), uint64 [E][N] for N <= 64;
 * for N > 64, W = ceil(N / 64) words per aircraft, word-major [W][E][N] (bit j: word j >> 6, bit j & 63). */
int bsg_set_state(bsg_handle *h, const double *ac_host, const uint64_t *reso_host, const int32_t *nac_host,
                  const double *ev_host);
int bsg_get_state(bsg_handle *h, double *ac_host, uint64_t *reso_host, int32_t *nac_host, double *ev_host);

/* device-side scenario generation (reset path of the env layer) for envs with mask_dev[e] != 0 (NULL = all) */
int bsg_reset(bsg_handle *h, const int32_t *mask_dev, uint64_t seed, void *stream);
/* Gymnasium "next-step" auto-reset: when enabled, every bsg_step is followed on the same stream by a
 * bsg_reset of the envs it reported done, and the NEXT bsg_step returns for those envs the first
 * observation of the new episode with reward 0 / terminated 0 / truncated 0 (their action is ignored).
 * enable = 2 is the same-step flavour SB3's VecEnv uses: the reset still follows the step, but the next step
 * integrates the fresh episode normally (the caller fetches its first observation with bsg_observe). */
int bsg_set_autoreset(bsg_handle *h, int32_t enable, uint64_t seed);
/* observation of the current state (what reset() returns) */
int bsg_observe(bsg_handle *h, void *obs_dev, void *stream);
/* one env step for every env: action intake -> substeps x (autopilot, CD, MVP, limits, kinematics) -> obs/reward */
int bsg_step(bsg_handle *h, const void *actions_dev /* real [E][act_dim] */, const bsg_step_out *out, void *stream);
/* same through host buffers (pinned staging, copies inside): the end-to-end call of the plugin boundary */
/* The pinned host block bsg_step_host fills (obs | reward | info[4] | terminated | truncated), as pointers: a caller
 * that passes NULL output pointers to bsg_step_host reads the results here without another host copy (the block of the
 * step waited for last).  Valid until the second next bsg_step_host(_begin) / bsg_destroy on this handle. */
int bsg_host_views(bsg_handle *h, void **obs, void **reward, void **info, int32_t **terminated, int32_t **truncated);
/* bsg_step_host moves the data in one of two ways, same results bit for bit: copy mode (actions H2D, one launch, one
 * D2H of the output block) or direct mode (the kernel reads the actions from and stores the outputs to the pinned
 * block over PCIe itself; default when the launch has >= 4096 CTAs, never with observation noise on).  The
 * environment variable BSG_HOST_DIRECT=0/1 forces one of them. */
int bsg_step_host(bsg_handle *h, const void *actions_host, void *obs_host, void *reward_host,
                  int32_t *terminated_host, int32_t *truncated_host, void *info_host);
/* The same call split in two, for a host that has other work while the step is in flight (a second handle / GPU to
 * feed, the previous step's bookkeeping): bsg_step_host_begin copies the actions into one of the handle's TWO pinned
 * blocks and enqueues H2D + the launch + D2H on the handle's own stream, then returns; bsg_step_host_wait blocks until
 * the OLDEST step in flight has landed and returns pointers into its pinned block (valid until two more begins).  At
 * most two steps may be in flight per handle; bsg_step_host = begin + wait + the optional copies. */
int bsg_step_host_begin(bsg_handle *h, const void *actions_host);
int bsg_step_host_wait(bsg_handle *h, void **obs, void **reward, void **info, int32_t **terminated, int32_t **truncated);
/* The same two calls for bindings whose per-call marshalling matters (ctypes): bsg_step_host_block = begin + wait, both
 * return the index (0 or 1) of the pinned block that holds the step's results (the pointers of a block are constant
 * between changes of the observation size: ask bsg_host_views once per block), or a negative bsg_status. */
int bsg_step_host_block(bsg_handle *h, const void *actions_host);
int bsg_step_host_wait_block(bsg_handle *h);

/* pack fp32 CD records (8 floats per aircraft) from SoA device arrays [E][N]; origin = (lat0, lon0) degrees */
int bsg_cd_pack(int32_t num_envs, int32_t num_ac, const double *lat_dev, const double *lon_dev,
                const float *alt_dev, const float *trk_dev, const float *gs_dev, const float *vs_dev,
                double lat0, double lon0, float *rec_dev /* [E][N][8] */, void *stream);
/* bsg_cd_pack + bsg_cd_sym_init of the same E x N rows in ONE launch (the peer sweep's step: one kernel fewer) */
int bsg_cd_pack_init(int32_t num_envs, int32_t num_ac, const double *lat_dev, const double *lon_dev, const float *alt_dev,
                     const float *trk_dev, const float *gs_dev, const float *vs_dev, double lat0, double lon0, float *rec_dev,
                     int32_t *nconf_dev, int32_t *nlos_dev, float *tcpamax_dev, int64_t *key_dev, void *stream);
int bsg_cd_sweep(const bsg_cd_params *p, const float *rec_dev, uint8_t *inconf_dev, int32_t *nconf_dev,
                 int32_t *nlos_dev, int32_t *near_idx_dev, float *near_tcpa_dev, float *near_dcpa_dev,
                 float *near_tin_dev, float *tcpamax_dev, void *stream);

/* Symmetric form of the same sweep: every unordered pair is evaluated once (the fp32 pair form is exactly
 * symmetric, so the row outputs are bit-identical to bsg_cd_sweep's).  Three stages on caller-owned device buffers
 * [E][N]: init, sweep (accumulates with order-independent integer atomics; tile_rank / tile_stride deal the tiles of
 * the upper triangle to GPUs round-robin), finalize.  With tile_stride > 1 the caller combines the partial outputs
 * across GPUs between sweep and finalize: nconf, nlos SUM (int32); tcpamax MAX (its float bits as int32: values are
 * non-negative); key MIN (int64).  row0 / row1 of bsg_cd_params are ignored (all rows are produced). */
int bsg_cd_sym_init(const bsg_cd_params *p, int32_t *nconf_dev, int32_t *nlos_dev, float *tcpamax_dev,
                    int64_t *key_dev, void *stream);
int bsg_cd_sym_sweep(const bsg_cd_params *p, const float *rec_dev, int32_t tile_rank, int32_t tile_stride,
                     int32_t *nconf_dev, int32_t *nlos_dev, float *tcpamax_dev, int64_t *key_dev, void *stream);
int bsg_cd_sym_finalize(const bsg_cd_params *p, const float *rec_dev, const int32_t *nconf_dev, const int64_t *key_dev,
                        uint8_t *inconf_dev, int32_t *near_idx_dev, float *near_tcpa_dev, float *near_dcpa_dev,
                        float *near_tin_dev, void *stream);

/* No counterpart in the reference (nor in BlueSky): this entry point exists for hosts that may not allocate GPU memory.
 * The three-stage sweep above needs four accumulator arrays from its caller (nconf, nlos, tcpamax and the int64 key,
 * 20 bytes per aircraft row).  Some users of this library cannot provide them:
 *   - a trainer or simulator that gave all of the GPU's memory to its own pool at start-up (torch's caching allocator, a
 *     fixed arena): a cudaMalloc made behind the pool's back fails although the pool has room, and a plain C caller
 *     has no way to borrow from the pool;
 *   - a conflict-detection call inside a fixed-rate tick: cudaMalloc / cudaFree stall the whole device for a time
 *     nobody bounds, which a per-step budget cannot absorb;
 *   - a binding (the ctypes stub of INTEGRATION.md, a C plugin) that is handed the record and output buffers by its
 *     host and owns no device allocator of its own.
 * bsg_cd_sym_static is bsg_cd_sym_init + bsg_cd_sym_sweep + bsg_cd_sym_finalize in one call, with the accumulators in
 * a device array compiled into the library.  Inputs and outputs are bsg_cd_sweep's (all rows; row0 / row1 are ignored;
 * any output pointer may be NULL) and the results are bit-identical to it.  The call itself allocates nothing: no
 * cudaMalloc, no cudaFree, no allocating library routine.
 *   Limit: num_envs * num_ac <= BSG_CD_STATIC_MAX_ROWS (262 144 rows; the giant-airspace config is 160 000).  The array
 * costs 5.25 MB on EVERY GPU that loads the library, from load until the process exits, used or not.  A larger problem
 * is refused with BSG_ERR_TOOLARGE before anything is launched -- use the three-stage calls with your own buffers.
 *   First call on a device: it looks the array's device address up (cudaGetSymbolAddress), once per device, and that
 * is also the moment the CUDA runtime loads the library's code onto that device.  Make that first call during your
 * start-up, outside any stream capture; later calls only launch kernels on `stream`.
 *   The device that is current when you call is the one whose array is used: rec_dev and the outputs must live there
 * (a record pointer of another device is refused with BSG_ERR_ARG).
 *   One array per GPU, shared by every call in the process: it still holds the previous call's values when the next one
 * starts (each call re-initialises the rows it reads, you never see stale data), and two calls whose kernels overlap
 * on the GPU -- different streams, different host threads -- race on it and corrupt each other's results.  Keep the
 * calls of one GPU on one stream, or order them with events. */
#define BSG_CD_STATIC_MAX_ROWS (1 << 18)
int bsg_cd_sym_static(const bsg_cd_params *p, const float *rec_dev, uint8_t *inconf_dev, int32_t *nconf_dev,
                      int32_t *nlos_dev, int32_t *near_idx_dev, float *near_tcpa_dev, float *near_dcpa_dev,
                      float *near_tin_dev, float *tcpamax_dev, void *stream);

/* The same symmetric sweep over PEER memory (one process per GPU, NVLink): aircraft j of every env lives on rank
 * j / rows_per_rank; rec / nconf / nlos / tcpamax / key [g] are device pointers to rank g's [E][rows_per_rank] buffers
 * as mapped into THIS process (symmetric-memory peer mappings; the local ones for g == own rank).  The sweep reads
 * the records from their owners and flushes the (rare) conflict updates to the owners' outputs with system-scope
 * atomics: no all-gather and no all-reduce, only a cross-GPU barrier before (records packed, outputs initialised
 * with bsg_cd_sym_init on the own block) and after the sweep.  finalize writes the outputs of the OWN rows
 * [E][rows_per_rank]. */
#define BSG_MAX_PEERS 8
typedef struct bsg_cd_peers {
    int32_t num_ranks, rows_per_rank;
    const float *rec[BSG_MAX_PEERS];
    int32_t *nconf[BSG_MAX_PEERS], *nlos[BSG_MAX_PEERS];
    float *tcpamax[BSG_MAX_PEERS];
    int64_t *key[BSG_MAX_PEERS];
} bsg_cd_peers;
int bsg_cd_sym_sweep_peers(const bsg_cd_params *p, const bsg_cd_peers *peers, int32_t tile_rank, int32_t tile_stride,
                           void *stream);
int bsg_cd_sym_finalize_peers(const bsg_cd_params *p, const bsg_cd_peers *peers, int32_t rank, uint8_t *inconf_dev,
                              int32_t *near_idx_dev, float *near_tcpa_dev, float *near_dcpa_dev, float *near_tin_dev,
                              void *stream);

/* ND-map (4D spatial-temporal conflict grid over straight-line look-ahead trajectories; builder-owned spec):
 * rec_dev float [E][6][n] (x, y, z [m], u, v, w [m/s]); occ uint8 [E][nt][nz][ny][nx] = min(count, 255);
 * conf_bits = 1 bit per cell, [E][nt][cells / 8] bytes, cell c of a slab <-> bit (c & 7) of byte (c >> 3) (c = (z ny + y) nx
 * + x): set where the 3x3x3 neighbourhood of the cell, clipped at the faces, holds >= 2 aircraft; ncell int32 [E][nt] =
 * number of set bits.  nx in {32, 64, 128}; nx ny nz a multiple of 128 and <= 77 K cells (one slab lives in shared
 * memory).  bsg_ndmap_unpack expands the bits to one 0/1 byte per cell. */
int bsg_ndmap(const bsg_ndmap_params *p, const float *rec_dev, uint8_t *occ_dev, uint8_t *conf_bits_dev,
              int32_t *ncell_dev, void *stream);
int bsg_ndmap_unpack(const bsg_ndmap_params *p, const uint8_t *conf_bits_dev, uint8_t *conf_bytes_dev, void *stream);
/* The same map over the LIVE state of a handle (p->num_envs / num_ac = the handle's): the records are the aircraft's
 * current position, flat-earth offsets from (lat0, lon0) scaled by the caller's metres per degree, altitude, and the
 * ground-speed vector / vertical speed; slots beyond an env's live aircraft are dropped.  rec_out_dev (may be NULL)
 * receives the packed records [E][6][n] the map was built from. */
typedef struct bsg_ndmap_geo { double lat0, lon0, m_per_deg_lat, m_per_deg_lon; } bsg_ndmap_geo;
int bsg_ndmap_state(bsg_handle *h, const bsg_ndmap_params *p, const bsg_ndmap_geo *geo, uint8_t *occ_dev,
                    uint8_t *conf_bits_dev, int32_t *ncell_dev, float *rec_out_dev, void *stream);

/* FMA-chain micro-benchmarks: the measured FP32 / FP64 (non-tensor) roofline denominators */
int bsg_peak_fp32(int device, int iters, double *tflops, double *ms);
int bsg_peak_fp64(int device, int iters, double *tflops, double *ms);

#ifdef __cplusplus
}
#endif
#endif
