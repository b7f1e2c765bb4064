// bsg_env.cuh -- observation / reward / termination / info assembly, the epilogue of the fused step.
// Restates the reference env layer for n-1 intruders (reference: 5):
//   Horizontal _get_obs/_get_reward/_check_*  bluesky_gym/envs/horizontal_cr_env.py:150-270
//   Vertical   _get_obs/_get_reward           bluesky_gym/envs/vertical_cr_env.py:113-240
//   Sector     _get_observation/_get_reward/_check_inside_airspace  bluesky_gym/envs/sector_cr_env.py:134-139,227-339
//   Merge      _get_obs/_get_reward/_check_*  bluesky_gym/envs/merge_env.py:160-283            (EXT kernels)
//   wind_u/v   WindFieldWrapper._get_wind_observation  bluesky_gym/wrappers/wind.py:55-64       (EXT kernels)
//   noise      NoisyObservationWrapper.add_noise  bluesky_gym/wrappers/uncertainty.py:19-32     (EXT kernels)
// Flat obs layouts (documented in DESIGN.md) follow the key order of the reference's Dict spaces.
#pragma once
#include "bsg_step.cuh"
#include "bsg_reset.cuh"

namespace bsg {

__host__ __device__ constexpr int env_obs_dim(int kind, int N) {
    return kind == BSG_KIND_HORIZONTAL ? 5 * (N - 1) + 3 : kind == BSG_KIND_VERTICAL ? 4 + 7 * (N - 1)
           : kind == BSG_KIND_SECTOR ? 31 : kind == BSG_KIND_MERGE ? 40 : 0;
}

// x / c for a literal constant c: the literal division in the fp64 parity path, a multiplication by the rounded reciprocal
// in the fp32 fast path (one instruction instead of the IEEE division sequence; <= 1 ulp apart)
template <typename R> __device__ __forceinline__ R divc(R x, double c) {
    if constexpr (sizeof(R) == 8) return x / R(c); else return x * float(1.0 / c);
}

template <typename R, bool EXT = false>
__device__ __forceinline__ void env_layer(const StepArgs<R> &A, const AcReg<R> &a, EnvShared<R> *es, R *sdist,
                                          int le, int i, int e, int nl, bool valid, bool live, bool skip, double *ev) {
    const int N = A.N, K = N - 1, kind = A.kind;
    double *sev = es[valid ? le : 0].ev;                               // staged scalar slots (shared memory)
    if (kind == BSG_KIND_RAW) return;                                  // uniform: no barrier skipped
    R *obs = A.out.obs ? (R *)A.out.obs : nullptr;
    const int od0 = env_obs_dim(kind, N);
    bool wind = false;
    if constexpr (EXT) wind = A.W.n > 0;
    const int od = od0 + ((EXT && A.W.obs) ? 2 : 0);
    if (obs) obs += (size_t)(valid ? e : 0) * od;
    // bs.traf.gs: the ground speed under wind (Horizontal / Vertical speed-difference slots)
    R my_gs = a.tas;
    if (wind && (kind == BSG_KIND_HORIZONTAL || kind == BSG_KIND_VERTICAL)) my_gs = sqrt(a.gsn * a.gsn + a.gse * a.gse);
    if (live && i == 0) {
        es[le].own_lat = a.lat; es[le].own_lon = a.lon; es[le].own_hdg = a.hdg; es[le].own_tas = my_gs;
        es[le].own_alt = a.alt; es[le].own_vs = a.vs; es[le].intr = 0;
    }
    __syncthreads();
    R px = R(0), py = R(0);                                            // sector: own-relative plane coords [m]
    if (valid) {
        const double olat = es[le].own_lat, olon = es[le].own_lon;
        const R ohdg = es[le].own_hdg, otas = es[le].own_tas;
        if (kind == BSG_KIND_HORIZONTAL || kind == BSG_KIND_VERTICAL) {
            if (i >= 1) {
                const int kk = i - 1;
                R o_dist = R(0), o_c = R(0), o_s = R(0), o_x = R(0), o_y = R(0), o_da = R(0), o_z = R(0);
                if (live) {
                    R qdr, d;
                    env_qdrdist<R>(olat, olon, a.lat, a.lon, qdr, d);
                    o_dist = kind == BSG_KIND_HORIZONTAL ? divc<R>(d * R(1.852), 150) : divc<R>(d * R(1.852), 200);
                    rsincos_deg<R>(bound180<R>(ohdg - qdr), o_s, o_c);
                    R sh, ch;
                    rsincos_deg<R>(ohdg - a.hdg, sh, ch);
                    o_x = divc<R>(-ch * my_gs, 150);
                    o_y = divc<R>(otas - sh * my_gs, 150);
                    bool hit = d < R(5);
                    if (kind == BSG_KIND_VERTICAL) {
                        o_da = divc<R>(a.alt - es[le].own_alt, 3000);
                        o_z = a.vs - es[le].own_vs;
                        hit = hit && fabs(es[le].own_alt - a.alt) < R(1000 * 0.3048);
                    }
                    if (hit) atomicAdd(&es[le].intr, 1);
                }
                if (obs) {
                    R *b = obs + (kind == BSG_KIND_VERTICAL ? 4 : 0);
                    b[kk] = o_dist; b[K + kk] = o_c; b[2 * K + kk] = o_s;
                    if (kind == BSG_KIND_HORIZONTAL) { b[3 * K + kk] = o_x; b[4 * K + kk] = o_y; }
                    else { b[3 * K + kk] = o_da; b[4 * K + kk] = o_x; b[5 * K + kk] = o_y; b[6 * K + kk] = o_z; }
                }
            } else if (live) {
                if (kind == BSG_KIND_HORIZONTAL) {
                    R qdr, d;
                    env_qdrdist<R>(olat, olon, sev[BSG_EV_WPT_LAT], sev[BSG_EV_WPT_LON], qdr, d);
                    R drift = bound180<R>(ohdg - qdr), sd, cd;
                    rsincos_deg<R>(drift, sd, cd);
                    es[le].drift = drift; es[le].wdist = d * R(1.852);
                    sev[BSG_EV_AC_HDG] = (double)a.hdg;
                    if (obs) { obs[5 * K] = divc<R>(d * R(1.852), 150); obs[5 * K + 1] = cd; obs[5 * K + 2] = sd; }
                } else {
                    R qdr, d;
                    env_qdrdist<R>(52.0, 4.0, olat, olon, qdr, d);
                    es[le].wdist = R(200) - d * R(1.852);              // runway distance
                    if (obs) {
                        obs[0] = divc<R>(a.alt - R(1500), 3000); obs[1] = divc<R>(a.vs - R(0), 5);
                        obs[2] = divc<R>(R(sev[BSG_EV_TARGET_ALT]) - R(1500), 3000);
                        obs[3] = divc<R>(es[le].wdist - R(100), 200);
                    }
                }
            }
        } else if (EXT && kind == BSG_KIND_MERGE) {   // -------------------------------------- Merge
            if (live && i >= 1) {                      // kwikdist_matrix row of the agent (merge_env.py:195), [NM]
                env_qdrdist<R>(olat, olon, a.lat, a.lon, px, py);        // px = bearing [deg], py = distance [NM]
                sdist[le * N + i] = py;
                if (py < R(4)) atomicAdd(&es[le].intr, 1);               // merge_env.py:275-283
            } else if (live) {
                const bool pre = sev[BSG_EV_WPT_REACH] == 0.0;           // drift to the FAF, then to the runway
                R qdr, d;
                env_qdrdist<R>(olat, olon, pre ? sev[BSG_EV_WPT_LAT] : sev[BSG_EV_WPT_X],
                               pre ? sev[BSG_EV_WPT_LON] : sev[BSG_EV_WPT_Y], qdr, d);
                R drift = bound180<R>(ohdg - qdr), sd, cd;
                rsincos_deg<R>(drift, sd, cd);
                es[le].drift = drift; es[le].wdist = d;
                if (obs) {
                    obs[0] = cd; obs[1] = sd; obs[2] = a.tas; obs[3] = d / R(250); obs[4] = R(sev[BSG_EV_WPT_REACH]);
                    for (int s = (nl - 1 < 0 ? 0 : nl - 1); s < 5; s++)
                        for (int q = 0; q < 7; q++) obs[5 + q * 5 + s] = R(0);
                }
            }
        } else {   // ---------------------------------------------------------------- Sector
            const double ccl = cos(kCenterLat * kD2R);
            // latlong_to_nm(CENTER, .) * NM2KM * 1000 (functions.py:130-131), own-relative
            if constexpr (sizeof(R) == 8) {
                double x0 = (olat - kCenterLat) * 60 * 1.852 * 1000, y0 = (olon - kCenterLon) * 60 * ccl * 1.852 * 1000;
                double x = (a.lat - kCenterLat) * 60 * 1.852 * 1000, y = (a.lon - kCenterLon) * 60 * ccl * 1.852 * 1000;
                px = x - x0; py = y - y0;
            } else {
                px = (float)((a.lat - olat) * (60 * 1.852 * 1000)); py = (float)((a.lon - olon) * (60 * ccl * 1.852 * 1000));
            }
            if (live && i >= 1) {
                sdist[le * N + i] = sqrt(px * px + py * py);
                R qdr, d;
                env_qdrdist<R>(olat, olon, a.lat, a.lon, qdr, d);
                if (d < R(5)) atomicAdd(&es[le].intr, 1);
            } else if (live) {
                const double wlat = kCenterLat + sev[BSG_EV_WPT_X] / 60, wlon = kCenterLon + sev[BSG_EV_WPT_Y] / (60 * ccl);
                R qdr, d;
                env_qdrdist<R>(olat, olon, wlat, wlon, qdr, d);
                R drift = bound180<R>(ohdg - qdr), sd, cd;
                rsincos_deg<R>(drift, sd, cd);
                es[le].drift = drift;
                if (obs) {
                    obs[0] = cd; obs[1] = sd; obs[2] = divc<R>(a.tas - R(150), 6);
                    for (int s = (nl - 1 < 0 ? 0 : nl - 1); s < 4; s++)
                        for (int q = 0; q < 7; q++) obs[3 + q * 4 + s] = R(0);
                }
            }
        }
    }
    __syncthreads();
    if (valid && kind == BSG_KIND_SECTOR && live && i >= 1) {          // the 4 nearest by (distance, index)
        const R mine = sdist[le * N + i];
        int rank = 0;
        for (int m = 1; m < nl; m++) {
            const R dm = sdist[le * N + m];
            rank += (dm < mine || (dm == mine && m < i)) ? 1 : 0;
        }
        if (rank < 4 && obs) {
            R sh, ch, so, co;
            rsincos_deg<R>(a.hdg, sh, ch);
            rsincos_deg<R>(es[le].own_hdg, so, co);
            const R vx = co * es[le].own_tas, vy = so * es[le].own_tas;
            const R vxi = ch * a.tas, vyi = sh * a.tas;
            const R trk = atan2(vyi - vy, vxi - vx);
            R *o = obs + 3;
            o[0 * 4 + rank] = divc<R>(px, 13000); o[1 * 4 + rank] = divc<R>(py, 13000);
            o[2 * 4 + rank] = divc<R>(vxi - vx, 32); o[3 * 4 + rank] = divc<R>(vyi - vy, 66);
            o[4 * 4 + rank] = cos(trk); o[5 * 4 + rank] = sin(trk);
            o[6 * 4 + rank] = divc<R>(mine - R(50000), 15000);
        }
    }
    if (EXT && valid && kind == BSG_KIND_MERGE && live && i >= 1) {    // the 5 nearest by (distance, index)
        const R mine = sdist[le * N + i];
        int rank = 0;
        for (int m = 1; m < nl; m++) {
            const R dm = sdist[le * N + m];
            rank += (dm < mine || (dm == mine && m < i)) ? 1 : 0;
        }
        if (rank < 5 && obs) {
            R sh, ch, so, co, sb, cb;
            rsincos_deg<R>(a.hdg, sh, ch);
            rsincos_deg<R>(es[le].own_hdg, so, co);
            rsincos_deg<R>(px, sb, cb);
            const R vx = co * es[le].own_tas, vy = so * es[le].own_tas;
            const R vxi = ch * a.tas, vyi = sh * a.tas;
            const R trk = atan2(vyi - vy, vxi - vx);
            R *o = obs + 5;
            o[0 * 5 + rank] = (py * R(1.852) * R(1000)) * cb / R(1000000);
            o[1 * 5 + rank] = (py * R(1.852) * R(1000)) * sb / R(1000000);
            o[2 * 5 + rank] = (vxi - vx) / R(150); o[3 * 5 + rank] = (vyi - vy) / R(150);
            o[4 * 5 + rank] = cos(trk); o[5 * 5 + rank] = sin(trk);
            o[6 * 5 + rank] = mine / R(250);
        }
    }
    if constexpr (EXT) {
        if (A.W.obs && live && i == 0 && obs) {        // wrappers/wind.py:55-64
            R wn, we, sh, ch;
            wind_at<R>(A.W, a.lat, a.lon, 0.0, wn, we);          // (the wrapper's scalar altitude: the field at altitude 0)
            rsincos_deg<R>(a.hdg, sh, ch);
            obs[od0] = (wn * ch + we * sh) / R(50);
            obs[od0 + 1] = (-wn * sh + we * ch) / R(50);
        }
        if (A.noise_sigma > R(0)) {                    // wrappers/uncertainty.py:19-32 (every slot, every step)
            __syncthreads();
            if (valid && obs) {
                for (int q = i; q < od; q += N) {
                    Philox g(A.noise_seed, (uint32_t)e, A.noise_ctr, 0x4E4F4953u + (uint32_t)q);
                    obs[q] += A.noise_sigma * (R)g.normal(0.0, 1.0);
                }
            }
        }
    }
    // ------------------------------------------------------------------ reward / termination / info (agent thread)
    if (live && i == 0 && skip && !A.observe_only) {                 // first observation of a fresh episode
        if (A.out.reward) ((R *)A.out.reward)[e] = R(0);
        if (A.out.terminated) A.out.terminated[e] = 0;
        if (A.out.truncated) A.out.truncated[e] = 0;
        if (A.out.done) A.out.done[e] = 0;
        if (A.done_next) A.done_next[e] = 0;
        if (A.out.info) {
            R *inf = (R *)A.out.info + (size_t)e * 4;
            inf[0] = R(0); inf[1] = R(0); inf[2] = kind == BSG_KIND_VERTICAL ? R(0) : R(NAN); inf[3] = R(0);
        }
    }
    if (live && i == 0 && !skip && !A.observe_only) {
        const int intr = es[le].intr;
        double rew = 0.0;
        int term = 0, trunc = 0;
        if (kind == BSG_KIND_HORIZONTAL) {
            double reach = 0.0;
            if (es[le].wdist < R(5) && sev[BSG_EV_WPT_REACH] != 1.0) { sev[BSG_EV_WPT_REACH] = 1.0; reach = 1.0; }
            const double dr = fabs((double)es[le].drift * kD2R);
            sev[BSG_EV_DRIFT_SUM] += dr; sev[BSG_EV_DRIFT_CNT] += 1.0;
            rew = reach + dr * -0.1 + (double)(-intr);
            term = sev[BSG_EV_WPT_REACH] == 1.0;
        } else if (kind == BSG_KIND_VERTICAL) {
            const double rwy = (double)es[le].wdist, alt = (double)a.alt;
            double pen = 0.0;
            if (rwy > 0 && alt > 0) pen = fabs(sev[BSG_EV_TARGET_ALT] - alt) * (-5.0 / 3000);
            else if (alt <= 0) { pen = -100; sev[BSG_EV_FINAL_ALT] = -100; term = 1; }
            else if (rwy <= 0) { pen = alt * (-50.0 / 3000); sev[BSG_EV_FINAL_ALT] = alt; term = 1; }
            rew = pen + (double)(-50 * intr);
        } else if (EXT && kind == BSG_KIND_MERGE) {   // merge_env.py:247-283
            double reach = 0.0;
            if (es[le].wdist < R(10) && sev[BSG_EV_WPT_REACH] != 1.0) {
                sev[BSG_EV_WPT_REACH] = 1.0; sev[BSG_EV_FINAL_ALT] = 1.0; reach = 1.0;
            } else if (es[le].wdist < R(20) && sev[BSG_EV_WPT_REACH] == 1.0) { sev[BSG_EV_FINAL_ALT] = 2.0; term = 1; }
            const double dr = fabs((double)es[le].drift * kD2R);
            sev[BSG_EV_DRIFT_SUM] += dr; sev[BSG_EV_DRIFT_CNT] += 1.0;
            rew = reach + dr * -0.1 + (double)(-intr);
        } else {
            const double dr = fabs((double)es[le].drift * kD2R);
            sev[BSG_EV_DRIFT_SUM] += dr; sev[BSG_EV_DRIFT_CNT] += 1.0;
            rew = dr * -0.1 + (double)(-intr);
            trunc = poly_inside(ev + BSG_EV_POLY, (int)sev[BSG_EV_NVERT], a.lat, a.lon) ? 0 : 1;
        }
        sev[BSG_EV_TOTAL_INTR] += (double)intr;
        sev[BSG_EV_TOTAL_REWARD] += rew;
        const double steps = sev[BSG_EV_STEPS] + 1.0;
        sev[BSG_EV_STEPS] = steps;
        if (A.max_steps > 0 && steps >= (double)A.max_steps) trunc = 1;
        if (A.out.reward) ((R *)A.out.reward)[e] = (R)rew;
        if (A.out.terminated) A.out.terminated[e] = term;
        if (A.out.truncated) A.out.truncated[e] = trunc;
        if (A.out.done) A.out.done[e] = term | trunc;
        if (A.done_next) A.done_next[e] = term | trunc;      // auto-reset: the next launch's `skip` flags
        if (A.out.info) {
            R *inf = (R *)A.out.info + (size_t)e * 4;
            inf[0] = (R)sev[BSG_EV_TOTAL_REWARD]; inf[1] = (R)sev[BSG_EV_TOTAL_INTR];
            inf[2] = kind == BSG_KIND_VERTICAL ? (R)sev[BSG_EV_FINAL_ALT] : (R)(sev[BSG_EV_DRIFT_SUM] / sev[BSG_EV_DRIFT_CNT]);
            inf[3] = kind == BSG_KIND_MERGE ? (R)sev[BSG_EV_FINAL_ALT] : (R)steps;
        }
    }
    if (live && i == 0) {                                              // write the staged slots back (fire and forget)
#pragma unroll
        for (int s = 0; s < BSG_EV_POLY; s++) ev[s] = sev[s];
    }
}

}  // namespace bsg
