// bsg_reset.cuh -- device-side scenario generation shared by the stand-alone reset kernel (bsg_reset) and the
// fused next-step auto-reset inside step_kernel.  Restates the reset path of the reference env layer:
//   Horizontal  bluesky_gym/envs/horizontal_cr_env.py:82-148  (bs.traf.cre + creconfs x K + waypoint)
//   Vertical    bluesky_gym/envs/vertical_cr_env.py:183-200,258-281
//   Sector      bluesky_gym/envs/sector_cr_env.py:87-217      (polygon growth, density draw, perimeter
//               waypoints, bounding-box rejection sampling, get_hdg)
//   Merge       bluesky_gym/envs/merge_env.py:100-158         (spawn radials towards the FAF, LNAV route for intruders)
// Same value laws as the reference; the random stream is Philox4x32-10 keyed by (seed) with counter
// (env, episode, draw index, stream) instead of numpy's global legacy generator (the global
// draw order cannot be reproduced on a GPU; parity runs inject oracle state through bsg_set_state).
// Stages per env: reset_env_level (one thread: env-level draws, polygon into the ev slots), for Sector the sorted
// perimeter waypoints (reset_wpt_draw / reset_wpt_rank, one thread per aircraft, a barrier between them), and
// reset_aircraft (one thread per aircraft).  All are __noinline__: rare, register-hungry fp64 code that
// must not inflate the register allocation of the step kernel's hot loop.
#pragma once
#include "bsg_step.cuh"

namespace bsg {

struct Philox {
    uint32_t k0, k1, c0, c1, c2, n;     // n = draw-block counter
    uint32_t buf[4];
    int have;
    __device__ Philox(uint64_t seed, uint32_t env, uint32_t episode, uint32_t stream)
        : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)), c0(env), c1(episode), c2(stream), n(0), have(0) {}
    __device__ void refill() {
        uint32_t a = c0, b = c1, c = c2, d = n++, x = k0, y = k1;
#pragma unroll
        for (int r = 0; r < 10; r++) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, a), lo0 = 0xD2511F53u * a;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c), lo1 = 0xCD9E8D57u * c;
            a = hi1 ^ b ^ x; b = lo1; c = hi0 ^ d ^ y; d = lo0;
            x += 0x9E3779B9u; y += 0xBB67AE85u;
        }
        buf[0] = a; buf[1] = b; buf[2] = c; buf[3] = d;
        have = 4;
    }
    __device__ uint32_t u32() { if (!have) refill(); return buf[--have]; }
    __device__ double uniform() {                      // [0, 1) with 53 bits
        const uint64_t hi = u32(), lo = u32();
        return (double)(((hi << 32) | lo) >> 11) * (1.0 / 9007199254740992.0);
    }
    __device__ int randint(int lo, int hi) { return lo + (int)(uniform() * (double)(hi - lo)); }   // [lo, hi)
    __device__ double normal(double mu, double sigma) {
        const double u1 = 1.0 - uniform(), u2 = uniform();
        return mu + sigma * sqrt(-2.0 * log(u1)) * cos(2.0 * kPi * u2);
    }
};

struct AcInit { double lat, lon, alt, hdg, cas, tas, vs, selalt; int lnav; };

// Traffic.creconfs geometry for one intruder of a target at (lat0, lon0), track trk0, speed gs0.
// Same construction with the inverse trigonometry folded away: the bearing enters only through its sine and
// cosine, which are the normalised components (hypot = |v_rel| because (rd, rx) is a unit vector), and
// atan2(gse, gsn) of a speed vector built from trk is trk itself wrapped to (-180, 180].
__device__ inline void creconfs(double lat0, double lon0, double coslat0, double trk0_deg, double gs0, double dpsi,
                                double cpa_nm, double tlosh, double &lat, double &lon, double &hdg) {
    const double trkref = trk0_deg * kD2R, cpa = cpa_nm * kNm, pzr = 5.0 * kNm;
    double st0, ct0, st, ct;
    sincos(trkref, &st0, &ct0);
    sincos(trkref + dpsi * kD2R, &st, &ct);
    const double vreln = gs0 * (ct0 - ct), vrele = gs0 * (st0 - st);
    const double vrel = sqrt(vreln * vreln + vrele * vrele);
    const double drelcpa = tlosh * vrel + (cpa > pzr ? 0.0 : sqrt(pzr * pzr - cpa * cpa));
    const double dist = sqrt(drelcpa * drelcpa + cpa * cpa);
    const double rd = drelcpa / dist, rx = cpa / dist;
    const double k = (dist / kNm) / vrel;                     // dnm * (sin, cos)(brn) = k * (y, x)
    const double dx = k * (-rx * vreln + rd * vrele), dy = k * (rd * vreln + rx * vrele);
    lat = lat0 + dy / 60.0;
    double lo = lon0 + dx / fmax(0.01, 60.0 * coslat0);
    lo = lo >= 180.0 ? lo - 360.0 : lo;
    lon = lo < -180.0 ? lo + 360.0 : lo;
    double h = trk0_deg + dpsi;
    h -= 360.0 * floor((h + 180.0) * (1.0 / 360.0));          // (-180, 180] like degrees(arctan2(gse, gsn))
    hdg = h <= -180.0 ? h + 360.0 : h;
}

template <typename R>
__device__ inline void write_ac(double *pos, R *st, size_t EN, size_t k, const AcInit &a) {
    pos[k] = a.lat; pos[EN + k] = a.lon;
    for (int f = 0; f < S_NS; f++) st[(size_t)f * EN + k] = R(0);
    st[S_ALT * EN + k] = (R)a.alt; st[S_HDG * EN + k] = (R)a.hdg; st[S_TAS * EN + k] = (R)a.tas;
    st[S_VS * EN + k] = (R)a.vs; st[S_CAS * EN + k] = (R)a.cas; st[S_SELSPD * EN + k] = (R)a.cas;
    st[S_SELALT * EN + k] = (R)a.selalt; st[S_APTRK * EN + k] = (R)a.hdg; st[S_AX * EN + k] = (R)kKts;
    double sh, ch;                                               // Traffic.cre: gsnorth / gseast from tas and hdg
    sincos(a.hdg * kD2R, &sh, &ch);
    st[S_GSN * EN + k] = (R)(a.tas * ch); st[S_GSE * EN + k] = (R)(a.tas * sh); st[S_LNAV * EN + k] = (R)a.lnav;
    for (int f = S_C_ALT; f < S_NS; f++) st[(size_t)f * EN + k] = (R)NAN;
}

struct ResetEnv {                 // env-level results the per-aircraft stage needs
    double alt0, target, gs0, minx, miny, maxx, maxy, perim;
    int nv, nac;
};

__device__ inline bool inside_poly_ll(const double *poly, int nv, double tx, double ty) {
    bool inside = false;
    double x0 = poly[2 * (nv - 1)], y0 = poly[2 * (nv - 1) + 1];
    bool f0 = y0 >= ty;
    for (int k = 0; k < nv; k++) {
        const double x1 = poly[2 * k], y1 = poly[2 * k + 1];
        const bool f1 = y1 >= ty;
        if (f0 != f1)
            if ((((y1 - ty) * (x0 - x1)) >= ((x1 - tx) * (y0 - y1))) == f1) inside = !inside;
        f0 = f1; x0 = x1; y0 = y1;
    }
    return inside;
}

// stage 1 (one thread per env): clears the scalar ev slots, draws the env-level quantities, grows the polygon
static __device__ __noinline__ void reset_env_level(int kind, int fixed_nac, int density_mode, int N, int e, uint64_t seed, uint32_t ep,
                                                    double *v, ResetEnv *Sp, const ResetConst &RC) {
    ResetEnv S = {};
    Philox g(seed, (uint32_t)e, ep, 0u);
    const double ccl = RC.ccl;
    for (int s = 0; s < BSG_EV_POLY; s++) v[s] = 0.0;
    S.nac = N;
    if (kind == BSG_KIND_HORIZONTAL) {
        S.gs0 = RC.tas_h;
        // functions.py:24-42 at bearing 0: asin(sin(lat1) cos d + cos(lat1) sin d) = lat1 + d
        const double d = (double)g.randint(100, 150) / 6371.0, lat1 = 52.0 * kD2R;
        v[BSG_EV_WPT_LAT] = (lat1 + d) * kR2D;
        v[BSG_EV_WPT_LON] = 4.0;
    } else if (kind == BSG_KIND_VERTICAL) {
        S.alt0 = (double)g.randint(2000, 4000);
        S.target = S.alt0 + (double)g.randint(-500, 500);
        S.gs0 = vcas2tas(150.0, S.alt0);
        v[BSG_EV_TARGET_ALT] = S.target;
    } else if (kind == BSG_KIND_SECTOR) {
        // random points on the circle of area 3750 NM^2, kept sorted by polar angle, until the polygon they span
        // covers >= 2400 NM^2 (sector_cr_env.py:141-160, functions.py:44-114).  Vertex coordinates are computed
        // once per draw and carried through the insertion; the shoelace sum then needs no trigonometry.
        const double Rr = sqrt(3750.0 / kPi);
        double ang[BSG_MAXV], px[BSG_MAXV], py[BSG_MAXV];
        int nv = 0;
        double area = 0.0;
        for (;;) {
            const double al = 2 * kPi * g.uniform();
            double sa, ca;
            sincos(al, &sa, &ca);
            const double t = al > kPi ? al - 2 * kPi : al;         // atan2(sin, cos): polar angle in (-pi, pi]
            int b = nv - 1;
            while (b >= 0 && ang[b] > t) { ang[b + 1] = ang[b]; px[b + 1] = px[b]; py[b + 1] = py[b]; b--; }
            ang[b + 1] = t; px[b + 1] = Rr * ca; py[b + 1] = Rr * sa;
            nv++;
            if (nv < 3) continue;
            area = 0.0;
            for (int k = 0; k < nv; k++) {
                const int m = k + 1 == nv ? 0 : k + 1;
                area += px[k] * py[m] - py[k] * px[m];
            }
            area = fabs(area) / 2.0;
            if (area >= 2400.0 || nv >= BSG_MAXV) break;
        }
        S.nv = nv;
        S.gs0 = RC.tas_s;
        // sector_cr_env.py:98-103: ac_density_mode "normal" N(0.005, 0.001) or "uniform" U(0.003, 0.007) aircraft per NM^2
        const double dens = density_mode ? 0.003 + 0.004 * g.uniform() : g.normal(0.005, 0.001);
        const int na = (int)fmax(ceil(dens * area), 5.0);
        S.nac = (fixed_nac || na > N) ? N : na;
        S.minx = S.maxx = px[0]; S.miny = S.maxy = py[0];
        S.perim = 0.0;
        for (int k = 0; k < nv; k++) {
            const int m = k + 1 == nv ? 0 : k + 1;
            const double dx = px[m] - px[k], dy = py[m] - py[k];
            S.perim += sqrt(dx * dx + dy * dy);
            S.minx = fmin(S.minx, px[k]); S.maxx = fmax(S.maxx, px[k]);
            S.miny = fmin(S.miny, py[k]); S.maxy = fmax(S.maxy, py[k]);
            v[BSG_EV_POLY + 2 * k] = kCenterLat + px[k] / 60;
            v[BSG_EV_POLY + 2 * k + 1] = kCenterLon + py[k] / (60 * ccl);
        }
        v[BSG_EV_NVERT] = nv;
    } else if (kind == BSG_KIND_MERGE) {      // the route every intruder gets: FAF, then the runway (merge_env.py:42-45)
        S.gs0 = RC.tas_m;
        v[BSG_EV_WPT_LAT] = RC.faf_lat; v[BSG_EV_WPT_LON] = RC.faf_lon;
        v[BSG_EV_WPT_X] = RC.rwy_lat; v[BSG_EV_WPT_Y] = RC.rwy_lon;
    }
    *Sp = S;
}

// Sector waypoints (sector_cr_env.py:162-176): num_ac perimeter distances U(0, perimeter), SORTED, and aircraft i is
// handed the i-th smallest -- so the agent (i = 0) always flies to the first waypoint past polygon vertex 0.
// Two stages with a barrier between them: every aircraft draws one distance (its own Philox stream), then finds
// the rank of its draw among the env's (ties by index) and files it under that rank; aircraft i reads wsort[i].
static __device__ __noinline__ void reset_wpt_draw(int e, int i, uint64_t seed, uint32_t ep, const ResetEnv *Sp, double *wdraw) {
    if (i >= Sp->nac) return;
    Philox g(seed, (uint32_t)e, ep, 0x40000000u + (uint32_t)i);
    wdraw[i] = g.uniform() * Sp->perim;
}
static __device__ __noinline__ void reset_wpt_rank(int i, const ResetEnv *Sp, const double *wdraw, double *wsort) {
    const int n = Sp->nac;
    if (i >= n) return;
    const double mine = wdraw[i];
    int r = 0;
    for (int m = 0; m < n; m++) { const double o = wdraw[m]; r += (o < mine || (o == mine && m < i)) ? 1 : 0; }
    wsort[r] = mine;
}

// last stage (one thread per aircraft slot i of env e): creates the aircraft in the device state
template <typename R>
static __device__ __noinline__ void reset_aircraft(int kind, int E, int N, int e, int i, uint64_t seed, uint32_t ep,
                                                   const ResetEnv *Sp, const double *wsort, double *v, double *pos, R *st,
                                                   unsigned long long *reso, const ResetConst &RC) {
    const ResetEnv S = *Sp;
    const size_t EN = (size_t)E * N, k = (size_t)e * N + i;
    const double ccl = RC.ccl;
    for (int w = 0; w < (N + 63) / 64; w++) reso[(size_t)w * EN + k] = 0ull;       // resopairs row, all its words
    AcInit a = {};
    a.lnav = -1;
    Philox g(seed, (uint32_t)e, ep, 1u + (uint32_t)i);
    if (kind == BSG_KIND_HORIZONTAL || kind == BSG_KIND_VERTICAL) {
        const double alt_own = kind == BSG_KIND_VERTICAL ? S.alt0 : 0.0;
        if (i == 0) {
            a.lat = 52.0; a.lon = 4.0; a.alt = alt_own; a.hdg = 0.0; a.cas = 150.0; a.tas = S.gs0; a.selalt = alt_own;
        } else {
            const double dpsi = g.randint(45, 315), cpa = g.randint(0, 5);
            double tlosh, dH = 0.0;
            if (kind == BSG_KIND_HORIZONTAL) tlosh = g.randint(100, 1000);
            else {
                tlosh = g.randint(100, (int)(180000.0 / S.gs0));
                const double avg_tod = 200000.0 / S.gs0 - 2 * S.target / 12.5;
                if (tlosh > avg_tod) dH = g.randint((int)(-alt_own + 500), (int)((S.target - alt_own) + 100));
                else dH = g.randint((int)((S.target - alt_own) - 500), (int)((S.target - alt_own) + 500));
            }
            creconfs(52.0, 4.0, RC.coslat_hv, 0.0, S.gs0, dpsi, cpa, tlosh, a.lat, a.lon, a.hdg);
            a.alt = alt_own + dH;
            // creconfs hands cre() the CAS of the target's ground speed at the intruder's altitude; the TAS that
            // cre() derives back from it is that ground speed again (vcas2tas o vtas2cas = id)
            a.cas = kind == BSG_KIND_HORIZONTAL ? RC.cas_h : vtas2cas(S.gs0, a.alt);
            a.tas = S.gs0;
            a.selalt = a.alt;
        }
    } else if (kind == BSG_KIND_SECTOR && i < S.nac) {
        // perimeter waypoint: the i-th smallest of the env's draws (reset_wpt_rank).  Vertices come back from the ev slots.
        const double *poly = v + BSG_EV_POLY;
        double d = wsort[i], wx = 0.0, wy = 0.0;
        for (int s2 = 0; s2 < S.nv; s2++) {
            const int m = (s2 + 1) % S.nv;
            const double x0 = (poly[2 * s2] - kCenterLat) * 60, y0 = (poly[2 * s2 + 1] - kCenterLon) * (60 * ccl);
            const double x1 = (poly[2 * m] - kCenterLat) * 60, y1 = (poly[2 * m + 1] - kCenterLon) * (60 * ccl);
            const double len = sqrt((x1 - x0) * (x1 - x0) + (y1 - y0) * (y1 - y0));
            if (d <= len || s2 == S.nv - 1) { const double fr = fmin(d / len, 1.0); wx = x0 + fr * (x1 - x0); wy = y0 + fr * (y1 - y0); break; }
            d -= len;
        }
        double x, y;
        do {
            x = S.minx + g.uniform() * (S.maxx - S.minx);
            y = S.miny + g.uniform() * (S.maxy - S.miny);
        } while (!inside_poly_ll(poly, S.nv, kCenterLat + x / 60, kCenterLon + y / (60 * ccl)));
        a.lat = kCenterLat + x / 60; a.lon = kCenterLon + y / (60 * ccl);
        const double wlat = (kCenterLat + wx / 60) * kD2R, wlon = (kCenterLon + wy / (60 * ccl)) * kD2R;
        const double l1 = a.lat * kD2R, o1 = a.lon * kD2R;                        // functions.get_hdg
        const double hx = sin(wlon - o1) * cos(wlat);
        const double hy = cos(l1) * sin(wlat) - sin(l1) * cos(wlat) * cos(wlon - o1);
        a.hdg = pymod(atan2(hx, hy) * kR2D + 360, 360);
        a.alt = 350.0; a.cas = 150.0; a.tas = S.gs0; a.selalt = 350.0;
        if (i == 0) { v[BSG_EV_WPT_X] = wx; v[BSG_EV_WPT_Y] = wy; }
    } else if (kind == BSG_KIND_MERGE) {
        // merge_env.py:110-115,148-156: a radial within +-15 deg of north from the FAF, 50-200 km out for the agent and
        // 20-500 km for the intruders (functions.get_point_at_distance, km on the 6371 km sphere), heading back along it
        const double brg = -15.0 + 30.0 * g.uniform();
        const double dkm = i == 0 ? 50.0 + 150.0 * g.uniform() : 20.0 + 480.0 * g.uniform();
        const double lat1 = RC.faf_lat * kD2R, lon1 = RC.faf_lon * kD2R, al = brg * kD2R, dr = dkm / 6371.0;
        const double lat2 = asin(sin(lat1) * cos(dr) + cos(lat1) * sin(dr) * cos(al));
        const double lon2 = lon1 + atan2(sin(al) * sin(dr) * cos(lat1), cos(dr) - sin(lat1) * sin(lat2));
        a.lat = lat2 * kR2D; a.lon = lon2 * kR2D;
        a.hdg = brg - 180.0;                                       // handed to cre() as is (negative)
        a.alt = 10000.0; a.cas = 100.0; a.tas = S.gs0; a.selalt = 10000.0;
        a.lnav = i == 0 ? -1 : 0;                                  // ADDWPT activates LNAV towards the FAF
    }
    write_ac<R>(pos, st, EN, k, a);
}

}  // namespace bsg
