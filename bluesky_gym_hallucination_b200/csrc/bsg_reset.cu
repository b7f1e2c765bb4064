// bsg_reset.cu -- device-side scenario generation: the reset path of the reference env layer, for the
// envs selected by a device mask (auto-reset without a host round trip).
//   Horizontal  bluesky_gym/envs/horizontal_cr_env.py:82-148  (bs.traf.cre + creconfs x K + waypoint)
//   Vertical    bluesky_gym/envs/vertical_cr_env.py:183-200,258-281
//   Sector      bluesky_gym/envs/sector_cr_env.py:87-217      (polygon growth, density draw, perimeter
//               waypoints, bounding-box rejection sampling, get_hdg)
// Same value laws as the reference; the random stream is Philox4x32-10 keyed by (seed) with counter
// (env, episode, draw index, stream) instead of numpy's global legacy generator (the
// global draw order cannot be reproduced on a GPU; parity runs inject oracle state through bsg_set_state).
// One CTA per env: thread 0 draws the env-level quantities into shared memory, the threads then create
// one aircraft each.
#include "bsg_handle.h"
#include "bsg_step.cuh"

using namespace bsg;

namespace {

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

struct AcInit { double lat, lon, alt, hdg, cas, tas, vs, selalt; };

// Traffic.creconfs geometry for one intruder of a target at (lat0, lon0), track trk0, speed gs0
__device__ void creconfs(double lat0, double lon0, double trk0_deg, double gs0, double dpsi, double cpa_nm, double tlosh,
                         double &lat, double &lon, double &hdg) {
    const double trkref = trk0_deg * kD2R, cpa = cpa_nm * kNm, pzr = 5.0 * kNm;
    const double trk = trkref + dpsi * kD2R;
    const double gsn = gs0 * cos(trk), gse = gs0 * sin(trk);
    const double vreln = gs0 * cos(trkref) - gsn, vrele = gs0 * sin(trkref) - gse;
    const double vrel = sqrt(vreln * vreln + vrele * vrele);
    const double drelcpa = tlosh * vrel + (cpa > pzr ? 0.0 : sqrt(pzr * pzr - cpa * cpa));
    const double dist = sqrt(drelcpa * drelcpa + cpa * cpa);
    const double rd = drelcpa / dist, rx = cpa / dist;
    const double brn = atan2(-rx * vreln + rd * vrele, rd * vreln + rx * vrele) * kR2D;
    const double dnm = dist / kNm;
    const double dx = dnm * sin(brn * kD2R), dy = dnm * cos(brn * kD2R);
    lat = lat0 + dy / 60.0;
    lon = pymod((lon0 + dx / fmax(0.01, 60.0 * cos(lat0 * kD2R))) + 180, 360) - 180;
    hdg = atan2(gse, gsn) * kR2D;
}

template <typename R>
__device__ void write_ac(double *pos, R *st, size_t EN, size_t k, const AcInit &a) {
    pos[k] = a.lat; pos[EN + k] = a.lon;
    for (int f = 0; f < S_NS; f++) st[(size_t)f * EN + k] = R(0);
    st[S_ALT * EN + k] = (R)a.alt; st[S_HDG * EN + k] = (R)a.hdg; st[S_TAS * EN + k] = (R)a.tas;
    st[S_VS * EN + k] = (R)a.vs; st[S_CAS * EN + k] = (R)a.cas; st[S_SELSPD * EN + k] = (R)a.cas;
    st[S_SELALT * EN + k] = (R)a.selalt; st[S_APTRK * EN + k] = (R)a.hdg; st[S_AX * EN + k] = (R)kKts;
    for (int f = S_C_ALT; f < S_NS; f++) st[(size_t)f * EN + k] = (R)NAN;
}

struct ResetShared {
    double alt0, target, gs0;
    int nv, nac;
    double px[BSG_MAXV], py[BSG_MAXV], cum[BSG_MAXV + 1], area;
    double minx, miny, maxx, maxy;
};

__device__ bool inside_poly_nm(const ResetShared &S, double ccl, double x, double y) {
    // the reference tests the lat/lon image of the point against the lat/lon polygon (areafilter.checkInside)
    const double tx = kCenterLat + x / 60, ty = kCenterLon + y / (60 * ccl);
    bool inside = false;
    double x0 = kCenterLat + S.px[S.nv - 1] / 60, y0 = kCenterLon + S.py[S.nv - 1] / (60 * ccl);
    bool f0 = y0 >= ty;
    for (int k = 0; k < S.nv; k++) {
        const double x1 = kCenterLat + S.px[k] / 60, y1 = kCenterLon + S.py[k] / (60 * ccl);
        const bool f1 = y1 >= ty;
        if (f0 != f1)
            if ((((y1 - ty) * (x0 - x1)) >= ((x1 - tx) * (y0 - y1))) == f1) inside = !inside;
        f0 = f1; x0 = x1; y0 = y1;
    }
    return inside;
}

template <typename R>
__global__ void reset_kernel(int kind, int E, int N, const int32_t *mask, uint64_t seed, uint32_t *episode,
                             double *pos, R *st, unsigned long long *reso, int *nac, double *ev) {
    const int e = blockIdx.x;
    if (mask && !mask[e]) return;                       // uniform per CTA
    __shared__ ResetShared S;
    const uint32_t ep = episode[e];
    const size_t EN = (size_t)E * N;
    double *v = ev + (size_t)e * BSG_NE;
    const double ccl = cos(kCenterLat * kD2R);
    if (threadIdx.x == 0) {
        Philox g(seed, (uint32_t)e, ep, 0u);
        for (int s = 0; s < BSG_EV_POLY; s++) v[s] = 0.0;
        S.nac = N;
        if (kind == BSG_KIND_HORIZONTAL) {
            S.gs0 = vcas2tas(150.0, 0.0);
            const double d = (double)g.randint(100, 150) / 6371.0, lat1 = 52.0 * kD2R;   // functions.py:24-42, bearing 0
            v[BSG_EV_WPT_LAT] = asin(sin(lat1) * cos(d) + cos(lat1) * sin(d)) * kR2D;
            v[BSG_EV_WPT_LON] = 4.0;
        } else if (kind == BSG_KIND_VERTICAL) {
            S.alt0 = (double)g.randint(2000, 4000);
            S.target = S.alt0 + (double)g.randint(-500, 500);
            S.gs0 = vcas2tas(150.0, S.alt0);
            v[BSG_EV_TARGET_ALT] = S.target;
        } else if (kind == BSG_KIND_SECTOR) {
            const double Rr = sqrt(3750.0 / kPi);
            double ang[BSG_MAXV];
            int nv = 3;
            for (int k = 0; k < 3; k++) { const double al = 2 * kPi * g.uniform(); ang[k] = atan2(Rr * sin(al), Rr * cos(al)); }
            for (;;) {
                for (int a = 1; a < nv; a++) {            // insertion sort by polar angle (sort_points_clockwise)
                    const double t = ang[a]; int b = a - 1;
                    while (b >= 0 && ang[b] > t) { ang[b + 1] = ang[b]; b--; }
                    ang[b + 1] = t;
                }
                double area = 0.0;
                for (int k = 0; k < nv; k++) { S.px[k] = Rr * cos(ang[k]); S.py[k] = Rr * sin(ang[k]); }
                for (int k = 0; k < nv; k++) { const int m = (k + 1) % nv; area += S.px[k] * S.py[m] - S.py[k] * S.px[m]; }
                S.area = fabs(area) / 2.0;
                if (S.area >= 2400.0 || nv >= BSG_MAXV) break;
                const double al = 2 * kPi * g.uniform();
                ang[nv++] = atan2(Rr * sin(al), Rr * cos(al));
            }
            S.nv = nv;
            const double dens = g.normal(0.005, 0.001);
            int na = (int)fmax(ceil(dens * S.area), 5.0);
            S.nac = na > N ? N : na;
            S.cum[0] = 0.0;
            S.minx = S.maxx = S.px[0]; S.miny = S.maxy = S.py[0];
            for (int k = 0; k < nv; k++) {
                const int m = (k + 1) % nv;
                const double dx = S.px[m] - S.px[k], dy = S.py[m] - S.py[k];
                S.cum[k + 1] = S.cum[k] + sqrt(dx * dx + dy * dy);
                S.minx = fmin(S.minx, S.px[k]); S.maxx = fmax(S.maxx, S.px[k]);
                S.miny = fmin(S.miny, S.py[k]); S.maxy = fmax(S.maxy, S.py[k]);
                v[BSG_EV_POLY + 2 * k] = kCenterLat + S.px[k] / 60;
                v[BSG_EV_POLY + 2 * k + 1] = kCenterLon + S.py[k] / (60 * ccl);
            }
            v[BSG_EV_NVERT] = nv;
        }
        nac[e] = S.nac;
        episode[e] = ep + 1;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        const size_t k = (size_t)e * N + i;
        reso[k] = 0ull;
        AcInit a = {};
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
                creconfs(52.0, 4.0, 0.0, S.gs0, dpsi, cpa, tlosh, a.lat, a.lon, a.hdg);
                a.alt = alt_own + dH;
                a.cas = vtas2cas(S.gs0, a.alt);
                a.tas = vcas2tas(a.cas, a.alt);
                a.selalt = a.alt;
            }
        } else if (kind == BSG_KIND_SECTOR && i < S.nac) {
            // perimeter waypoint: the reference sorts its draws and hands waypoint i to aircraft i; any
            // assignment of i.i.d. draws has the same law, so aircraft i keeps its own draw
            const double d = g.uniform() * S.cum[S.nv];
            int sgm = 0;
            while (sgm < S.nv - 1 && d > S.cum[sgm + 1]) sgm++;
            const int m = (sgm + 1) % S.nv;
            const double fr = (d - S.cum[sgm]) / (S.cum[sgm + 1] - S.cum[sgm]);
            const double wx = S.px[sgm] + fr * (S.px[m] - S.px[sgm]), wy = S.py[sgm] + fr * (S.py[m] - S.py[sgm]);
            double x, y;
            do {
                x = S.minx + g.uniform() * (S.maxx - S.minx);
                y = S.miny + g.uniform() * (S.maxy - S.miny);
            } while (!inside_poly_nm(S, ccl, x, y));
            a.lat = kCenterLat + x / 60; a.lon = kCenterLon + y / (60 * ccl);
            const double wlat = (kCenterLat + wx / 60) * kD2R, wlon = (kCenterLon + wy / (60 * ccl)) * kD2R;
            const double l1 = a.lat * kD2R, o1 = a.lon * kD2R;                        // functions.get_hdg
            const double hx = sin(wlon - o1) * cos(wlat);
            const double hy = cos(l1) * sin(wlat) - sin(l1) * cos(wlat) * cos(wlon - o1);
            a.hdg = pymod(atan2(hx, hy) * kR2D + 360, 360);
            a.alt = 350.0; a.cas = 150.0; a.tas = vcas2tas(150.0, 350.0); a.selalt = 350.0;
            if (i == 0) { v[BSG_EV_WPT_X] = wx; v[BSG_EV_WPT_Y] = wy; }
        }
        write_ac<R>(pos, st, EN, k, a);
    }
}

}  // namespace

extern "C" int bsg_reset(bsg_handle *h, const int32_t *mask_dev, uint64_t seed, void *stream) {
    if (!h) return BSG_ERR_ARG;
    const bsg_config &g = h->cfg;
    if (g.env_kind == BSG_KIND_RAW) { h->err = "bsg_reset: RAW handles take their traffic through bsg_set_state"; return BSG_ERR_UNSUPPORTED; }
    CK(h, cudaSetDevice(g.device));
    if (!h->episode) {
        CK(h, cudaMalloc(&h->episode, g.num_envs * sizeof(uint32_t)));
        CK(h, cudaMemset(h->episode, 0, g.num_envs * sizeof(uint32_t)));
    }
    cudaStream_t s = (cudaStream_t)stream;
    if (g.precision == BSG_FP64)
        reset_kernel<double><<<g.num_envs, 64, 0, s>>>(g.env_kind, g.num_envs, g.max_ac, mask_dev, seed, h->episode, h->pos,
                                                        (double *)h->st, h->reso, h->nac, h->ev);
    else
        reset_kernel<float><<<g.num_envs, 64, 0, s>>>(g.env_kind, g.num_envs, g.max_ac, mask_dev, seed, h->episode, h->pos,
                                                       (float *)h->st, h->reso, h->nac, h->ev);
    h->launches++;
    CK(h, cudaGetLastError());
    return BSG_OK;
}
