"""Host-side (numpy, vectorised over envs) scenario generation in the core's fp64 interchange layout.
It restates the reset path of the reference env layer for N aircraft per env:
  Horizontal  bluesky_gym/envs/horizontal_cr_env.py:82-148 (own at BlueSky's default spawn + creconfs intruders)
  Vertical    bluesky_gym/envs/vertical_cr_env.py:183-200,258-281
  Sector      bluesky_gym/envs/sector_cr_env.py:87-217 (rejection-grown polygon, density draw, rejection sampling)
  Merge       bluesky_gym/envs/merge_env.py:100-158 (spawn radials towards the FAF; intruders on the LNAV route)
with the same value laws but numpy's Generator instead of the global legacy stream.
Used for seeded synthetic traffic (tests, bench) and as the host reset of the vector env; the
device-side generator (bsg_reset) draws from the same laws with Philox.
"""
import numpy as np
from . import _lib as L

NM, FT, RE = 1852.0, 0.3048, 6371000.0
CENTER = np.array([51.990426702297746, 4.376124857109851])
_R, _P0, _RHO0 = 287.05287, 101325.0, 1.225


def _vatmos(h):
    T = np.maximum(288.15 - 0.0065 * h, 216.65)
    rho = 1.225 * (T / 288.15) ** 4.256848030018761 * np.exp(-np.maximum(0.0, h - 11000.0) / 6341.552161)
    return rho * _R * T, rho, T


def vcas2tas(cas, h):
    p, rho, _ = _vatmos(h)
    qdyn = _P0 * ((1.0 + _RHO0 * cas * cas / (7.0 * _P0)) ** 3.5 - 1.0)
    return np.sqrt(7.0 * p / rho * ((1.0 + qdyn / p) ** (2.0 / 7.0) - 1.0))


def vtas2cas(tas, h):
    p, rho, _ = _vatmos(h)
    qdyn = p * ((1.0 + rho * tas * tas / (7.0 * p)) ** 3.5 - 1.0)
    return np.sqrt(7.0 * _P0 / _RHO0 * ((qdyn / _P0 + 1.0) ** (2.0 / 7.0) - 1.0))


def _blank(E, N):
    ac = np.zeros((E, L.NF, N)); reso = np.zeros((E, N), np.uint64)
    nac = np.full(E, N, np.int32); ev = np.zeros((E, L.NE))
    return ac, reso, nac, ev


def _fill(ac, idx, lat, lon, alt, hdg, cas, vs=0.0, selalt=None):
    """BlueSky Traffic.cre bookkeeping for aircraft column(s) idx."""
    F = L.F
    tas = vcas2tas(np.asarray(cas, dtype=np.float64) + 0 * alt, alt)
    ac[:, F['lat'], idx] = lat; ac[:, F['lon'], idx] = lon; ac[:, F['alt'], idx] = alt
    ac[:, F['hdg'], idx] = hdg; ac[:, F['tas'], idx] = tas; ac[:, F['cas'], idx] = cas
    ac[:, F['vs'], idx] = vs; ac[:, F['selspd'], idx] = cas
    ac[:, F['selalt'], idx] = alt if selalt is None else selalt
    ac[:, F['aptrk'], idx] = hdg
    ac[:, F['ax'], idx] = 0.514444
    ac[:, F['gsn'], idx] = tas * np.cos(np.radians(hdg)); ac[:, F['gse'], idx] = tas * np.sin(np.radians(hdg))
    ac[:, F['lnav'], idx] = -1.0


def _creconfs(rng, shape, lat0, lon0, trk0_deg, gs0, tlosh, dpsi, cpa_nm):
    """Vectorised Traffic.creconfs geometry: intruder position + heading."""
    trkref = np.radians(trk0_deg)
    cpa = cpa_nm * NM
    pzr = 5.0 * NM
    trk = trkref + np.radians(dpsi)
    gsn, gse = gs0 * np.cos(trk), gs0 * np.sin(trk)
    vreln, vrele = gs0 * np.cos(trkref) - gsn, gs0 * np.sin(trkref) - gse
    vrel = np.sqrt(vreln * vreln + vrele * vrele)
    drelcpa = tlosh * vrel + np.where(cpa > pzr, 0.0, np.sqrt(np.maximum(pzr * pzr - cpa * cpa, 0.0)))
    dist = np.sqrt(drelcpa * drelcpa + cpa * cpa)
    rd, rx = drelcpa / dist, cpa / dist
    brn = np.degrees(np.arctan2(-rx * vreln + rd * vrele, rd * vreln + rx * vrele))
    dnm = dist / NM
    dx, dy = dnm * np.sin(np.radians(brn)), dnm * np.cos(np.radians(brn))
    lat = lat0 + dy / 60.0
    lon = ((lon0 + dx / np.maximum(0.01, 60.0 * np.cos(np.radians(lat0)))) + 180) % 360 - 180
    hdg = np.degrees(np.arctan2(gse, gsn))
    return lat, lon, hdg


def horizontal(E, N, seed=0):
    """Own at (52, 4), hdg 0, alt 0, CAS 150; N-1 creconfs intruders; waypoint 100-149 km due north."""
    rng = np.random.default_rng(seed)
    ac, reso, nac, ev = _blank(E, N)
    K = N - 1
    z = np.zeros(E)
    _fill(ac, 0, 52.0 + z, 4.0 + z, z, z, 150.0 + z)
    gs0 = ac[:, L.F['tas'], 0][:, None]
    dpsi = rng.integers(45, 315, (E, K)).astype(float)
    cpa = rng.integers(0, 5, (E, K)).astype(float)
    tlosh = rng.integers(100, 1000, (E, K)).astype(float)
    lat, lon, hdg = _creconfs(rng, (E, K), 52.0, 4.0, 0.0, gs0, tlosh, dpsi, cpa)
    zk = np.zeros((E, K))
    _fill(ac, slice(1, N), lat, lon, zk, hdg, vtas2cas(gs0 + zk, zk))
    d = rng.integers(100, 150, E).astype(float) / 6371.0            # functions.get_point_at_distance, bearing 0
    lat1 = np.radians(52.0)
    ev[:, L.EV['wpt_lat']] = np.degrees(np.arcsin(np.sin(lat1) * np.cos(d) + np.cos(lat1) * np.sin(d)))
    ev[:, L.EV['wpt_lon']] = 4.0
    ev[:, L.EV['ac_hdg']] = 0.0
    return ac, reso, nac, ev


def vertical(E, N, seed=0):
    rng = np.random.default_rng(seed)
    ac, reso, nac, ev = _blank(E, N)
    K = N - 1
    z = np.zeros(E)
    alt0 = rng.integers(2000, 4000, E).astype(float)
    target = alt0 + rng.integers(-500, 500, E)
    _fill(ac, 0, 52.0 + z, 4.0 + z, alt0, z, 150.0 + z)
    gs0 = ac[:, L.F['tas'], 0][:, None]
    dpsi = rng.integers(45, 315, (E, K)).astype(float)
    cpa = rng.integers(0, 5, (E, K)).astype(float)
    hi = np.floor(180000.0 / gs0).astype(int)
    tlosh = np.floor(100 + rng.random((E, K)) * (hi - 100)).astype(float)
    avg_tod = 200000.0 / gs0 - 2 * target[:, None] / 12.5
    a2, t2 = alt0[:, None], target[:, None]
    lo1, hi1 = np.trunc(-a2 + 500), np.trunc((t2 - a2) + 100)
    lo2, hi2 = np.trunc((t2 - a2) - 500), np.trunc((t2 - a2) + 500)
    lo = np.where(tlosh > avg_tod, lo1, lo2); hi_ = np.where(tlosh > avg_tod, hi1, hi2)
    dH = np.floor(lo + rng.random((E, K)) * np.maximum(hi_ - lo, 1))
    lat, lon, hdg = _creconfs(rng, (E, K), 52.0, 4.0, 0.0, gs0, tlosh, dpsi, cpa)
    alt = a2 + dH
    _fill(ac, slice(1, N), lat, lon, alt, hdg, vtas2cas(gs0 + 0 * alt, alt))
    ev[:, L.EV['target_alt']] = target
    return ac, reso, nac, ev


def _inside(vlat, vlon, tx, ty):
    """Crossing-number test, same edge rule as the CUDA core (matplotlib point_in_path)."""
    inside = np.zeros(np.shape(tx), dtype=bool)
    n = len(vlat)
    x0, y0 = vlat[n - 1], vlon[n - 1]
    for k in range(n):
        x1, y1 = vlat[k], vlon[k]
        f0, f1 = y0 >= ty, y1 >= ty
        cross = (f0 != f1) & ((((y1 - ty) * (x0 - x1)) >= ((x1 - tx) * (y0 - y1))) == f1)
        inside ^= cross
        x0, y0 = x1, y1
    return inside


def _polygon(rng):
    """sector_cr_env.py:141-160: points on a circle of area 3750 NM^2 until the polygon area >= 2400 NM^2."""
    Rr = np.sqrt(3750 / np.pi)
    ang = list(2 * np.pi * rng.random(3))
    while True:
        a = np.sort(np.arctan2(Rr * np.sin(ang), Rr * np.cos(ang)))
        p = np.stack([Rr * np.cos(a), Rr * np.sin(a)], 1)
        area = 0.5 * abs(np.sum(p[:, 0] * np.roll(p[:, 1], -1) - p[:, 1] * np.roll(p[:, 0], -1)))
        if area >= 2400 or len(ang) >= L.MAXV:
            return p, area
        ang.append(2 * np.pi * rng.random())


def sector(E, N, seed=0, n_poly=64, fixed_n=True):
    """Sector scenarios.  n_poly distinct reference-style polygons are shared round-robin by the envs;
    fixed_n=True gives every env N live aircraft (throughput configs), else the reference's density draw."""
    rng = np.random.default_rng(seed)
    ac, reso, nac, ev = _blank(E, N)
    ccl = np.cos(np.radians(CENTER[0]))
    polys = [_polygon(rng) for _ in range(min(n_poly, E))]
    for pi, (p, area) in enumerate(polys):
        envs = np.arange(pi, E, len(polys))
        ne = len(envs)
        vlat = CENTER[0] + p[:, 0] / 60; vlon = CENTER[1] + p[:, 1] / (60 * ccl)
        nv = len(p)
        ev[envs, L.EV['nvert']] = nv
        ev[np.ix_(envs, L.EV['poly'] + 2 * np.arange(nv))] = vlat
        ev[np.ix_(envs, L.EV['poly'] + 2 * np.arange(nv) + 1)] = vlon
        if not fixed_n:
            dens = rng.normal(0.005, 0.001, ne)
            nac[envs] = np.clip(np.maximum(np.ceil(dens * area), 5), 5, N).astype(np.int32)
        # waypoints on the perimeter (sector_cr_env.py:162-188), sorted per env like the reference's d_list
        seg = np.linalg.norm(np.roll(p, -1, 0) - p, axis=1)
        cum = np.concatenate([[0], np.cumsum(seg)])
        d = rng.random((ne, N)) * cum[-1]
        d[np.arange(N)[None, :] >= nac[envs][:, None]] = np.inf          # only the num_ac live aircraft draw (and sort)
        d = np.sort(d, axis=1)
        d[~np.isfinite(d)] = 0.0
        k = np.clip(np.searchsorted(cum, d, side='right') - 1, 0, nv - 1)
        frac = (d - cum[k]) / seg[k]
        wp = p[k] + frac[..., None] * (np.roll(p, -1, 0)[k] - p[k])
        # positions: rejection sampling in the bounding box (sector_cr_env.py:190-204)
        lo, hi = p.min(0), p.max(0)
        pos = np.zeros((ne, N, 2)); need = np.ones((ne, N), dtype=bool)
        while need.any():
            c = lo + rng.random((ne, N, 2)) * (hi - lo)
            la = CENTER[0] + c[..., 0] / 60; lo_ = CENTER[1] + c[..., 1] / (60 * ccl)
            ok = _inside(vlat, vlon, la, lo_) & need
            pos[ok] = c[ok]; need &= ~ok
        lat = CENTER[0] + pos[..., 0] / 60; lon = CENTER[1] + pos[..., 1] / (60 * ccl)
        wlat = np.radians(CENTER[0] + wp[..., 0] / 60); wlon = np.radians(CENTER[1] + wp[..., 1] / (60 * ccl))
        l1, o1 = np.radians(lat), np.radians(lon)          # functions.get_hdg
        x = np.sin(wlon - o1) * np.cos(wlat)
        y = np.cos(l1) * np.sin(wlat) - np.sin(l1) * np.cos(wlat) * np.cos(wlon - o1)
        hdg = (np.degrees(np.arctan2(x, y)) + 360) % 360
        sub = ac[envs]
        _fill(sub, slice(0, N), lat, lon, np.full((ne, N), 350.0), hdg, np.full((ne, N), 150.0))
        ac[envs] = sub
        ev[envs, L.EV['wpt_x']] = wp[:, 0, 0]; ev[envs, L.EV['wpt_y']] = wp[:, 0, 1]
    return ac, reso, nac, ev


RWY = (52.36239301495972, 4.713195734579777)            # merge_env.py:39-40


def point_at_distance(lat1, lon1, d_km, bearing, R=6371.0):
    """functions.get_point_at_distance (bluesky_gym/envs/common/functions.py:24-42)."""
    lat1, lon1, a = np.radians(lat1), np.radians(lon1), np.radians(bearing)
    lat2 = np.arcsin(np.sin(lat1) * np.cos(d_km / R) + np.cos(lat1) * np.sin(d_km / R) * np.cos(a))
    lon2 = lon1 + np.arctan2(np.sin(a) * np.sin(d_km / R) * np.cos(lat1), np.cos(d_km / R) - np.sin(lat1) * np.sin(lat2))
    return np.degrees(lat2), np.degrees(lon2)


FAF = point_at_distance(RWY[0], RWY[1], 200.0, 0.0)     # merge_env.py:42-45


def merge(E, N=20, seed=0):
    """Agent 50-200 km and N-1 intruders 20-500 km out on radials within +-15 deg of north from the FAF, all heading
    back along their radial at CAS 100 m/s, 10 km altitude; intruders fly the LNAV route FAF -> runway."""
    rng = np.random.default_rng(seed)
    ac, reso, nac, ev = _blank(E, N)
    brg = rng.uniform(-15.0, 15.0, (E, N))
    dist = rng.uniform(20.0, 500.0, (E, N))
    dist[:, 0] = rng.uniform(50.0, 200.0, E)
    lat, lon = point_at_distance(FAF[0], FAF[1], dist, brg)
    _fill(ac, slice(0, N), lat, lon, np.full((E, N), 10000.0), brg - 180.0, np.full((E, N), 100.0))
    ac[:, L.F['lnav'], 1:] = 0.0
    ev[:, L.EV['wpt_lat']] = FAF[0]; ev[:, L.EV['wpt_lon']] = FAF[1]
    ev[:, L.EV['wpt_x']] = RWY[0]; ev[:, L.EV['wpt_y']] = RWY[1]
    return ac, reso, nac, ev


def raw_traffic(E, N, seed=0, box_nm=200.0, lat0=52.0, lon0=4.0):
    """Uniform traffic for the stand-alone CD / ND-map configs (SURVEY.md 8d, C4/C5 value laws)."""
    rng = np.random.default_rng(seed)
    half = box_nm / 2 / 60.0
    lat = lat0 + rng.uniform(-half, half, (E, N))
    lon = lon0 + rng.uniform(-half, half, (E, N)) / np.cos(np.radians(lat0))
    alt = rng.uniform(10000 * FT, 40000 * FT, (E, N))
    trk = rng.uniform(0, 360, (E, N)); gs = rng.uniform(120, 250, (E, N))
    vs = rng.choice([-10.0, 0.0, 0.0, 10.0], (E, N))
    return lat, lon, alt, trk, gs, vs
