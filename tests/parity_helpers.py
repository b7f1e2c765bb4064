"""Enumeration helpers of the parity tests (BASELINE.json north_star: flags / indices / counts are bit-exact "except
pairs within a stated epsilon of the zone thresholds, which must be enumerated").

Every quantity here is computed from ORACLE / reference-golden data in fp64:
  * env_threshold_margin: per env, the smallest relative distance of the post-step state to a threshold of the
    reference's env layer (intrusion 5 NM / 1000 ft, waypoint 5 km, crash / runway, sector polygon edge) -- what can
    legitimately flip reward / terminated / truncated under fp32 rounding;
The CD / MVP side of the enumeration (near-threshold rows of the last detection, per-aircraft margins of the decisions
that steer a resolution, MVP's conditioning) comes from the oracle's own trace: oracle.step_trace (oracle/c/bsgo.h).
"""
import numpy as np

NM, FT = 1852.0, 0.3048
KCLAT, KCLON = 51.990426702297746, 4.376124857109851         # sector_cr_env.py:16


def kwikdist_nm(lata, lona, latb, lonb):
    """bluesky.tools.geo.kwikdist, fp64."""
    dlat = np.radians(latb - lata)
    dlon = np.radians(((lonb - lona) + 180.0) % 360.0 - 180.0)
    cav = np.cos(np.radians(lata + latb) * 0.5)
    return 6371000.0 * np.sqrt(dlat * dlat + dlon * dlon * cav * cav) / NM


def _edge_dist_nm(lat, lon, vlat, vlon):
    """Distance [NM] of a point to the nearest edge of the polygon (flat, sector centre latitude)."""
    c = np.cos(np.radians(KCLAT))
    px, py = (lat - KCLAT) * 60.0, (lon - KCLON) * 60.0 * c
    x, y = (vlat - KCLAT) * 60.0, (vlon - KCLON) * 60.0 * c
    best = np.inf
    for k in range(len(x)):
        m = (k + 1) % len(x)
        ex, ey = x[m] - x[k], y[m] - y[k]
        t = np.clip(((px - x[k]) * ex + (py - y[k]) * ey) / max(ex * ex + ey * ey, 1e-300), 0.0, 1.0)
        best = min(best, np.hypot(px - (x[k] + t * ex), py - (y[k] + t * ey)))
    return best


def env_threshold_margin(kind, ac1, ev, nac, F, EV):
    """[E] smallest relative margin of the post-step state ac1 to a reward / termination threshold of the env layer.
    kind: 0 Horizontal (horizontal_cr_env.py:237-270), 1 Vertical (vertical_cr_env.py:213-240), 2 Sector
    (sector_cr_env.py:129-139,324-339)."""
    E = ac1.shape[0]
    out = np.full(E, np.inf)
    for e in range(E):
        n = int(nac[e])
        lat, lon, alt = ac1[e, F['lat'], :n], ac1[e, F['lon'], :n], ac1[e, F['alt'], :n]
        if n < 1 or not np.isfinite(lat[0]):
            continue
        d = kwikdist_nm(lat[0], lon[0], lat[1:], lon[1:])
        m = np.inf
        if d.size:
            m = np.min(np.abs(d - 5.0) / 5.0)
            if kind == 1:
                m = min(m, np.min(np.abs(np.abs(alt[1:] - alt[0]) - 1000 * FT) / (1000 * FT)))
        if kind == 0:
            wd = kwikdist_nm(lat[0], lon[0], ev[e, EV['wpt_lat']], ev[e, EV['wpt_lon']]) * 1.852
            m = min(m, abs(wd - 5.0) / 5.0)
        elif kind == 1:
            rwy = 200.0 - kwikdist_nm(52.0, 4.0, lat[0], lon[0]) * 1.852
            m = min(m, abs(rwy) / 200.0, abs(alt[0]) / 3000.0)
        elif kind == 4:      # Merge (merge_env.py:247-283): intrusion at 4 NM, FAF reached at 10, runway at 20 (distance units of the env)
            if d.size:
                m = np.min(np.abs(d - 4.0) / 4.0)
            pre = ev[e, EV['wpt_reach']] != 1.0
            wd = kwikdist_nm(lat[0], lon[0], ev[e, EV['wpt_lat']] if pre else ev[e, EV['wpt_x']],
                             ev[e, EV['wpt_lon']] if pre else ev[e, EV['wpt_y']])
            m = min(m, abs(wd - 10.0) / 10.0, abs(wd - 20.0) / 20.0)
        elif kind == 2:
            nv = int(ev[e, EV['nvert']])
            vlat = ev[e, EV['poly']:EV['poly'] + 2 * nv:2]; vlon = ev[e, EV['poly'] + 1:EV['poly'] + 2 * nv:2]
            m = min(m, _edge_dist_nm(lat[0], lon[0], vlat, vlon) / 50.0)
        out[e] = m
    return out
