"""ORACLE (test infrastructure, not product code).

CPU restatement of the flat-earth helpers of ``bluesky/tools/geo.py``
(bluesky-simulator==1.0.3, un-vendored).  Reference call
sites: bluesky_gym/envs/horizontal_cr_env.py:169,190,265;
vertical_cr_env.py:135,158,235; sector_cr_env.py:260,334; merge_env.py:195.
parity unpinned vs upstream (source absent here).
"""
import numpy as np

nm = 1852.0
_RE = 6371000.0


def kwikqdrdist(lata, lona, latb, lonb):
    """Bearing [deg, 0..360) and distance [NM] from a to b, mean-latitude flat earth."""
    dlat = np.radians(latb - lata)
    dlon = np.radians(((lonb - lona) + 180) % 360 - 180)
    cavelat = np.cos(np.radians(lata + latb) * 0.5)
    dangle = np.sqrt(dlat * dlat + dlon * dlon * cavelat * cavelat)
    dist = _RE * dangle / nm
    qdr = np.degrees(np.arctan2(dlon * cavelat, dlat)) % 360.0
    return qdr, dist


def kwikdist(lata, lona, latb, lonb):
    dlat = np.radians(latb - lata)
    dlon = np.radians(((lonb - lona) + 180) % 360 - 180)
    cavelat = np.cos(np.radians(lata + latb) * 0.5)
    dangle = np.sqrt(dlat * dlat + dlon * dlon * cavelat * cavelat)
    return _RE * dangle / nm


def kwikqdrdist_matrix(lata, lona, latb, lonb):
    """All-pairs version: row i = from aircraft i of set a, column j = to aircraft j of set b."""
    lata = np.asarray(lata, dtype=np.float64).reshape(-1, 1)
    lona = np.asarray(lona, dtype=np.float64).reshape(-1, 1)
    latb = np.asarray(latb, dtype=np.float64).reshape(1, -1)
    lonb = np.asarray(lonb, dtype=np.float64).reshape(1, -1)
    return kwikqdrdist(lata, lona, latb, lonb)


def kwikdist_matrix(lata, lona, latb, lonb):
    """Distance [NM] with plain numpy broadcasting (``.T`` on the first operand), as upstream: a scalar origin
    against a vector gives a vector (merge_env.py:195), two numpy matrices give the all-pairs matrix."""
    lata = np.asarray(lata, dtype=np.float64); lona = np.asarray(lona, dtype=np.float64)
    latb = np.asarray(latb, dtype=np.float64); lonb = np.asarray(lonb, dtype=np.float64)
    dlat = np.radians(latb - lata.T)
    dlon = np.radians(((lonb - lona.T) + 180) % 360 - 180)
    cavelat = np.cos(np.radians(lata + latb.T) * 0.5)
    dangle = np.sqrt(dlat * dlat + dlon * dlon * cavelat * cavelat)
    return _RE * dangle / nm


def kwikpos(latd1, lond1, qdr, dist):
    """Position at bearing qdr [deg] and distance dist [NM] from (latd1, lond1)."""
    dx = dist * np.sin(np.radians(qdr))
    dy = dist * np.cos(np.radians(qdr))
    dlat = dy / 60.0
    dlon = dx / (np.maximum(0.01, 60.0 * np.cos(np.radians(latd1))))
    latd2 = latd1 + dlat
    lond2 = ((lond1 + dlon) + 180) % 360 - 180
    return latd2, lond2


def rwgs84(latd):
    """Earth radius [m] of the WGS-84 ellipsoid at latitude latd [deg]."""
    lat = np.radians(latd)
    a = 6378137.0
    b = 6356752.314245
    coslat = np.cos(lat); sinlat = np.sin(lat)
    an = a * a * coslat; bn = b * b * sinlat
    ad = a * coslat; bd = b * sinlat
    return np.sqrt((an * an + bn * bn) / (ad * ad + bd * bd))


def qdrdist(latd1, lond1, latd2, lond2):
    """Great-circle bearing [deg, -180..180] and distance [NM] with the WGS-84 radius of the mean latitude
    (``geo.qdrdist``: what the LNAV autopilot steers by; reference call path merge_env.py:155-156
    ``addwpt``/``dest`` -> autopilot.update)."""
    res1 = rwgs84(0.5 * (latd1 + latd2))
    a = 6378137.0
    r1 = rwgs84(latd1); r2 = rwgs84(latd2)
    res2 = 0.5 * (np.abs(latd1) * (r1 + a) + np.abs(latd2) * (r2 + a)) / \
        np.maximum(0.000001, np.abs(latd1) + np.abs(latd2))
    sw = (latd1 * latd2 >= 0.0)
    r = sw * res1 + (1 - sw) * res2
    lat1 = np.radians(latd1); lon1 = np.radians(lond1)
    lat2 = np.radians(latd2); lon2 = np.radians(lond2)
    coslat1 = np.cos(lat1); coslat2 = np.cos(lat2)
    d = r * np.arccos(np.clip(coslat1 * coslat2 * np.cos(lon2 - lon1) + np.sin(lat1) * np.sin(lat2), -1.0, 1.0))
    qdr = np.degrees(np.arctan2(np.sin(lon2 - lon1) * coslat2,
                                coslat1 * np.sin(lat2) - np.sin(lat1) * coslat2 * np.cos(lon2 - lon1)))
    return qdr, d / nm
