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
    return kwikqdrdist_matrix(lata, lona, latb, lonb)[1]


def kwikpos(latd1, lond1, qdr, dist):
    """Position at bearing qdr [deg] and distance dist [NM] from (latd1, lond1)."""
    dx = dist * np.sin(np.radians(qdr))
    dy = dist * np.cos(np.radians(qdr))
    dlat = dy / 60.0
    dlon = dx / (np.maximum(0.01, 60.0 * np.cos(np.radians(latd1))))
    latd2 = latd1 + dlat
    lond2 = ((lond1 + dlon) + 180) % 360 - 180
    return latd2, lond2
