"""ORACLE (test infrastructure, not product code).

CPU restatement of ``bluesky/tools/areafilter.py`` POLY areas.
Upstream delegates the point test to ``matplotlib.path.Path.contains_points``
(matplotlib is absent here); the crossing-number test below restates the
published algorithm of matplotlib's ``point_in_path`` (Franklin pnpoly variant
with implicit path closure).  Reference call sites:
bluesky_gym/envs/sector_cr_env.py:89,136,160,203.  parity unpinned vs upstream.
"""
import numpy as np

_areas = {}


def _crossing_inside(vx, vy, tx, ty):
    n = len(vx)
    inside = False
    x0, y0 = vx[n - 1], vy[n - 1]
    yflag0 = y0 >= ty
    for k in range(n):
        x1, y1 = vx[k], vy[k]
        yflag1 = y1 >= ty
        if yflag0 != yflag1:
            if ((y1 - ty) * (x0 - x1) >= (x1 - tx) * (y0 - y1)) == yflag1:
                inside = not inside
        yflag0 = yflag1
        x0, y0 = x1, y1
    return inside


class Poly:
    def __init__(self, name, coordinates, top=1e9, bottom=-1e9):
        self.name = name
        c = np.asarray(coordinates, dtype=np.float64).reshape(-1, 2)
        self.vlat = c[:, 0].copy()
        self.vlon = c[:, 1].copy()
        self.top = top
        self.bottom = bottom

    def checkInside(self, lat, lon, alt):
        lat = np.atleast_1d(lat)
        lon = np.atleast_1d(lon)
        alt = np.atleast_1d(alt)
        hit = np.array([_crossing_inside(self.vlat, self.vlon, la, lo) for la, lo in zip(lat, lon)], dtype=bool)
        return np.all((hit, self.bottom <= alt, alt <= self.top), axis=0)


def defineArea(name, shape, coordinates, top=1e9, bottom=-1e9):
    if shape.upper() != 'POLY':
        raise NotImplementedError("oracle shim supports POLY areas only")
    _areas[name] = Poly(name, coordinates, top, bottom)
    return True


def deleteArea(name):
    _areas.pop(name, None)


def hasArea(name):
    return name in _areas


def checkInside(name, lat, lon, alt):
    if name not in _areas:
        return np.zeros(np.size(lat), dtype=bool)
    return _areas[name].checkInside(lat, lon, alt)


def reset():
    _areas.clear()
