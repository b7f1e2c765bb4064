"""ORACLE (test infrastructure, not product code).

Restatement of ``bluesky/traffic/windfield.py`` (bluesky-simulator==1.0.3, un-vendored) for the surface
the reference's WindFieldWrapper drives (bluesky_gym/wrappers/wind.py:28 ``addpointvne``, :57 ``getdata``):
wind vectors defined at points, no altitude profile (the wrapper's ``alt=None`` default):
  winddim 0  no wind
  winddim 1  one point: the same vector everywhere
  winddim 2  several points: horizontal inverse-distance-squared weighting on the mean-latitude flat earth
A 3-D field (altitude profiles, winddim 3) is outside this restatement and raises.
parity unpinned vs upstream (source absent here).
"""
import numpy as np


class Windfield:
    def __init__(self):
        self.clear()

    def clear(self):
        self.winddim = 0
        self.lat = np.zeros(0)
        self.lon = np.zeros(0)
        self.vnorth = np.zeros(0)
        self.veast = np.zeros(0)

    def addpointvne(self, lat, lon, vnorth, veast, windalt=None):
        if windalt is not None:
            raise NotImplementedError("3-D wind profiles are outside the oracle's restatement")
        lat = np.atleast_1d(np.asarray(lat, dtype=np.float64))
        lon = np.atleast_1d(np.asarray(lon, dtype=np.float64))
        vn = np.broadcast_to(np.atleast_1d(np.asarray(vnorth, dtype=np.float64)), lat.shape)
        ve = np.broadcast_to(np.atleast_1d(np.asarray(veast, dtype=np.float64)), lat.shape)
        self.lat = np.append(self.lat, lat)
        self.lon = np.append(self.lon, lon)
        self.vnorth = np.append(self.vnorth, vn)
        self.veast = np.append(self.veast, ve)
        self.winddim = min(2, len(self.lat))

    def getdata(self, userlat, userlon, useralt=0.0):
        scalar = np.ndim(userlat) == 0
        lat = np.atleast_1d(np.asarray(userlat, dtype=np.float64))
        lon = np.atleast_1d(np.asarray(userlon, dtype=np.float64))
        if self.winddim == 0:
            vn = np.zeros(lat.shape); ve = np.zeros(lat.shape)
        elif self.winddim == 1:
            vn = np.ones(lat.shape) * self.vnorth[0]
            ve = np.ones(lat.shape) * self.veast[0]
        else:
            eps = 1e-20
            la = lat.reshape(-1, 1); lo = lon.reshape(-1, 1)
            cavelat = np.cos(np.radians(0.5 * (la + self.lat.reshape(1, -1))))
            dy = np.radians(la - self.lat.reshape(1, -1))
            dx = cavelat * np.radians(((lo - self.lon.reshape(1, -1)) + 180.0) % 360.0 - 180.0)
            invd2 = 1.0 / (eps + dx * dx + dy * dy)
            horfact = invd2 / invd2.sum(axis=1, keepdims=True)
            vn = (horfact * self.vnorth.reshape(1, -1)).sum(axis=1)
            ve = (horfact * self.veast.reshape(1, -1)).sum(axis=1)
        if scalar:
            return vn[0], ve[0]
        return vn, ve
