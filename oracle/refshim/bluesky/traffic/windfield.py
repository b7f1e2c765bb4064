"""ORACLE (test infrastructure, not product code).

Restatement of ``bluesky/traffic/windfield.py`` (bluesky-simulator==1.0.3, un-vendored) for the surface
the reference's WindFieldWrapper drives (bluesky_gym/wrappers/wind.py:28 ``addpointvne``, :57 ``getdata``):
  winddim 0  no wind
  winddim 1  one point: the same vector everywhere
  winddim 2  several points: horizontal inverse-distance-squared weighting on the mean-latitude flat earth
  winddim 3  a point given with an altitude profile (the wrapper's ``alt=`` argument): its vectors are resampled onto a
             fixed altitude axis (0 .. 45 000 ft in 100 ft steps, ``np.interp``: constant beyond the profile's ends) and
             ``getdata`` interpolates linearly between the two axis levels around the requested altitude.
As recalled from upstream: ``getdata`` takes the altitude from ndarray / list / Python-float arguments only -- any other
scalar type (a numpy float64 such as ``bs.traf.alt[idx]``: the wrapper's observation, ``selhdgcmd``) reads the field at
altitude 0.  Restated supported surface: ONE profile point (what a single ``addpointvne`` call with ``windalt`` adds).
parity unpinned vs upstream (source absent here).
"""
import numpy as np


FT = 0.3048


class Windfield:
    def __init__(self):
        self.altmax = 45000.0 * FT
        self.altstep = 100.0 * FT
        self.altaxis = np.arange(0.0, self.altmax + self.altstep, self.altstep)
        self.clear()

    def clear(self):
        self.winddim = 0
        self.prof_vn = None        # [nalt] profile of the (single) profile point on the altitude axis
        self.prof_ve = None
        self.lat = np.zeros(0)
        self.lon = np.zeros(0)
        self.vnorth = np.zeros(0)
        self.veast = np.zeros(0)

    def addpointvne(self, lat, lon, vnorth, veast, windalt=None):
        if windalt is not None:
            if len(self.lat) or np.ndim(lat) != 0:
                raise NotImplementedError("the oracle restates ONE wind point with an altitude profile")
            alttab = np.asarray(windalt, dtype=np.float64)
            self.prof_vn = np.interp(self.altaxis, alttab, np.asarray(vnorth, dtype=np.float64))
            self.prof_ve = np.interp(self.altaxis, alttab, np.asarray(veast, dtype=np.float64))
            self.lat = np.append(self.lat, float(lat)); self.lon = np.append(self.lon, float(lon))
            self.vnorth = np.append(self.vnorth, self.prof_vn[0]); self.veast = np.append(self.veast, self.prof_ve[0])
            self.winddim = 3
            return
        if self.winddim == 3:
            raise NotImplementedError("the oracle restates ONE wind point with an altitude profile")
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
        if self.winddim == 3:
            # altitude: ndarray / list / Python float only (see the module docstring); one point: no horizontal weights
            if type(useralt) == np.ndarray:
                alt = useralt.astype(np.float64)
            elif type(useralt) == list:
                alt = np.array(useralt, dtype=np.float64)
            elif type(useralt) == float:
                alt = useralt * np.ones(lat.shape)
            else:
                alt = np.zeros(lat.shape)
            idxalt = np.maximum(0.0, np.minimum(self.altaxis[-1] - 1e-20, alt)) / self.altstep
            ialt = np.floor(idxalt).astype(int)
            falt = idxalt - ialt
            vn = (1.0 - falt) * self.prof_vn[ialt] + falt * self.prof_vn[ialt + 1]
            ve = (1.0 - falt) * self.prof_ve[ialt] + falt * self.prof_ve[ialt + 1]
        elif self.winddim == 0:
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
