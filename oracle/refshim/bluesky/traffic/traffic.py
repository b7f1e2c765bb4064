"""ORACLE (test infrastructure, not product code).

CPU restatement (numpy, fp64, BlueSky-style per-env arrays) of the subset of
``bluesky/traffic/traffic.py`` + ``autopilot.py`` + ``aporasas.py`` the reference
envs drive: ``cre``, ``creconfs``, ``delete``, ``reset``,
``id2idx``, ``update`` (autopilot -> ASAS -> AP-or-ASAS select -> performance
limits -> airspeed/heading/vs -> ground speed -> position), ``ap.selaltcmd``.
Reference call sites: bluesky_gym/envs/horizontal_cr_env.py:85,91,109,133;
vertical_cr_env.py:198-200,252-256,270-271,289; sector_cr_env.py:88,121,211,217.
bluesky-simulator==1.0.3 is un-vendored: parity unpinned vs upstream.
"""
from math import radians, degrees, cos, sin, atan2, sqrt
import numpy as np
from ..tools import geo
from ..tools.aero import (fpm, kts, ft, nm, g0, Rearth, vatmos, vcasormach, vcasormach2tas,
                          vtas2cas, vtas2mach, tas2cas)
from .perf import Perf
from .asas import ConflictDetection
from .mvp import MVP
from .windfield import Windfield

_FLOAT_FIELDS = ('lat', 'lon', 'alt', 'hdg', 'trk', 'tas', 'gs', 'gsnorth', 'gseast', 'cas', 'M',
                 'vs', 'selspd', 'selalt', 'selvs', 'ax', 'bank', 'apvsdef')
_BOOL_FIELDS = ('swlnav', 'swvnav', 'swvnavspd', 'swhdgsel', 'swaltsel')


def _degto180(a):
    return (a + 180.0) % 360.0 - 180.0


class Route:
    """Per-aircraft flight plan: the slice of ``bluesky/traffic/route.py`` that ``ACID ADDWPT lat lon`` and
    ``ACID DEST lat lon`` exercise (merge_env.py:155-156).  Waypoints are fly-by positions without altitude or
    speed constraints that LNAV acts on (VNAV is never switched on by the reference)."""

    def __init__(self):
        self.wplat = []
        self.wplon = []
        self.iactwp = -1
        self.has_dest = False

    @property
    def nwp(self):
        return len(self.wplat)


class Autopilot:
    def __init__(self, traf):
        self.traf = traf
        self.reset()

    _ARR = ('trk', 'tas', 'alt', 'vs', 'wplat', 'wplon', 'next_qdr', 'swlastwp')

    def reset(self):
        self.trk = np.zeros(0); self.tas = np.zeros(0); self.alt = np.zeros(0); self.vs = np.zeros(0)
        # active waypoint (bs.traf.actwp): position, direction of the leg after it (-999 = none), last-wp switch
        self.wplat = np.zeros(0); self.wplon = np.zeros(0); self.next_qdr = np.zeros(0)
        self.swlastwp = np.zeros(0, dtype=bool)
        self.route = []

    def create(self, n):
        t = self.traf
        self.trk = np.append(self.trk, t.trk[-n:])
        self.tas = np.append(self.tas, t.tas[-n:])
        self.alt = np.append(self.alt, t.alt[-n:])
        self.vs = np.append(self.vs, np.zeros(n))
        self.wplat = np.append(self.wplat, np.full(n, 89.99))
        self.wplon = np.append(self.wplon, np.zeros(n))
        self.next_qdr = np.append(self.next_qdr, np.full(n, -999.0))
        self.swlastwp = np.append(self.swlastwp, np.zeros(n, dtype=bool))
        self.route.extend(Route() for _ in range(n))

    def delete(self, idx):
        for k in self._ARR:
            setattr(self, k, np.delete(getattr(self, k), idx))
        self.route.pop(idx)

    # ---------------------------------------------------------------- flight plan (LNAV)
    def addwpt(self, idx, lat, lon, dest=False):
        """``ADDWPT`` / ``DEST`` with a lat/lon position.  A destination stays the last waypoint; other waypoints
        are inserted before it.  The first waypoint of a route becomes the active one and switches LNAV on
        (``Route.direct``); the active leg keeps the next-leg direction it had when it was activated."""
        r = self.route[idx]
        if dest:
            if r.has_dest:
                r.wplat[-1] = lat; r.wplon[-1] = lon
            else:
                r.wplat.append(lat); r.wplon.append(lon); r.has_dest = True
        else:
            pos = r.nwp - 1 if r.has_dest else r.nwp
            r.wplat.insert(pos, lat); r.wplon.insert(pos, lon)
            if r.iactwp >= pos:
                r.iactwp += 1
        if r.nwp == 1:
            r.iactwp = 0
            self.wplat[idx] = lat; self.wplon[idx] = lon
            self.next_qdr[idx] = -999.0
            self.swlastwp[idx] = False
            self.traf.swlnav[idx] = True

    def _calcturn(self, tas, bank, wpqdr, next_wpqdr):
        turnrad = tas * tas / (np.maximum(0.01, np.tan(bank)) * g0)
        turndist = np.abs(turnrad * np.tan(np.radians(0.5 * np.abs(_degto180(wpqdr % 360.0 - next_wpqdr % 360.0)))))
        return turndist, turnrad

    def _update_fms(self):
        """Waypoint switching (``ActiveWaypoint.Reached`` + ``Autopilot.update_fms``): a fly-by waypoint counts as
        reached inside the turn-anticipation distance, or when the aircraft flies away from it within 1.02 turn
        radii; past the last waypoint LNAV switches off and the aircraft keeps its track."""
        t = self.traf
        qdr, distnm = geo.qdrdist(t.lat, t.lon, self.wplat, self.wplon)
        dist = distnm * nm
        next_qdr = np.where(self.next_qdr < -900.0, qdr, self.next_qdr)
        turndist, turnrad = self._calcturn(t.tas, t.bank, qdr, next_qdr)
        away = np.abs(_degto180(t.trk % 360.0 - qdr % 360.0)) > 90.0
        circling = away & (dist < turnrad * 1.02)
        reached = t.swlnav & ((dist < turndist) | circling)
        for i in np.where(reached)[0]:
            r = self.route[i]
            if self.swlastwp[i] or r.iactwp + 1 >= r.nwp:
                t.swlnav[i] = False
                continue
            r.iactwp += 1
            self.wplat[i] = r.wplat[r.iactwp]; self.wplon[i] = r.wplon[r.iactwp]
            self.swlastwp[i] = (r.iactwp == r.nwp - 1)
            if r.iactwp + 1 < r.nwp:
                q, _ = geo.qdrdist(r.wplat[r.iactwp], r.wplon[r.iactwp], r.wplat[r.iactwp + 1], r.wplon[r.iactwp + 1])
                self.next_qdr[i] = q
            else:
                self.next_qdr[i] = -999.0
            qdr[i], _ = geo.qdrdist(t.lat[i], t.lon[i], self.wplat[i], self.wplon[i])
        return qdr % 360.0

    def update(self):
        """Targets follow the selected values; with LNAV on the commanded track is the bearing to the active
        waypoint.  VNAV is never on in the reference's envs."""
        t = self.traf
        if t.swlnav.any():
            qdr2wp = self._update_fms()
            self.trk = np.where(t.swlnav, qdr2wp, self.trk)
        self.tas = vcasormach2tas(t.selspd, t.alt)
        self.alt = t.selalt.copy()
        self.vs = np.where(np.abs(t.selvs) > 0.1, t.selvs, t.apvsdef)

    def selaltcmd(self, idx, alt, vspd=None):
        t = self.traf
        t.selalt[idx] = alt
        t.swvnav[idx] = False
        if vspd:
            t.selvs[idx] = vspd
        else:
            delalt = alt - t.alt[idx]
            if t.selvs[idx] * delalt < 0.0 and abs(t.selvs[idx]) > 0.01:
                t.selvs[idx] = 0.0

    def selhdgcmd(self, idx, hdg):
        """``HDG``: with a wind field the selected heading becomes the track it gives in the current wind
        (above 50 ft; on the ground the heading itself)."""
        t = self.traf
        t.swlnav[idx] = False
        if t.wind.winddim > 0 and t.alt[idx] > 50.0 * ft:
            vnwnd, vewnd = t.wind.getdata(t.lat[idx], t.lon[idx], t.alt[idx])
            gsnorth = t.tas[idx] * np.cos(np.radians(hdg)) + vnwnd
            gseast = t.tas[idx] * np.sin(np.radians(hdg)) + vewnd
            self.trk[idx] = np.degrees(np.arctan2(gseast, gsnorth)) % 360.0
        else:
            self.trk[idx] = hdg

    def selspdcmd(self, idx, casmach):
        self.traf.selspd[idx] = casmach
        self.traf.swvnavspd[idx] = False


class APorASAS:
    def __init__(self, traf):
        self.traf = traf
        self.hdg = np.zeros(0); self.tas = np.zeros(0); self.alt = np.zeros(0); self.vs = np.zeros(0)

    def update(self):
        t = self.traf
        act = t.cr.active if t.cr.enabled else np.zeros(t.ntraf, dtype=bool)
        trk = np.where(act, t.cr.trk, t.ap.trk)
        if t.wind.winddim > 0:      # the ASAS speed command is a ground speed: back to TAS through the wind
            vwn, vwe = t.wind.getdata(t.lat, t.lon, t.alt)
            asastasnorth = t.cr.tas * np.cos(np.radians(t.cr.trk)) - vwn
            asastaseast = t.cr.tas * np.sin(np.radians(t.cr.trk)) - vwe
            asastas = np.sqrt(asastasnorth ** 2 + asastaseast ** 2)
        else:
            asastas = t.cr.tas
        self.tas = np.where(act, asastas, t.ap.tas)
        self.alt = np.where(act, t.cr.alt, t.ap.alt)
        self.vs = np.abs(np.where(act, t.cr.vs, t.ap.vs))
        if t.wind.winddim > 0:      # heading that holds the commanded track in the wind
            Vw = np.sqrt(vwn * vwn + vwe * vwe)
            winddir = np.arctan2(vwe, vwn)
            drift = np.radians(trk) - winddir
            steer = np.arcsin(np.minimum(1.0, np.maximum(-1.0, Vw * np.sin(drift) / np.maximum(0.001, t.tas))))
            self.hdg = (trk + np.degrees(steer)) % 360.0
        else:
            self.hdg = trk % 360.0


class Traffic:
    def __init__(self, settings, sim):
        self.settings = settings
        self.sim = sim
        self.wind = Windfield()
        self.ap = Autopilot(self)
        self.perf = Perf(self)
        self.cd = ConflictDetection(self, settings)
        self.cr = MVP(self, settings)
        self.aporasas = APorASAS(self)
        self.reset()

    # ------------------------------------------------------------ bookkeeping
    def reset(self):
        self.ntraf = 0
        self.id = []
        self.type = []
        for k in _FLOAT_FIELDS:
            setattr(self, k, np.zeros(0))
        for k in _BOOL_FIELDS:
            setattr(self, k, np.zeros(0, dtype=bool))
        self.ap.reset(); self.perf.reset(); self.cd.reset(); self.cr.reset()
        self.wind.clear()

    def id2idx(self, acid):
        if not isinstance(acid, str):
            return [self.id2idx(a) for a in acid]
        try:
            return self.id.index(acid.upper())
        except ValueError:
            return -1

    def delete(self, idx):
        self.id.pop(idx); self.type.pop(idx)
        for k in _FLOAT_FIELDS + _BOOL_FIELDS:
            setattr(self, k, np.delete(getattr(self, k), idx))
        self.ap.delete(idx); self.perf.delete(idx); self.cd.delete(idx); self.cr.delete(idx)
        self.ntraf -= 1
        return True

    def cre(self, acid, actype="B744", aclat=52.0, aclon=4.0, achdg=None, acalt=0, acspd=0):
        """Direct call: no unit conversion (acalt [m], acspd [m/s CAS or Mach])."""
        acid = acid.upper()
        if acid in self.id:
            return False, acid + " already exists."
        n = 1
        for k in _FLOAT_FIELDS:
            setattr(self, k, np.append(getattr(self, k), np.zeros(n)))
        for k in _BOOL_FIELDS:
            setattr(self, k, np.append(getattr(self, k), np.zeros(n, dtype=bool)))
        self.ntraf += n
        self.id.append(acid); self.type.append(actype)
        aclon = float(aclon)
        if aclon > 180.0:
            aclon -= 360.0
        if aclon < -180.0:
            aclon += 360.0
        achdg = 0.0 if achdg is None else achdg
        self.lat[-1] = aclat; self.lon[-1] = aclon; self.alt[-1] = acalt
        self.hdg[-1] = achdg; self.trk[-1] = achdg
        tas, cas, M = vcasormach(np.float64(acspd), np.float64(acalt))
        self.tas[-1] = tas; self.cas[-1] = cas; self.M[-1] = M
        self.gs[-1] = self.tas[-1]
        hdgrad = np.radians(achdg)
        self.gsnorth[-1] = self.tas[-1] * np.cos(hdgrad)
        self.gseast[-1] = self.tas[-1] * np.sin(hdgrad)
        self.apvsdef[-1] = 1500.0 * fpm
        self.ax[-1] = kts
        self.bank[-1] = np.radians(25.0)
        self.selspd[-1] = self.cas[-1]
        self.selalt[-1] = self.alt[-1]
        self.ap.create(n); self.perf.create(n); self.cd.create(n); self.cr.create(n)
        return True

    def creconfs(self, acid, actype, targetidx, dpsi, dcpa, tlosh, dH=None, tlosv=None, spd=None):
        """Create an aircraft in conflict with the target."""
        latref = self.lat[targetidx]; lonref = self.lon[targetidx]; altref = self.alt[targetidx]
        trkref = radians(self.trk[targetidx])
        gsref = self.gs[targetidx]; vsref = self.vs[targetidx]
        cpa = dcpa * nm
        pzr = self.settings.asas_pzr * nm
        pzh = self.settings.asas_pzh * ft
        trk = trkref + radians(dpsi)
        if dH is None:
            acalt = altref
            acvs = 0.0
        else:
            acalt = altref + dH
            tlosv = tlosh if tlosv is None else tlosv
            acvs = vsref - np.sign(dH) * (abs(dH) - pzh) / tlosv
        gsn, gse = gsref * cos(trk), gsref * sin(trk)
        vreln, vrele = gsref * cos(trkref) - gsn, gsref * sin(trkref) - gse
        vrel = sqrt(vreln * vreln + vrele * vrele)
        drelcpa = tlosh * vrel + (0 if cpa > pzr else sqrt(pzr * pzr - cpa * cpa))
        dist = sqrt(drelcpa * drelcpa + cpa * cpa)
        rd = drelcpa / dist
        rx = cpa / dist
        brn = degrees(atan2(-rx * vreln + rd * vrele, rd * vreln + rx * vrele))
        aclat, aclon = geo.kwikpos(latref, lonref, brn, dist / nm)
        tasn, tase = gsn, gse
        acspd = tas2cas(sqrt(tasn * tasn + tase * tase), acalt)
        achdg = degrees(atan2(tase, tasn))
        self.cre(acid, actype, aclat, aclon, achdg, acalt, acspd)
        self.ap.selaltcmd(len(self.lat) - 1, altref, acvs)
        self.vs[-1] = acvs

    # ------------------------------------------------------------------ update
    def update(self):
        if self.ntraf == 0:
            return
        self.ap.update()
        self.cd.update()
        self.cr.update(self.cd)
        self.aporasas.update()
        self.perf.update()
        self.aporasas.tas, self.aporasas.vs, self.aporasas.alt = self.perf.limits(
            self.aporasas.tas, self.aporasas.vs, self.aporasas.alt, self.ax)
        self.update_airspeed()
        self.update_groundspeed()
        self.update_pos()

    def update_airspeed(self):
        dt = self.sim.simdt
        axmax = self.perf.axmax
        delta_spd = self.aporasas.tas - self.tas
        need_ax = np.abs(delta_spd) > np.abs(dt * axmax)
        self.ax = need_ax * np.sign(delta_spd) * axmax
        self.tas = np.where(need_ax, self.tas + self.ax * dt, self.aporasas.tas)
        self.cas = vtas2cas(self.tas, self.alt)
        self.M = vtas2mach(self.tas, self.alt)
        eps = 1e-8
        turnrate = np.degrees(g0 * np.tan(self.bank) / np.maximum(self.tas, eps))
        delhdg = (self.aporasas.hdg - self.hdg + 180) % 360 - 180
        self.swhdgsel = np.abs(delhdg) > np.abs(dt * turnrate)
        self.hdg = np.where(self.swhdgsel, self.hdg + dt * turnrate * np.sign(delhdg),
                            self.aporasas.hdg) % 360.0
        delta_alt = self.aporasas.alt - self.alt
        self.swaltsel = np.abs(delta_alt) > 1.05 * np.maximum(np.abs(dt * self.aporasas.vs),
                                                               np.abs(dt * self.vs))
        target_vs = self.swaltsel * np.sign(delta_alt) * np.abs(self.aporasas.vs)
        delta_vs = target_vs - self.vs
        need_az = np.abs(delta_vs) > 300 * fpm
        az = need_az * np.sign(delta_vs) * (300 * fpm)
        self.vs = np.where(need_az, self.vs + az * dt, target_vs)
        self.vs = np.where(np.isfinite(self.vs), self.vs, 0)

    def update_groundspeed(self):
        if self.wind.winddim == 0:
            self.gsnorth = self.tas * np.cos(np.radians(self.hdg))
            self.gseast = self.tas * np.sin(np.radians(self.hdg))
            self.gs = self.tas
            self.trk = self.hdg
        else:                       # wind acts above 50 ft
            applywind = self.alt > 50.0 * ft
            vnwnd, vewnd = self.wind.getdata(self.lat, self.lon, self.alt)
            self.gsnorth = self.tas * np.cos(np.radians(self.hdg)) + vnwnd * applywind
            self.gseast = self.tas * np.sin(np.radians(self.hdg)) + vewnd * applywind
            self.gs = np.logical_not(applywind) * self.tas + \
                applywind * np.sqrt(self.gsnorth ** 2 + self.gseast ** 2)
            self.trk = np.logical_not(applywind) * self.hdg + \
                applywind * np.degrees(np.arctan2(self.gseast, self.gsnorth)) % 360.0

    def update_pos(self):
        dt = self.sim.simdt
        self.alt = np.where(self.swaltsel, np.round(self.alt + self.vs * dt, 6), self.aporasas.alt)
        self.lat = self.lat + np.degrees(dt * self.gsnorth / Rearth)
        coslat = np.cos(np.deg2rad(self.lat))
        self.lon = self.lon + np.degrees(dt * self.gseast / coslat / Rearth)
