"""ORACLE (test infrastructure, not product code).

CPU restatement of BlueSky's MVP resolver (``bluesky/traffic/asas/mvp.py``:
``MVP.resolve`` / ``MVP.MVP``) and of ``ConflictResolution.resumenav``
(``bluesky/traffic/asas/resolution.py``), SURVEY.md A.7.  Defaults restated:
resofach = resofacv = 1.05, horizontal (HDG+SPD) resolution, no priority rules.
The resolution-method switches of the stack commands RMETHH / RMETHV
(``swresohoriz``, ``swresovert``) select horizontal (default), vertical-only
or combined resolutions; the vertical term, ``timesolveV`` and the resolution
altitude follow the tail of ``MVP.resolve``.  parity unpinned vs upstream (bluesky-simulator==1.0.3 un-vendored).
Pairs are (ownship index, intruder index); aircraft are never deleted mid-episode
in the hot envs, so indices stand in for callsigns.
"""
import numpy as np

_RE = 6371000.0


class MVP:
    def __init__(self, traf, settings):
        self.traf = traf
        self.settings = settings
        self.enabled = False
        self.resofach = settings.asas_marh
        self.resofacv = settings.asas_marv
        self.swresohoriz = True      # RMETHH BOTH (the default): heading + speed
        self.swresovert = False      # RMETHV V/S: vertical speed only
        self.reset()

    def reset(self):
        z = np.zeros(0)
        self.active = np.zeros(0, dtype=bool)
        self.trk = z.copy(); self.tas = z.copy(); self.vs = z.copy(); self.alt = z.copy()
        self.resooffac = np.zeros(0, dtype=bool)
        self.noresoac = np.zeros(0, dtype=bool)
        self.resopairs = set()

    def create(self, n):
        self.active = np.append(self.active, np.zeros(n, dtype=bool))
        for k in ('trk', 'tas', 'vs', 'alt'):
            setattr(self, k, np.append(getattr(self, k), np.zeros(n)))
        self.resooffac = np.append(self.resooffac, np.zeros(n, dtype=bool))
        self.noresoac = np.append(self.noresoac, np.zeros(n, dtype=bool))

    def delete(self, idx):
        for k in ('active', 'trk', 'tas', 'vs', 'alt', 'resooffac', 'noresoac'):
            setattr(self, k, np.delete(getattr(self, k), idx))
        self.resopairs = set()

    # ------------------------------------------------------------------ update
    def update(self, cd):
        if not self.enabled or self.traf.ntraf == 0:
            return
        if cd.confpairs:
            self.trk, self.tas, self.vs, self.alt = self.resolve(cd)
        self.resumenav(cd)

    def pair_mvp(self, cd, qdr, dist, tcpa, tLOS, i1, i2):
        t = self.traf
        rpz_m = max(cd.rpz[i1], cd.rpz[i2]) * self.resofach
        hpz_m = max(cd.hpz[i1], cd.hpz[i2]) * self.resofacv
        dtlook = cd.dtlookahead[i1]
        q = np.radians(qdr)
        drel = np.array([np.sin(q) * dist, np.cos(q) * dist, t.alt[i2] - t.alt[i1]])
        v1 = np.array([t.gseast[i1], t.gsnorth[i1], t.vs[i1]])
        v2 = np.array([t.gseast[i2], t.gsnorth[i2], t.vs[i2]])
        vrel = v2 - v1
        dcpa = drel + vrel * tcpa
        dabsH = np.sqrt(dcpa[0] * dcpa[0] + dcpa[1] * dcpa[1])
        iH = rpz_m - dabsH
        if dabsH <= 10.0:
            dabsH = 10.0
            dcpa[0] = drel[1] / dist * dabsH
            dcpa[1] = -drel[0] / dist * dabsH
        if rpz_m < dist and dabsH < dist:
            erratum = np.cos(np.arcsin(rpz_m / dist) - np.arcsin(dabsH / dist))
            dv1 = ((rpz_m / erratum - dabsH) * dcpa[0]) / (abs(tcpa) * dabsH)
            dv2 = ((rpz_m / erratum - dabsH) * dcpa[1]) / (abs(tcpa) * dabsH)
        else:
            dv1 = (iH * dcpa[0]) / (abs(tcpa) * dabsH)
            dv2 = (iH * dcpa[1]) / (abs(tcpa) * dabsH)
        if abs(vrel[2]) > 0.0:
            iV = hpz_m
            tsolV = abs(drel[2] / vrel[2])
        else:
            iV = hpz_m - abs(drel[2])
            tsolV = tLOS
        if tsolV > dtlook:
            tsolV = tLOS
            iV = hpz_m
        if abs(vrel[2]) > 0.0:
            dv3 = (iV / tsolV) * (-vrel[2] / abs(vrel[2]))
        else:
            dv3 = iV / tsolV
        return np.array([dv1, dv2, dv3]), tsolV

    def resolve(self, cd):
        t = self.traf
        n = t.ntraf
        dv = np.zeros((n, 3))
        timesolveV = np.ones(n) * 1e9
        for (i1, i2), qdr, dist, tcpa, tLOS in zip(cd.confpairs, cd.qdr, cd.dist, cd.tcpa, cd.tLOS):
            dv_mvp, tsolV = self.pair_mvp(cd, qdr, dist, tcpa, tLOS, i1, i2)
            if tsolV < timesolveV[i1]:
                timesolveV[i1] = tsolV
            dv_mvp[2] = 0.5 * dv_mvp[2]
            dv[i1] = dv[i1] - dv_mvp
            if self.noresoac[i2]:
                dv[i1] = dv[i1] + dv_mvp
            if self.resooffac[i1]:
                dv[i1] = 0.0
        dv = dv.T
        v = np.array([t.gseast, t.gsnorth, t.vs])
        newv = v + dv
        if self.swresohoriz:          # horizontal, heading + speed (the default method)
            newtrack = (np.arctan2(newv[0, :], newv[1, :]) * 180 / np.pi) % 360
            newgs = np.sqrt(newv[0, :] ** 2 + newv[1, :] ** 2)
            newvs = t.vs
        elif self.swresovert:         # vertical speed only
            newtrack = t.trk.copy()
            newgs = t.gs.copy()
            newvs = newv[2, :]
        else:                         # horizontal + vertical
            newtrack = (np.arctan2(newv[0, :], newv[1, :]) * 180 / np.pi) % 360
            newgs = np.sqrt(newv[0, :] ** 2 + newv[1, :] ** 2)
            newvs = newv[2, :]
        newgscapped = np.maximum(t.perf.vmin, np.minimum(t.perf.vmax, newgs))
        vscapped = np.maximum(t.perf.vsmin, np.minimum(t.perf.vsmax, newvs))
        # Resolution altitude: follow the autopilot's level-off altitude when it also solves the conflict, else
        # the altitude reached after timesolveV at the capped vertical speed
        asasalttemp = vscapped * timesolveV + t.alt
        signdvs = np.sign(vscapped - t.ap.vs * np.sign(t.selalt - t.alt))
        signalt = np.sign(asasalttemp - t.selalt)
        alt = np.where(np.logical_or(signdvs == 0, signdvs == signalt), asasalttemp, t.selalt)
        altcond = np.logical_and(timesolveV < cd.dtlookahead, np.abs(dv[2, :]) > 0.0)
        alt = np.where(altcond, asasalttemp, alt)
        # horizontal-only resolutions: ASAS altitude equals the selected altitude
        if self.swresohoriz:
            alt = t.selalt.copy()
        self.dv_last = dv.T.copy()
        self.timesolveV = timesolveV
        return newtrack, newgscapped, vscapped, alt

    def resumenav(self, cd):
        t = self.traf
        self.resopairs.update(cd.confpairs)
        delpairs = set()
        changeactive = dict()
        for (i1, i2) in sorted(self.resopairs):
            dist = _RE * np.array([np.radians(t.lon[i2] - t.lon[i1]) *
                                   np.cos(0.5 * np.radians(t.lat[i2] + t.lat[i1])),
                                   np.radians(t.lat[i2] - t.lat[i1])])
            vrel = np.array([t.gseast[i2] - t.gseast[i1], t.gsnorth[i2] - t.gsnorth[i1]])
            past_cpa = np.dot(dist, vrel) > 0.0
            rpz = max(cd.rpz[i1], cd.rpz[i2])
            hdist = np.linalg.norm(dist)
            hor_los = hdist < rpz
            is_bouncing = abs(t.trk[i1] - t.trk[i2]) < 30.0 and hdist < rpz * self.resofach
            if (not past_cpa) or hor_los or is_bouncing:
                changeactive[i1] = True
            else:
                changeactive[i1] = changeactive.get(i1, False)
                delpairs.add((i1, i2))
        for idx, active in changeactive.items():
            self.active[idx] = active
        self.resopairs -= delpairs
