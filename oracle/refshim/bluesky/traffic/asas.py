"""ORACLE (test infrastructure, not product code).

CPU restatement (numpy, fp64) of BlueSky's state-based conflict detection
(``bluesky/traffic/asas/statebased.py: StateBased.detect``) and
the MVP resolver (``bluesky/traffic/asas/mvp.py`` + ``resolution.py:resumenav``,
SURVEY.md A.7).  bluesky-simulator==1.0.3 is un-vendored: parity unpinned vs
upstream.  The reference envs never enable ASAS; this is the feature you get
with ``ASAS ON`` / ``RESO MVP``.
"""
import numpy as np
from ..tools import geo
from ..tools.aero import nm, ft


def detect(lat, lon, alt, trk, gs, vs, rpz, hpz, dtlookahead):
    """All-pairs state-based CD.  Returns a dict of full N x N matrices + row outputs."""
    n = len(lat)
    I = np.eye(n)
    qdr, dist = geo.kwikqdrdist_matrix(lat, lon, lat, lon)
    qdr = np.asarray(qdr)
    dist = np.asarray(dist) * nm + 1e9 * I
    qdrrad = np.radians(qdr)
    dx = dist * np.sin(qdrrad)
    dy = dist * np.cos(qdrrad)
    trkrad = np.radians(trk)
    u = gs * np.sin(trkrad)
    v = gs * np.cos(trkrad)
    # upstream: du = ownu(1xN) - intu.T(Nx1)  =>  du[i][j] = u_j - u_i (velocity of j relative to i)
    du = u.reshape(1, -1) - u.reshape(-1, 1)
    dv = v.reshape(1, -1) - v.reshape(-1, 1)
    dv2 = du * du + dv * dv
    dv2 = np.where(np.abs(dv2) < 1e-6, 1e-6, dv2)
    vrel = np.sqrt(dv2)
    tcpa = -(du * dx + dv * dy) / dv2 + 1e9 * I
    dcpa2 = np.abs(dist * dist - tcpa * tcpa * dv2)
    rpzm = np.maximum(rpz.reshape(-1, 1), rpz.reshape(1, -1))
    R2 = rpzm * rpzm
    swhorconf = dcpa2 < R2
    dtinhor = np.sqrt(np.maximum(0.0, R2 - dcpa2)) / vrel
    tinhor = np.where(swhorconf, tcpa - dtinhor, 1e8)
    touthor = np.where(swhorconf, tcpa + dtinhor, -1e8)
    dalt = alt.reshape(1, -1) - alt.reshape(-1, 1) + 1e9 * I   # alt_j - alt_i
    dvs = vs.reshape(1, -1) - vs.reshape(-1, 1)                # vs_j - vs_i
    dvs = np.where(np.abs(dvs) < 1e-6, 1e-6, dvs)
    hpzm = np.maximum(hpz.reshape(-1, 1), hpz.reshape(1, -1))
    tcrosshi = (dalt + hpzm) / -dvs
    tcrosslo = (dalt - hpzm) / -dvs
    tinver = np.minimum(tcrosshi, tcrosslo)
    toutver = np.maximum(tcrosshi, tcrosslo)
    tinconf = np.maximum(tinver, tinhor)
    toutconf = np.minimum(toutver, touthor)
    swconfl = (swhorconf * (tinconf <= toutconf) * (toutconf > 0.0)
               * (tinconf < np.asarray(dtlookahead).reshape(-1, 1)) * (1.0 - I)).astype(bool)
    inconf = np.any(swconfl, 1)
    tcpamax = np.max(tcpa * swconfl, 1)
    swlos = (dist < rpzm) * (np.abs(dalt) < hpzm)
    swlos = swlos.astype(bool)
    return dict(qdr=qdr, dist=dist, dx=dx, dy=dy, du=du, dv=dv, dv2=dv2, tcpa=tcpa, dcpa2=dcpa2,
                R2=R2, tinhor=tinhor, touthor=touthor, tinver=tinver, toutver=toutver,
                tinconf=tinconf, toutconf=toutconf, swhorconf=swhorconf, swconfl=swconfl,
                swlos=swlos, inconf=inconf, tcpamax=tcpamax, dalt=dalt, dvs=dvs, hpzm=hpzm,
                rpzm=rpzm)


class ConflictDetection:
    """Holds per-aircraft zones and the last detection result (bs.traf.cd)."""

    def __init__(self, traf, settings):
        self.traf = traf
        self.settings = settings
        self.enabled = False
        self.reset()

    def reset(self):
        self.rpz = np.zeros(0); self.hpz = np.zeros(0); self.dtlookahead = np.zeros(0)
        self.clear()

    def clear(self):
        n = len(self.rpz)
        self.confpairs = []; self.lospairs = []
        self.qdr = np.zeros(0); self.dist = np.zeros(0); self.dcpa = np.zeros(0)
        self.tcpa = np.zeros(0); self.tLOS = np.zeros(0)
        self.inconf = np.zeros(n, dtype=bool); self.tcpamax = np.zeros(n)
        self.last = None

    def create(self, n):
        s = self.settings
        self.rpz = np.append(self.rpz, np.full(n, s.asas_pzr * nm))
        self.hpz = np.append(self.hpz, np.full(n, s.asas_pzh * ft))
        self.dtlookahead = np.append(self.dtlookahead, np.full(n, s.asas_dtlookahead))
        self.inconf = np.append(self.inconf, np.zeros(n, dtype=bool))
        self.tcpamax = np.append(self.tcpamax, np.zeros(n))

    def delete(self, idx):
        for k in ('rpz', 'hpz', 'dtlookahead', 'inconf', 'tcpamax'):
            setattr(self, k, np.delete(getattr(self, k), idx))

    def update(self):
        t = self.traf
        if not self.enabled or t.ntraf == 0:
            return
        r = detect(t.lat, t.lon, t.alt, t.trk, t.gs, t.vs, self.rpz, self.hpz, self.dtlookahead)
        self.last = r
        ii, jj = np.where(r['swconfl'])
        self.confpairs = list(zip(ii.tolist(), jj.tolist()))      # index pairs (row-major order)
        li, lj = np.where(r['swlos'])
        self.lospairs = list(zip(li.tolist(), lj.tolist()))
        self.qdr = r['qdr'][ii, jj]; self.dist = r['dist'][ii, jj]
        self.dcpa = np.sqrt(r['dcpa2'][ii, jj]); self.tcpa = r['tcpa'][ii, jj]
        self.tLOS = r['tinconf'][ii, jj]
        self.inconf = r['inconf']; self.tcpamax = r['tcpamax']
