"""ORACLE (test infrastructure, not product code).

ctypes front-end of oracle/_build/libbsgo.so (the plain-C fp64 restatement, see
oracle/c/bsgo.h).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, '_build', 'libbsgo.so')

FIELDS = ['lat', 'lon', 'alt', 'hdg', 'tas', 'vs', 'cas', 'selspd', 'selalt', 'selvs', 'aptrk', 'ax',
          'asas_trk', 'asas_tas', 'asas_vs', 'asas_alt', 'asas_active', 'phase', 'gsn', 'gse', 'lnav']
F = {k: i for i, k in enumerate(FIELDS)}
NF = len(FIELDS)
MAXV = 32
EV = {k: i for i, k in enumerate(['wpt_lat', 'wpt_lon', 'wpt_reach', 'ac_hdg', 'target_alt', 'wpt_x', 'wpt_y',
                                  'nvert', 'total_reward', 'total_intr', 'drift_sum', 'drift_cnt', 'final_alt',
                                  'steps', 'poly'])}
NE = EV['poly'] + 2 * MAXV
KIND_H, KIND_V, KIND_S, KIND_RAW, KIND_M = 0, 1, 2, 3, 4
MAXWIND = 8
NM, FT = 1852.0, 0.3048


class Cfg(C.Structure):
    _fields_ = [('kind', C.c_int32), ('n', C.c_int32), ('substeps', C.c_int32), ('asas_on', C.c_int32),
                ('reso_on', C.c_int32), ('reso_off_own', C.c_int32), ('simdt', C.c_double), ('rpz', C.c_double),
                ('hpz', C.c_double), ('dtlook', C.c_double), ('resofach', C.c_double), ('resofacv', C.c_double),
                ('nwind', C.c_int32), ('wind_obs', C.c_int32), ('wlat', C.c_double * MAXWIND),
                ('wlon', C.c_double * MAXWIND), ('wvn', C.c_double * MAXWIND), ('wve', C.c_double * MAXWIND),
                ('reso_mode', C.c_int32), ('nknot', C.c_int32), ('walt', C.c_double * MAXWIND)]


class CdOut(C.Structure):
    _fields_ = [('inconf', C.c_void_p), ('nconf', C.c_void_p), ('nlos', C.c_void_p), ('near_idx', C.c_void_p),
                ('near_tcpa', C.c_void_p), ('near_dcpa', C.c_void_p), ('near_tin', C.c_void_p),
                ('tcpamax', C.c_void_p)]


class NdCfg(C.Structure):
    _fields_ = [('nx', C.c_int32), ('ny', C.c_int32), ('nz', C.c_int32), ('nt', C.c_int32),
                ('x0', C.c_float), ('y0', C.c_float), ('z0', C.c_float), ('inv_cx', C.c_float),
                ('inv_cy', C.c_float), ('inv_cz', C.c_float), ('dt', C.c_float)]


def build():
    """Compile the C restatement (building the checker is not using it)."""
    subprocess.check_call(['make', '-s', '-C', _HERE])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.bsgo_obs_dim.restype = C.c_int
        _lib.bsgo_obs_dim_cfg.restype = C.c_int
        for f in ('bsgo_vcas2tas', 'bsgo_vtas2cas', 'bsgo_vcasormach2tas'):
            getattr(_lib, f).restype = C.c_double
            getattr(_lib, f).argtypes = [C.c_double, C.c_double]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_cfg(kind, n, substeps=None, simdt=None, asas_on=0, reso_on=0, reso_off_own=0,
             rpz=5 * NM, hpz=1000 * FT, dtlook=300.0, resofach=1.05, resofacv=1.05, wind=None, wind_obs=0, reso_mode=0):
    """wind = (lat[], lon[], vnorth[], veast[]) of the WindFieldWrapper's points (scalars allowed), or None.
    reso_mode: MVP resolution method, 0 horizontal (default) / 1 vertical only / 2 horizontal + vertical."""
    d_sub = {KIND_H: 10, KIND_V: 30, KIND_S: 5, KIND_RAW: 1, KIND_M: 10}
    d_dt = {KIND_H: 5.0, KIND_V: 1.0, KIND_S: 1.0, KIND_RAW: 1.0, KIND_M: 5.0}
    c = Cfg(kind, n, d_sub[kind] if substeps is None else substeps, asas_on, reso_on, reso_off_own,
            d_dt[kind] if simdt is None else simdt, rpz, hpz, dtlook, resofach, resofacv)
    c.reso_mode = int(reso_mode)
    if wind is not None and len(wind) == 5:
        # WindFieldWrapper(env, lat, lon, vnorth[], veast[], alt=[]): ONE point with an altitude profile (knots)
        wl, wo, vn, ve, wa = [np.atleast_1d(np.asarray(x, dtype=np.float64)) for x in wind]
        assert len(wl) == 1 and 1 <= len(wa) <= MAXWIND and len(vn) == len(ve) == len(wa)
        c.nwind = 1; c.wind_obs = int(wind_obs); c.nknot = len(wa)
        c.wlat[0] = wl[0]; c.wlon[0] = wo[0]
        for k in range(len(wa)):
            c.walt[k] = wa[k]; c.wvn[k] = vn[k]; c.wve[k] = ve[k]
    elif wind is not None:
        wl, wo, vn, ve = [np.atleast_1d(np.asarray(x, dtype=np.float64)) for x in wind]
        vn = np.broadcast_to(vn, wl.shape); ve = np.broadcast_to(ve, wl.shape)
        assert 1 <= len(wl) <= MAXWIND
        c.nwind = len(wl); c.wind_obs = int(wind_obs)
        for k in range(len(wl)):
            c.wlat[k] = wl[k]; c.wlon[k] = wo[k]; c.wvn[k] = vn[k]; c.wve[k] = ve[k]
    return c


def obs_dim(kind, n):
    return lib().bsgo_obs_dim(kind, n)


def obs_dim_cfg(cfg):
    return lib().bsgo_obs_dim_cfg(C.byref(cfg))


def new_cdout(E, n):
    return dict(inconf=np.zeros((E, n), np.uint8), nconf=np.zeros((E, n), np.int32), nlos=np.zeros((E, n), np.int32),
                near_idx=np.full((E, n), -1, np.int32), near_tcpa=np.zeros((E, n)), near_dcpa=np.zeros((E, n)),
                near_tin=np.zeros((E, n)), tcpamax=np.zeros((E, n)))


def _cdstruct(cd):
    return CdOut(*[_p(cd[k]) for k in ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa', 'near_dcpa', 'near_tin',
                                       'tcpamax')])


def step(cfg, ac, reso, nac, ev, actions, nthreads=1):
    """In-place K substeps + env layer.  ac [E,NF,N] f64, reso [E,N] u64, nac [E] i32, ev [E,NE] f64."""
    E = ac.shape[0]
    od = obs_dim_cfg(cfg)
    obs = np.zeros((E, od)); rew = np.zeros(E); term = np.zeros(E, np.int32); trunc = np.zeros(E, np.int32)
    cd = new_cdout(E, cfg.n)
    act = None if actions is None else np.ascontiguousarray(actions, dtype=np.float64)
    lib().bsgo_step(C.byref(cfg), C.c_int(E), _p(ac), _p(reso), _p(nac), _p(ev), _p(act), _p(obs), _p(rew),
                    _p(term), _p(trunc), _cdstruct(cd), C.c_int(nthreads))
    return obs, rew, term, trunc, cd


def step_trace(cfg, ac, reso, nac, ev, actions, eps, nthreads=1):
    """step() plus the trace the parity tests enumerate from: returns (obs, rew, term, trunc, cd, trace) with
    trace = dict(near_conf, near_los [E,N] bool, res_margin, kappa [E,N]) -- see oracle/c/bsgo.h."""
    E = ac.shape[0]
    od = obs_dim_cfg(cfg)
    obs = np.zeros((E, od)); rew = np.zeros(E); term = np.zeros(E, np.int32); trunc = np.zeros(E, np.int32)
    cd = new_cdout(E, cfg.n)
    nc = np.zeros((E, cfg.n), np.uint8); nl = np.zeros((E, cfg.n), np.uint8)
    rm = np.zeros((E, cfg.n)); kp = np.zeros((E, cfg.n))
    act = None if actions is None else np.ascontiguousarray(actions, dtype=np.float64)
    lib().bsgo_step_trace(C.byref(cfg), C.c_int(E), _p(ac), _p(reso), _p(nac), _p(ev), _p(act), _p(obs), _p(rew),
                          _p(term), _p(trunc), _cdstruct(cd), C.c_double(eps), _p(nc), _p(nl), _p(rm), _p(kp),
                          C.c_int(nthreads))
    return obs, rew, term, trunc, cd, dict(near_conf=nc.astype(bool), near_los=nl.astype(bool), res_margin=rm, kappa=kp)


def observe(cfg, ac, nac, ev):
    E = ac.shape[0]
    obs = np.zeros((E, obs_dim_cfg(cfg)))
    lib().bsgo_observe(C.byref(cfg), C.c_int(E), _p(ac), _p(nac), _p(ev), _p(obs))
    return obs


def cd_rows(lat, lon, alt, trk, gs, vs, rpz=5 * NM, hpz=1000 * FT, dtlook=300.0, row0=0, row1=None, nthreads=1):
    """Row outputs of one state-based CD sweep; inputs [E,N] f64."""
    E, n = lat.shape
    row1 = n if row1 is None else row1
    cd = new_cdout(E, n)
    arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (lat, lon, alt, trk, gs, vs)]
    lib().bsgo_cd_rows(C.c_int(E), C.c_int(n), *[_p(a) for a in arrs], C.c_double(rpz), C.c_double(hpz),
                       C.c_double(dtlook), C.c_int(row0), C.c_int(row1), _cdstruct(cd), C.c_int(nthreads))
    return cd


def cd_rows_near(lat, lon, alt, trk, gs, vs, eps, rpz=5 * NM, hpz=1000 * FT, dtlook=300.0, nthreads=1):
    """cd_rows plus the near-threshold enumeration: (cd, near_conf [E,N] u8, near_los [E,N] u8) -- row i holds a pair
    whose relative margin to a conflict / LoS threshold is below eps (no dense N x N matrices: giant airspaces)."""
    E, n = lat.shape
    cd = new_cdout(E, n)
    nc = np.zeros((E, n), np.uint8); nl = np.zeros((E, n), np.uint8)
    arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (lat, lon, alt, trk, gs, vs)]
    lib().bsgo_cd_rows_near(C.c_int(E), C.c_int(n), *[_p(a) for a in arrs], C.c_double(rpz), C.c_double(hpz),
                            C.c_double(dtlook), C.c_int(0), C.c_int(n), _cdstruct(cd), C.c_double(eps), _p(nc), _p(nl),
                            C.c_int(nthreads))
    return cd, nc, nl


def detect_dense(lat, lon, alt, trk, gs, vs, rpz=5 * NM, hpz=1000 * FT, dtlook=300.0):
    """Dense N x N matrices for one env (small N)."""
    n = len(lat)
    out = dict(tcpa=np.zeros((n, n)), dcpa2=np.zeros((n, n)), tinconf=np.zeros((n, n)), toutconf=np.zeros((n, n)),
               swconfl=np.zeros((n, n), np.uint8), swlos=np.zeros((n, n), np.uint8),
               margin_conf=np.zeros((n, n)), margin_los=np.zeros((n, n)))
    arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (lat, lon, alt, trk, gs, vs)]
    lib().bsgo_detect_dense(C.c_int(n), *[_p(a) for a in arrs], C.c_double(rpz), C.c_double(hpz), C.c_double(dtlook),
                            *[_p(out[k]) for k in ('tcpa', 'dcpa2', 'tinconf', 'toutconf', 'swconfl', 'swlos',
                                                   'margin_conf', 'margin_los')])
    return out


def ndmap(rec, nx=64, ny=64, nz=8, nt=30, x0=0.0, y0=0.0, z0=0.0, cx=1.0, cy=1.0, cz=1.0, dt=10.0, nthreads=1):
    """rec [E,6,n] f32 (x,y,z,u,v,w).  Returns occ u8, conf u8 [E,nt,nz,ny,nx], ncell i32 [E,nt]."""
    rec = np.ascontiguousarray(rec, dtype=np.float32)
    E, _, n = rec.shape
    g = NdCfg(nx, ny, nz, nt, x0, y0, z0, np.float32(1.0) / np.float32(cx), np.float32(1.0) / np.float32(cy),
              np.float32(1.0) / np.float32(cz), dt)
    occ = np.zeros((E, nt, nz, ny, nx), np.uint8); conf = np.zeros_like(occ); ncell = np.zeros((E, nt), np.int32)
    lib().bsgo_ndmap(C.byref(g), C.c_int(E), C.c_int(n), _p(rec), _p(occ), _p(conf), _p(ncell), C.c_int(nthreads))
    return occ, conf, ncell


def philox(k0, k1, c0, c1, c2, c3):
    out = (C.c_uint32 * 4)()
    lib().bsgo_philox(C.c_uint32(k0), C.c_uint32(k1), C.c_uint32(c0), C.c_uint32(c1), C.c_uint32(c2),
                      C.c_uint32(c3), out)
    return list(out)
