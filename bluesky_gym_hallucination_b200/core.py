"""Host-side wrapper of one libbsgcore handle (one GPU).  PyTorch is plumbing only: it owns the
device buffers handed to the C-ABI (`tensor.data_ptr()`) and the CUDA stream."""
import ctypes as C
import numpy as np
import torch
from . import _lib as L

OBS_KEYS = {
    # key order of the reference's Dict spaces (horizontal_cr_env.py:51-58, vertical_cr_env.py:70-81,
    # sector_cr_env.py:54-63, merge_env.py:59-72); value = number of entries as a function of K = intruders
    L.KIND_HORIZONTAL: [('intruder_distance', 'K'), ('cos_difference_pos', 'K'), ('sin_difference_pos', 'K'),
                        ('x_difference_speed', 'K'), ('y_difference_speed', 'K'), ('waypoint_distance', 1),
                        ('cos_drift', 1), ('sin_drift', 1)],
    L.KIND_VERTICAL: [('altitude', 1), ('vz', 1), ('target_altitude', 1), ('runway_distance', 1),
                      ('intruder_distance', 'K'), ('cos_difference_pos', 'K'), ('sin_difference_pos', 'K'),
                      ('altitude_difference', 'K'), ('x_difference_speed', 'K'), ('y_difference_speed', 'K'),
                      ('z_difference_speed', 'K')],
    L.KIND_SECTOR: [('cos(drift)', 1), ('sin(drift)', 1), ('airspeed', 1), ('x_r', 4), ('y_r', 4), ('vx_r', 4),
                    ('vy_r', 4), ('cos(track)', 4), ('sin(track)', 4), ('distances', 4)],
    L.KIND_MERGE: [('cos(drift)', 1), ('sin(drift)', 1), ('airspeed', 1), ('waypoint_dist', 1), ('faf_reached', 1),
                   ('x_r', 5), ('y_r', 5), ('vx_r', 5), ('vy_r', 5), ('cos(track)', 5), ('sin(track)', 5),
                   ('distances', 5)],
}
WIND_KEYS = [('wind_u', 1), ('wind_v', 1)]      # WindFieldWrapper(augment_obs=True), wrappers/wind.py:21-24


def obs_slices(kind, n_ac, wind_obs=False):
    """[(key, start, stop)] of the flat observation vector."""
    K, out, p = n_ac - 1, [], 0
    for key, w in OBS_KEYS[kind] + (WIND_KEYS if wind_obs else []):
        w = K if w == 'K' else w
        out.append((key, p, p + w))
        p += w
    return out


def _ptr(t):
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        return t.ctypes.data_as(C.c_void_p)
    return C.c_void_p(t.data_ptr())


class BsgCore:
    """One batched simulation core: E envs x N aircraft slots on one CUDA device."""

    def __init__(self, kind, num_envs, max_ac, precision='fp32', device=0, substeps=0, simdt=0.0, asas_on=False,
                 reso_on=False, reso_off_own=False, max_episode_steps=0, rpz_m=0.0, hpz_m=0.0, dtlookahead=0.0,
                 cd_outputs=False, fixed_nac=False, density_mode='normal', reso_mode=0):
        """reso_mode: MVP resolution method (BlueSky RMETHH / RMETHV) -- 0 / 'horizontal' (heading + speed, the default),
        1 / 'vertical' (vertical speed only), 2 / 'both' (horizontal + vertical); needs reso_on."""
        self.lib = L.load()
        reso_mode = {'horizontal': 0, 'vertical': 1, 'both': 2}.get(reso_mode, reso_mode)
        if not torch.cuda.is_available():
            raise L.BsgError("libbsgcore needs a CUDA device (no CPU fallback)")
        self.kind, self.E, self.N = kind, num_envs, max_ac
        self.fp64 = precision in ('fp64', L.FP64)
        self.device = torch.device('cuda', device)
        self.rdtype = torch.float64 if self.fp64 else torch.float32
        self.npdtype = np.float64 if self.fp64 else np.float32
        cfg = L.Config(C.sizeof(L.Config), kind, num_envs, max_ac, L.FP64 if self.fp64 else L.FP32, substeps,
                       int(asas_on), int(reso_on), int(reso_off_own), max_episode_steps, device, int(fixed_nac),
                       {'normal': 0, 'uniform': 1}[density_mode], int(reso_mode), simdt, rpz_m,
                       hpz_m, dtlookahead, 0.0, 0.0)
        self.h = C.c_void_p()
        rc = self.lib.bsg_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            h, self.h = self.h, C.c_void_p()
            msg = self.lib.bsg_last_error(h).decode() if h else ''
            if h:
                self.lib.bsg_destroy(h)
            raise L.BsgError(f"bsg_create: {self.lib.bsg_status_string(rc).decode()} {msg}")
        self.cd = None
        self._cd_outputs = cd_outputs
        self.wind_obs = False
        self._alloc_outputs()

    def _alloc_outputs(self):
        """(Re)allocate the caller-owned output tensors; obs_dim follows bsg_set_wind(augment_obs)."""
        cd_outputs = self._cd_outputs
        self._views = None
        self._blocks = [None, None]                     # numpy views of the two pinned host blocks (re-made with the outputs)
        self.__dict__.pop('_view_cache', None)
        self.obs_dim = self.lib.bsg_obs_dim(self.h)
        self.act_dim = self.lib.bsg_act_dim(self.h)
        E, N, dev, rd = self.E, self.N, self.device, self.rdtype
        self.obs = torch.zeros((E, max(self.obs_dim, 1)), dtype=rd, device=dev)
        self.reward = torch.zeros(E, dtype=rd, device=dev)
        self.terminated = torch.zeros(E, dtype=torch.int32, device=dev)
        self.truncated = torch.zeros(E, dtype=torch.int32, device=dev)
        self.info = torch.zeros((E, 4), dtype=rd, device=dev)
        self.done = torch.zeros(E, dtype=torch.int32, device=dev)
        self.cd = None
        if cd_outputs:
            self.cd = dict(inconf=torch.zeros((E, N), dtype=torch.uint8, device=dev),
                           nconf=torch.zeros((E, N), dtype=torch.int32, device=dev),
                           nlos=torch.zeros((E, N), dtype=torch.int32, device=dev),
                           near_idx=torch.zeros((E, N), dtype=torch.int32, device=dev),
                           near_tcpa=torch.zeros((E, N), dtype=rd, device=dev),
                           near_dcpa=torch.zeros((E, N), dtype=rd, device=dev),
                           near_tin=torch.zeros((E, N), dtype=rd, device=dev),
                           tcpamax=torch.zeros((E, N), dtype=rd, device=dev))
        cdp = [(_ptr(self.cd[k]) if self.cd else None) for k in ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa',
                                                                   'near_dcpa', 'near_tin', 'tcpamax')]
        self._out = L.StepOut(_ptr(self.obs) if self.obs_dim else None, _ptr(self.reward), _ptr(self.terminated),
                              _ptr(self.truncated), _ptr(self.info), *cdp, _ptr(self.done))

    # ------------------------------------------------------------------ wrapper features (EXT kernels)
    def set_wind(self, lat=None, lon=None, vnorth=None, veast=None, alt=None, augment_obs=False):
        """WindFieldWrapper (bluesky_gym/wrappers/wind.py:9-29): wind vectors [m/s] at points; None switches it off.
        alt = altitudes [m] of a profile: ONE point (scalar lat, lon) whose vnorth / veast are given at those altitudes."""
        if lat is not None and alt is not None:
            al = np.ascontiguousarray(np.atleast_1d(np.asarray(alt, dtype=np.float64)))
            vn = np.ascontiguousarray(np.broadcast_to(np.atleast_1d(np.asarray(vnorth, dtype=np.float64)), al.shape))
            ve = np.ascontiguousarray(np.broadcast_to(np.atleast_1d(np.asarray(veast, dtype=np.float64)), al.shape))
            if np.ndim(lat) != 0 or np.ndim(lon) != 0:
                raise ValueError("a wind altitude profile belongs to ONE point: scalar lat / lon")
            L.check(self.lib.bsg_set_wind_profile(self.h, float(lat), float(lon), len(al), _ptr(al), _ptr(vn), _ptr(ve),
                                                  int(augment_obs)), self.h)
        elif lat is None:
            L.check(self.lib.bsg_set_wind(self.h, 0, None, None, None, None, 0), self.h)
        else:
            la = np.atleast_1d(np.asarray(lat, dtype=np.float64)).copy()
            lo = np.ascontiguousarray(np.broadcast_to(np.atleast_1d(np.asarray(lon, dtype=np.float64)), la.shape))
            vn = np.ascontiguousarray(np.broadcast_to(np.atleast_1d(np.asarray(vnorth, dtype=np.float64)), la.shape))
            ve = np.ascontiguousarray(np.broadcast_to(np.atleast_1d(np.asarray(veast, dtype=np.float64)), la.shape))
            L.check(self.lib.bsg_set_wind(self.h, len(la), _ptr(la), _ptr(lo), _ptr(vn), _ptr(ve), int(augment_obs)), self.h)
        self.wind_obs = bool(augment_obs) and lat is not None
        if self.lib.bsg_obs_dim(self.h) != self.obs_dim:
            self._alloc_outputs()

    def set_obs_noise(self, sigma, seed=0):
        """NoisyObservationWrapper (bluesky_gym/wrappers/uncertainty.py:5-8): obs + N(0, sigma) on every slot."""
        L.check(self.lib.bsg_set_obs_noise(self.h, float(sigma), int(seed)), self.h)

    def close(self):
        h = getattr(self, 'h', None)
        if h:
            self.h = None
            self.lib.bsg_destroy(h)

    __del__ = close

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------ state interchange (fp64, host)
    def set_state(self, ac=None, reso=None, nac=None, ev=None):
        ac = None if ac is None else np.ascontiguousarray(ac, dtype=np.float64)
        reso = None if reso is None else np.ascontiguousarray(reso, dtype=np.uint64)
        nac = None if nac is None else np.ascontiguousarray(nac, dtype=np.int32)
        ev = None if ev is None else np.ascontiguousarray(ev, dtype=np.float64)
        if ac is not None:
            assert ac.shape == (self.E, L.NF, self.N), ac.shape
        if ev is not None:
            assert ev.shape == (self.E, L.NE), ev.shape
        if reso is not None:
            assert reso.size == ((self.N + 63) // 64) * self.E * self.N, reso.shape      # [E, N] or [W, E, N] (N > 64)
        torch.cuda.synchronize(self.device)
        L.check(self.lib.bsg_set_state(self.h, _ptr(ac), _ptr(reso), _ptr(nac), _ptr(ev)), self.h)

    def live_aircraft(self):
        """nac[E]: live aircraft per env, read back from the device."""
        nac = np.zeros(self.E, np.int32)
        L.check(self.lib.bsg_get_state(self.h, None, None, _ptr(nac), None), self.h)
        return nac

    def get_state(self):
        # resopairs rows: [E, N] uint64 for N <= 64; beyond, W = ceil(N / 64) words per aircraft, word-major [W, E, N]
        W = (self.N + 63) // 64
        ac = np.zeros((self.E, L.NF, self.N)); reso = np.zeros((self.E, self.N) if W == 1 else (W, self.E, self.N), np.uint64)
        nac = np.zeros(self.E, np.int32); ev = np.zeros((self.E, L.NE))
        torch.cuda.synchronize(self.device)
        L.check(self.lib.bsg_get_state(self.h, _ptr(ac), _ptr(reso), _ptr(nac), _ptr(ev)), self.h)
        return ac, reso, nac, ev

    # ------------------------------------------------------------------ hot path
    def step(self, actions):
        """actions: CUDA tensor [E, act_dim] of the handle's real dtype.  Asynchronous on the current stream;
        results land in self.obs / reward / terminated / truncated / info (and self.cd)."""
        if self.act_dim:
            assert actions.is_cuda and actions.dtype == self.rdtype and actions.is_contiguous()
            assert actions.numel() == self.E * self.act_dim
        L.check(self.lib.bsg_step(self.h, _ptr(actions) if self.act_dim else None, C.byref(self._out), self._stream()),
                self.h)
        return self.obs, self.reward, self.terminated, self.truncated, self.info

    def observe(self):
        L.check(self.lib.bsg_observe(self.h, _ptr(self.obs), self._stream()), self.h)
        return self.obs

    def reset(self, mask=None, seed=0):
        L.check(self.lib.bsg_reset(self.h, _ptr(mask), C.c_uint64(seed), self._stream()), self.h)

    def set_autoreset(self, enable=True, seed=0):
        L.check(self.lib.bsg_set_autoreset(self.h, int(enable), C.c_uint64(seed)), self.h)   # 0 off, 1 next-step, 2 same-step

    def step_host(self, actions, obs, reward, terminated, truncated, info):
        """End-to-end call with HOST numpy buffers (H2D + kernel + D2H inside)."""
        L.check(self.lib.bsg_step_host(self.h, _ptr(actions), _ptr(obs), _ptr(reward), _ptr(terminated),
                                       _ptr(truncated), _ptr(info)), self.h)

    def host_views(self):
        """numpy views of the pinned host block the last step_host / step_host_wait filled: (obs [E, obs_dim], reward [E],
        terminated [E], truncated [E], info [E, 4]).  The handle alternates between two blocks: valid until the second
        next step_host / close."""
        p = [C.c_void_p() for _ in range(5)]
        L.check(self.lib.bsg_host_views(self.h, *[C.byref(x) for x in p]), self.h)
        cache = self.__dict__.setdefault('_view_cache', {})
        if p[0].value not in cache:
            cache[p[0].value] = self._views_of(p)
        return cache[p[0].value]

    def _views_of(self, p):
        ct = C.c_double if self.fp64 else C.c_float
        arr = lambda ptr, cty, *shape: np.ctypeslib.as_array((cty * int(np.prod(shape))).from_address(ptr.value)).reshape(shape)
        E = self.E
        return (arr(p[0], ct, E, max(self.obs_dim, 1)), arr(p[1], ct, E), arr(p[3], C.c_int32, E),
                arr(p[4], C.c_int32, E), arr(p[2], ct, E, 4))

    def step_host_begin(self, actions):
        """Enqueue one end-to-end step (host actions -> pinned H2D -> launch -> D2H into a pinned block) and return."""
        L.check(self.lib.bsg_step_host_begin(self.h, _ptr(actions)), self.h)

    def step_host_wait(self):
        """Wait for the oldest step begun; returns (obs, reward, terminated, truncated, info) views of its pinned block
        (valid until two more begins)."""
        return self._block_views(self.lib.bsg_step_host_wait_block(self.h))

    def _block_views(self, q):
        """views of pinned block q (the value bsg_step_host_block / _wait_block returned), built once per block"""
        if q < 0:
            L.check(q, self.h)
        v = self._blocks[q]
        if v is None:                                   # the block just filled is the one bsg_host_views describes
            p = [C.c_void_p() for _ in range(5)]
            L.check(self.lib.bsg_host_views(self.h, *[C.byref(x) for x in p]), self.h)
            v = self._blocks[q] = self._views_of(p)
        return v

    def step_host_views(self, actions):
        """End-to-end call with a HOST action array; results are returned as views of the pinned block
        (no second host copy): H2D actions + kernel + D2H obs/reward/flags/info inside."""
        # one ctypes call without pointer marshalling: the C side returns which of the two pinned blocks it filled
        if not (isinstance(actions, np.ndarray) and actions.dtype == self.npdtype and actions.flags.c_contiguous):
            actions = np.ascontiguousarray(np.asarray(actions), dtype=self.npdtype)
        return self._block_views(self.lib.bsg_step_host_block(self.h, actions.ctypes.data))

    def ndmap(self, nx=64, ny=64, nz=8, nt=30, x0=0.0, y0=0.0, z0=0.0, cx=1.0, cy=1.0, cz=1.0, dt=10.0, lat0=52.0, lon0=4.0,
              out=None, want_rec=False):
        """ND-map of the LIVE state (bsg_ndmap_state): straight-line look-ahead of every live aircraft from where it is
        now, flat-earth metres from (lat0, lon0).  Returns (occ, conf_bits, ncell[, rec]) like core.ndmap."""
        E, N, dev = self.E, self.N, self.device
        if out is None:
            out = (torch.empty((E, nt, nz, ny, nx), dtype=torch.uint8, device=dev),
                   torch.empty((E, nt, nz * ny * nx // 8), dtype=torch.uint8, device=dev),
                   torch.empty((E, nt), dtype=torch.int32, device=dev))
        rec = torch.empty((E, 6, N), dtype=torch.float32, device=dev) if want_rec else None
        p = L.NdmapParams(E, N, nx, ny, nz, nt, x0, y0, z0, cx, cy, cz, dt, 0)
        m_lat = float(np.radians(1.0) * 6371000.0)
        geo = L.NdmapGeo(lat0, lon0, m_lat, m_lat * float(np.cos(np.radians(lat0))))
        L.check(self.lib.bsg_ndmap_state(self.h, C.byref(p), C.byref(geo), _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), _ptr(rec),
                                         self._stream()), self.h)
        return out + ((rec,) if want_rec else ())

    def launch_count(self):
        n = C.c_int64(0)
        L.check(self.lib.bsg_launch_count(self.h, C.byref(n)), self.h)
        return n.value


# ---------------------------------------------------------------------- stand-alone kernels
def cd_pack(lat, lon, alt, trk, gs, vs, lat0, lon0):
    """SoA CUDA tensors [E,N] (lat/lon f64, others f32) -> packed fp32 CD records [E,N,8]."""
    lib = L.load()
    E, N = lat.shape
    rec = torch.empty((E, N, 8), dtype=torch.float32, device=lat.device)
    s = C.c_void_p(torch.cuda.current_stream(lat.device).cuda_stream)
    L.check(lib.bsg_cd_pack(E, N, _ptr(lat), _ptr(lon), _ptr(alt), _ptr(trk), _ptr(gs), _ptr(vs), float(lat0),
                            float(lon0), _ptr(rec), s))
    return rec


def cd_sweep(rec, row0=0, row1=None, rpz_m=5 * 1852.0, hpz_m=1000 * 0.3048, dtlookahead=300.0, out=None):
    """All-pairs state-based CD row outputs for rows [row0,row1) of every env.  rec: [E,N,8] f32 CUDA."""
    lib = L.load()
    E, N, _ = rec.shape
    row1 = N if row1 is None else row1
    dev = rec.device
    if out is None:
        out = dict(inconf=torch.zeros((E, N), dtype=torch.uint8, device=dev),
                   nconf=torch.zeros((E, N), dtype=torch.int32, device=dev),
                   nlos=torch.zeros((E, N), dtype=torch.int32, device=dev),
                   near_idx=torch.full((E, N), -1, dtype=torch.int32, device=dev),
                   near_tcpa=torch.zeros((E, N), dtype=torch.float32, device=dev),
                   near_dcpa=torch.zeros((E, N), dtype=torch.float32, device=dev),
                   near_tin=torch.zeros((E, N), dtype=torch.float32, device=dev),
                   tcpamax=torch.zeros((E, N), dtype=torch.float32, device=dev))
    p = L.CdParams(E, N, row0, row1, rpz_m, hpz_m, dtlookahead, 0)
    s = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    L.check(lib.bsg_cd_sweep(C.byref(p), _ptr(rec), *[_ptr(out[k]) for k in ('inconf', 'nconf', 'nlos', 'near_idx',
                                                                                 'near_tcpa', 'near_dcpa', 'near_tin',
                                                                                 'tcpamax')], s))
    return out


def cd_sweep_sym(rec, rpz_m=5 * 1852.0, hpz_m=1000 * 0.3048, dtlookahead=300.0, out=None, group=None, rank=0, world=1):
    """Symmetric all-pairs sweep (every unordered pair once): same row outputs as cd_sweep for ALL rows.
    world > 1: the tiles of the upper triangle are dealt round-robin to the ranks of `group` and the partial outputs
    are combined with three all-reduces (sum / max / min) before the finalize stage."""
    lib = L.load()
    E, N, _ = rec.shape
    dev = rec.device
    if out is None:
        out = dict(inconf=torch.empty((E, N), dtype=torch.uint8, device=dev),
                   cnt=torch.empty((2, E, N), dtype=torch.int32, device=dev),        # nconf, nlos (one SUM all-reduce)
                   near_idx=torch.empty((E, N), dtype=torch.int32, device=dev),
                   near_tcpa=torch.empty((E, N), dtype=torch.float32, device=dev),
                   near_dcpa=torch.empty((E, N), dtype=torch.float32, device=dev),
                   near_tin=torch.empty((E, N), dtype=torch.float32, device=dev),
                   tcpamax=torch.empty((E, N), dtype=torch.float32, device=dev),
                   key=torch.empty((E, N), dtype=torch.int64, device=dev))
        out['nconf'], out['nlos'] = out['cnt'][0], out['cnt'][1]
    p = L.CdParams(E, N, 0, N, rpz_m, hpz_m, dtlookahead, 0)
    s = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    L.check(lib.bsg_cd_sym_init(C.byref(p), _ptr(out['nconf']), _ptr(out['nlos']), _ptr(out['tcpamax']), _ptr(out['key']), s))
    L.check(lib.bsg_cd_sym_sweep(C.byref(p), _ptr(rec), rank, world, _ptr(out['nconf']), _ptr(out['nlos']),
                                 _ptr(out['tcpamax']), _ptr(out['key']), s))
    if world > 1:
        from . import sharding
        sharding.reduce_sym_outputs(out['cnt'], out['tcpamax'], out['key'], group)
    L.check(lib.bsg_cd_sym_finalize(C.byref(p), _ptr(rec), _ptr(out['nconf']), _ptr(out['key']), _ptr(out['inconf']),
                                    _ptr(out['near_idx']), _ptr(out['near_tcpa']), _ptr(out['near_dcpa']),
                                    _ptr(out['near_tin']), s))
    return out


def cd_sweep_sym_static(rec, out, rpz_m=5 * 1852.0, hpz_m=1000 * 0.3048, dtlookahead=300.0):
    """bsg_cd_sym_static: the symmetric sweep for a host that must not allocate GPU memory (memory handed to a pool,
    fixed-rate tick, binding without an allocator -- include/bsgcore.h).  `out` is the caller's dict of output tensors
    (cd_sweep's keys, [E,N]; a missing key is not written); the row accumulators live in an array inside the library, so
    E * N <= _lib.CD_STATIC_MAX_ROWS and calls on one GPU must not overlap.  Same results as cd_sweep, bit for bit."""
    lib = L.load()
    E, N, _ = rec.shape
    p = L.CdParams(E, N, 0, N, rpz_m, hpz_m, dtlookahead, 0)
    s = C.c_void_p(torch.cuda.current_stream(rec.device).cuda_stream)
    with torch.cuda.device(rec.device):
        L.check(lib.bsg_cd_sym_static(C.byref(p), _ptr(rec), *[_ptr(out[k]) if k in out else None for k in
                                      ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa', 'near_dcpa', 'near_tin', 'tcpamax')], s))
    return out


def ndmap(rec, nx=64, ny=64, nz=8, nt=30, x0=0.0, y0=0.0, z0=0.0, cx=1.0, cy=1.0, cz=1.0, dt=10.0, out=None):
    """rec: [E,6,n] f32 CUDA (x,y,z,u,v,w).  Returns occ u8 [E,nt,nz,ny,nx], conf_bits u8 [E,nt,nz*ny*nx/8] (1 bit per
    cell: cell c of a slab <-> bit c & 7 of byte c >> 3; ndmap_unpack gives the dense 0/1 bytes), ncell i32 [E,nt]."""
    lib = L.load()
    E, _, n = rec.shape
    dev = rec.device
    if out is None:
        out = (torch.empty((E, nt, nz, ny, nx), dtype=torch.uint8, device=dev),
               torch.empty((E, nt, nz * ny * nx // 8), dtype=torch.uint8, device=dev),
               torch.empty((E, nt), dtype=torch.int32, device=dev))
    p = L.NdmapParams(E, n, nx, ny, nz, nt, x0, y0, z0, cx, cy, cz, dt, 0)
    s = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    L.check(lib.bsg_ndmap(C.byref(p), _ptr(rec), _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), s))
    return out


def ndmap_unpack(conf_bits, nx=64, ny=64, nz=8):
    """conf_bits u8 [E,nt,cells/8] -> dense u8 0/1 [E,nt,nz,ny,nx]."""
    lib = L.load()
    E, nt, _ = conf_bits.shape
    dev = conf_bits.device
    out = torch.empty((E, nt, nz, ny, nx), dtype=torch.uint8, device=dev)
    p = L.NdmapParams(E, 1, nx, ny, nz, nt, 0.0, 0.0, 0.0, 1.0, 1.0, 1.0, 1.0, 0)
    s = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    L.check(lib.bsg_ndmap_unpack(C.byref(p), _ptr(conf_bits), _ptr(out), s))
    return out


def peak_fp32(device=0, iters=4096):
    lib = L.load()
    tf, ms = C.c_double(0), C.c_double(0)
    L.check(lib.bsg_peak_fp32(device, iters, C.byref(tf), C.byref(ms)))
    return tf.value, ms.value


def peak_fp64(device=0, iters=2048):
    lib = L.load()
    tf, ms = C.c_double(0), C.c_double(0)
    L.check(lib.bsg_peak_fp64(device, iters, C.byref(tf), C.byref(ms)))
    return tf.value, ms.value


class PeerSweep:
    """Giant-airspace symmetric sweep over NVLink peer memory (one process per GPU).

    Every rank owns aircraft rows [rank*rows, (rank+1)*rows) of every env.  Its packed records and its row-output
    accumulators live in ONE symmetric-memory block that every peer maps; the sweep kernel reads the intruder
    records straight from their owners and flushes the (rare) conflict updates to the owners' accumulators with
    system-scope atomics, so a step is: pack + init (local) -> barrier -> sweep -> barrier -> finalize (own rows).
    No all-gather, no all-reduce.  world == 1 uses plain device memory and no barrier."""

    def __init__(self, E, N, rank=0, world=1, group=None, device=None):
        import torch.distributed as dist
        if N % world:
            raise ValueError(f"{N} aircraft do not split into {world} equal row blocks")
        self.lib = L.load()
        self.E, self.N, self.rank, self.world, self.rows = E, N, rank, world, N // world
        dev = torch.device('cuda', torch.cuda.current_device()) if device is None else device
        n = E * self.rows
        # block layout (bytes): rec [2][n][8] f32 | nconf [n] i32 | nlos [n] i32 | tcpamax [n] f32 | key [n] i64
        # TWO record regions, alternated by step parity: finalize of step k reads the winner's record from its owner
        # (no cross-GPU barrier follows it), so a faster rank's pack of step k+1 must not land in the region a slower
        # rank is still reading; it lands in the other one, which nobody reads before the next barrier.
        off = {'rec': 0, 'rec1': 32 * n, 'nconf': 64 * n, 'nlos': 68 * n, 'tcpamax': 72 * n, 'key': 80 * n}   # key 8-aligned
        nbytes = 88 * n
        self.h = None
        if world > 1:
            import torch.distributed._symmetric_memory as symm
            self.block = symm.empty((nbytes,), dtype=torch.uint8, device=dev)
            self.h = symm.rendezvous(self.block, (group or dist.group.WORLD).group_name)
            bases = list(self.h.buffer_ptrs)
        else:
            self.block = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
            bases = [self.block.data_ptr()]
        view = lambda key, dt, *shape: self.block[off[key]:off[key] + int(np.prod(shape)) * torch.empty((), dtype=dt).element_size()].view(dt).view(*shape)
        self.recs = (view('rec', torch.float32, E, self.rows, 8), view('rec1', torch.float32, E, self.rows, 8))
        self.rec = self.recs[0]
        self.parity = 0
        self.debug_finalize_delay = 0
        self.nconf = view('nconf', torch.int32, E, self.rows)
        self.nlos = view('nlos', torch.int32, E, self.rows)
        self.tcpamax = view('tcpamax', torch.float32, E, self.rows)
        self.key = view('key', torch.int64, E, self.rows)
        self.peer_tabs = (L.CdPeers(), L.CdPeers())          # one peer table per record region
        for par, tab in enumerate(self.peer_tabs):
            tab.num_ranks, tab.rows_per_rank = world, self.rows
            for g in range(world):
                tab.rec[g] = bases[g] + off['rec1' if par else 'rec']
                for f in ('nconf', 'nlos', 'tcpamax', 'key'):
                    getattr(tab, f)[g] = bases[g] + off[f]
        self.peers = self.peer_tabs[0]
        self.out = dict(inconf=torch.empty((E, self.rows), dtype=torch.uint8, device=dev),
                        near_idx=torch.empty((E, self.rows), dtype=torch.int32, device=dev),
                        near_tcpa=torch.empty((E, self.rows), dtype=torch.float32, device=dev),
                        near_dcpa=torch.empty((E, self.rows), dtype=torch.float32, device=dev),
                        near_tin=torch.empty((E, self.rows), dtype=torch.float32, device=dev),
                        nconf=self.nconf, nlos=self.nlos, tcpamax=self.tcpamax)

    def step(self, lat, lon, alt, trk, gs, vs, lat0, lon0, rpz_m=5 * 1852.0, hpz_m=1000 * 0.3048, dtlookahead=300.0):
        """SoA CUDA tensors [E, rows] of the OWN aircraft -> row outputs [E, rows] of the own rows."""
        lib, E, dev = self.lib, self.E, self.block.device
        s = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        full = L.CdParams(E, self.N, 0, self.N, rpz_m, hpz_m, dtlookahead, 0)
        self.rec, self.peers = self.recs[self.parity], self.peer_tabs[self.parity]
        self.parity ^= 1
        # records packed and row accumulators reset by ONE launch
        L.check(lib.bsg_cd_pack_init(E, self.rows, _ptr(lat), _ptr(lon), _ptr(alt), _ptr(trk), _ptr(gs), _ptr(vs), float(lat0),
                                     float(lon0), _ptr(self.rec), _ptr(self.nconf), _ptr(self.nlos), _ptr(self.tcpamax),
                                     _ptr(self.key), s))
        if self.h is not None:
            self.h.barrier(channel=0)                  # every rank's records packed, accumulators initialised
        L.check(lib.bsg_cd_sym_sweep_peers(C.byref(full), C.byref(self.peers), self.rank, self.world, s))
        if self.h is not None:
            self.h.barrier(channel=1)                  # every rank's atomics have landed
        if self.debug_finalize_delay:                  # test hook: widen the window between the barrier and finalize
            torch.cuda._sleep(int(self.debug_finalize_delay))
        o = self.out
        L.check(lib.bsg_cd_sym_finalize_peers(C.byref(full), C.byref(self.peers), self.rank, _ptr(o['inconf']),
                                              _ptr(o['near_idx']), _ptr(o['near_tcpa']), _ptr(o['near_dcpa']),
                                              _ptr(o['near_tin']), s))
        return o
