"""ctypes binding of libbsgcore.so (include/bsgcore.h).  The CUDA core is the only compute path:
if the shared library is missing or cannot be loaded this module raises -- there is no CPU fallback."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get('BSG_LIB_PATH') or os.path.join(_HERE, 'libbsgcore.so')   # override: A/B of kernel variants (tools/ab.sh)

ABI_VERSION = 4
KIND_HORIZONTAL, KIND_VERTICAL, KIND_SECTOR, KIND_RAW, KIND_MERGE = 0, 1, 2, 3, 4
MAX_WIND = 8
FP32, FP64 = 0, 1
FIELDS = ['lat', 'lon', 'alt', 'hdg', 'tas', 'vs', 'cas', 'selspd', 'selalt', 'selvs', 'aptrk', 'ax',
          'asas_trk', 'asas_tas', 'asas_vs', 'asas_alt', 'asas_active', 'phase', 'gsn', 'gse', 'lnav']
F = {k: i for i, k in enumerate(FIELDS)}
NF = len(FIELDS)
MAXV = 32
EV = {k: i for i, k in enumerate(['wpt_lat', 'wpt_lon', 'wpt_reach', 'ac_hdg', 'target_alt', 'wpt_x', 'wpt_y', 'nvert',
                                  'total_reward', 'total_intr', 'drift_sum', 'drift_cnt', 'final_alt', 'steps',
                                  'poly'])}
NE = EV['poly'] + 2 * MAXV


class Config(C.Structure):
    _fields_ = [('struct_size', C.c_int32), ('env_kind', C.c_int32), ('num_envs', C.c_int32), ('max_ac', C.c_int32),
                ('precision', C.c_int32), ('substeps', C.c_int32), ('asas_on', C.c_int32), ('reso_on', C.c_int32),
                ('reso_off_own', C.c_int32), ('max_episode_steps', C.c_int32), ('device', C.c_int32),
                ('fixed_nac', C.c_int32), ('density_mode', C.c_int32), ('reso_mode', C.c_int32),
                ('simdt', C.c_double), ('rpz_m', C.c_double), ('hpz_m', C.c_double),
                ('dtlookahead', C.c_double), ('resofach', C.c_double), ('resofacv', C.c_double)]


class StepOut(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ('obs', 'reward', 'terminated', 'truncated', 'info', 'cd_inconf', 'cd_nconf',
                                          'cd_nlos', 'cd_near_idx', 'cd_near_tcpa', 'cd_near_dcpa', 'cd_near_tin',
                                          'cd_tcpamax', 'done')]


class CdParams(C.Structure):
    _fields_ = [('num_envs', C.c_int32), ('num_ac', C.c_int32), ('row0', C.c_int32), ('row1', C.c_int32),
                ('rpz_m', C.c_float), ('hpz_m', C.c_float), ('dtlookahead', C.c_float), ('reserved', C.c_int32)]


MAX_PEERS = 8
CD_STATIC_MAX_ROWS = 1 << 18     # BSG_CD_STATIC_MAX_ROWS
ERR_TOOLARGE = -6


class CdPeers(C.Structure):
    _fields_ = [('num_ranks', C.c_int32), ('rows_per_rank', C.c_int32), ('rec', C.c_void_p * MAX_PEERS),
                ('nconf', C.c_void_p * MAX_PEERS), ('nlos', C.c_void_p * MAX_PEERS), ('tcpamax', C.c_void_p * MAX_PEERS),
                ('key', C.c_void_p * MAX_PEERS)]


class NdmapParams(C.Structure):
    _fields_ = [('num_envs', C.c_int32), ('num_ac', C.c_int32), ('nx', C.c_int32), ('ny', C.c_int32),
                ('nz', C.c_int32), ('nt', C.c_int32), ('x0', C.c_float), ('y0', C.c_float), ('z0', C.c_float),
                ('cell_x', C.c_float), ('cell_y', C.c_float), ('cell_z', C.c_float), ('dt', C.c_float),
                ('reserved', C.c_int32)]


class NdmapGeo(C.Structure):
    _fields_ = [('lat0', C.c_double), ('lon0', C.c_double), ('m_per_deg_lat', C.c_double), ('m_per_deg_lon', C.c_double)]


# every symbol include/bsgcore.h declares: (name, restype, argtypes)
_VP, _I, _I64P = C.c_void_p, C.c_int, C.POINTER(C.c_int64)
SYMBOLS = {
    'bsg_abi_version': (_I, []),
    'bsg_status_string': (C.c_char_p, [_I]),
    'bsg_device_count': (_I, [C.POINTER(C.c_int)]),
    'bsg_create': (_I, [C.POINTER(Config), C.POINTER(_VP)]),
    'bsg_destroy': (_I, [_VP]),
    'bsg_last_error': (C.c_char_p, [_VP]),
    'bsg_obs_dim': (_I, [_VP]),
    'bsg_act_dim': (_I, [_VP]),
    'bsg_launch_count': (_I, [_VP, _I64P]),
    'bsg_set_wind': (_I, [_VP, C.c_int32, _VP, _VP, _VP, _VP, C.c_int32]),
    'bsg_set_wind_profile': (_I, [_VP, C.c_double, C.c_double, C.c_int32, _VP, _VP, _VP, C.c_int32]),
    'bsg_set_obs_noise': (_I, [_VP, C.c_double, C.c_uint64]),
    'bsg_set_state': (_I, [_VP, _VP, _VP, _VP, _VP]),
    'bsg_get_state': (_I, [_VP, _VP, _VP, _VP, _VP]),
    'bsg_reset': (_I, [_VP, _VP, C.c_uint64, _VP]),
    'bsg_set_autoreset': (_I, [_VP, C.c_int32, C.c_uint64]),
    'bsg_observe': (_I, [_VP, _VP, _VP]),
    'bsg_step': (_I, [_VP, _VP, C.POINTER(StepOut), _VP]),
    'bsg_step_host': (_I, [_VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    'bsg_host_views': (_I, [_VP, _VP, _VP, _VP, _VP, _VP]),
    'bsg_step_host_begin': (_I, [_VP, _VP]),
    'bsg_step_host_wait': (_I, [_VP, _VP, _VP, _VP, _VP, _VP]),
    'bsg_step_host_block': (_I, [_VP, _VP]),
    'bsg_step_host_wait_block': (_I, [_VP]),
    'bsg_cd_pack': (_I, [C.c_int32, C.c_int32, _VP, _VP, _VP, _VP, _VP, _VP, C.c_double, C.c_double, _VP, _VP]),
    'bsg_cd_pack_init': (_I, [C.c_int32, C.c_int32, _VP, _VP, _VP, _VP, _VP, _VP, C.c_double, C.c_double, _VP, _VP, _VP, _VP, _VP, _VP]),
    'bsg_cd_sweep': (_I, [C.POINTER(CdParams), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    'bsg_cd_sym_init': (_I, [C.POINTER(CdParams), _VP, _VP, _VP, _VP, _VP]),
    'bsg_cd_sym_sweep': (_I, [C.POINTER(CdParams), _VP, C.c_int32, C.c_int32, _VP, _VP, _VP, _VP, _VP]),
    'bsg_cd_sym_static': (_I, [C.POINTER(CdParams), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    'bsg_cd_sym_sweep_peers': (_I, [C.POINTER(CdParams), C.POINTER(CdPeers), C.c_int32, C.c_int32, _VP]),
    'bsg_cd_sym_finalize_peers': (_I, [C.POINTER(CdParams), C.POINTER(CdPeers), C.c_int32, _VP, _VP, _VP, _VP, _VP, _VP]),
    'bsg_cd_sym_finalize': (_I, [C.POINTER(CdParams), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    'bsg_ndmap': (_I, [C.POINTER(NdmapParams), _VP, _VP, _VP, _VP, _VP]),
    'bsg_ndmap_unpack': (_I, [C.POINTER(NdmapParams), _VP, _VP, _VP]),
    'bsg_ndmap_state': (_I, [_VP, C.POINTER(NdmapParams), C.POINTER(NdmapGeo), _VP, _VP, _VP, _VP, _VP]),
    'bsg_peak_fp32': (_I, [_I, _I, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    'bsg_peak_fp64': (_I, [_I, _I, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
}

_lib = None


def load():
    """Load libbsgcore.so and bind every entry point; raises if the CUDA extension is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export the symbol
        fn.restype = res
        fn.argtypes = args
    if lib.bsg_abi_version() != ABI_VERSION:
        raise ImportError(f"libbsgcore.so ABI {lib.bsg_abi_version()} != binding ABI {ABI_VERSION}")
    _lib = lib
    return lib


class BsgError(RuntimeError):
    pass


def check(rc, handle=None):
    if rc != 0:
        lib = load()
        msg = lib.bsg_status_string(rc).decode()
        if handle:
            msg += ': ' + lib.bsg_last_error(handle).decode()
        raise BsgError(f"libbsgcore: {msg} (status {rc})")
