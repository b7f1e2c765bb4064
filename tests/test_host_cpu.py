"""CPU tests (no GPU): the C-ABI library loads and exports every symbol include/bsgcore.h declares, the
host-side mirror of the reference interface (registration, spaces, obs layout), the host scenario
generators, and the multi-process (gloo, world_size 2) sharding logic."""
import os
import re
import sys
import numpy as np
import pytest
from conftest import ROOT


def test_library_exports_every_declared_symbol():
    from bluesky_gym_hallucination_b200 import _lib
    hdr = open(os.path.join(ROOT, 'include', 'bsgcore.h')).read()
    declared = set(re.findall(r'^(?:int|const char \*)\s*(bsg_\w+)\(', hdr, flags=re.M))
    assert len(declared) >= 20
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    lib = _lib.load()                       # binds all of them; raises on a missing symbol
    assert lib.bsg_abi_version() == _lib.ABI_VERSION == 4
    assert lib.bsg_status_string(-4).decode() == 'unsupported configuration'


def test_struct_layouts_match_header():
    import ctypes as C
    from bluesky_gym_hallucination_b200 import _lib
    assert C.sizeof(_lib.Config) == 14 * 4 + 6 * 8
    assert C.sizeof(_lib.StepOut) == 14 * 8
    assert C.sizeof(_lib.CdParams) == 32 and C.sizeof(_lib.NdmapParams) == 14 * 4
    assert C.sizeof(_lib.CdPeers) == 8 + 5 * 8 * _lib.MAX_PEERS
    assert _lib.NF == 21 and _lib.NE == 14 + 64


def test_no_cpu_fallback_without_device():
    """The product path fails loudly when no CUDA device is present (never routes to the oracle)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from bluesky_gym_hallucination_b200 import _lib
    from bluesky_gym_hallucination_b200.core import BsgCore
    with pytest.raises(_lib.BsgError):
        BsgCore(0, 4, 6)
    src = ''.join(open(os.path.join(ROOT, 'bluesky_gym_hallucination_b200', f)).read()
                  for f in os.listdir(os.path.join(ROOT, 'bluesky_gym_hallucination_b200')) if f.endswith('.py'))
    assert 'import oracle' not in src and 'from oracle' not in src


def test_registration_and_spaces_mirror_the_reference():
    import bluesky_gym_hallucination_b200 as pkg
    from bluesky_gym_hallucination_b200 import vec_env, core, _lib as L
    pkg.register_envs()
    assert [r[0] for r in pkg._REGISTRATIONS] == ['HorizontalCREnv-v0', 'VerticalCREnv-v0', 'SectorCREnv-v0', 'MergeEnv-v0']
    assert [r[2] for r in pkg._REGISTRATIONS] == [300, 300, 200, 50]      # bluesky_gym/__init__.py:21,27,33,45
    obs, act = vec_env.single_spaces(L.KIND_MERGE, 20, wind_obs=True)     # merge_env.py:57-76 + wrappers/wind.py:21-24
    assert list(obs.spaces) == ['cos(drift)', 'sin(drift)', 'airspeed', 'waypoint_dist', 'faf_reached', 'x_r', 'y_r',
                                'vx_r', 'vy_r', 'cos(track)', 'sin(track)', 'distances', 'wind_u', 'wind_v']
    assert sum(s.shape[0] for s in obs.spaces.values()) == 42 and act.shape == (2,) and obs['x_r'].shape == (5,)
    obs, act = vec_env.single_spaces(L.KIND_HORIZONTAL, 6)
    assert list(obs.spaces) == ['intruder_distance', 'cos_difference_pos', 'sin_difference_pos', 'x_difference_speed',
                                'y_difference_speed', 'waypoint_distance', 'cos_drift', 'sin_drift']
    assert obs['intruder_distance'].shape == (5,) and obs['cos_drift'].shape == (1,) and act.shape == (1,)
    assert obs['intruder_distance'].dtype == np.float64
    obs, act = vec_env.single_spaces(L.KIND_VERTICAL, 6)
    assert len(obs.spaces) == 11 and sum(s.shape[0] for s in obs.spaces.values()) == 39
    obs, act = vec_env.single_spaces(L.KIND_SECTOR, 32)
    # the shipped SB3 policies pin these shapes: (1,) x3, (4,) x7, action (2,)
    assert [s.shape[0] for s in obs.spaces.values()] == [1, 1, 1, 4, 4, 4, 4, 4, 4, 4] and act.shape == (2,)
    assert 'cos(track)' in obs.spaces and obs['airspeed'].high[0] == 1
    assert core.obs_slices(L.KIND_HORIZONTAL, 20)[-1] == ('sin_drift', 97, 98)
    assert vec_env.INFO_KEYS[L.KIND_VERTICAL] == ('total_reward', 'total_intrusions', 'final_altitude')


def test_host_scenarios_follow_the_reference_laws(orc):
    from bluesky_gym_hallucination_b200 import scenario as sc, _lib as L
    ac, reso, nac, ev = sc.horizontal(64, 20, seed=1)
    assert np.all(ac[:, L.F['lat'], 0] == 52.0) and np.all(nac == 20)
    d = orc.detect_dense(ac[0, L.F['lat']], ac[0, L.F['lon']], ac[0, L.F['alt']], ac[0, L.F['hdg']], ac[0, L.F['tas']],
                         ac[0, L.F['vs']], dtlook=1e9)
    assert d['swconfl'][0, 1:].all() and np.all((d['tinconf'][0, 1:] > 95) & (d['tinconf'][0, 1:] < 1005))
    ac, reso, nac, ev = sc.sector(32, 40, seed=2, fixed_n=False)
    assert np.all((nac >= 5) & (nac <= 40))
    for e in range(0, 32, 5):
        nv = int(ev[e, L.EV['nvert']])
        vlat = ev[e, L.EV['poly']:L.EV['poly'] + 2 * nv:2]; vlon = ev[e, L.EV['poly'] + 1:L.EV['poly'] + 2 * nv:2]
        assert sc._inside(vlat, vlon, ac[e, L.F['lat'], :nac[e]], ac[e, L.F['lon'], :nac[e]]).all()


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from bluesky_gym_hallucination_b200 import sharding
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    E, N = 3, 40
    full = torch.arange(E * N * 8, dtype=torch.float32).reshape(E, N, 8)         # every rank knows the truth
    r0, r1 = sharding.row_block(N, rank, world)
    got = sharding.gather_records(full[:, r0:r1].contiguous(), N, world)
    e0, e1 = sharding.env_shard(10, rank, world)
    # symmetric sweep: partial outputs of two ranks -> sum / max (float bits) / min (packed key)
    cnt = torch.full((2, E, N), rank + 1, dtype=torch.int32)
    tmax = torch.full((E, N), 10.0 * (rank + 1)); tmax[0, 0] = 0.0
    key = torch.full((E, N), (100 - rank) << 32 | 7, dtype=torch.int64); key[1, 1] = -(5 + rank) << 32
    sharding.reduce_sym_outputs(cnt, tmax, key)
    ok = bool((cnt == 3).all()) and float(tmax[0, 0]) == 0.0 and float(tmax[1, 1]) == 20.0 \
        and int(key[0, 0]) == (99 << 32 | 7) and int(key[1, 1]) == -(6 << 32)
    q.put((rank, bool(torch.equal(got, full)) and ok, (r0, r1), (e0, e1)))
    dist.destroy_process_group()


def test_row_block_allgather_gloo_world2():
    """N > 1 path on CPU: 2 gloo ranks, row-block partition + the single all-gather restore global order."""
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 1000
    ps = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=120) for _ in ps)
    [p.join(60) for p in ps]
    assert res[0][1] and res[1][1]
    assert res[0][2] == (0, 20) and res[1][2] == (20, 40)
    assert res[0][3] == (0, 5) and res[1][3] == (5, 10)
