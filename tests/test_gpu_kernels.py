"""GPU tests of the stand-alone kernels behind the C-ABI: tiled CD sweep (giant airspace), ND-map,
FMA peak probes.  Checker: the CPU oracle (oracle/)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
EPS32 = 2e-4


def _pack(lat, lon, alt, trk, gs, vs):
    from bluesky_gym_hallucination_b200 import core
    dev = torch.device('cuda', 0)
    t64 = lambda a: torch.as_tensor(a, dtype=torch.float64, device=dev)
    t32 = lambda a: torch.as_tensor(a, dtype=torch.float32, device=dev)
    return core.cd_pack(t64(lat), t64(lon), t32(alt), t32(trk), t32(gs), t32(vs), 52.0, 4.0)


@pytest.mark.parametrize('E,N,box', [(3, 1500, 200.0), (2, 1031, 120.0), (1, 37, 60.0)])
def test_cd_sweep_vs_oracle(orc, E, N, box):
    """Row outputs of the TMA-tiled sweep == oracle rows; every mismatching row is explained by a pair the
    oracle places within EPS32 of a zone threshold (those pairs are enumerated)."""
    from bluesky_gym_hallucination_b200 import core, scenario as sc
    lat, lon, alt, trk, gs, vs = sc.raw_traffic(E, N, seed=2, box_nm=box)
    ref = orc.cd_rows(lat, lon, alt, trk, gs, vs, nthreads=8)
    out = core.cd_sweep(_pack(lat, lon, alt, trk, gs, vs))
    torch.cuda.synchronize()
    got = {k: v.cpu().numpy() for k, v in out.items()}
    assert ref['nconf'].sum() > 0
    near_rows = 0
    for e in range(E):
        d = orc.detect_dense(lat[e], lon[e], alt[e], trk[e], gs[e], vs[e])
        off = ~np.eye(N, dtype=bool)
        near_c = ((d['margin_conf'] < EPS32) & off).any(1)
        near_l = ((d['margin_los'] < EPS32) & off).any(1)
        bad_c = got['nconf'][e] != ref['nconf'][e]
        bad_l = got['nlos'][e] != ref['nlos'][e]
        assert not (bad_c & ~near_c).any(), f"conflict-count mismatch off-threshold, env {e}: rows {np.where(bad_c & ~near_c)[0][:8]}"
        assert not (bad_l & ~near_l).any(), f"LoS-count mismatch off-threshold, env {e}"
        near_rows += int(near_c.sum() + near_l.sum())
        ok = ~bad_c & ~near_c
        assert np.array_equal(got['inconf'][e][ok], ref['inconf'][e][ok])
        same = ok & (got['near_idx'][e] == ref['near_idx'][e]) & (ref['near_idx'][e] >= 0)
        # the soonest conflict of a row is an argmin over tCPA: a different index is accepted only where the oracle
        # itself holds a TIE -- the intruder the GPU chose is a conflict of the row whose tCPA is within 1e-5 (relative,
        # floor 1 s) of the row's minimum; every such row is enumerated
        diff = np.where(ok & (ref['near_idx'][e] >= 0) & (got['near_idx'][e] != ref['near_idx'][e]))[0]
        ties = []
        for i in diff:
            jg, jr = int(got['near_idx'][e][i]), int(ref['near_idx'][e][i])
            assert jg >= 0 and d['swconfl'][i, jg], f"env {e} row {i}: GPU's soonest intruder {jg} is no conflict of the oracle's row"
            gap = abs(d['tcpa'][i, jg] - d['tcpa'][i, jr]) / max(abs(d['tcpa'][i, jr]), 1.0)
            assert gap < 1e-5, f"env {e} row {i}: intruder {jg} vs {jr}, tCPA gap {gap:.2e}"
            ties.append((int(i), jg, jr, float(gap)))
        if ties:
            print(f"env {e}: soonest-conflict ties enumerated (row, GPU intruder, oracle intruder, relative tCPA gap): {ties[:6]}")
        for k, tol in (('near_tcpa', 1e-5), ('near_tin', 2e-5), ('tcpamax', 1e-5)):
            err = np.abs(got[k][e][same] - ref[k][e][same]) / np.maximum(np.abs(ref[k][e][same]), 1.0)
            assert err.max(initial=0) < tol * 10, (k, err.max())
        # dCPA: 1e-5 relative, floored at the protected-zone radius (what it is compared against)
        err = np.abs(got['near_dcpa'][e][same] - ref['near_dcpa'][e][same]) / np.maximum(ref['near_dcpa'][e][same], 5 * 1852.0)
        assert err.max(initial=0) < 1e-5, err.max()
    print(f"near-threshold rows enumerated: {near_rows}")


def test_giant_airspace_full_env_vs_oracle(orc):
    """BASELINE.json configs[4] at its FULL per-env size: one env of 10 000 aircraft (1e8 ordered pairs), symmetric
    sweep (the default C5 kernel) vs the oracle's rows; every mismatching row is enumerated by the oracle as holding a
    pair within EPS32 of a zone threshold (oracle.cd_rows_near: no dense matrices)."""
    from bluesky_gym_hallucination_b200 import core, scenario as sc
    E, N = 1, 10000
    lat, lon, alt, trk, gs, vs = sc.raw_traffic(E, N, seed=3, box_nm=1000.0)
    ref, near_c, near_l = orc.cd_rows_near(lat, lon, alt, trk, gs, vs, EPS32, nthreads=16)
    out = core.cd_sweep_sym(_pack(lat, lon, alt, trk, gs, vs))
    torch.cuda.synchronize()
    got = {k: v.cpu().numpy() for k, v in out.items()}
    assert ref['nconf'].sum() > 1000
    near_c, near_l = near_c.astype(bool), near_l.astype(bool)
    bad_c = got['nconf'] != ref['nconf']
    bad_l = got['nlos'] != ref['nlos']
    print(f"C5 full env: {int(ref['nconf'].sum())} conflict entries, {int(bad_c.sum())} conflict-count / {int(bad_l.sum())} LoS-count "
          f"row mismatches, near-threshold rows enumerated {int(near_c.sum())} + {int(near_l.sum())}")
    assert not (bad_c & ~near_c).any(), np.argwhere(bad_c & ~near_c)[:8]
    assert not (bad_l & ~near_l & ~near_c).any(), np.argwhere(bad_l & ~near_l)[:8]
    ok = ~bad_c & ~near_c
    assert np.array_equal(got['inconf'][ok] != 0, ref['inconf'][ok] != 0)
    same = ok & (got['near_idx'] == ref['near_idx']) & (ref['near_idx'] >= 0)
    # a different soonest intruder only at a tie of the oracle's own tCPA values (three-aircraft oracle evaluation of
    # the row's aircraft with both candidates); every such row is enumerated
    ties = []
    for e, i in np.argwhere(ok & (ref['near_idx'] >= 0) & (got['near_idx'] != ref['near_idx'])):
        jg, jr = int(got['near_idx'][e, i]), int(ref['near_idx'][e, i])
        assert jg >= 0
        sel = [i, jg, jr]
        d3 = orc.detect_dense(lat[e][sel], lon[e][sel], alt[e][sel], trk[e][sel], gs[e][sel], vs[e][sel])
        assert d3['swconfl'][0, 1], f"row {i}: GPU's soonest intruder {jg} is no conflict of the oracle's row"
        gap = abs(d3['tcpa'][0, 1] - d3['tcpa'][0, 2]) / max(abs(d3['tcpa'][0, 2]), 1.0)
        assert gap < 1e-4, f"row {i}: intruder {jg} vs {jr}, tCPA gap {gap:.2e}"      # (the bound of the time comparison below)
        ties.append((int(i), jg, jr, float(gap)))
    print(f"C5 full env: soonest-conflict ties enumerated: {len(ties)} {ties[:5]}")
    # times: fp32 records in a 1000 NM box resolve positions to 0.125 m, i.e. tCPA to ~0.1 m / |v_rel| -- an absolute
    # floor of 0.02 s beside the 1e-4 relative bound
    for k in ('near_tcpa', 'tcpamax'):
        d = np.abs(got[k][same] - ref[k][same])
        assert np.all(d <= np.maximum(1e-4 * np.abs(ref[k][same]), 2e-2)), (k, d.max())


def test_cd_sweep_row_block_equals_full():
    """Row-block partition of the pair matrix (the multi-GPU sharding) == the full sweep, bitwise."""
    from bluesky_gym_hallucination_b200 import core, scenario as sc
    lat, lon, alt, trk, gs, vs = sc.raw_traffic(2, 1200, seed=4)
    rec = _pack(lat, lon, alt, trk, gs, vs)
    full = core.cd_sweep(rec)
    parts = core.cd_sweep(rec, 0, 500)
    parts = core.cd_sweep(rec, 500, 1200, out=parts)
    torch.cuda.synchronize()
    for k in full:
        assert torch.equal(full[k], parts[k]), k


@pytest.mark.parametrize('E,N,box', [(2, 1200, 300.0), (3, 257, 60.0), (1, 20, 30.0), (2, 700, 150.0)])
def test_cd_sweep_symmetric_equals_ordered(E, N, box):
    """Symmetric sweep (every unordered pair once, integer atomics) == the ordered sweep, bitwise, on dense and
    sparse traffic; and the tile-dealt partial sweeps (the multi-GPU split) combine to the same result."""
    from bluesky_gym_hallucination_b200 import core, scenario as sc, _lib as L
    import ctypes as C
    lat, lon, alt, trk, gs, vs = sc.raw_traffic(E, N, seed=11, box_nm=box)
    rec = _pack(lat, lon, alt, trk, gs, vs)
    full = core.cd_sweep(rec)
    sym = core.cd_sweep_sym(rec)
    torch.cuda.synchronize()
    assert int(full['nconf'].sum()) > 0
    for k in ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa', 'near_dcpa', 'near_tin', 'tcpamax'):
        assert torch.equal(full[k], sym[k]), k
    # three "ranks": partial outputs combined by sum / max / min like the all-reduce would
    lib = L.load()
    p = L.CdParams(E, N, 0, N, 5 * 1852.0, 1000 * 0.3048, 300.0, 0)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    parts = []
    for r in range(3):
        o = dict(nconf=torch.empty((E, N), dtype=torch.int32, device='cuda'), nlos=torch.empty((E, N), dtype=torch.int32, device='cuda'),
                 tcpamax=torch.empty((E, N), dtype=torch.float32, device='cuda'), key=torch.empty((E, N), dtype=torch.int64, device='cuda'))
        assert lib.bsg_cd_sym_init(C.byref(p), ptr(o['nconf']), ptr(o['nlos']), ptr(o['tcpamax']), ptr(o['key']), None) == 0
        assert lib.bsg_cd_sym_sweep(C.byref(p), ptr(rec), r, 3, ptr(o['nconf']), ptr(o['nlos']), ptr(o['tcpamax']), ptr(o['key']), None) == 0
        parts.append(o)
    torch.cuda.synchronize()
    nconf = sum(o['nconf'] for o in parts)
    key = torch.stack([o['key'] for o in parts]).min(0).values
    tmax = torch.stack([o['tcpamax'].view(torch.int32) for o in parts]).max(0).values.view(torch.float32)
    assert torch.equal(nconf, full['nconf']) and torch.equal(sum(o['nlos'] for o in parts), full['nlos'])
    assert torch.equal(tmax, full['tcpamax']) and torch.equal(key, sym['key'])


def test_peer_sweep_single_rank_equals_ordered():
    """PeerSweep (peer-table form of the symmetric sweep; world 1 = plain device memory) == ordered sweep, bitwise."""
    from bluesky_gym_hallucination_b200 import core, scenario as sc
    E, N = 2, 900
    lat, lon, alt, trk, gs, vs = sc.raw_traffic(E, N, seed=21, box_nm=250.0)
    rec = _pack(lat, lon, alt, trk, gs, vs)
    full = core.cd_sweep(rec)
    t64 = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float64, device='cuda')
    t32 = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32, device='cuda')
    ps = core.PeerSweep(E, N)
    out = ps.step(t64(lat), t64(lon), t32(alt), t32(trk), t32(gs), t32(vs), 52.0, 4.0)
    torch.cuda.synchronize()
    assert int(full['nconf'].sum()) > 0
    for k in ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa', 'near_dcpa', 'near_tin', 'tcpamax'):
        assert torch.equal(full[k], out[k]), k


@pytest.mark.parametrize('E,n', [(3, 500), (1, 7)])
def test_ndmap_bit_exact(orc, E, n):
    from bluesky_gym_hallucination_b200 import core
    rng = np.random.default_rng(9)
    rec = np.zeros((E, 6, n), np.float32)
    rec[:, 0] = rng.uniform(-5000, 300000, (E, n)); rec[:, 1] = rng.uniform(-5000, 300000, (E, n))
    rec[:, 2] = rng.uniform(2000, 13000, (E, n))
    rec[:, 3] = rng.uniform(-250, 250, (E, n)); rec[:, 4] = rng.uniform(-250, 250, (E, n))
    rec[:, 5] = rng.choice([-10.0, 0.0, 10.0], (E, n))
    kw = dict(nx=64, ny=64, nz=8, nt=30, x0=0.0, y0=0.0, z0=3048.0, cx=4630.0, cy=4630.0, cz=304.8 * 4, dt=10.0)
    occ, conf, ncell = orc.ndmap(rec, nthreads=8, **kw)
    g_occ, g_bits, g_ncell = core.ndmap(torch.as_tensor(rec, device='cuda'), **kw)
    g_conf = core.ndmap_unpack(g_bits)
    torch.cuda.synchronize()
    assert occ.sum() > 0 and (conf.sum() > 0 or n < 100)
    assert np.array_equal(g_occ.cpu().numpy(), occ)
    assert np.array_equal(g_conf.cpu().numpy(), conf)
    assert np.array_equal(g_ncell.cpu().numpy(), ncell)
    # the bitmask itself: cell c of a slab <-> bit c & 7 of byte c >> 3 (numpy little-endian bit order)
    assert np.array_equal(g_bits.cpu().numpy(), np.packbits(conf.reshape(E, 30, -1), axis=-1, bitorder='little'))


@pytest.mark.parametrize('nx,ny,nz,n', [(32, 16, 4, 40), (128, 24, 4, 300), (64, 64, 8, 700), (64, 32, 2, 300)])
def test_ndmap_grids_and_saturation(orc, nx, ny, nz, n):
    """Other grid shapes (rows of 1 / 2 / 4 words), more aircraft than CTA threads (records re-read from memory), and the
    saturating counters: > 255 aircraft parked in ONE cell must give occ == 255 there and leave its neighbours alone."""
    from bluesky_gym_hallucination_b200 import core
    E, nt = 2, 5
    rng = np.random.default_rng(3)
    rec = np.zeros((E, 6, n), np.float32)
    rec[:, 0] = rng.uniform(0, nx * 1000.0, (E, n)); rec[:, 1] = rng.uniform(0, ny * 1000.0, (E, n))
    rec[:, 2] = rng.uniform(0, nz * 300.0, (E, n))
    rec[:, 3] = rng.uniform(-200, 200, (E, n)); rec[:, 4] = rng.uniform(-200, 200, (E, n))
    if n >= 300:                        # 280 aircraft flying in formation inside one cell, away from the faces
        rec[0, 0, :280] = 5500.0; rec[0, 1, :280] = 5500.0; rec[0, 2, :280] = 450.0 if nz > 2 else 150.0
        rec[0, 3, :280] = 10.0; rec[0, 4, :280] = -10.0; rec[0, 5, :280] = 0.0
    kw = dict(nx=nx, ny=ny, nz=nz, nt=nt, x0=0.0, y0=0.0, z0=0.0, cx=1000.0, cy=1000.0, cz=300.0, dt=10.0)
    occ, conf, ncell = orc.ndmap(rec, nthreads=4, **kw)
    g_occ, g_bits, g_ncell = core.ndmap(torch.as_tensor(rec, device='cuda'), **kw)
    g_conf = core.ndmap_unpack(g_bits, nx=nx, ny=ny, nz=nz)
    torch.cuda.synchronize()
    if n >= 300:
        assert occ.max() == 255
    assert np.array_equal(g_occ.cpu().numpy(), occ)
    assert np.array_equal(g_conf.cpu().numpy(), conf)
    assert np.array_equal(g_ncell.cpu().numpy(), ncell)


def test_ndmap_of_live_state(orc):
    """bsg_ndmap_state: the map of a handle's LIVE traffic (Sector envs with the density law: dead slots are dropped)
    equals the oracle's map of the records the call reports, and those records are the state's positions / velocities."""
    from bluesky_gym_hallucination_b200.core import BsgCore
    from bluesky_gym_hallucination_b200 import core, _lib as L
    E, N = 48, 50
    c = BsgCore(L.KIND_SECTOR, E, N, precision='fp32', asas_on=True)
    c.reset(seed=3)
    c.step(torch.zeros((E, 2), dtype=torch.float32, device='cuda'))
    kw = dict(nx=64, ny=64, nz=8, nt=30, x0=-110e3, y0=-90e3, z0=-1000.0, cx=3000.0, cy=3000.0, cz=500.0, dt=10.0)
    g_occ, g_bits, g_ncell, rec = c.ndmap(lat0=51.990426702297746, lon0=4.376124857109851, want_rec=True, **kw)
    torch.cuda.synchronize()
    ac, _, nac, _ = c.get_state()
    rec = rec.cpu().numpy()
    F = L.F
    m_lat = np.radians(1.0) * 6371000.0
    for e in range(0, E, 5):
        n = nac[e]
        assert np.allclose(rec[e, 1, :n], (ac[e, F['lat'], :n] - 51.990426702297746) * m_lat, rtol=1e-6, atol=0.05)
        assert np.allclose(rec[e, 0, :n], (ac[e, F['lon'], :n] - 4.376124857109851) * m_lat * np.cos(np.radians(51.990426702297746)), rtol=1e-6, atol=0.05)
        assert np.allclose(rec[e, 2, :n], ac[e, F['alt'], :n]) and np.allclose(rec[e, 3, :n], ac[e, F['gse'], :n], atol=1e-3)
        assert np.allclose(rec[e, 4, :n], ac[e, F['gsn'], :n], atol=1e-3) and (rec[e, 0, n:] < -1e30).all()
    occ, conf, ncell = orc.ndmap(rec, nthreads=8, **kw)
    assert occ.sum() > 0 and int(occ.reshape(E, 30, -1).sum(-1).max()) <= nac.max()
    assert np.array_equal(occ.reshape(E, 30, -1).sum(-1)[:, 0] <= nac, np.ones(E, bool))
    assert np.array_equal(g_occ.cpu().numpy(), occ)
    assert np.array_equal(core.ndmap_unpack(g_bits).cpu().numpy(), conf)
    assert np.array_equal(g_ncell.cpu().numpy(), ncell)
    c.close()


def test_fma_peaks_are_sane():
    from bluesky_gym_hallucination_b200 import core
    tf32, _ = core.peak_fp32(0, 2048)
    tf64, _ = core.peak_fp64(0, 1024)
    assert 20.0 < tf32 < 90.0, tf32      # B200: 148 SM x 128 lanes x 2 x clock
    assert 5.0 < tf64 < 50.0, tf64


def test_device_reset_properties(orc):
    """bsg_reset (Philox scenario generation on the device): value laws of the reference's reset path,
    determinism per (seed, env, episode), masked reset leaves other envs untouched, and the creconfs
    known-answer: every Horizontal intruder is detected in conflict with tLOS ~ its drawn tlosh."""
    from bluesky_gym_hallucination_b200.core import BsgCore
    from bluesky_gym_hallucination_b200 import scenario as sc, _lib as L
    E, N = 96, 20
    F, EV = L.F, L.EV
    c = BsgCore(L.KIND_HORIZONTAL, E, N, precision='fp64')
    c.reset(seed=77)
    ac, reso, nac, ev = c.get_state()
    assert (nac == N).all() and (reso == 0).all()
    assert np.all(ac[:, F['lat'], 0] == 52.0) and np.all(ac[:, F['lon'], 0] == 4.0)
    assert np.all((ev[:, EV['wpt_lat']] > 52.89) & (ev[:, EV['wpt_lat']] < 53.35))
    for e in (0, 5, 95):
        d = orc.detect_dense(ac[e, F['lat']], ac[e, F['lon']], ac[e, F['alt']], ac[e, F['hdg']], ac[e, F['tas']], ac[e, F['vs']],
                             dtlook=1e9)
        assert d['swconfl'][0, 1:].all()
        assert np.all((d['tinconf'][0, 1:] > 95) & (d['tinconf'][0, 1:] < 1005))
        assert np.all(np.sqrt(d['dcpa2'][0, 1:]) < 5 * 1852.0)
    c2 = BsgCore(L.KIND_HORIZONTAL, E, N, precision='fp64')
    c2.reset(seed=77)
    assert np.array_equal(c2.get_state()[0], ac)
    mask = torch.zeros(E, dtype=torch.int32, device='cuda'); mask[3] = 1
    c.reset(mask=mask, seed=78)                # another seed: env 3 alone draws a new scenario
    ac2 = c.get_state()[0]
    assert np.array_equal(np.delete(ac2, 3, 0), np.delete(ac, 3, 0)) and not np.array_equal(ac2[3], ac[3])
    c.reset(mask=mask, seed=77)                # the same seed again: the Gymnasium contract, reset(seed=s) reproduces
    assert np.array_equal(c.get_state()[0], ac)
    c.close(); c2.close()
    # Vertical
    c = BsgCore(L.KIND_VERTICAL, E, 6, precision='fp32')
    c.reset(seed=5)
    ac, _, _, ev = c.get_state()
    a0 = ac[:, F['alt'], 0]
    assert np.all((a0 >= 2000) & (a0 < 4000)) and np.all(np.abs(ev[:, EV['target_alt']] - a0) <= 500)
    assert np.all(ac[:, F['selalt'], 1:] == ac[:, F['alt'], 1:])
    c.close()
    # Sector: everything inside its polygon, reference density law, headings to the perimeter
    N = 50
    c = BsgCore(L.KIND_SECTOR, E, N, precision='fp64')
    c.reset(seed=9)
    ac, _, nac, ev = c.get_state()
    assert np.all((nac >= 5) & (nac <= N)) and 8 < nac.mean() < 25
    for e in range(0, E, 7):
        nv = int(ev[e, EV['nvert']])
        vlat = ev[e, EV['poly']:EV['poly'] + 2 * nv:2]; vlon = ev[e, EV['poly'] + 1:EV['poly'] + 2 * nv:2]
        n = nac[e]
        assert 3 <= nv <= L.MAXV
        assert sc._inside(vlat, vlon, ac[e, F['lat'], :n], ac[e, F['lon'], :n]).all()
        assert np.all((ac[e, F['hdg'], :n] >= 0) & (ac[e, F['hdg'], :n] < 360))
    # a reset env steps without truncating immediately
    act = torch.zeros((E, 2), dtype=torch.float64, device='cuda')
    c.step(act)
    torch.cuda.synchronize()
    assert c.truncated.sum().item() <= E // 10
    c.close()


def test_cd_sweep_equal_velocity_pairs(orc):
    """Pairs with identical velocity vectors (|v_rel|^2 under the reference's 1e-6 clamp): dCPA must equal the
    current distance, not 0 -- the cross-product form needs its clamp correction term."""
    from bluesky_gym_hallucination_b200 import core, scenario as sc
    lat, lon, alt, trk, gs, vs = sc.raw_traffic(2, 400, seed=8, box_nm=150.0)
    trk[:] = np.round(trk / 45.0) * 45.0       # 8 headings
    gs[:] = 150.0                              # one speed -> thousands of equal-velocity pairs
    ref, near_c, near_l = orc.cd_rows_near(lat, lon, alt, trk, gs, vs, EPS32, nthreads=8)
    out = core.cd_sweep(_pack(lat, lon, alt, trk, gs, vs))
    torch.cuda.synchronize()
    # every row that differs is enumerated and must hold a pair the oracle places within EPS32 of a zone threshold
    bad = out['nconf'].cpu().numpy() != ref['nconf']
    print(f"equal-velocity pairs: rows differing {np.argwhere(bad).tolist()}, near-threshold rows {int(near_c.astype(bool).sum())}")
    assert not (bad & ~near_c.astype(bool)).any(), np.argwhere(bad & ~near_c.astype(bool))
    bad_l = out['nlos'].cpu().numpy() != ref['nlos']
    assert not (bad_l & ~near_l.astype(bool) & ~near_c.astype(bool)).any(), np.argwhere(bad_l)
    assert int(ref['nconf'].sum()) > 100


def test_giant_airspace_full_size_properties():
    """BASELINE.json C5 size (16 envs x 10 000 aircraft, 1000 x 1000 NM): symmetric sweep == ordered sweep bitwise,
    row-block halves == full, and the properties of a symmetric conflict matrix: sum of the conflict counts is even
    per env, an aircraft is in conflict iff its count is positive, the nearest intruder of a conflicting aircraft is
    itself in conflict, LoS counts never exceed conflict counts."""
    from bluesky_gym_hallucination_b200 import core, scenario as sc
    E, N = 16, 10000
    lat, lon, alt, trk, gs, vs = sc.raw_traffic(E, N, seed=3, box_nm=1000.0)
    rec = _pack(lat, lon, alt, trk, gs, vs)
    full = core.cd_sweep(rec)
    sym = core.cd_sweep_sym(rec)
    parts = core.cd_sweep(rec, 0, N // 2)
    parts = core.cd_sweep(rec, N // 2, N, out=parts)
    torch.cuda.synchronize()
    for k in ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa', 'near_dcpa', 'near_tin', 'tcpamax'):
        assert torch.equal(full[k], sym[k]), k
        assert torch.equal(full[k], parts[k]), k
    nconf, nlos, inconf, near = (full[k].cpu().numpy() for k in ('nconf', 'nlos', 'inconf', 'near_idx'))
    assert nconf.sum() > 1000
    assert (nconf.sum(1) % 2 == 0).all() and (nlos.sum(1) % 2 == 0).all()
    assert np.array_equal(inconf != 0, nconf > 0)
    assert (nlos <= nconf).all()
    e_idx, i_idx = np.nonzero(nconf > 0)
    j_idx = near[e_idx, i_idx]
    assert (j_idx >= 0).all() and (j_idx != i_idx).all() and (nconf[e_idx, j_idx] > 0).all()
    assert (near[nconf == 0] == -1).all()


def test_ndmap_full_size_properties(orc):
    """BASELINE.json C4 size (1024 envs x 500 aircraft x 30 look-ahead steps, 64 x 64 x 8 grid): envs across the batch are
    bit-exact against the CPU restatement; for the whole batch ncell == number of conflict cells, every slab holds at
    most n aircraft, and a conflict cell has an occupied cell in its 3x3x3 neighbourhood."""
    from bluesky_gym_hallucination_b200 import core
    E, n = 1024, 500
    rng = np.random.default_rng(12)
    rec = np.zeros((E, 6, n), np.float32)
    rec[:, 0] = rng.uniform(0, 370400, (E, n)); rec[:, 1] = rng.uniform(0, 370400, (E, n))     # 200 x 200 NM box
    rec[:, 2] = rng.uniform(3048, 12192, (E, n))
    hdg = rng.uniform(0, 2 * np.pi, (E, n)); spd = rng.uniform(120, 250, (E, n))
    rec[:, 3] = spd * np.sin(hdg); rec[:, 4] = spd * np.cos(hdg)
    rec[:, 5] = rng.choice([-10.0, 0.0, 0.0, 10.0], (E, n))
    kw = dict(nx=64, ny=64, nz=8, nt=30, x0=0.0, y0=0.0, z0=3048.0, cx=370400.0 / 64, cy=370400.0 / 64, cz=304.8 * 4, dt=10.0)
    g_occ, g_bits, g_ncell = core.ndmap(torch.as_tensor(rec, device='cuda'), **kw)
    g_conf = core.ndmap_unpack(g_bits)
    torch.cuda.synchronize()
    pick = [0, 1, 511, 1022, 1023]
    occ, conf, ncell = orc.ndmap(np.ascontiguousarray(rec[pick]), nthreads=8, **kw)
    assert occ.sum() > 0 and conf.sum() > 0
    assert np.array_equal(g_occ[pick].cpu().numpy(), occ)
    assert np.array_equal(g_conf[pick].cpu().numpy(), conf)
    assert np.array_equal(g_ncell[pick].cpu().numpy(), ncell)
    assert torch.equal(g_conf.view(E, 30, -1).sum(-1, dtype=torch.int32), g_ncell)
    assert g_bits.numel() * 8 == g_conf.numel()                 # 1 bit per cell leaves the device: 1.125 B per cell in all
    assert int(g_occ.view(E, 30, -1).sum(-1, dtype=torch.int32).max()) <= n
    for e0 in range(0, E, 128):                       # dilation property, 128 envs at a time (memory)
        o = (g_occ[e0:e0 + 128] > 0).float().view(-1, 1, 8, 64, 64)
        nb = torch.nn.functional.max_pool3d(o, 3, stride=1, padding=1).view(128, 30, 8, 64, 64) > 0
        assert bool((nb | (g_conf[e0:e0 + 128] == 0)).all())


def test_reset_seed_reproduces_after_episodes():
    """reset(seed=s) draws the same scenario whenever it is called -- after construction and after several auto-reset
    episodes -- and the auto-reset episodes that follow it repeat as well (the episode counter that keys
    Philox is re-based by every stand-alone reset)."""
    from bluesky_gym_hallucination_b200.core import BsgCore
    from bluesky_gym_hallucination_b200 import _lib as L
    E, N = 64, 6
    c = BsgCore(L.KIND_HORIZONTAL, E, N, precision='fp32', max_episode_steps=4)
    act = torch.zeros((E, 1), dtype=torch.float32, device='cuda')

    def run():
        c.reset(seed=31)
        c.set_autoreset(True, seed=32)
        first = c.get_state()[0].copy()
        for _ in range(11):                     # TimeLimit 4: every env passes through two auto-resets
            c.step(act)
        torch.cuda.synchronize()
        return first, c.get_state()[0].copy()
    a0, a1 = run()
    b0, b1 = run()
    assert np.array_equal(a0, b0, equal_nan=True) and np.array_equal(a1, b1, equal_nan=True)
    assert not np.array_equal(a0, a1, equal_nan=True)
    c.close()


def test_sector_waypoints_follow_the_sorted_law():
    """sector_cr_env.py:162-176: num_ac perimeter draws are SORTED and aircraft i is handed the i-th smallest, so the
    agent's waypoint is the minimum of num_ac uniforms (mean perimeter / (num_ac + 1)) -- not a uniform point -- and
    the arc-length positions increase with the aircraft index.  Device generator vs the host twin (scenario.py)."""
    from bluesky_gym_hallucination_b200.core import BsgCore
    from bluesky_gym_hallucination_b200 import scenario as sc, _lib as L
    E, N = 2048, 32
    F, EV = L.F, L.EV
    c = BsgCore(L.KIND_SECTOR, E, N, precision='fp64', fixed_nac=True)
    c.reset(seed=5)
    ac, _, nac, ev = c.get_state()
    c.close()
    assert (nac == N).all()
    ccl = np.cos(np.radians(51.990426702297746))

    def arc_fraction(ev_row, wx, wy):
        nv = int(ev_row[EV['nvert']])
        x = (ev_row[EV['poly']:EV['poly'] + 2 * nv:2] - 51.990426702297746) * 60
        y = (ev_row[EV['poly'] + 1:EV['poly'] + 2 * nv:2] - 4.376124857109851) * 60 * ccl
        seg = np.hypot(np.roll(x, -1) - x, np.roll(y, -1) - y)
        best, s0 = (np.inf, 0.0), 0.0
        for k in range(nv):
            m = (k + 1) % nv
            t = np.clip(((wx - x[k]) * (x[m] - x[k]) + (wy - y[k]) * (y[m] - y[k])) / seg[k] ** 2, 0, 1)
            d = np.hypot(wx - (x[k] + t * (x[m] - x[k])), wy - (y[k] + t * (y[m] - y[k])))
            if d < best[0]:
                best = (d, s0 + t * seg[k])
            s0 += seg[k]
        assert best[0] < 1e-6
        return best[1] / seg.sum()
    frac = np.array([arc_fraction(ev[e], ev[e, EV['wpt_x']], ev[e, EV['wpt_y']]) for e in range(E)])
    expect = 1.0 / (nac + 1.0)                                  # E[min of n uniforms]
    assert abs(frac.mean() - expect.mean()) < 4 * np.sqrt((expect ** 2).mean() / E), (frac.mean(), expect.mean())
    assert frac.mean() < 0.15                                   # a uniform perimeter point would give 0.5
    h_ac, _, h_nac, h_ev = sc.sector(E, N, seed=6)
    h_frac = np.array([arc_fraction(h_ev[e], h_ev[e, EV['wpt_x']], h_ev[e, EV['wpt_y']]) for e in range(E)])
    from scipy import stats
    assert stats.ks_2samp(frac * (nac + 1), h_frac * (h_nac + 1)).pvalue > 1e-3


@pytest.mark.parametrize('kind,N', [(2, 50), (0, 20), (2, 64), (2, 33)])
def test_pair_work_list_equals_owner_walk(kind, N):
    """Resolution kernels (fp32, N <= 64): the env's unordered conflicting / resolving pairs are evaluated once each from
    a work list and both owners pick the result up (ownship j: the exact negative).  The fallback in which every owner
    evaluates its own row (BSG_FORCE_WALK; also taken when an env overflows the list) must give the same bits -- state,
    resopairs rows, observations and CD rows, over chained steps with MVP active."""
    import os
    from bluesky_gym_hallucination_b200.core import BsgCore
    E = 192
    rng = np.random.default_rng(5)
    cores = [BsgCore(kind, E, N, precision='fp32', asas_on=True, reso_on=True, reso_off_own=True, cd_outputs=True) for _ in range(2)]
    for c in cores:
        c.reset(seed=17)
    na = 2 if kind == 2 else 1
    nres = 0
    try:
        for s in range(12):
            act = torch.as_tensor(rng.uniform(-1, 1, (E, na)).astype(np.float32), device='cuda')
            outs = []
            for w, c in enumerate(cores):
                if w: os.environ['BSG_FORCE_WALK'] = '1'
                else: os.environ.pop('BSG_FORCE_WALK', None)
                c.step(act)
                torch.cuda.synchronize()
                ac, reso, nac, ev = c.get_state()
                outs.append((ac, reso, c.obs.cpu().numpy(), {k: v.cpu().numpy() for k, v in c.cd.items()}))
            (a0, r0, o0, cd0), (a1, r1, o1, cd1) = outs
            assert np.array_equal(r0, r1), f"step {s}: resopairs rows differ"
            assert np.array_equal(a0, a1, equal_nan=True), f"step {s}: aircraft state differs"
            assert np.array_equal(o0, o1, equal_nan=True), f"step {s}: observations differ"
            for k in cd0:
                assert np.array_equal(cd0[k], cd1[k], equal_nan=True), (s, k)
            nres += int((r0 != 0).sum())
    finally:
        os.environ.pop('BSG_FORCE_WALK', None)
        for c in cores:
            c.close()
    assert nres > 0, "no pair in resolution: the test did not exercise MVP"


def test_step_host_begin_wait_pipeline():
    """bsg_step_host_begin / bsg_step_host_wait (two pinned blocks per handle, two steps in flight) return, step by
    step, exactly what the blocking bsg_step_host returns for the same action sequence; a third begin is refused."""
    from bluesky_gym_hallucination_b200.core import BsgCore
    from bluesky_gym_hallucination_b200 import _lib as L
    E, N, T = 256, 20, 7
    rng = np.random.default_rng(2)
    acts = [rng.uniform(-1, 1, (E, 1)).astype(np.float32) for _ in range(T)]
    cores = [BsgCore(L.KIND_HORIZONTAL, E, N, precision='fp32', asas_on=True) for _ in range(2)]
    for c in cores:
        c.reset(seed=8); c.set_autoreset(True, seed=9)
    ref = []
    for a in acts:
        o, r, te, tr, info = cores[0].step_host_views(a)
        ref.append((o.copy(), r.copy(), te.copy(), tr.copy(), info.copy()))
    c = cores[1]
    got = []
    c.step_host_begin(acts[0])
    for k in range(1, T):
        c.step_host_begin(acts[k])                       # two in flight
        if k == 1:
            with pytest.raises(L.BsgError):
                c.step_host_begin(acts[k])
        got.append(tuple(x.copy() for x in c.step_host_wait()))
    got.append(tuple(x.copy() for x in c.step_host_wait()))
    with pytest.raises(L.BsgError):
        c.step_host_wait()
    for k in range(T):
        for x, y in zip(ref[k], got[k]):
            assert np.array_equal(x, y, equal_nan=True), k
    for c in cores:
        c.close()
