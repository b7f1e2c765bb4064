"""GPU test of bsg_cd_sym_static, the entry point for hosts that may not allocate GPU memory (include/bsgcore.h):
a trainer whose memory pool already owns the whole GPU, a fixed-rate tick that cannot absorb a cudaMalloc stall, a
binding that is handed its buffers and has no allocator.  Its sweep accumulators live in an array inside the library
that every call shares, so the test makes two calls in a row on DIFFERENT traffic and requires each to equal the
existing bsg_cd_sweep on the same records bit for bit (the tolerance test_cd_sweep_symmetric_equals_ordered uses):
a row the second call failed to re-initialise would carry the first call's conflicts.

The file is named test_zz_* so that pytest collects it last: `-x` stops at the first failure, and a failure here must
not hide the other tests.  It needs nothing from them (the first call on the device does the entry point's set-up)."""
import ctypes as C
import pytest
import torch

pytestmark = pytest.mark.gpu
KEYS = ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa', 'near_dcpa', 'near_tin', 'tcpamax')
DTYPES = dict(inconf=torch.uint8, nconf=torch.int32, nlos=torch.int32, near_idx=torch.int32, near_tcpa=torch.float32,
              near_dcpa=torch.float32, near_tin=torch.float32, tcpamax=torch.float32)


def _records(E, N, seed, box):
    from bluesky_gym_hallucination_b200 import core, scenario as sc
    dev = torch.device('cuda', 0)
    lat, lon, alt, trk, gs, vs = sc.raw_traffic(E, N, seed=seed, box_nm=box)
    t64 = lambda a: torch.as_tensor(a, dtype=torch.float64, device=dev)
    t32 = lambda a: torch.as_tensor(a, dtype=torch.float32, device=dev)
    return core.cd_pack(t64(lat), t64(lon), t32(alt), t32(trk), t32(gs), t32(vs), 52.0, 4.0)


def test_static_scratch_sweep_equals_cd_sweep_on_two_calls_in_a_row():
    from bluesky_gym_hallucination_b200 import core, _lib as L
    lib = L.load()
    ptr = lambda t: C.c_void_p(t.data_ptr())
    # call 1: dense traffic, many conflicts in every row block; call 2: fewer, sparser aircraft and fewer envs, so
    # the rows it uses overlap the first call's leftovers; call 3: the first problem again
    cases = [(3, 1500, 21, 200.0), (2, 700, 22, 150.0), (3, 1500, 21, 200.0)]
    recs = [_records(*c) for c in cases]
    want = [core.cd_sweep(r) for r in recs]                       # the existing entry point (ordered sweep)
    # the caller owns inputs and outputs; they are made before the calls and pre-filled with a pattern
    outs = [{k: torch.full(r.shape[:2], 77, dtype=DTYPES[k], device=r.device) for k in KEYS} for r in recs]
    torch.cuda.synchronize()
    assert int(want[0]['nconf'].sum()) > 0 and int(want[1]['nconf'].sum()) > 0
    for n, (rec, out) in enumerate(zip(recs, outs)):
        E, N, _ = rec.shape
        p = L.CdParams(E, N, 0, N, 5 * 1852.0, 1000 * 0.3048, 300.0, 0)
        if n > 0:                                                 # set-up (first call on the device) is behind us
            torch.cuda.synchronize()
            free0, alloc0 = torch.cuda.mem_get_info(0)[0], torch.cuda.memory_allocated(0)
        rc = lib.bsg_cd_sym_static(C.byref(p), ptr(rec), *[ptr(out[k]) for k in KEYS], None)
        assert rc == 0, lib.bsg_status_string(rc)
        torch.cuda.synchronize()
        if n > 0:                                                 # the call took no device memory, from CUDA or from torch
            assert torch.cuda.mem_get_info(0)[0] >= free0 and torch.cuda.memory_allocated(0) == alloc0
        for k in KEYS:
            assert torch.equal(out[k], want[n][k]), (n, k)
    # the Python wrapper, with outputs the caller leaves out (NULL pointers are skipped)
    part = {k: torch.empty_like(want[1][k]) for k in ('nconf', 'near_idx')}
    core.cd_sweep_sym_static(recs[1], part)
    torch.cuda.synchronize()
    assert torch.equal(part['nconf'], want[1]['nconf']) and torch.equal(part['near_idx'], want[1]['near_idx'])


def test_static_scratch_sweep_refuses_what_does_not_fit():
    """More rows than the built-in array holds: a clear status, nothing launched, the outputs untouched."""
    from bluesky_gym_hallucination_b200 import _lib as L
    lib = L.load()
    rec = _records(1, 64, 3, 60.0)
    nconf = torch.full((1, 64), 77, dtype=torch.int32, device='cuda')
    E = 2 * L.CD_STATIC_MAX_ROWS // 4096
    p = L.CdParams(E, 4096, 0, 4096, 5 * 1852.0, 1000 * 0.3048, 300.0, 0)
    assert E * 4096 > L.CD_STATIC_MAX_ROWS
    rc = lib.bsg_cd_sym_static(C.byref(p), C.c_void_p(rec.data_ptr()), None, C.c_void_p(nconf.data_ptr()), None, None,
                               None, None, None, None, None)
    assert rc == L.ERR_TOOLARGE and b'BSG_CD_STATIC_MAX_ROWS' in lib.bsg_status_string(rc)
    with pytest.raises(L.BsgError, match='built-in scratch'):
        L.check(rc)
    torch.cuda.synchronize()
    assert bool((nconf == 77).all())
    # exactly the limit is accepted by the size check (a host pointer is then refused as an argument error, unlaunched)
    p = L.CdParams(L.CD_STATIC_MAX_ROWS // 4096, 4096, 0, 4096, 5 * 1852.0, 1000 * 0.3048, 300.0, 0)
    host = (C.c_float * 8)()
    assert lib.bsg_cd_sym_static(C.byref(p), C.cast(host, C.c_void_p), *([None] * 9)) == -1
