#!/usr/bin/env python
"""bench.py -- BASELINE.json metric: env-steps/sec (and CPA pair-evals/sec) of the batched BlueSky-Gym core.

  python bench.py --gpus N --steps K --warmup W                (our arm; torchrun launches N ranks for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W   (the reference's numpy/BlueSky CPU path)

Headline line (config.workload "c2"): BASELINE.json configs[1] -- HorizontalCREnv-v0, 4096 envs x 20 aircraft per GPU,
state-based CD every substep + CPA obs, fp32 fast path, next-step auto-reset on the device.  A "step" = one fused
bsg_step launch (auto-reset inside) over the whole batch = 4096 env-steps.  For N > 1 the envs shard across ranks with
no data-path collective (weak scaling).
Without --workload the same JSON line also carries the two SHARDED configs of BASELINE.json, measured in the same run at
the same N: key "c3" (SectorCREnv 65 536 envs x 50 aircraft with MVP, the 65 536 envs split over the N ranks: strong
scaling) and key "c5" (giant airspace 16 x 10 000, pair matrix split over the N ranks through NVLink peer memory); at
N = 1 also key "v" (VerticalCREnv batch) and the CPU rows of the measurement table (cpu_baseline, "c1", numpy detect).
--workload X times one config alone (c2 | c3 | c3w | v | merge | merge_cd | c4 | c5).
One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_PAIR = 74.0          # algorithmic FP32 flop per ordered CD pair
EXEC_FLOP_PER_PAIR = 42.0     # executed by the canonical fp32 pair form per UNORDERED pair (csrc/bsg_pair.cuh)
WORKLOADS = {
    # name: (kind, envs per GPU, aircraft, asas, reso, description)
    'c2': (0, 4096, 20, 1, 0, "HorizontalCREnv-v0 4096 envs x 20 aircraft, state-based CD + CPA obs"),
    'c3': (2, 65536, 50, 1, 1, "SectorCREnv-v0 65536 envs x 50 aircraft, CD + MVP (intruders) + reward"),
    'v': (1, 4096, 6, 0, 0, "VerticalCREnv-v0 4096 envs x 6 aircraft, 30 substeps of 1 s, ASAS off as in the reference"),
    # SURVEY 8(f) rows, the EXT instantiations of the step kernel (not the driver's line):
    'merge': (4, 4096, 20, 0, 0, "MergeEnv-v0 4096 envs x 20 aircraft, LNAV route following, ASAS off as in the reference"),
    'merge_cd': (4, 4096, 20, 1, 0, "MergeEnv-v0 4096 envs x 20 aircraft, LNAV route following + state-based CD"),
    'c3w': (2, 65536, 50, 1, 1, "SectorCREnv-v0 65536 envs x 50 aircraft, CD + MVP, 3-point wind field + wind obs + obs noise"),
}
SUBSTEPS = {0: 10, 1: 30, 2: 5, 4: 10}


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits',
                                          '-i', str(self.idx), '-lms', '20'], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            p = [x.strip() for x in r.split(',')]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2]))
            except ValueError:
                continue
            for n, v in zip(names, p[5:9]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        if not sm:
            why = "nvidia-smi unavailable" if self.proc is None else "no sample inside the region"
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [why], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def ncu_traffic(workload):
    """DRAM bytes of one launch of the dominant kernel from the committed ncu capture (profiles/ncu_traffic.json)."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')) as f:
            return json.load(f).get(workload, {}).get('bytes')
    except (OSError, ValueError):
        return None


def cpu_baselines(n_ac, asas, budget_s=12.0):
    """Rank 0, N = 1 only: the oracle timed on the host cores on a bounded sample of the same workload."""
    import numpy as np
    from oracle import oracle as orc, envs_np
    from bluesky_gym_hallucination_b200 import scenario
    cores = len(os.sched_getaffinity(0))
    # (a) reference-style path: numpy BlueSky stand-in + Python env loops, one simulator per process
    per_proc = 24
    v_np, total_np, wall_np = envs_np.timed_sample(n_ac - 1, bool(asas), per_proc, cores)
    # (b) plain-C fp64 port, OpenMP over envs
    E = 256
    ac, reso, nac, ev = scenario.horizontal(E, n_ac, seed=99)
    cfg = orc.make_cfg(orc.KIND_H, n_ac, asas_on=asas)
    act = np.random.default_rng(0).uniform(-1, 1, (E, 1))
    orc.step(cfg, ac, reso, nac, ev, act, nthreads=cores)
    t0 = time.perf_counter(); n = 0
    while time.perf_counter() - t0 < budget_s / 3:
        orc.step(cfg, ac, reso, nac, ev, act, nthreads=cores); n += E
    v_c = n / (time.perf_counter() - t0)
    return ({"value": v_np, "unit": "env-steps/s", "cores": cores, "kind": "port",
             "sample": f"{cores} processes x {per_proc} env-steps of HorizontalCREnv with {n_ac} aircraft, "
                       f"ASAS {'ON' if asas else 'OFF'}; numpy BlueSky stand-in (oracle/refshim) + Python env port "
                       f"(oracle/envs_np.py), the reference's SubprocVecEnv style"},
            {"value": v_c, "unit": "env-steps/s", "cores": cores, "kind": "port",
             "sample": f"plain-C fp64 port (oracle/c), OpenMP over {E} envs x {n_ac} aircraft, {n} env-steps"})


def env_config(workload, E, N, kind, asas, reso, desc, world, strong):
    """config dict of an env-batch workload -- the SAME keys and values on our arm and on the reference arm."""
    cfg = {"workload": workload, "description": desc, "envs_per_gpu": E, "aircraft": N, "substeps": SUBSTEPS[kind],
           "asas": bool(asas), "mvp": bool(reso), "autoreset": "next-step",
           "l2": "256 MB buffer zeroed between timed iterations (L2 flush)",
           "sharding": ("envs split over the ranks (strong), no collective" if strong else "envs per rank fixed (weak), no collective")}
    return cfg


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path = numpy BlueSky-style simulator + per-step
    Python env loops, one simulator singleton per process, all host cores (the reference's own scaling method,
    scripts/multi_processing_example.py:51-53).  /root/reference is Python and cannot travel to the GPU box, so this
    times the oracle port of it (oracle/refshim + oracle/envs_np.py, replay-checked against the unmodified classes).
    A bench step = every worker process advances its env by `per_proc` env-steps (a bounded sample of the 4096-env
    batch); exactly --steps steps are timed after --warmup."""
    rank = env_int('RANK', 0)
    if rank != 0:
        return
    from oracle import envs_np
    wl = args.workload or 'c2'
    kind, E, N, asas, reso, desc = WORKLOADS[wl]
    cores = len(os.sched_getaffinity(0))
    K, W = args.steps, args.warmup
    # the reference's own classes when tools/make_oracle_ref.sh has shipped them (oracle/_ref), else the port of them
    use_ref = kind == 0 and envs_np.ref_available() and not os.environ.get('BSG_REF_PORT')      # BSG_REF_PORT=1: time the port
    pool = envs_np.WorkerPool(cores, N - 1, bool(asas), use_ref=use_ref)
    t1 = pool.step(1)                                    # one env-step per process: sizes the sample
    per_proc = max(1, min(32, int(90.0 / max(K + W, 1) / max(t1, 1e-4))))      # whole run ~1.5 min
    for _ in range(W):
        pool.step(per_proc)
    busy = 0.0
    for _ in range(K):
        busy += pool.step(per_proc)
    pool.close()
    tot = K * cores * per_proc
    value = tot / busy
    pairs = value * SUBSTEPS[kind] * N * (N - 1) * (1 if asas else 0)
    line = {"impl": "reference", "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
            "steps": K, "warmup": W, "ms_per_step": 1e3 * busy / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": env_config(wl, E, N, kind, asas, reso, desc, 1, False),
            # the config above names the WORKLOAD (same dict as our arm's); a bench step of this arm is a bounded sample of it:
            "envs_per_bench_step": cores * per_proc, "env_steps_per_process_per_bench_step": per_proc,
            "pair_evals_per_sec": pairs,
            "classes": ("unmodified reference env class (oracle/_ref/bluesky_gym HorizontalCREnv, NUM_INTRUDERS set on the module) on "
                        "the numpy BlueSky stand-in" if use_ref else "port of the reference env class (oracle/envs_np.py)"),
            "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": cores, "kind": "port",
                             "sample": f"{cores} persistent processes x {per_proc} env-steps per bench step "
                                       f"({cores * per_proc} of the batch's {E} envs), numpy BlueSky stand-in (oracle/refshim: "
                                       f"Tier B is a port) + " + ("the UNMODIFIED reference env class (oracle/_ref)" if use_ref
                                                                  else "Python env port (oracle/envs_np.py)")},
            "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


class Ctx:
    """ranks / device / process group of this bench process (one group for every workload of the run)."""
    def __init__(self):
        import torch
        import torch.distributed as dist
        self.rank, self.world, self.local = env_int('RANK', 0), env_int('WORLD_SIZE', 1), env_int('LOCAL_RANK', 0)
        # one slice of the host cores per rank (linear in the local rank, which follows the GPUs' NUMA order on an HGX
        # board): the ranks' host loops, their CUDA / NCCL helper threads and the pinned blocks they touch first stay on
        # their own cores instead of migrating over each other (the e2e leg is a host loop per rank)
        cores = sorted(os.sched_getaffinity(0))
        self.cores = len(cores)
        if self.world > 1 and len(cores) >= 2 * self.world and not os.environ.get('BSG_NO_PIN'):
            per = len(cores) // self.world
            mine = cores[self.local * per:(self.local + 1) * per]
            try:
                os.sched_setaffinity(0, mine)
                self.cores = len(mine)
            except OSError:
                pass
        torch.cuda.set_device(self.local)
        self.dev = torch.device('cuda', self.local)
        if self.world > 1:
            dist.init_process_group('nccl', device_id=self.dev)
        self.flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=self.dev)       # > 126 MB L2
        self.peak32 = None

    def peak(self):
        if self.peak32 is None:
            from bluesky_gym_hallucination_b200 import core as bc
            self.peak32, _ = bc.peak_fp32(self.local, 4096)
        return self.peak32

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()

    def max_over_ranks(self, x):
        import torch
        import torch.distributed as dist
        t = torch.tensor([x], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x):
        import torch
        import torch.distributed as dist
        t = torch.tensor([x], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


def run_envs(ctx, workload, K, W, strong=False, envs=0, cpu=False, precision='fp32'):
    """One env-batch workload: K fused step launches timed on the device (max over ranks), then the same metric end to
    end through host buffers.  strong: the workload's envs are split over the ranks (total fixed)."""
    import numpy as np
    import torch
    from bluesky_gym_hallucination_b200 import core as bc
    rank, world, local, dev = ctx.rank, ctx.world, ctx.local, ctx.dev
    kind, E, N, asas, reso, desc = WORKLOADS[workload]
    if envs:
        E = envs
    if strong:
        E = E // world
    W = max(W, 3)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()                      # running from before the warm-up: short timed regions still get samples
    c = bc.BsgCore(kind, E, N, precision=precision, device=local, asas_on=bool(asas), reso_on=bool(reso),
                   reso_off_own=True, max_episode_steps={0: 300, 1: 300, 2: 200, 4: 50}[kind], cd_outputs=True,
                   fixed_nac=True)    # every aircraft slot live ("N fixed, mask all-active")
    if workload == 'c3w':             # WindFieldWrapper(augment_obs=True) + NoisyObservationWrapper(0.1) on device
        c.set_wind([51.5, 52.6, 52.0], [3.5, 4.2, 5.5], [10.0, -4.0, 6.0], [-8.0, 12.0, 3.0], augment_obs=True)
        c.set_obs_noise(0.1, seed=5 + rank)
    c.reset(seed=1234 + rank)
    c.set_autoreset(True, seed=99 + rank)
    gen = torch.Generator(device=dev); gen.manual_seed(7 + rank)
    acts = [(torch.rand((E, c.act_dim), device=dev, generator=gen) * 2 - 1).to(c.rdtype) for _ in range(16)]
    flush = ctx.flush
    peak32 = ctx.peak()
    for w in range(W):
        c.step(acts[w % 16])
    torch.cuda.synchronize()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(K)]
    ctx.barrier()
    torch.cuda.synchronize()
    l0 = c.launch_count()
    t_wall0 = time.perf_counter()
    for k in range(K):
        flush.zero_()                                   # L2 flush between timed iterations (not timed)
        ev[k][0].record()
        c.step(acts[k % 16])                            # ONE launch: step kernel with the fused next-step auto-reset
        ev[k][1].record()
    torch.cuda.synchronize()
    ctx.barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = c.launch_count() - l0
    dev_ms = sum(ev[k][0].elapsed_time(ev[k][1]) for k in range(K))
    # the dominant kernel: with the next-step auto-reset fused, a step IS one step_kernel launch, so its average
    # launch duration is the event time of the timed region itself (steady-state traffic, resets included)
    kernel_ms = dev_ms / K
    dev_ms_max = ctx.max_over_ranks(dev_ms)
    # envs that start an episode in a step draw a scenario instead of integrating: they run no pair sweep and earn no
    # pair credit.  Their steady-state share, counted over an untimed replay of the same loop:
    fresh, nrep = 0, 64
    for k in range(nrep):
        c.step(acts[k % 16])
        fresh += int((c.terminated | c.truncated).sum().item())
    fresh_frac = fresh / float(nrep * E)
    # ---------------------------------------------------------------- end-to-end through host buffers
    Ke = max(200, min(K, 1000))                          # a floor of timed iterations whatever --steps says
    a_host = [a.cpu().numpy().copy() for a in acts]
    sink = 0.0
    for k in range(3):
        c.step_host_views(a_host[k])
    ctx.barrier()
    t0 = time.perf_counter()
    for k in range(Ke):
        # host actions -> pinned H2D -> ONE launch -> D2H of obs/reward/flags/info into pinned host memory, which the
        # caller reads in place (views); the reward read below is the step's device->host result read
        obs_v, rew_v, term_v, trunc_v, info_v = c.step_host_views(a_host[k % 16])
        sink += float(rew_v[0])
    t_e2e = time.perf_counter() - t0
    e2e_value = world * E * Ke / ctx.max_over_ranks(t_e2e)
    # the same loop with TWO steps in flight (bsg_step_host_begin / _wait: SB3's step_async / step_wait over two pinned
    # blocks): the D2H of step k overlaps the launch of step k + 1.  A reported extra, not the headline: a policy that needs
    # obs k before it can choose action k + 1 runs the synchronous loop above (or two env batches in this one).
    ctx.barrier()
    t0 = time.perf_counter()
    c.step_host_begin(a_host[0])
    for k in range(1, Ke + 1):
        if k < Ke:
            c.step_host_begin(a_host[k % 16])
        obs_v, rew_v, term_v, trunc_v, info_v = c.step_host_wait()
        sink += float(rew_v[0])
    t_pipe = time.perf_counter() - t0
    e2e_pipe = world * E * Ke / ctx.max_over_ranks(t_pipe)
    clocks = sampler.stop() if rank == 0 else None          # sampled across warm-up, the device-timed and the e2e regions
    nac = c.live_aircraft().astype(np.float64)                  # counted from the device state, not assumed
    assert nac.min() == N and nac.max() == N, (nac.min(), nac.max())
    obs_dim, act_dim = c.obs_dim, c.act_dim
    c.close()
    if rank != 0:
        return None
    sub = SUBSTEPS[kind]
    pairs_per_launch = float(sub * np.sum(nac * (nac - 1))) * (1.0 - fresh_frac) if asas else 0.0   # ordered CD entries of one launch
    value = world * E * K / (dev_ms_max * 1e-3)
    achieved = FLOP_PER_PAIR * pairs_per_launch / (kernel_ms * 1e-3) / 1e12
    bytes_per_envstep = 128.0 * N + 4 * obs_dim + 4 * act_dim + 20
    cfg = env_config(workload, E, N, kind, asas, reso, desc, world, strong)       # identical on the reference arm
    line = {
        "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dev_ms_max / K, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": cfg,
        "pair_evals_per_sec": world * pairs_per_launch * K / (dev_ms_max * 1e-3),
        "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak32, "unit": "TFLOP/s", "frac": achieved / peak32,
                     "traffic": ncu_traffic(workload), "kernel": "bsg::step_kernel<float>" + (", EXT" if kind == 4 or workload == 'c3w' else ""), "kernel_ms": kernel_ms,
                     "flop_per_pair": FLOP_PER_PAIR, "pairs_per_launch": pairs_per_launch,
                     "pairs_note": "ordered CD entries (every unordered pair is evaluated once and credited for both aircraft), "
                                   "every substep of the step; envs that start an episode in the step are not credited",
                     "peak_source": "bsg_peak_fp32 FMA chain measured in this run (burst); MEASURED_PEAKS.json holds no FP32 figure",
                     "executed": {"flop_per_unordered_pair": EXEC_FLOP_PER_PAIR, "frac": EXEC_FLOP_PER_PAIR * 0.5 * pairs_per_launch / (kernel_ms * 1e-3) / 1e12 / peak32,
                                  "note": "what the SMs execute for the sweep alone: every unordered pair once, 9 FMA + 15 add/mul + 2 MUFU + max3 (42 flop)"},
                     "hbm_achieved_gbs": bytes_per_envstep * E / (kernel_ms * 1e-3) / 1e9,
                     "hbm_peak_gbs": load_peaks().get("hbm_gbs")},
        "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": int(act_dim * E * 4),
                "d2h_bytes_per_step": int(E * (obs_dim + 5) * 4 + E * 8), "steps": Ke,
                "api": "bsg_step_host + bsg_host_views (host actions -> pinned block, read by the kernel over PCIe; one launch; D2H of obs/reward/info/flags into pinned host memory read in place; the launch + copy replayed as one CUDA graph)",
                "two_steps_in_flight": {"value": e2e_pipe, "unit": "env-steps/s",
                                        "api": "bsg_step_host_begin / bsg_step_host_wait, same copies per step; reported extra (the headline e2e is the synchronous loop)"}},
        "gpu_launches": int(launches), "clocks": clocks, "wall_s_timed_region": t_wall,
        "traffic_state": {"live_aircraft_per_env": {"min": int(nac.min()), "mean": float(nac.mean()), "max": int(nac.max())},
                          "fresh_episode_env_share": fresh_frac},
    }
    if not asas:     # no pair sweep in the launch (the reference leaves ASAS off here): state / obs traffic bounds it
        hb, hp = line["roofline"]["hbm_achieved_gbs"], line["roofline"]["hbm_peak_gbs"] or 6552.3
        line["roofline"].update({"bound": "hbm", "achieved": hb, "peak": hp, "unit": "GB/s", "frac": hb / hp})
    if cpu and world == 1 and kind != 4:
        np_b, c_b = cpu_baselines(N, asas)
        line["cpu_baseline"] = np_b
        line["cpu_baseline_c_openmp"] = c_b
    return line


def cpu_table():
    """The CPU rows of BASELINE.md's measurement table that no GPU workload covers (rank 0, N = 1)."""
    from oracle import envs_np
    v, eps, dt = envs_np.vertical_1000_steps(seed=0, steps=1000)
    det = envs_np.detect_rates(sizes=(20, 50, 500), budget_s=2.0)
    return ({"value": v, "unit": "env-steps/s", "cores": 1, "kind": "port", "episodes": eps, "seconds": dt,
             "sample": "BASELINE.json configs[0]: ONE VerticalCREnv-v0, random actions, 1000 env steps (30 sim steps each) on "
                       "one core; numpy BlueSky stand-in + Python env port (oracle/envs_np.VerticalNp, replay-checked "
                       "against the unmodified reference class)"},
            {"unit": "ordered pair-evals/s", "cores": 1, "kind": "port", "by_aircraft": {str(k): x for k, x in det.items()},
             "sample": "numpy state-based detection alone (oracle/refshim bluesky.traffic.asas.detect = BlueSky "
                       "StateBased.detect restated), all N x N pairs, 2 s per size"})


def run_giant(ctx, K, W, mode):
    """c5: giant airspace, 16 envs x 10 000 aircraft, one all-pairs CD sweep per step.  N > 1: the pair matrix is
    sharded over the ranks (modes below)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from bluesky_gym_hallucination_b200 import core as bc, scenario, sharding
    rank, world, local, dev = ctx.rank, ctx.world, ctx.local, ctx.dev
    E, N = 16, 10000
    W = max(W, 3)
    lat, lon, alt, trk, gs, vs = scenario.raw_traffic(E, N, seed=3, box_nm=1000.0)       # same traffic on every rank
    r0, r1 = sharding.row_block(N, rank, world)
    t64 = lambda a: torch.as_tensor(np.ascontiguousarray(a[:, r0:r1]), dtype=torch.float64, device=dev)
    t32 = lambda a: torch.as_tensor(np.ascontiguousarray(a[:, r0:r1]), dtype=torch.float32, device=dev)
    mine = (t64(lat), t64(lon), t32(alt), t32(trk), t32(gs), t32(vs))
    rec_all = torch.empty((world, E, r1 - r0, 8), dtype=torch.float32, device=dev) if world > 1 else None
    out = None
    flush = ctx.flush
    peak32 = ctx.peak()

    # sweep modes: "peer" (default) = symmetric sweep over NVLink peer memory, loads / atomics to the owners inside the
    # kernel, two cross-GPU barriers per step; "allreduce" = symmetric sweep + NCCL all-gather + 3 all-reduces;
    # "ordered" = ordered row block + NCCL all-gather (the baseline design)
    sym = mode != 'ordered'
    grp = dist.group.WORLD if world > 1 else None
    peer = bc.PeerSweep(E, N, rank, world, grp, dev) if mode == 'peer' else None

    def one():
        nonlocal out
        if mode == 'peer':
            out = peer.step(*mine, 52.0, 4.0)
            return
        rec = bc.cd_pack(*mine, 52.0, 4.0)                                   # [E, rows, 8]
        full = sharding.gather_records(rec, N, world, buf=rec_all)
        if sym:      # every unordered pair once; tiles dealt round-robin to the ranks, outputs all-reduced
            out = bc.cd_sweep_sym(full, out=out, group=grp, rank=rank, world=world)
        else:        # ordered row block per rank, row outputs local
            out = bc.cd_sweep(full, r0, r1, out=out)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(W):
        one()
    torch.cuda.synchronize()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(K)]
    ctx.barrier()
    torch.cuda.synchronize()
    for k in range(K):
        flush.zero_()
        ev[k][0].record(); one(); ev[k][1].record()
    torch.cuda.synchronize()
    ctx.barrier()
    dev_ms = sum(ev[k][0].elapsed_time(ev[k][1]) for k in range(K))
    clocks = sampler.stop() if rank == 0 else None
    ms = ctx.max_over_ranks(dev_ms) / K
    if rank != 0:
        return None
    pairs = float(E) * N * (N - 1)
    ach = FLOP_PER_PAIR * pairs / world / (ms * 1e-3) / 1e12
    # what the SMs execute: the symmetric sweep evaluates every UNORDERED pair once (42 flop); the ordered one every entry
    exe = (42.0 * pairs / 2 if sym else FLOP_PER_PAIR * pairs) / world / (ms * 1e-3) / 1e12
    return {
        "metric": "pair-evals/sec", "value": pairs / (ms * 1e-3), "unit": "pair-evals/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "c5", "description": "giant airspace 16 envs x 10000 aircraft all-pairs CPA, pair matrix sharded over the ranks",
                   "sweep": "symmetric: every unordered pair evaluated once, ordered entries reported" if sym else "ordered row block",
                   "mode": mode,
                   "collective": ("none" if world == 1 else
                                  {"peer": "none: records read from / conflict updates flushed to the owning GPU over NVLink peer memory inside the sweep kernel; 2 symmetric-memory barriers per step",
                                   "allreduce": "one NCCL all-gather of packed state + 3 all-reduces of the row outputs (sum/max/min) per step",
                                   "ordered": "one NCCL all-gather of packed state per step, no reduce"}[mode]),
                   "l2": "256 MB flush between timed iterations"},
        "roofline": {"bound": "fp32", "achieved": ach, "peak": peak32, "unit": "TFLOP/s", "frac": ach / peak32,
                     "frac_is": "SURVEY 8(d) convention: 74 flop credited per ORDERED entry; the symmetric sweep evaluates each unordered "
                                "pair once, so this figure can exceed 1 -- executed.frac is the fraction of the FP32 peak",
                     "executed": {"flop_per_unordered_pair": 42.0 if sym else FLOP_PER_PAIR, "tflops": exe, "frac": exe / peak32},
                     "traffic": ncu_traffic('c5'), "kernel": ("cd_sym_sweep_kernel" if sym else "cd_sweep_kernel") + " (+ pack, collectives, init/finalize inside the timed step)"},
        # kernels per step: peer = pack+init | sweep | finalize (+ 2 symmetric-memory barrier kernels at N > 1);
        # allreduce = pack | init | sweep | finalize; ordered = pack | sweep
        "gpu_launches": (3 if mode == "peer" else 4 if sym else 2) * K + (2 * K if (mode == "peer" and world > 1) else 0),
        # what crosses NVLink per step, algorithmic: every rank reads the packed records (32 B per aircraft) it does not own
        "nvlink_bytes_per_step": {"per_rank": 32.0 * E * N * (world - 1) / world, "all_ranks": 32.0 * E * N * (world - 1),
                                  "note": "algorithmic: remote 32-byte records read once per sweep (tile reuse through L2 / shared "
                                          "memory), plus the rare conflict updates flushed to their owners"},
        "clocks": clocks}


def run_ndmap(ctx, K, W, envs=0):
    """c4: ND-map, 1024 envs x 500 aircraft x 30 look-ahead steps, 64x64x8 grid (HBM-write bound)."""
    import numpy as np
    import torch
    from bluesky_gym_hallucination_b200 import core as bc
    local, dev, world = ctx.local, ctx.dev, ctx.world
    E, n, W = envs or 1024 // world, 500, max(W, 3)           # the config's 1024 envs are split over the ranks (strong)
    rng = np.random.default_rng(4 + ctx.rank)
    rec = np.zeros((E, 6, n), np.float32)
    box = 200 * 1852.0
    rec[:, 0] = rng.uniform(0, box, (E, n)); rec[:, 1] = rng.uniform(0, box, (E, n))
    rec[:, 2] = rng.uniform(3048, 12192, (E, n))
    hdg = rng.uniform(0, 2 * np.pi, (E, n)); gs = rng.uniform(120, 250, (E, n))
    rec[:, 3] = gs * np.sin(hdg); rec[:, 4] = gs * np.cos(hdg); rec[:, 5] = rng.choice([-10.0, 0.0, 0.0, 10.0], (E, n))
    d_rec = torch.as_tensor(rec, device=dev)
    kw = dict(nx=64, ny=64, nz=8, nt=30, x0=0.0, y0=0.0, z0=3048.0, cx=box / 64, cy=box / 64, cz=9144.0 / 8, dt=10.0)
    sampler = ClockSampler(local)
    if ctx.rank == 0:
        sampler.start()
    out = bc.ndmap(d_rec, **kw)
    for _ in range(W):
        bc.ndmap(d_rec, out=out, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        bc.ndmap(d_rec, out=out, **kw)                   # 1.13 GB written per 1024-env build >> L2
    e1.record(); torch.cuda.synchronize()
    ctx.barrier()
    ms = ctx.max_over_ranks(e0.elapsed_time(e1)) / K
    clocks = sampler.stop() if ctx.rank == 0 else None
    if ctx.rank != 0:
        return None
    # the kernel only WRITES: the ceiling of a write-only stream on this GPU, measured here with a device memset of 1 GiB
    # (best of 5), beside the COPY figure of MEASURED_PEAKS.json (read + write bytes) the contract's `peak` is
    wbuf = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    wbest = 1e9
    for _ in range(6):
        w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0.record(); wbuf.zero_(); w1.record(); torch.cuda.synchronize()
        wbest = min(wbest, w0.elapsed_time(w1))
    write_peak = (1 << 30) / (wbest * 1e-3) / 1e9
    del wbuf
    cells = E * 30 * 64 * 64 * 8
    # algorithmic bytes: occ 1 B per cell + conflict BITMASK 1/8 B per cell + the aircraft records + ncell
    nd_bytes = 1.125 * cells + 24.0 * n * E + 4 * 30 * E
    gbs = nd_bytes / (ms * 1e-3) / 1e9
    peak = load_peaks().get("hbm_gbs", 6650.0)
    return {
        "metric": "ndmap-samples/sec", "value": world * E * n * 30 / (ms * 1e-3), "unit": "trajectory samples/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak" if (envs or world == 1) else "strong",
        "vs_baseline": None,
        "dtype": "f32+u8", "data": "synthetic",
        "config": {"workload": "c4", "description": f"ND-map {world * E} envs x 500 aircraft x 30 steps, grid 64x64x8", "envs_per_gpu": E,
                   "l2": "1.13 GB written per build (>> 126 MB L2)"},
        "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": ncu_traffic('c4'),
                     "kernel": "ndmap_kernel", "bytes_per_build": nd_bytes,
                     "bytes_per_cell": 1.125, "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy: read + write bytes)",
                     "write_only_peak_gbs": write_peak, "frac_of_write_only_peak": gbs / write_peak,
                     "write_only_note": "the kernel is a pure write stream; a device memset of 1 GiB measured in this run is the "
                                        "ceiling of such a stream (about 0.59 of the copy figure on B200)"},
        "gpu_launches": K, "clocks": clocks}


def compact(line):
    """the sub-record of a secondary workload inside the headline line"""
    if line is None:
        return None
    keep = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "scaling", "config", "pair_evals_per_sec",
            "roofline", "e2e", "gpu_launches", "nvlink_bytes_per_step")
    return {k: line[k] for k in keep if k in line}


def summary(line):
    """{config: {value, ms, frac[, xfrac = executed fraction][, e2e]}} of the headline line and its sub-records, 4 digits"""
    r4 = lambda x: None if x is None else float(f"{x:.4g}")
    out = {}
    for key, rec in (("c2", line), ("c3", line.get("c3")), ("c4", line.get("c4")), ("c5", line.get("c5")), ("v", line.get("v"))):
        if not rec:
            continue
        roof = rec.get("roofline", {})
        d = {"value": r4(rec["value"]), "ms": r4(rec["ms_per_step"]), "frac": r4(roof.get("frac"))}
        if "executed" in roof:
            d["xfrac"] = r4(roof["executed"].get("frac"))
        if "e2e" in rec:
            d["e2e"] = r4(rec["e2e"]["value"])
        if "nccl_ordered" in rec:
            d["nccl_ordered"] = r4(rec["nccl_ordered"]["value"])
        out[key] = d
    if "c2_fp64_parity_mode" in line:
        out["c2_f64"] = {"value": r4(line["c2_fp64_parity_mode"]["value"]), "ms": r4(line["c2_fp64_parity_mode"]["ms_per_step"])}
    return out


def load_peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {"hbm_gbs": 6650.0, "source": "fallback"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--c5-mode', default='peer', choices=['peer', 'allreduce', 'ordered'])
    ap.add_argument('--workload', default=None, choices=sorted(WORKLOADS) + ['c4', 'c5'])
    ap.add_argument('--envs', type=int, default=0, help='override envs per GPU')
    ap.add_argument('--no-cpu', action='store_true', help='skip the CPU legs (cpu_baseline, c1, numpy detect)')
    ap.add_argument('--no-extras', action='store_true', help='headline workload only (no c3 / c5 / v sub-records)')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
        return
    ctx = Ctx()
    K, W = args.steps, args.warmup
    if args.workload == 'c5':
        line = run_giant(ctx, K, W, args.c5_mode)
    elif args.workload == 'c4':
        line = run_ndmap(ctx, K, W, args.envs)
    elif args.workload:
        line = run_envs(ctx, args.workload, K, W, strong=(args.workload in ('c3', 'c3w') and ctx.world > 1 and not args.envs),
                        envs=args.envs, cpu=not args.no_cpu)
    else:
        line = run_envs(ctx, 'c2', K, W, envs=args.envs, cpu=not args.no_cpu)
        if not args.no_extras:
            # the sharded configs of BASELINE.json at the same N, bounded so that the default run stays within minutes
            c3 = run_envs(ctx, 'c3', min(K, 200), W, strong=True)
            c5 = run_giant(ctx, min(K, 100), W, args.c5_mode)
            c4 = run_ndmap(ctx, min(K, 100), W)
            # what the fp64 parity mode (the reference's own arithmetic type) costs on the headline config
            f64 = run_envs(ctx, 'c2', min(K, 100), W, precision='fp64') if ctx.world == 1 else None
            # the north_star's literal C5 design (ordered row block per rank + ncclAllGather of the packed state) beside
            # the fused peer sweep: what reading the records over NVLink inside the kernel buys
            c5o = run_giant(ctx, min(K, 50), W, 'ordered') if (ctx.world > 1 and args.c5_mode == 'peer') else None
            v = run_envs(ctx, 'v', min(K, 500), W) if ctx.world == 1 else None
            # the same kernel on a launch four times C2's size: what the 4096-env batch leaves on the table is the fixed
            # latency of one launch (about 22 us: launch, cold prologue, the two integration chains, epilogue), not the sweep
            big = run_envs(ctx, 'c2', min(K, 200), W, envs=16384) if ctx.world == 1 else None
            if ctx.rank == 0:
                line["c3"] = compact(c3); line["c5"] = compact(c5); line["c4"] = compact(c4)
                if c5o is not None:
                    line["c5"]["nccl_ordered"] = {"value": c5o["value"], "unit": c5o["unit"], "ms_per_step": c5o["ms_per_step"],
                                                  "collective": c5o["config"]["collective"]}
                if f64 is not None:
                    line["c2_fp64_parity_mode"] = {"value": f64["value"], "unit": f64["unit"], "ms_per_step": f64["ms_per_step"],
                                                   "dtype": "f64", "e2e": f64["e2e"]["value"], "steps": f64["steps"],
                                                   "vs_fp32": f64["value"] / line["value"]}
                if v is not None:
                    line["v"] = compact(v)
                if big is not None:
                    line["c2_16384_envs"] = compact(big)
                    line["roofline"]["frac_at_16384_envs"] = big["roofline"]["frac"]
                if ctx.world == 1 and not args.no_cpu:
                    line["c1"], line["numpy_detect"] = cpu_table()
    if ctx.rank == 0 and line is not None:
        if not args.workload and not args.no_extras:
            # numbers only, last key of the line: the driver's record keeps the tail of stdout, so every BASELINE config of
            # this N is in it whatever the length of the sub-records above
            line["summary"] = summary(line)
        print(json.dumps(line), flush=True)
    ctx.close()


if __name__ == '__main__':
    main()
