"""torchrun --nproc-per-node G tools/check_peer_sweep.py : PeerSweep over G GPUs == ordered sweep, bitwise, per rank."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from bluesky_gym_hallucination_b200 import core as bc, scenario, sharding
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local); dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
ok = True
for E, N, box in ((3, 1024 * world // 2 + 8 * world, 200.0), (2, 250 * world, 60.0), (16, 10000, 1000.0)):
    lat, lon, alt, trk, gs, vs = scenario.raw_traffic(E, N, seed=5, box_nm=box)
    r0, r1 = sharding.row_block(N, rank, world)
    t64 = lambda a: torch.as_tensor(np.ascontiguousarray(a[:, r0:r1]), dtype=torch.float64, device=dev)
    t32 = lambda a: torch.as_tensor(np.ascontiguousarray(a[:, r0:r1]), dtype=torch.float32, device=dev)
    mine = (t64(lat), t64(lon), t32(alt), t32(trk), t32(gs), t32(vs))
    ps = bc.PeerSweep(E, N, rank, world, dist.group.WORLD, dev)
    # Repeated steps reuse the symmetric block and the traffic CHANGES between them (every aircraft moves along its
    # track).  Odd ranks delay their finalize by ~50 ms (test hook) while the even ranks run ahead and pack the next
    # step's records: with a single record region the late finalize would read records of the wrong step.
    ps.debug_finalize_delay = 100_000_000 if rank % 2 == 1 else 0
    steps, outs = [], []
    for it in range(4):
        moved = (mine[0] + it * 2e-3 * torch.cos(torch.deg2rad(mine[3].double())),
                 mine[1] + it * 3e-3 * torch.sin(torch.deg2rad(mine[3].double()))) + mine[2:]
        out = ps.step(*moved, 52.0, 4.0)
        steps.append(moved); outs.append({k: v.clone() for k, v in out.items()})
    torch.cuda.synchronize()
    for it, (moved, out) in enumerate(zip(steps, outs)):
        rec = bc.cd_pack(*moved, 52.0, 4.0)
        full = sharding.gather_records(rec, N, world)
        ref = bc.cd_sweep(full, r0, r1)
        torch.cuda.synchronize()
        for k in ('inconf', 'nconf', 'nlos', 'near_idx', 'near_tcpa', 'near_dcpa', 'near_tin', 'tcpamax'):
            same = torch.equal(ref[k][:, r0:r1], out[k])
            ok &= same
            if not same:
                print(f"rank {rank} E {E} N {N} step {it}: {k} differs in {(ref[k][:, r0:r1] != out[k]).sum().item()} entries", flush=True)
    print(f"rank {rank} E {E} N {N}: conflicts in own rows {int(out['nconf'].sum())}, match {ok}", flush=True)
    dist.barrier()
t = torch.tensor([int(ok)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("PEER SWEEP CHECK", "OK" if t.item() else "FAILED", flush=True)
dist.destroy_process_group()
