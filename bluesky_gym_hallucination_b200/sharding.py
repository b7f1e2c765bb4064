"""Multi-GPU partitioning of the path (one process per GPU, torch.distributed for the plumbing).

* env sharding (C2/C3/C4): envs are independent (no cross-env term anywhere in the path), every rank owns a
  contiguous env range and its own libbsgcore handle -- no data-path collective.
* giant airspace (C5): the N x N pair matrix is row-block partitioned; rank g owns ownship rows
  [g*N/G, (g+1)*N/G) of every env, packs the CD records of its own aircraft and ONE all-gather of the packed
  state (32 B per aircraft) gives every rank all intruder columns.  Row reductions are local: no reduce step.
* giant airspace, symmetric sweep: every unordered pair is evaluated once; the tiles of the upper triangle are dealt
  round-robin to the ranks (tile_of_rank), each rank accumulates partial row outputs for ALL rows and three
  all-reduces (reduce_sym_outputs: sum / max / min, all order-independent integer operations) combine them.
"""
import torch
import torch.distributed as dist


def env_shard(num_envs, rank, world):
    """Contiguous env range [e0, e1) of this rank."""
    return rank * num_envs // world, (rank + 1) * num_envs // world


def row_block(num_ac, rank, world):
    """Ownship rows [r0, r1) of this rank (every env)."""
    return rank * num_ac // world, (rank + 1) * num_ac // world


def gather_records(rec_local, num_ac, world, buf=None):
    """rec_local: [E, rows_of_this_rank, 8] packed CD records.  Returns [E, num_ac, 8] with the aircraft in
    global order on every rank.  Needs equal row blocks (num_ac % world == 0) for the single all-gather."""
    if world == 1:
        return rec_local
    E, rows, w = rec_local.shape
    if rows * world != num_ac:
        raise ValueError(f"row blocks must be equal for the single all-gather: {num_ac} aircraft over {world} ranks")
    if buf is None:
        buf = torch.empty((world, E, rows, w), dtype=rec_local.dtype, device=rec_local.device)
    try:
        dist.all_gather_into_tensor(buf, rec_local.contiguous())
    except (RuntimeError, NotImplementedError):          # backends without the flat variant
        parts = [torch.empty_like(rec_local) for _ in range(world)]
        dist.all_gather(parts, rec_local.contiguous())
        buf = torch.stack(parts, 0)
    return buf.permute(1, 0, 2, 3).reshape(E, num_ac, w).contiguous()     # rank-major -> aircraft-major


def reduce_sym_outputs(cnt, tcpamax, key, group=None):
    """Combine the partial outputs of the symmetric sweep across ranks, in place:
    cnt [2, E, N] int32 (nconf, nlos) SUM; tcpamax [E, N] float32 >= 0, MAX through its int32 bit pattern (order
    preserving for non-negative floats); key [E, N] int64 = (order-preserving int32 of tCPA) << 32 | intruder, MIN."""
    dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(tcpamax.view(torch.int32), op=dist.ReduceOp.MAX, group=group)
    dist.all_reduce(key, op=dist.ReduceOp.MIN, group=group)
    return cnt, tcpamax, key
