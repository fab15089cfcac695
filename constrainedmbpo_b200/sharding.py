"""Row sharding of the rollout batch across the GPUs of one box and the all-gather of generated
transitions into the replicated model pool.

The reference has no multi-device path (mpi4py at world size 1, softlearning/policies/safe_utils/
mpi_tools.py:7-93), so the parity statement is: the rank-order concatenation of the per-rank
transition lists equals the single-device result (same kernels, same row -> Philox counter map via
``row_offset``).  Everything here is device-agnostic torch plumbing (NCCL on CUDA tensors, gloo on
CPU tensors in the tests); no arithmetic of the path happens in this file.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

RECORD_FIELDS = ("obs", "act", "next_obs", "rew", "cost", "dyn_error", "term")


def shard_bounds(n_rows: int, world: int, rank: int):
    """Contiguous row block of ``rank``: the first ``n_rows % world`` ranks get one extra row."""
    base, rem = divmod(int(n_rows), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pack_records(pool: dict, row_offset: int = 0):
    """Populated transitions of a ``[B,T,.]`` pool (modelbuffer.py:45-114, filtered by
    ``populated_mask`` like ``ModelBuffer.get``, modelbuffer.py:348-405) as one fp32 matrix
    ``[n, 2*obs+act+6]``: obs | act | next_obs | rew | cost | dyn_error | term | path | t, in
    [path, t] order; ``path`` is the GLOBAL start-state index (exact in fp32 below 2**24 rows per
    box; the int64 twin ``coords`` carries it losslessly)."""
    mask = pool["populated"].bool()
    path, t = mask.nonzero(as_tuple=True)
    cols = []
    for k in RECORD_FIELDS:
        v = pool[k][mask]
        cols.append(v.reshape(v.shape[0], -1).to(torch.float32))
    rec = torch.cat(cols, dim=1)
    coords = torch.stack([path + int(row_offset), t], dim=1).to(torch.int64)
    return rec, coords


def unpack_records(rec: torch.Tensor, obs_dim: int, act_dim: int):
    o = 0
    out = {}
    for k, wdt in (("obs", obs_dim), ("act", act_dim), ("next_obs", obs_dim), ("rew", 1), ("cost", 1),
                   ("dyn_error", 1), ("term", 1)):
        v = rec[:, o:o + wdt]
        out[k] = v if wdt > 1 else v[:, 0]
        o += wdt
    out["term"] = out["term"] != 0
    return out


def _all_gather_rows(x: torch.Tensor, group=None):
    """Variable-length all-gather along dim 0 (counts exchanged first, payload padded to the max)."""
    world = dist.get_world_size(group)
    n = torch.tensor([x.shape[0]], dtype=torch.int64, device=x.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    m = max(counts) if counts else 0
    pad = torch.zeros((m,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    pad[: x.shape[0]] = x
    if x.is_cuda:
        full = torch.empty((world * m,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(full, pad, group=group)
        parts = [full[r * m: r * m + counts[r]] for r in range(world)]
    else:
        bufs = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(bufs, pad, group=group)
        parts = [bufs[r][: counts[r]] for r in range(world)]
    return torch.cat(parts, dim=0), counts


def gather_transitions(pool: dict, row_offset: int = 0, group=None):
    """All ranks' populated transitions, concatenated in rank order: ``(records, coords, counts)``."""
    rec, coords = pack_records(pool, row_offset)
    rec_all, counts = _all_gather_rows(rec, group)
    coords_all, _ = _all_gather_rows(coords, group)
    return rec_all, coords_all, counts


class PoolGatherer:
    """Per-rollout all-gather of the transitions a ``ModelRollout`` generated (one NCCL all-gather
    of the packed records + one of the coordinates)."""

    def __init__(self, rollout, world, group=None):
        self.ro, self.world, self.group = rollout, world, group
        self.last = None

    def launch(self):
        self.last = gather_transitions(self.ro.pool, self.ro.row_offset, self.group)

    def wait(self):
        return self.last
