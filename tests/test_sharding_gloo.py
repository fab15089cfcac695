"""N>1 host logic on CPU: world_size-2 gloo run of the row sharding + transition all-gather.  Each rank
rolls out its row block with the ORACLE (standing in for the CUDA step, which needs a GPU) and the
gathered records must equal the single-process rollout of all rows (SURVEY §8e parity statement)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.helpers import make_inputs, make_params, orc


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _oracle_pool(rows_lo, rows_hi, B_total):
    """Oracle rollout of rows [lo, hi) with per-row deterministic elite draws keyed by GLOBAL row id."""
    p = make_params("antsafe", (32, 32), parity_scalers=False)
    obs, _ = make_inputs("antsafe", B_total)
    P = np.random.default_rng(4).standard_normal((29, 8)) / np.sqrt(29)
    pol = lambda o: np.tanh(o @ P)
    T = 6
    ids = {"alive": np.arange(rows_lo, rows_hi)}

    def step_fn(cur, a, step):
        gid = ids["alive"]
        eidx = np.asarray(p.elites)[(gid * 7 + step) % len(p.elites)]   # counter-based, shard-invariant
        out = orc.fake_env_step(p, "antsafe", cur, a, elite_idx=eidx, cares_about_cost=True)
        return out
    # track which global rows are alive through the same masks the sampler applies
    pool = orc.rollout(p, "antsafe", obs[rows_lo:rows_hi].astype(np.float64), pol, T, "schedule", step_fn=_Tracking(step_fn, ids))
    return {k: torch.from_numpy(np.asarray(v)) for k, v in pool.items()
            if k in ("obs", "act", "next_obs", "rew", "cost", "dyn_error", "term", "populated")}


class _Tracking:
    """Keeps the global ids of the alive rows in sync with the sampler's compaction (terminal mask)."""
    def __init__(self, fn, ids):
        self.fn, self.ids = fn, ids

    def __call__(self, cur, a, step):
        out = self.fn(cur, a, step)
        term = np.squeeze(out[2], -1)
        self.ids["alive"] = self.ids["alive"][~term]
        return out


def _worker(rank, world, port, B, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from constrainedmbpo_b200 import sharding
    lo, hi = sharding.shard_bounds(B, world, rank)
    pool = _oracle_pool(lo, hi, B)
    rec, coords, counts = sharding.gather_transitions(pool, row_offset=lo)
    if rank == 0:
        q.put((rec.numpy(), coords.numpy(), counts))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds_partition():
    from constrainedmbpo_b200.sharding import shard_bounds
    for n, w in ((10, 3), (8, 8), (5, 8), (1_000_000, 8), (0, 2)):
        b = [shard_bounds(n, w, r) for r in range(w)]
        assert b[0][0] == 0 and b[-1][1] == n
        assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
        sizes = [hi - lo for lo, hi in b]
        assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(300)
def test_two_rank_gather_equals_single_process():
    from constrainedmbpo_b200 import sharding
    B, world = 101, 2   # odd: ragged shards
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, B, q)) for r in range(world)]
    for p in procs:
        p.start()
    rec, coords, counts = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = _oracle_pool(0, B, B)
    ref_rec, ref_coords = sharding.pack_records(full, 0)
    assert sum(counts) == ref_rec.shape[0] and len(counts) == 2
    assert np.array_equal(coords, ref_coords.numpy())
    assert np.array_equal(rec, ref_rec.numpy())
    un = sharding.unpack_records(torch.from_numpy(rec), 29, 8)
    assert un["obs"].shape == (rec.shape[0], 29) and un["term"].dtype == torch.bool
