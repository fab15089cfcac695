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


def _norm_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from constrainedmbpo_b200.gae import normalize_advantages
    from constrainedmbpo_b200.sharding import shard_bounds
    adv, cadv, ret, cret, pop = _norm_data()
    lo, hi = shard_bounds(adv.shape[0], world, rank)
    a, c = torch.from_numpy(adv[lo:hi].copy()), torch.from_numpy(cadv[lo:hi].copy())
    stats = normalize_advantages(a, c, torch.from_numpy(pop[lo:hi].astype(np.uint8)), ret=torch.from_numpy(ret[lo:hi]),
                                 cret=torch.from_numpy(cret[lo:hi]), group=dist.group.WORLD)
    q.put((rank, a.numpy(), c.numpy(), stats.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def _norm_data():
    r = np.random.default_rng(5)
    B, T = 90, 7
    lens = r.integers(0, T, B)
    pop = np.arange(T)[None, :] < lens[:, None]
    f = lambda s: (r.standard_normal((B, T)) * s + 0.3).astype(np.float32)
    return f(2.0), f(0.5), f(1.0), f(1.0), pop


@pytest.mark.timeout(300)
def test_two_rank_advantage_normalisation_is_global():
    """ModelBuffer.get normalises with GLOBAL statistics (mpi_statistics_scalar, mpi_tools.py:71-93): two ranks holding
    row shards must produce what one process produces on the whole pool (oracle.normalize_advantages)."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_norm_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=240) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    adv, cadv, ret, cret, pop = _norm_data()
    ra, rc, am, asd, cm_ = orc.normalize_advantages(adv, cadv, pop)
    a = np.concatenate([g[1] for g in got]); c = np.concatenate([g[2] for g in got])
    assert np.allclose(a[pop], ra[pop], rtol=2e-6, atol=2e-6) and np.allclose(c[pop], rc[pop], rtol=2e-6, atol=2e-6)
    assert np.array_equal(a[~pop], adv[~pop])                     # unpopulated entries untouched
    for g in got:                                                 # every rank reports the same global statistics
        assert np.allclose(g[3][:3], [am, asd, cm_], rtol=1e-6) and g[3][5] == pop.sum()
        assert np.isclose(g[3][3], ret[pop].mean(), rtol=1e-6)


def test_sharded_rollout_refuses_max_samples():
    import constrainedmbpo_b200 as cm

    class E:
        device = "cpu"
    with pytest.raises(NotImplementedError, match="global row order"):
        cm.ModelRollout(E(), 10, rollout_mode="schedule", max_samples=100, sharded=True)
    with pytest.raises(NotImplementedError):
        cm.ModelRollout(E(), 10, rollout_mode="schedule", max_samples=100, row_offset=500)
    assert cm.ModelRollout(E(), 10, rollout_mode="schedule", max_samples=100).max_samples == 100
