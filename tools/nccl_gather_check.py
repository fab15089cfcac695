#!/usr/bin/env python
"""N-GPU check of the transition all-gather (run under torchrun, one rank per GPU):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/nccl_gather_check.py
Each rank rolls out its row block (global row ids via row_offset); the gathered records must equal, bit for bit,
the rollout of ALL rows on one GPU (SURVEY §8e parity statement) -- for ragged ('uncertainty') and equal
('schedule') shard sizes, pipelined over several rollouts."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import constrainedmbpo_b200 as cm  # noqa: E402
from constrainedmbpo_b200 import sharding  # noqa: E402
from tests.helpers import build_cuda_env, make_inputs, make_params  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    env, hidden, B = "antsafe", (200, 200, 200, 200), 20_001
    p = make_params(env, hidden, parity_scalers=False, bias=False)
    obs_all, _ = make_inputs(env, B)
    ok = True
    for mode, T, lim in (("uncertainty", 10, 3e-5), ("schedule", 5, None), ("iv_gae", 5, None)):
        fe = cm.FakeEnv(cm.SimpleEnv(29, 8), env, num_networks=7, num_elites=5, hidden_dims=hidden,
                        cares_about_cost=True, device=str(dev), seed=3)
        if rank == 0:      # only rank 0 holds the "trained" state; the others receive it (weights replicated, §8e)
            fe._model.load_state([w.astype(np.float32) for w in p.weights], [b.astype(np.float32) for b in p.biases],
                                 scaler_in=(p.mu_in, p.var_in), scaler_out=(p.mu_out, p.var_out), elites=p.elites)
        sharding.broadcast_model_state(fe._model, src=0)
        lo, hi = sharding.shard_bounds(B, world, rank)
        obs = torch.from_numpy(obs_all[lo:hi]).to(dev)
        full, counts = _run(fe, obs, lo, T, mode, lim, world, 3)
        ref, _ = _run(fe, torch.from_numpy(obs_all).to(dev), 0, T, mode, lim, None, 1)
        same = full.shape == ref.shape and torch.equal(full.view(torch.int32), ref.view(torch.int32))
        ok = ok and same
        if rank == 0:
            print(f"{mode}: gathered {tuple(full.shape)} counts={counts} equal_to_single_gpu={same}", flush=True)
    # dense gather of the pool arenas themselves (sharding.DensePoolGatherer): equal shards, double-buffered rollouts
    Bd = 20_000 // world * world
    for mode, T, lim in (("schedule", 6, None), ("uncertainty", 8, 3e-5), ("iv_gae", 4, None)):
        fe = cm.FakeEnv(cm.SimpleEnv(29, 8), env, num_networks=7, num_elites=5, hidden_dims=hidden,
                        cares_about_cost=True, device=str(dev), seed=3)
        if rank == 0:
            fe._model.load_state([w.astype(np.float32) for w in p.weights], [b.astype(np.float32) for b in p.biases],
                                 scaler_in=(p.mu_in, p.var_in), scaler_out=(p.mu_out, p.var_out), elites=p.elites)
        sharding.broadcast_model_state(fe._model, src=0)
        same = _run_dense(fe, obs_all[:Bd], T, mode, lim, rank, world, dev)
        ok = ok and same
        if rank == 0:
            print(f"{mode}: dense pool gather dense_equal={same}", flush=True)
    t = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(t) == 1 else 1)


def _run(fe, obs, lo, T, mode, lim, world, n):
    perms = [np.random.default_rng(100 + t).permutation(7) for t in range(T + 1)]
    ro = cm.ModelRollout(fe, T, rollout_mode=mode, dkl_lim=lim, row_offset=lo, perm_fn=lambda t: perms[t])
    pol = cm.StubPolicy(fe.obs_dim, fe.act_dim, device=fe.device)
    G = sharding.P2PPoolGatherer if os.environ.get("CMBPO_TRANSPORT") == "p2p" else sharding.PoolGatherer
    g = G(ro, world) if world else None
    for _ in range(n):
        fe._rng_offset = 0
        ro.rng_base.zero_()             # same Philox streams in every repetition and in the single-GPU reference
        ro.reset(obs)
        ro.run(pol)
        if g:
            g.launch()
    if g:
        return g.wait()
    torch.cuda.synchronize()
    rec, coords = sharding.pack_records(ro.pool, lo)
    return torch.cat([rec, coords.to(torch.int32).view(torch.float32)], dim=1), None


def _run_dense(fe, obs_all, T, mode, lim, rank, world, dev):
    perms = [np.random.default_rng(100 + t).permutation(7) for t in range(T + 1)]
    pol = cm.StubPolicy(fe.obs_dim, fe.act_dim, device=fe.device)
    Bs = obs_all.shape[0] // world
    lo = rank * Bs
    alloc = sharding.symmetric_pool_alloc()
    ros = [cm.ModelRollout(fe, T, rollout_mode=mode, dkl_lim=lim, row_offset=lo, perm_fn=lambda t: perms[t], pool_alloc=alloc)
           for _ in range(2)]
    obs = torch.from_numpy(obs_all[lo:lo + Bs]).to(dev)
    for r in ros:                       # first reset allocates (collective: same order on every rank)
        fe._rng_offset = 0
        r.reset(obs)
    g = sharding.DensePoolGatherer(ros, world)
    for k in range(4):                  # pipelined: the pulls of rollout k overlap rollout k+1 in the other arena
        i = k & 1
        g.before_rollout(i)
        fe._rng_offset = 0
        ros[i].rng_base.zero_()
        ros[i].reset(obs)
        ros[i].run(pol)
        g.launch(i)
    g.wait()
    ref = cm.ModelRollout(fe, T, rollout_mode=mode, dkl_lim=lim, row_offset=0, perm_fn=lambda t: perms[t])
    fe._rng_offset = 0
    ref.rng_base.zero_()
    ref.reset(torch.from_numpy(obs_all).to(dev))
    ref.run(pol)
    torch.cuda.synchronize()
    same = True
    for i in (0, 1):
        got = g.pools(i)
        for name, v in ref.pool.items():
            per_member = mode == "iv_gae" and name not in ("dyn_error", "populated")     # [E, B, T, ...]: B is axis 1
            x = torch.cat(got[name], dim=1 if per_member else 0)
            same = same and x.shape == v.shape and bool(torch.equal(x, v))
    return same


if __name__ == "__main__":
    main()
