#!/usr/bin/env python
"""Benchmark of the B200-native model-rollout path (BASELINE.json metric: rollout transitions/sec).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the CPU implementation of the path

One "step" = one full batched model rollout of the workload's start states: ``ModelSampler.reset``
followed by ``ModelSampler.sample`` until no path is alive (cmbpo.py:335-348), i.e. up to
``max_path_length-1`` passes of policy -> FakeEnv.step -> cutoff -> model-pool write.  A
"transition" is one (row, step) that reaches the model pool.

Default workload = BASELINE.json configs[2], the configuration the metric's targets are quoted on
(>= 60 % of tensor peak at 1 M-row batches, 1/2/4/8 GPUs): HalfCheetah-shaped 7-member / 5-elite
4x200 swish ensemble, 'schedule' rollouts of length 10, 1 M start states.  For N > 1 the SAME 1 M start
states are split over the ranks (strong scaling), weights replicated, and the generated transitions are
all-gathered into the replicated model pool.  The other BASELINE configs are ``--workload`` choices
(configs[0] ``antsafe_predict_10k``, configs[1] ``antsafe_uncertainty_H30_100k``, configs[3]
``humanoidsafe_iv_gae_H12_1M``, configs[4] sweep points).

Prints ONE JSON line on rank 0 (see DESIGN.md "Measurement" for every field).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H4 = (200, 200, 200, 200)
WORKLOADS = {
    # rows = start states of the WHOLE job when scaling == "strong", per GPU when "weak"
    # BASELINE configs[0]: one BNN.predict on 10 k start states (bnn.py:671-717), reference on CPU beside it
    "antsafe_predict_10k": dict(kind="predict", env="antsafe", obs=29, act=8, hidden=H4, E=7, K=5, mode="predict",
                                T=2, rows=10_000, dkl_depth=0, cost=False, scaling="weak"),
    # configs[1]: 'uncertainty' rollouts up to H = 30 on ONE B200 (N > 1: every rank its own 100 k)
    "antsafe_uncertainty_H30_100k": dict(kind="rollout", env="antsafe", obs=29, act=8, hidden=H4, E=7, K=5,
                                         mode="uncertainty", T=30, rows=100_000, dkl_depth=5, cost=True, scaling="weak"),
    # configs[2]: 1 M start states sharded across 1/2/4/8 B200, all-gather into the model pool
    "halfcheetah_schedule_L10_1M": dict(kind="rollout", env="halfcheetah", obs=17, act=6, hidden=H4, E=7, K=5,
                                        mode="schedule", T=10, rows=1_000_000, dkl_depth=0, cost=False, scaling="strong"),
    # configs[3]: Humanoid-shaped ensemble (MBPO-truncated 45-dim obs, 17 act), 'iv_gae' rollouts (every member rolls
    # its own state, 3-D FakeEnv.step) with the static cost head, 1 M start states over the GPUs, maxroll 12
    "humanoidsafe_iv_gae_H12_1M": dict(kind="rollout", env="humanoidsafe", obs=45, act=17, hidden=H4, E=7, K=5,
                                       mode="iv_gae", T=12, rows=1_000_000, dkl_depth=0, cost=True, scaling="strong"),
    "humanoidsafe_iv_gae_H12_125k": dict(kind="rollout", env="humanoidsafe", obs=45, act=17, hidden=H4, E=7, K=5,
                                         mode="iv_gae", T=12, rows=125_000, dkl_depth=0, cost=True, scaling="weak"),
    # the shape every shipped CMBPO config uses (examples/config/antsafe/cmbpo_uc.py:36)
    "antsafe_shipped_2x512_uncertainty_H30_100k": dict(kind="rollout", env="antsafe", obs=29, act=8, hidden=(512, 512),
                                                       E=7, K=5, mode="uncertainty", T=30, rows=100_000, dkl_depth=5,
                                                       cost=True, scaling="weak"),
    # configs[4] sweep points: one sampler step (T = 2) at a fixed batch
    "antsafe_step_1M": dict(kind="rollout", env="antsafe", obs=29, act=8, hidden=H4, E=7, K=5, mode="schedule", T=2,
                            rows=1_000_000, dkl_depth=0, cost=True, scaling="weak"),
    "antsafe_step_4M": dict(kind="rollout", env="antsafe", obs=29, act=8, hidden=H4, E=7, K=5, mode="schedule", T=2,
                            rows=4_000_000, dkl_depth=0, cost=True, scaling="weak"),
    "antsafe_step_1M_4x256": dict(kind="rollout", env="antsafe", obs=29, act=8, hidden=(256,) * 4, E=7, K=5,
                                  mode="schedule", T=2, rows=1_000_000, dkl_depth=0, cost=True, scaling="weak"),
    "antsafe_step_1M_4x512": dict(kind="rollout", env="antsafe", obs=29, act=8, hidden=(512,) * 4, E=7, K=5,
                                  mode="schedule", T=2, rows=1_000_000, dkl_depth=0, cost=True, scaling="weak"),
    "antsafe_step_256k_4x1024": dict(kind="rollout", env="antsafe", obs=29, act=8, hidden=(1024,) * 4, E=7, K=5,
                                     mode="schedule", T=2, rows=262_144, dkl_depth=0, cost=True, scaling="weak"),
}
DEFAULT_WORKLOAD = "halfcheetah_schedule_L10_1M"
METRIC = "rollout transitions/sec (7-member ensemble step)"
MODEL_SEED = 5


def flops_per_row(w):
    """F = 2*E*(in*H + (L-1)*H^2 + H*2D), unpadded, GEMMs only."""
    dims = [w["obs"] + w["act"]] + list(w["hidden"]) + [2 * (w["obs"] + 1)]
    return 2 * w["E"] * sum(a * b for a, b in zip(dims[:-1], dims[1:]))


def shard_rows(w, world, rank):
    """(first global row, rows) of this rank."""
    if w["scaling"] == "strong":
        base, rem = divmod(w["rows"], world)
        lo = rank * base + min(rank, rem)
        return lo, base + (1 if rank < rem else 0)
    return rank * w["rows"], w["rows"]


def config_of(wname, w, n_gpus):
    """What defines the workload -- identical for both arms (the driver compares them)."""
    total = w["rows"] if w["scaling"] == "strong" else w["rows"] * n_gpus
    return {"workload": wname, "kind": w["kind"], "env": w["env"], "members": w["E"], "elites": w["K"],
            "hidden": list(w["hidden"]), "mode": w["mode"], "max_path_length": w["T"], "start_states_total": total,
            "scaling": w["scaling"], "init": f"fc.py:154-185 truncated normal, seed {MODEL_SEED}; identity scalers",
            "l2": "no explicit flush: the per-step working set (model pool + head scratch, GBs at the named size) "
                  "exceeds the 126 MB L2"}


def synth_start_states(w, n, seed, lo=0):
    """SURVEY §8d synthetic start states: N(0,1) with env overrides so thresholds are crossed.  Rows [lo, lo+n) of
    the stream of `seed` (strong scaling: every rank cuts its block out of the same global batch)."""
    rng = np.random.default_rng(seed)
    tot = lo + n
    obs = rng.standard_normal((tot, w["obs"]), dtype=np.float32)
    if w["env"].startswith("ant"):
        obs[:, 0] = rng.uniform(0.3, 0.9, tot)
        q = rng.standard_normal((tot, 4))
        obs[:, 1:5] = q / np.linalg.norm(q, axis=-1, keepdims=True)
        obs[:, -1] = rng.uniform(-4, 4, tot)
    if w["env"].startswith("humanoid"):
        obs[:, 0] = rng.uniform(1.1, 1.9, tot)
    return np.ascontiguousarray(obs[lo:])


def draw_weights(w, seed=MODEL_SEED):
    """The fp32 initial weights ``BNN.finalize`` draws for this shape and seed (constrainedmbpo_b200/bnn.py finalize:
    torch CPU generator, truncated normal std 1/(2 sqrt(in)), fc.py:154-185): lets the CPU legs time the SAME ensemble
    the GPU arm runs without touching the CUDA library."""
    import torch
    gen = torch.Generator(device="cpu")
    gen.manual_seed(int(seed))
    dims = [w["obs"] + w["act"]] + list(w["hidden"]) + [2 * (w["obs"] + 1)]
    W, B = [], []
    for k, n in zip(dims[:-1], dims[1:]):
        std = 1.0 / (2.0 * np.sqrt(k))
        t = torch.empty(w["E"], k, n, dtype=torch.float32)
        torch.nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=gen)
        W.append(t.numpy())
        B.append(np.zeros((w["E"], 1, n), np.float32))
    return W, B


def cpu_params(w):
    """oracle.EnsembleParams of the benchmark ensemble (CPU legs only: cpu_baseline, --impl reference)."""
    from oracle import cmbpo_oracle as orc
    W, B = draw_weights(w)
    one = lambda d, v: np.full((1, d), v, np.float32)
    i, o = w["obs"] + w["act"], w["obs"] + 1
    return orc.EnsembleParams(W, B, ["swish"] * len(w["hidden"]) + [None], mu_in=one(i, 0), var_in=one(i, 1),
                              mu_out=one(o, 0), var_out=one(o, 1), probabilistic=True, elites=list(range(w["K"])))


def stub_policy_matrix(w, seed=4):
    import torch
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(w["obs"], w["act"], generator=g) / np.sqrt(w["obs"])).numpy()


def committed_traffic(wname, rows_per_launch):
    """dram bytes per forward launch from the committed ``ncu --set full`` capture of THIS workload's forward launch
    (profiles/traffic.json, written by tools/summarize_profiles.py from the .ncu-rep); None if the capture was taken
    at another row count."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None, None
    d = json.load(open(path)).get(wname)
    if not d or abs(d["rows_per_launch"] - rows_per_launch) > 0.01 * rows_per_launch:
        return None, None
    return d["dram_bytes_per_launch"], d["source"]


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(bf16_burst=d.get("bf16_tflops"), bf16_sustained=d.get("bf16_tflops_sustained"),
                    hbm=d.get("hbm_gbs"), source="measured")
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback")


def bind_to_gpu_numa_node(gpu_index):
    """Pin this rank's host threads (and so its pinned staging buffers, first-touched locally) to the CPUs NVML
    reports as local to the GPU.  Best effort: returns the CPU count bound, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = [64 * w + b for w, word in enumerate(mask) for b in range(64) if (word >> b) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


def pick_peak(peaks, clk):
    """Roofline denominator from the clocks record: the burst figure when the timed region ran unthrottled at (near)
    the maximum SM clock, the sustained one when the power cap was active or the clock sagged."""
    try:
        burst = clk and not clk.get("reasons") and clk["sm_mhz"] >= 0.9 * clk["sm_max_mhz"]
    except Exception:
        burst = False
    return ("burst", peaks["bf16_burst"]) if burst else ("sustained", peaks["bf16_sustained"])


# ----------------------------------------------------------------------------------------------
# CPU arm (cpu_baseline leg and --impl reference): the fp32 op-for-op port, all host threads
# ----------------------------------------------------------------------------------------------
def cpu_sample(w, rows, steps=1, warmup=0, dkl_lim=None, seed=1):
    """Times the CPU port (oracle/cpu_baseline.py) on `rows` start states of the same workload and ensemble."""
    import torch
    from oracle import cpu_baseline as cb
    ens = cb.TorchEnsemble(cpu_params(w), threads=os.cpu_count())
    P = stub_policy_matrix(w)
    policy = lambda o: np.tanh(np.asarray(o, np.float32) @ P)
    start = synth_start_states(w, rows, seed)
    times, counts = [], []
    if w["kind"] == "predict":
        act = policy(start)
        x = np.concatenate([start, act], -1)
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            ens.predict_factored(x)
            if i >= warmup:
                times.append(time.perf_counter() - t0); counts.append(rows)
    else:
        if w["mode"] == "uncertainty" and dkl_lim is None:
            dkl_lim = cb.compute_dynamics_dkl(ens, w["env"], synth_start_states(w, 5000, 99), policy, w["dkl_depth"])
        for i in range(warmup + steps):
            n, dt, _ = cb.rollout(ens, w["env"], start, policy, w["T"], w["mode"], dkl_lim=dkl_lim,
                                  cares_about_cost=w["cost"])
            if i >= warmup:
                times.append(dt); counts.append(n)
    return dict(transitions=int(sum(counts)), seconds=float(sum(times)), dkl_lim=dkl_lim,
                threads=torch.get_num_threads())


def default_cpu_rows(w):
    """Bounded sample: ~4 s of CPU work per step on a 16-core box (the port does ~40 k transitions/s), so that the
    reference arm's K steps and the 3-step cpu_baseline leg (~10-15 s) both finish within minutes."""
    if w["kind"] == "predict":
        return w["rows"]
    per_row = max(w["T"] - 1, 1) * (w["E"] if w["mode"] == "iv_gae" else 1)
    return int(min(w["rows"], max(2048, 150_000 // per_row)))


def bytes_per_transition(w):
    """fp32 record that reaches the model pool: obs, act, next_obs, rew, cost, dyn_error + term byte
    ('iv_gae': one obs/act/next_obs/rew/cost/term per member, modelbuffer.py:48-51)."""
    if w["mode"] == "iv_gae":
        return w["E"] * (4 * (2 * w["obs"] + w["act"] + 2) + 1) + 4
    return 4 * (2 * w["obs"] + w["act"] + 3) + 1


def run_reference(args, w, wname):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rows = args.cpu_rows or default_cpu_rows(w)
    r = cpu_sample(w, rows, steps=args.steps, warmup=min(args.warmup, 1))
    val = r["transitions"] / r["seconds"]
    sample = (f"{args.steps} x {rows} start states of the workload's {w['rows']}, same ensemble (seed {MODEL_SEED}) / mode / H; "
              f"TF 1.14 is not installable here -> the fp32 op-for-op port is timed (torch BLAS GEMMs in the role of TF's "
              f"Eigen contraction + the reference's NumPy post-processing), {r['threads']} threads")
    out = {
        "impl": "reference", "metric": METRIC, "value": val,
        "unit": "transitions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["seconds"] / args.steps, "higher_is_better": True, "scaling": w["scaling"],
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_of(wname, w, args.gpus),
        "cpu_baseline": {"value": val, "unit": "transitions/s", "cores": r["threads"], "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "transitions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
def build_env(w, local_rank, operand_dtype="bf16"):
    import constrainedmbpo_b200 as cm
    fe = cm.FakeEnv(cm.SimpleEnv(w["obs"], w["act"]), w["env"], num_networks=w["E"], num_elites=w["K"],
                    hidden_dims=w["hidden"], cares_about_cost=w["cost"], device=f"cuda:{local_rank}", seed=MODEL_SEED,
                    operand_dtype=operand_dtype)
    fe._model.set_elites(list(range(w["K"])))      # fitted-model stand-in: the first K members are the elites
    return fe


def fwd_kernel_time(fe, fn, reps):
    """CUDA events around every ensemble-forward launch of `reps` eager calls of fn (cmbpo_profile)."""
    import ctypes as C
    import torch
    lib, h = fe._model._lib, fe._model._handle
    lib.cmbpo_profile(h, 1)
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    ms, n = C.c_double(), C.c_int64()
    lib.cmbpo_profile_read(h, C.byref(ms), C.byref(n))
    lib.cmbpo_profile(h, 0)
    return ms.value / reps, int(n.value) // reps


def run_gpu(args, w, wname):
    import torch
    import torch.distributed as dist
    import constrainedmbpo_b200 as cm

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else None   # N=1: all cores stay free for the cpu_baseline leg
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    row0, rows = shard_rows(w, world, rank)
    if args.rows:
        row0, rows = rank * args.rows, args.rows
    fe = build_env(w, local_rank, args.operand_dtype)
    model = fe._model
    policy = cm.StubPolicy(w["obs"], w["act"], device=dev, seed=4)
    start_host = torch.from_numpy(synth_start_states(w, rows, 1, lo=row0)).pin_memory()
    start_dev = start_host.to(dev)
    F = flops_per_row(w)
    peaks = measured_peaks()
    steps, warm = args.steps, max(args.warmup, 3)
    gather = dgather = None
    dkl_lim = None

    if w["kind"] == "predict":
        # ---- configs[0]: one BNN.predict per step ---------------------------------------------------------------
        act_dev = policy(start_dev).clone()
        x_dev = torch.cat([start_dev, act_dev], dim=-1).contiguous()
        x_host = x_dev.cpu().numpy()

        def one_step():
            model.predict_device([x_dev], 2, 0)
        per_step_units, rows_forwarded, launches_fwd = rows, rows, 1
        kernels_per_step = 3          # stage_inputs, ensemble_fwd, finalize_predict
        step_eager = one_step
    else:
        if w["mode"] == "uncertainty":  # cmbpo.py:283-285: 5000 archive states, depth = rollout_schedule[2]
            cal = torch.from_numpy(synth_start_states(w, 5000, 99)).to(dev)
            dkl_lim = cm.compute_dynamics_dkl(fe, policy, cal, depth=w["dkl_depth"])
        # N > 1: 'schedule' pools are dense (every path stores every step), so the pool arenas themselves are all-gathered
        # by copy engines (sharding.DensePoolGatherer, two rollouts alternating so the pulls overlap the next rollout);
        # ragged pools ('uncertainty', 'iv_gae' terminals) go as packed records (P2PPoolGatherer / PoolGatherer)
        fmt = args.gather_format if args.gather_format != "auto" else ("dense" if w["mode"] == "schedule" else "records")
        dense = world > 1 and args.gather != "none" and fmt == "dense"
        alloc = None
        if dense:
            from constrainedmbpo_b200.sharding import DensePoolGatherer, symmetric_pool_alloc
            # symmetric memory needs peer access between all ranks' GPUs; if the box cannot provide it, every rank falls
            # back (collectively) to packed records over NCCL rather than failing the run
            ok = torch.ones(1, device=dev)
            try:
                import torch.distributed._symmetric_memory as _symm
                probe = _symm.empty(1024, dtype=torch.uint8, device=dev)
                _symm.rendezvous(probe, dist.group.WORLD)
            except Exception as exc:    # noqa: BLE001
                ok.zero_()
                if rank == 0:
                    print(f"[bench] symmetric memory unavailable ({type(exc).__name__}: {exc}); using NCCL records", file=sys.stderr)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if float(ok) < 1:
                dense, fmt = False, "records"
                args.transport = "nccl"
            else:
                alloc = symmetric_pool_alloc()
        ros = [cm.ModelRollout(fe, w["T"], rollout_mode=w["mode"], dkl_lim=dkl_lim, row_offset=row0, pool_alloc=alloc)
               for _ in range(2 if dense else 1)]
        ro = ros[0]

        def step_eager():
            ro.reset(start_dev)
            ro.run(policy)
        if args.no_graph:
            step_fns = [(lambda r=r: (r.reset(start_dev), r.run(policy))) for r in ros]
        else:
            for r in ros:
                r.capture(start_dev, policy)
            step_fns = [r.replay for r in ros]
        step_fn = step_fns[0]
        dgather = None
        if dense:
            dgather = DensePoolGatherer(ros, world, pack_only=(args.gather == "pack"))
        elif world > 1 and args.gather != "none":
            from constrainedmbpo_b200.sharding import P2PPoolGatherer, PoolGatherer
            G = P2PPoolGatherer if args.transport == "p2p" else PoolGatherer
            group = None
            if args.nccl_max_ctas > 0:   # cap the SMs NCCL takes from the persistent forward kernel
                opts = dist.ProcessGroupNCCL.Options()
                opts.config.max_ctas = args.nccl_max_ctas
                group = dist.new_group(pg_options=opts)
            gather = G(ro, world, group=group, pack_only=(args.gather == "pack"))
        tick = [0]

        def one_step():
            if dgather is not None:
                i = tick[0] & 1
                tick[0] += 1
                dgather.before_rollout(i)
                step_fns[i]()
                dgather.launch(i)
                return
            step_fn()
            if gather is not None:
                gather.launch()

    for _ in range(warm):
        one_step()
    for g_ in (gather, dgather):
        if g_ is not None:
            g_.wait()
    torch.cuda.synchronize()
    model.check_status()
    if w["kind"] == "rollout":
        n_alive, ptr, per_step_units, finished = ro.counters()
        hist = ro.alive_hist.cpu().numpy().astype(np.int64)
        rows_forwarded = int(hist.sum())
        launches_fwd = int((hist > 0).sum())
        # forward launches per sampler step: 1 for the fused kernels, one per FC layer and 131072-row slab for wide_layer_kernel
        fwd_launches = (len(w["hidden"]) + 1) * ((rows + 131071) // 131072) if model._lib.cmbpo_forward_path(model._handle) == 1 else 1
        # per sampler step: stub policy, input staging, ensemble forward (fused: 1 launch; wide: one per FC layer and
        # 131072-row slab), step post, 3 bookkeeping kernels (decide+scan, store+scan, compact)
        kernels_per_step = (6 + fwd_launches) * (w["T"] - 1) + (3 if gather is not None else 0)   # dense gather: copy engines only

    # ---- timed region: K steps, device-resident inputs ------------------------------------------------------------
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(steps):
        one_step()
    for g_ in (gather, dgather):
        if g_ is not None:
            g_.wait()
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    clk = clocks.stop() if rank == 0 else None
    model.check_status()
    t = torch.tensor([ms, float(per_step_units)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, total_units = float(tmax[0]), float(tsum[1])
    else:
        total_units = float(per_step_units)
    value = total_units * steps / (ms * 1e-3)

    # ---- roofline of the dominant kernel (ensemble forward): CUDA events around every launch ------------------------
    reps = min(steps, 5)
    fwd_ms_per_step, _ = fwd_kernel_time(fe, step_eager, reps)
    achieved_tf = F * rows_forwarded / (fwd_ms_per_step * 1e-3) / 1e12

    # ---- e2e: HOST inputs in (pinned -> H2D every step), results out to pinned host memory, inside the timed region --
    # up to 40 steps: the last rollout's device -> host copy cannot hide under a following rollout (the clock stops only when
    # it has landed), so a short run charges that tail to few steps (10 steps: ~10 % of the region at the default workload)
    e2e_steps = max(1, min(steps, 40))
    if w["kind"] == "predict":
        for _ in range(2):
            model.predict(x_host, factored=True)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            mean_h, var_h = model.predict(x_host, factored=True)
        e2e_s = time.perf_counter() - t0
        n_units = float(rows * e2e_steps)
        h2d, d2h = int(x_host.nbytes), int(mean_h.nbytes + var_h.nbytes)
        assert np.isfinite(mean_h).all()
    else:
        from constrainedmbpo_b200.sharding import HostDrain
        drain = HostDrain(ro)

        # The start states of rollout k+1 travel host -> device on a side stream while rollout k runs (the graph reads a
        # fixed buffer, so they land in a staging tensor and a device-to-device copy hands them over): one pinned H2D copy
        # per step inside the timed region, off the critical path.
        side = torch.cuda.Stream(device=dev)
        stage = torch.empty_like(start_dev)
        h2d_done, taken = torch.cuda.Event(), torch.cuda.Event()
        with torch.cuda.stream(side):
            stage.copy_(start_host, non_blocking=True)
            h2d_done.record(side)

        def e2e_step():
            main = torch.cuda.current_stream(dev)
            main.wait_event(h2d_done)
            start_dev.copy_(stage, non_blocking=True)
            taken.record(main)
            with torch.cuda.stream(side):
                side.wait_event(taken)
                stage.copy_(start_host, non_blocking=True)      # next step's start states
                h2d_done.record(side)
            step_fn()
            drain.launch()
        for _ in range(2):
            e2e_step()
        drain.wait()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        n_units = 0
        for k in range(e2e_steps):
            e2e_step()
            if k > 0:
                n_units += drain.last[1][0]          # rollout k-1: its device->host copy was just enqueued
        host_rec, cnt = drain.wait()                 # the last rollout's records are on the host when this returns
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        n_units += cnt[0]
        assert host_rec.shape[0] == cnt[0] > 0 and float(host_rec[:, 0].abs().sum()) > 0
        assert drain.last_next is not None and float(drain.last_next.abs().sum()) > 0   # compact records + last next_obs
        h2d, d2h = int(start_host.numel() * 4), drain.bytes_per_rollout(per_step_units)
    te = torch.tensor([e2e_s, float(n_units)], dtype=torch.float64, device=dev)
    if world > 1:
        a = te.clone(); dist.all_reduce(a, op=dist.ReduceOp.MAX)
        b = te.clone(); dist.all_reduce(b, op=dist.ReduceOp.SUM)
        e2e_s, n_units = float(a[0]), float(b[1])
    e2e_val = n_units / e2e_s
    model.check_status()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- CPU baseline beside it (rank 0, N=1 only; bounded sample of the same workload and ensemble) -----------------
    cpu = None
    if args.gpus == 1 and not args.no_cpu:
        W_cpu, _ = draw_weights(w)
        assert all(np.array_equal(a, b.cpu().numpy()) for a, b in zip(W_cpu, model.weights)), "CPU leg ensemble != GPU ensemble"
        crow = args.cpu_rows or default_cpu_rows(w)
        r = cpu_sample(w, crow, steps=3, warmup=0, dkl_lim=dkl_lim)
        cpu = {"value": r["transitions"] / r["seconds"], "unit": "transitions/s", "cores": r["threads"],
               "kind": "port",
               "sample": f"3 x {crow} of the workload's {w['rows']} start states, same ensemble/mode/H/dkl_lim; "
                         f"{r['transitions']} transitions in {r['seconds']:.2f} s; TF 1.14 is not installable -> fp32 op-for-op "
                         f"port (torch BLAS GEMMs + NumPy post-processing, 64k-row chunks)"}

    traffic, traffic_src = committed_traffic(wname, rows_forwarded / max(launches_fwd, 1))
    fwd_kernel = ("ensemble_fwd_kernel", "wide_layer_kernel", "fused_wide_kernel")[model._lib.cmbpo_forward_path(model._handle)]
    which, peak = pick_peak(peaks, clk)
    cfg = config_of(wname, w, args.gpus)
    out = {
        "metric": METRIC,
        "value": value, "unit": "transitions/s", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": ms / steps, "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None,
        "dtype": args.operand_dtype, "data": "synthetic",
        "config": cfg,
        "details": {"start_states_per_gpu": rows, "dkl_lim": dkl_lim, "units_per_step_per_gpu": per_step_units,
                    "rows_forwarded_per_step": rows_forwarded, "cuda_graph": w["kind"] == "rollout" and not args.no_graph,
                    "timed_region_s": ms * 1e-3,
                    "parallelism": f"rows sharded x{args.gpus} ({w['scaling']} scaling), weights replicated" +
                                   (f", all-gather of transitions into the replicated pool (records, transport={args.transport}, "
                                    f"gather={args.gather})" if gather is not None else "") +
                                   (f", all-gather of the pool arenas into the replicated pool (dense: one copy-engine pull per "
                                    f"peer over NVLink peer mappings, double-buffered rollouts, gather={args.gather})" if dgather is not None else "")},
        "clocks": clk,
        "e2e": {"value": e2e_val, "unit": "transitions/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "host_cpus_bound": numa},
        "gpu_launches": int(kernels_per_step * steps),
        "roofline": {"bound": "tensor", "kernel": fwd_kernel,
                     "achieved": achieved_tf, "peak": peak, "unit": "TFLOP/s", "frac": achieved_tf / peak,
                     "peak_source": f"{peaks['source']} bf16 {which} (chosen from the clocks record)",
                     "frac_of_burst": achieved_tf / peaks["bf16_burst"], "frac_of_sustained": achieved_tf / peaks["bf16_sustained"],
                     "traffic": traffic, "traffic_source": traffic_src,
                     "flops_per_row": F, "launches_per_step": launches_fwd,
                     "kernel_ms_per_step": fwd_ms_per_step,
                     "kernel_share_of_step": fwd_ms_per_step / (ms / steps)},
        "cpu_baseline": cpu,
    }
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override start states per GPU")
    ap.add_argument("--cpu-rows", type=int, default=0, help="start states of the CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--gather", default="full", choices=["full", "pack", "none"],
                    help="N>1 diagnosis: full = pack + all-gather (the measured path), pack = packing only, none")
    ap.add_argument("--gather-format", default="auto", choices=["auto", "dense", "records"],
                    help="N>1: dense = pull whole pool arenas (default for 'schedule'), records = packed populated transitions")
    ap.add_argument("--transport", default="p2p", choices=["nccl", "p2p"],
                    help="N>1: NCCL all-gather kernels, or copy-engine pulls over symmetric-memory peer mappings")
    ap.add_argument("--nccl-max-ctas", type=int, default=0, help="N>1: NCCL max_ctas of the gather group (0 = default)")
    ap.add_argument("--operand-dtype", default="bf16", choices=["bf16", "tf32"],
                    help="tensor-core operand type of the ensemble GEMMs (tf32: layer-at-a-time kernel, rtol 1e-3)")
    ap.add_argument("--no-graph", action="store_true", help="launch kernels eagerly instead of replaying a CUDA graph")
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, w, args.workload)
    else:
        run_gpu(args, w, args.workload)


if __name__ == "__main__":
    main()
