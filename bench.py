#!/usr/bin/env python
"""Benchmark of the B200-native model-rollout path (BASELINE.json metric: rollout transitions/sec).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the CPU implementation of the path

One "step" = one full batched model rollout of the workload's start states: ``ModelSampler.reset``
followed by ``ModelSampler.sample`` until no path is alive (cmbpo.py:335-348), i.e. up to
``max_path_length-1`` passes of policy -> FakeEnv.step -> cutoff -> model-pool write.  A
"transition" is one (row, step) that reaches the model pool.  N=1 workload = BASELINE.json
configs[1]: antsafe-shaped 7-member / 5-elite 4x200 swish ensemble, 'uncertainty' rollout mode,
H=30, 100k start states.  For N>1 every rank rolls out its own 100k start states with replicated
weights (weak scaling) and the stored transitions are all-gathered into the replicated model pool
with NCCL.

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

WORKLOADS = {
    # name: env, hidden, E, elites, mode, T (max_path_length), rows per GPU, dkl depth
    "antsafe_uncertainty_H30_100k": dict(env="antsafe", obs=29, act=8, hidden=(200, 200, 200, 200), E=7, K=5,
                                         mode="uncertainty", T=30, rows=100_000, dkl_depth=5, cost=True),
    "halfcheetah_schedule_L10_1M": dict(env="halfcheetah", obs=17, act=6, hidden=(200, 200, 200, 200), E=7, K=5,
                                        mode="schedule", T=10, rows=1_000_000, dkl_depth=0, cost=False),
    # the shape every shipped CMBPO config uses (examples/config/antsafe/cmbpo_uc.py:36): two 512-wide hidden
    # layers -> the layer-at-a-time wide kernel (wide_fwd.cu)
    "antsafe_shipped_2x512_uncertainty_H30_100k": dict(env="antsafe", obs=29, act=8, hidden=(512, 512), E=7, K=5,
                                                       mode="uncertainty", T=30, rows=100_000, dkl_depth=5, cost=True),
    "antsafe_step_1M_4x512": dict(env="antsafe", obs=29, act=8, hidden=(512, 512, 512, 512), E=7, K=5,
                                  mode="schedule", T=2, rows=1_000_000, dkl_depth=0, cost=True),
    "antsafe_step_256k_4x1024": dict(env="antsafe", obs=29, act=8, hidden=(1024, 1024, 1024, 1024), E=7, K=5,
                                     mode="schedule", T=2, rows=262_144, dkl_depth=0, cost=True),
    # BASELINE.json configs[3]: Humanoid-shaped ensemble (MBPO-truncated 45-dim obs, 17 act), 'iv_gae' rollouts
    # (every member rolls its own state, 3-D FakeEnv.step) with the static cost head; 1M start states over 8 GPUs
    # = 125k per GPU, maxroll 12
    "humanoidsafe_iv_gae_H12_125k": dict(env="humanoidsafe", obs=45, act=17, hidden=(200, 200, 200, 200), E=7, K=5,
                                         mode="iv_gae", T=12, rows=125_000, dkl_depth=0, cost=True),
    "antsafe_step_1M": dict(env="antsafe", obs=29, act=8, hidden=(200, 200, 200, 200), E=7, K=5,
                            mode="schedule", T=2, rows=1_000_000, dkl_depth=0, cost=True),
}
DEFAULT_WORKLOAD = "antsafe_uncertainty_H30_100k"


def flops_per_row(w):
    """F = 2*E*(in*H + (L-1)*H^2 + H*2D), unpadded, GEMMs only."""
    dims = [w["obs"] + w["act"]] + list(w["hidden"]) + [2 * (w["obs"] + 1)]
    return 2 * w["E"] * sum(a * b for a, b in zip(dims[:-1], dims[1:]))


def bytes_per_transition(w):
    """fp32 record that reaches the model pool: obs, act, next_obs, rew, cost, dyn_error + term byte
    ('iv_gae': one obs/act/next_obs/rew/cost/term per member, modelbuffer.py:48-51)."""
    if w["mode"] == "iv_gae":
        return w["E"] * (4 * (2 * w["obs"] + w["act"] + 2) + 1) + 4
    return 4 * (2 * w["obs"] + w["act"] + 3) + 1


def synth_start_states(w, n, seed):
    """SURVEY §8d synthetic start states: N(0,1) with env overrides so thresholds are crossed."""
    rng = np.random.default_rng(seed)
    obs = rng.standard_normal((n, w["obs"]), dtype=np.float32)
    if w["env"].startswith("ant"):
        obs[:, 0] = rng.uniform(0.3, 0.9, n)
        q = rng.standard_normal((n, 4))
        obs[:, 1:5] = q / np.linalg.norm(q, axis=-1, keepdims=True)
        obs[:, -1] = rng.uniform(-4, 4, n)
    if w["env"].startswith("humanoid"):
        obs[:, 0] = rng.uniform(1.1, 1.9, n)
    return obs


def synth_params(w, seed=0):
    from oracle import cmbpo_oracle as orc
    rng = np.random.default_rng(seed)
    return orc.init_ensemble(rng, w["obs"] + w["act"], w["obs"] + 1, w["hidden"], w["E"], probabilistic=True,
                             use_scaler_in=True, use_scaler_out=True, elites=list(range(w["K"])), dtype=np.float32)


def stub_policy_matrix(w, seed=4):
    import torch
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(w["obs"], w["act"], generator=g) / np.sqrt(w["obs"])).numpy()


def committed_traffic(w, rows_per_launch):
    """dram bytes per forward launch from the committed ncu --set full capture of the same kernel and
    ensemble shape (profiles/*_fwd_1M_traffic.json: one 1M-row launch), scaled by rows per launch."""
    import glob
    if tuple(w["hidden"]) != (200, 200, 200, 200) or w["env"] != "antsafe" or w["mode"] == "iv_gae":
        return None, None
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_fwd_1M_traffic.json")))
    if not files:
        return None, None
    d = json.load(open(files[-1]))
    return d["bytes_per_row"] * rows_per_launch, d["source"]


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(bf16_burst=d.get("bf16_tflops"), bf16_sustained=d.get("bf16_tflops_sustained"),
                    hbm=d.get("hbm_gbs"), source="measured")
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback")


def bind_to_gpu_numa_node(gpu_index):
    """Pin this rank's host threads (and so its pinned staging buffers, first-touched locally) to the CPUs NVML
    reports as local to the GPU, so that 8 ranks streaming records to the host do not cross sockets on a multi-socket
    box (a no-op on the single-node pool this was measured on).  Best effort: returns the CPU count bound, or None."""
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
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region."""
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


# ----------------------------------------------------------------------------------------------
# CPU arm (cpu_baseline leg and --impl reference)
# ----------------------------------------------------------------------------------------------
def cpu_rollout_sample(w, rows, steps=1, warmup=0, dkl_lim=None, seed=1):
    """Times the CPU port (oracle/cpu_baseline.py) on `rows` start states of the same workload."""
    import torch
    from oracle import cpu_baseline as cb
    p = synth_params(w)
    ens = cb.TorchEnsemble(p, threads=os.cpu_count())
    P = stub_policy_matrix(w)
    policy = lambda o: np.tanh(np.asarray(o, np.float32) @ P)
    start = synth_start_states(w, rows, seed)
    if w["mode"] == "uncertainty" and dkl_lim is None:
        dkl_lim = cb.compute_dynamics_dkl(ens, w["env"], synth_start_states(w, 5000, 99), policy, w["dkl_depth"])
    times, counts = [], []
    for i in range(warmup + steps):
        n, dt, _ = cb.rollout(ens, w["env"], start, policy, w["T"], w["mode"], dkl_lim=dkl_lim,
                              cares_about_cost=w["cost"])
        if i >= warmup:
            times.append(dt); counts.append(n)
    return dict(transitions=int(sum(counts)), seconds=float(sum(times)), dkl_lim=dkl_lim,
                threads=torch.get_num_threads())


def run_reference(args, w, wname):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rows = args.cpu_rows or 8192
    r = cpu_rollout_sample(w, rows, steps=args.steps, warmup=min(args.warmup, 1))
    val = r["transitions"] / r["seconds"]
    sample = (f"{args.steps} rollouts of {rows} start states (of the workload's {w['rows']}), same ensemble/mode/"
              f"H; TF 1.14 unavailable on this box -> fp32 op-for-op port timed (torch BLAS GEMMs + NumPy post-processing)")
    out = {
        "impl": "reference", "metric": "rollout transitions/sec (7-member ensemble step)", "value": val,
        "unit": "transitions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["seconds"] / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wname, "rows_per_step": rows, "mode": w["mode"], "max_path_length": w["T"]},
        "cpu_baseline": {"value": val, "unit": "transitions/s", "cores": r["threads"], "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "transitions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
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

    rows = args.rows or w["rows"]
    p = synth_params(w)
    fe = cm.FakeEnv(cm.SimpleEnv(w["obs"], w["act"]), w["env"], num_networks=w["E"], num_elites=w["K"],
                    hidden_dims=w["hidden"], cares_about_cost=w["cost"], device=f"cuda:{local_rank}", seed=5)
    fe._model.load_state(p.weights, p.biases, scaler_in=(p.mu_in, p.var_in), scaler_out=(p.mu_out, p.var_out),
                         elites=p.elites)
    policy = cm.StubPolicy(w["obs"], w["act"], device=dev, seed=4)
    dkl_lim = None
    if w["mode"] == "uncertainty":  # cmbpo.py:283-285: 5000 archive states, depth = rollout_schedule[2]
        cal = torch.from_numpy(synth_start_states(w, 5000, 99)).to(dev)
        dkl_lim = cm.compute_dynamics_dkl(fe, policy, cal, depth=w["dkl_depth"])
    start_host = torch.from_numpy(synth_start_states(w, rows, 1 + rank)).pin_memory()
    start_dev = start_host.to(dev)
    ro = cm.ModelRollout(fe, w["T"], rollout_mode=w["mode"], dkl_lim=dkl_lim, row_offset=rank * rows)
    use_graph = not args.no_graph
    if use_graph:
        ro.capture(start_dev, policy)
        step_fn = ro.replay
    else:
        def step_fn():
            ro.reset(start_dev)
            ro.run(policy)
    gather = None
    if world > 1 and args.gather != "none":
        from constrainedmbpo_b200.sharding import P2PPoolGatherer, PoolGatherer
        if args.transport == "p2p":
            PoolGatherer = P2PPoolGatherer
        group = None
        if args.nccl_max_ctas > 0:   # cap the SMs NCCL takes from the persistent forward kernel
            opts = dist.ProcessGroupNCCL.Options()
            opts.config.max_ctas = args.nccl_max_ctas
            group = dist.new_group(pg_options=opts)
        gather = PoolGatherer(ro, world, group=group, pack_only=(args.gather == "pack"))

    def one_step():
        step_fn()
        if gather is not None:
            gather.launch()

    for _ in range(max(args.warmup, 3)):
        one_step()
    if gather is not None:
        gather.wait()
    torch.cuda.synchronize()
    n_alive, ptr, per_rollout, finished = ro.counters()
    hist = ro.alive_hist.cpu().numpy().astype(np.int64)
    rows_forwarded = int(hist.sum())
    launches_fwd = int((hist > 0).sum())

    # ---- timed region: K rollouts, device-resident inputs -------------------------------------
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(args.steps):
        one_step()
    if gather is not None:
        gather.wait()
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    clk = clocks.stop() if rank == 0 else None
    fe._model._lib.cmbpo_status(fe._model._handle, None)
    t = torch.tensor([ms, float(per_rollout)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, total_per_rollout = float(tmax[0]), float(tsum[1])
    else:
        total_per_rollout = float(per_rollout)
    value = total_per_rollout * args.steps / (ms * 1e-3)

    # ---- roofline of the dominant kernel (ensemble forward): CUDA events around every launch ----
    import ctypes as C
    lib, h = fe._model._lib, fe._model._handle
    lib.cmbpo_profile(h, 1)
    for _ in range(min(args.steps, 5)):
        ro.reset(start_dev)
        ro.run(policy)
    torch.cuda.synchronize()
    fwd_ms, fwd_n = C.c_double(), C.c_int64()
    lib.cmbpo_profile_read(h, C.byref(fwd_ms), C.byref(fwd_n))
    lib.cmbpo_profile(h, 0)
    reps = min(args.steps, 5)
    F = flops_per_row(w)
    fwd_ms_per_rollout = fwd_ms.value / reps
    achieved_tf = F * rows_forwarded / (fwd_ms_per_rollout * 1e-3) / 1e12
    # full-occupancy launches only (first step of every rollout: all rows alive)
    peaks = measured_peaks()
    peak = peaks["bf16_sustained"]

    # ---- e2e: host start states in (pinned -> H2D every step), populated transitions out to pinned host memory
    # (pack kernel + one D2H per rollout on a side stream, overlapping the next rollout), all inside the timed region
    from constrainedmbpo_b200.sharding import HostDrain
    drain = HostDrain(ro)

    def e2e_step():
        start_dev.copy_(start_host, non_blocking=True)
        step_fn()
        drain.launch()
    for _ in range(2):
        e2e_step()
    drain.wait()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e2e_steps = max(1, min(args.steps, 10))
    t0 = time.perf_counter()
    n_tr = 0
    for k in range(e2e_steps):
        e2e_step()
        if k > 0:
            n_tr += drain.last[1][0]          # rollout k-1: its device->host copy was just enqueued
    host_rec, cnt = drain.wait()              # the last rollout's records are on the host when this returns
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    n_tr += cnt[0]
    assert host_rec.shape[0] == cnt[0] > 0 and float(host_rec[:, 0].abs().sum()) > 0
    te = torch.tensor([e2e_s, float(n_tr)], dtype=torch.float64, device=dev)
    if world > 1:
        a = te.clone(); dist.all_reduce(a, op=dist.ReduceOp.MAX)
        b = te.clone(); dist.all_reduce(b, op=dist.ReduceOp.SUM)
        e2e_s, n_tr = float(a[0]), float(b[1])
    e2e_val = n_tr / e2e_s
    rec = drain.W * 4   # packed record: the transition payload + (path, t) coordinates

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- CPU baseline beside it (rank 0, N=1 only; bounded sample) ---------------------------------
    cpu = None
    if args.gpus == 1 and not args.no_cpu:
        crow = args.cpu_rows or 16384
        r = cpu_rollout_sample(w, crow, steps=1, warmup=0, dkl_lim=dkl_lim)
        cpu = {"value": r["transitions"] / r["seconds"], "unit": "transitions/s", "cores": r["threads"],
               "kind": "port",
               "sample": f"1 rollout of {crow} of the workload's {rows} start states, same ensemble/mode/H/dkl_lim; "
                         f"{r['transitions']} transitions in {r['seconds']:.2f} s; TF 1.14 unavailable -> fp32 op-for-op "
                         f"port (torch BLAS GEMMs + NumPy post-processing, 64k-row chunks)"}

    traffic, traffic_src = committed_traffic(w, rows_forwarded / max(launches_fwd, 1))
    wide = max(w["hidden"]) > 254
    fwd_kernel = "wide_layer_kernel" if wide else "ensemble_fwd_kernel"
    # stub policy, input staging, ensemble forward (fused: 1 launch; wide: one per FC layer and 131072-row slab),
    # step post, 3 bookkeeping kernels (decide+scan, store+scan, compact); N>1 adds 3 pack kernels per rollout
    fwd_launches = (len(w["hidden"]) + 1) * ((rows + 131071) // 131072) if wide else 1
    kernels_per_sampler_step = 6 + fwd_launches
    out = {
        "metric": "rollout transitions/sec (7-member ensemble step)",
        "value": value, "unit": "transitions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": wname, "env": w["env"], "members": w["E"], "elites": w["K"], "hidden": list(w["hidden"]),
                   "mode": w["mode"], "max_path_length": w["T"], "start_states_per_gpu": rows,
                   "dkl_lim": dkl_lim, "transitions_per_rollout_per_gpu": per_rollout,
                   "rows_forwarded_per_rollout": rows_forwarded, "cuda_graph": use_graph,
                   "l2": "no explicit flush: per-rollout working set (model pool + head scratch, > 0.9 GB at the named size) exceeds the 126 MB L2",
                   "parallelism": f"rows sharded x{args.gpus}, weights replicated" + (f", all-gather of transitions into the replicated pool (transport={args.transport}: "
                                    f"{'copy-engine pulls over symmetric-memory peer mappings' if args.transport == 'p2p' else 'NCCL'}, gather={args.gather})" if world > 1 else "")},
        "clocks": clk,
        "e2e": {"value": e2e_val, "unit": "transitions/s", "h2d_bytes_per_step": int(start_host.numel() * 4),
                "d2h_bytes_per_step": int(rec * per_rollout), "steps": e2e_steps, "host_cpus_bound": numa},
        "gpu_launches": int((kernels_per_sampler_step * (w["T"] - 1) + (3 if gather is not None else 0)) * args.steps),
        "roofline": {"bound": "tensor", "kernel": fwd_kernel, "achieved": achieved_tf, "peak": peak,
                     "unit": "TFLOP/s", "frac": achieved_tf / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": f"{peaks['source']} bf16 sustained (burst {peaks['bf16_burst']})",
                     "flops_per_row": F, "launches_per_rollout": launches_fwd,
                     "kernel_ms_per_rollout": fwd_ms_per_rollout,
                     "kernel_share_of_step": fwd_ms_per_rollout / (ms / args.steps)},
        "cpu_baseline": cpu,
    }
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override start states per GPU")
    ap.add_argument("--cpu-rows", type=int, default=0, help="start states of the CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--gather", default="full", choices=["full", "pack", "none"],
                    help="N>1 diagnosis: full = pack + NCCL all-gather (the measured path), pack = packing only, none")
    ap.add_argument("--transport", default="p2p", choices=["nccl", "p2p"],
                    help="N>1: NCCL all-gather kernels, or copy-engine pulls over symmetric-memory peer mappings")
    ap.add_argument("--nccl-max-ctas", type=int, default=0, help="N>1: NCCL max_ctas of the gather group (0 = default)")
    ap.add_argument("--no-graph", action="store_true", help="launch kernels eagerly instead of replaying a CUDA graph")
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, w, args.workload)
    else:
        run_gpu(args, w, args.workload)


if __name__ == "__main__":
    main()
