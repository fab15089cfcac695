"""Turns the raw ncu outputs of tools/gpu_evidence.sh (gpurun_out/<tag>_*) into the small text
summaries committed under profiles/.

  python tools/summarize_profiles.py r01
"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("cmbpo::", "")
    return name[:70]


def launch_list():
    src = os.path.join(G, f"{tag}_launches.csv")
    if not os.path.exists(src):
        return
    rows = [r for r in csv.reader(open(src)) if len(r) >= 15 and r[0].isdigit()]
    tot = collections.defaultdict(float)
    cnt = collections.Counter()
    ours = 0.0
    for r in rows:
        k = short(r[4])
        ns = float(r[14])
        tot[k] += ns
        cnt[k] += 1
    total = sum(tot.values())
    with open(os.path.join(P, f"{tag}_launches_summary.txt"), "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none -c 600: python bench.py --steps 2 --warmup 3 --no-cpu --no-graph\n")
        f.write(f"# first {len(rows)} launches (cold-cache, serialised: use the SHARES, not the absolute times)\n")
        f.write(f"{'kernel':72s} {'launches':>8s} {'total_us':>10s} {'avg_us':>9s} {'share':>7s}\n")
        for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
            f.write(f"{k:72s} {cnt[k]:8d} {v / 1e3:10.1f} {v / 1e3 / cnt[k]:9.2f} {100 * v / total:6.1f}%\n")
    # the raw list, trimmed to (id, kernel, grid, block, ns)
    with open(os.path.join(P, f"{tag}_launches.csv"), "w") as f:
        f.write("id,kernel,block,grid,gpu_time_ns\n")
        for r in rows:
            f.write(f"{r[0]},{short(r[4])},\"{r[7]}\",\"{r[8]}\",{r[14]}\n")


KEYS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__cycles_elapsed.avg.per_second",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__cluster_size",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
]


def to_bytes(val, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(val.replace(",", "")) * mult.get(unit, 1)


def full_capture(stem, title, traffic_rows=None):
    rep = os.path.join(G, f"{tag}_{stem}.ncu-rep")
    if not os.path.exists(rep):
        return
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    rows = [r for r in rows if len(r) > 20]
    hdr, units = rows[0], rows[1]
    with open(os.path.join(P, f"{tag}_{stem}_ncu_summary.txt"), "w") as f:
        f.write(f"# {title}\n# ncu --set full --clock-control none --import-source on (one launch); values per launch\n")
        for vals in rows[2:]:
            d = dict(zip(hdr, zip(units, vals)))
            if traffic_rows and "dram__bytes_read.sum" in d:
                rd, wr = to_bytes(d["dram__bytes_read.sum"][1], d["dram__bytes_read.sum"][0]), to_bytes(d["dram__bytes_write.sum"][1], d["dram__bytes_write.sum"][0])
                json.dump({"kernel": short(d.get("Kernel Name", ("", "?"))[1]), "rows": traffic_rows, "dram_bytes_read": rd,
                           "dram_bytes_write": wr, "bytes_per_row": (rd + wr) / traffic_rows,
                           "source": f"profiles/{tag}_{stem}_ncu_summary.txt (ncu --set full, one launch)"},
                          open(os.path.join(P, f"{tag}_{stem}_traffic.json"), "w"))
            f.write(f"kernel: {short(d.get('Kernel Name', ('', '?'))[1])}  grid {d.get('Grid Size', ('', '?'))[1]} block {d.get('Block Size', ('', '?'))[1]}\n")
            for k in KEYS:
                if k in d:
                    f.write(f"  {k:78s} {d[k][1]:>18s} {d[k][0]}\n")
            for h in hdr:
                if h.startswith("smsp__average_warp") and h.endswith("_per_issue_active.ratio") or h.startswith("smsp__average_warps_issue_stalled"):
                    try:
                        if float(d[h][1]) >= 0.3:
                            f.write(f"  {h:78s} {d[h][1]:>18s}\n")
                    except ValueError:
                        pass


def bench_lines():
    for stem in ("bench", "bench_reference", "bench_halfcheetah", "bench_step1M", "bench_wide512", "bench_step1M_4x512",
                 "bench_step256k_4x1024", "bench_iv_gae", "bench_n2", "bench_n2_nccl"):
        src = os.path.join(G, f"{tag}_{stem}.json")
        if os.path.exists(src) and os.path.getsize(src):
            line = open(src).read().strip().splitlines()[-1]
            try:
                json.loads(line)
            except ValueError:
                continue
            open(os.path.join(P, f"{tag}_{stem}.json"), "w").write(line + "\n")
    for stem in ("fwd_bench.log", "pytest_gpu.log", "sweep.log", "trace_report.txt", "gather_check.log", "pack_bench.log"):
        src = os.path.join(G, f"{tag}_{stem}")
        if os.path.exists(src):
            open(os.path.join(P, f"{tag}_{stem}"), "w").write(open(src).read())


launch_list()
full_capture("fwd_1M", "ensemble_fwd_kernel, antsafe 4x200, 1,000,000 rows (tools/fwd_bench.py 1000000)", traffic_rows=1_000_000)
full_capture("post_1M", "step_post_kernel, antsafe, 1,000,000 rows (tools/fwd_bench.py 1000000)")
full_capture("wide_512", "wide_layer_kernel, antsafe 4x512, 262,144 rows (one layer launch)")
full_capture("post3d", "step_post3d_kernel, humanoidsafe iv_gae, 125,000 rows x 7 members")
full_capture("pack", "k_pack_write, default workload pool (440k records x 72 columns)")
bench_lines()
print(sorted(os.listdir(P)))
