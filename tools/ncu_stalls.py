"""Summarise an ncu --set full report: per kernel duration, DRAM bytes, issue activity, top stall reasons, occupancy.
  python tools/ncu_stalls.py gpurun_out/x.ncu-rep [> profiles/r02_x_ncu_summary.txt]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, units = rows[0], rows[1]
keys = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct"]
for r in rows[2:]:
    d = dict(zip(h, r))
    print(f"kernel: {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
    for k in keys:
        if k in d:
            print(f"  {k:75s} {d[k]:>16s} {units[h.index(k)]}")
    st = sorted(((float(v.replace(',', '')), k) for k, v in d.items() if "issue_stalled" in k and "per_issue_active" in k and v), reverse=True)
    for v, k in st[:7]:
        print(f"  stall {k.split('issue_stalled_')[1].split('_per_issue')[0]:30s} {v:8.2f}")
