"""Prints the in-kernel timeline of the fused wide kernel written by a -DCMBPO_TRACE build (`make trace`):
  CMBPO_LIB=build/libcmbpo_b200_trace.so CMBPO_TRACE_OUT=gpurun_out/fw_trace.bin python tools/wide_bench.py --widths 512 --rows 100000 --layers 4 --reps 1
  python tools/fw_trace_report.py gpurun_out/fw_trace.bin [first event] [last event]
Tags.  MMA issuer (cluster 0): 10000 + 1000 l + 100 c top of N chunk c of layer l | 15000 + 1000 l + 100 c + g block group g
of the activation tile ready (A_RDY), its K blocks are issued next | 20000 + 1000 l + 100 c chunk issued.  Epilogue warps (leader CTA, quadrant 0, part 0 =
epi0 / part 3 = epi3): 30000 + l pass start | 31000 + l ACC0_FULL seen | 35000 + 10 l + g A_FREE[g] seen | 36000 + 10 l + g
group g arrived | 32000 + l early part done | 33000 + l LAYER_DONE seen | 34000 + l pass done (head: accumulator drained) |
37000 + l head stores done."""
import sys
import numpy as np

SLOTS = 2048
raw = np.fromfile(sys.argv[1], dtype=np.int64)
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3]) if len(sys.argv) > 3 else 400
roles = {}
for r, name in enumerate(["mma", "epi0", "epi3"]):
    a = raw[r * 2 * SLOTS:(r + 1) * 2 * SLOTS].reshape(-1, 2)
    roles[name] = a[a[:, 0] != 0]
t0 = min(a[0, 0] for a in roles.values() if len(a))
ev = sorted([(int(t - t0), name, int(tag)) for name, a in roles.items() for t, tag in a])
prev = {}
for t, name, tag in ev[lo:hi]:
    d = t - prev.get(name, t)
    prev[name] = t
    pad = {"mma": "", "epi0": " " * 28, "epi3": " " * 56}[name]
    print(f"{t:9d} {pad}{name} {tag:6d} (+{d})")
