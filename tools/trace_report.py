"""Prints the in-kernel timeline written by a -DCMBPO_TRACE build (see Makefile `trace`).

  CMBPO_LIB=build/libcmbpo_b200_trace.so CMBPO_TRACE_OUT=gpurun_out/trace.bin python tools/fwd_bench.py 1000000 1
  python tools/trace_report.py gpurun_out/trace.bin
Tags: MMA issuer 100+l top of layer | 110+10s+l before A_READY wait | 130+10s+l operands ready |
150+10s+l MMAs + commit issued.  Epilogue warp 4: 200+10s+l before D_FULL wait | 220+10s+l D_FULL seen |
240+10s+l pass done (A_READY arrive issued).
"""
import sys
import numpy as np

SLOTS = 2048
raw = np.fromfile(sys.argv[1], dtype=np.int64)
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3]) if len(sys.argv) > 3 else 400
roles = {}
for r, name in enumerate(["mma", "epiX", "epiY", "mmaY"]):
    a = raw[r * 2 * SLOTS:(r + 1) * 2 * SLOTS].reshape(-1, 2)
    a = a[a[:, 0] != 0]
    roles[name] = a
t0 = min(a[0, 0] for a in roles.values() if len(a))
ev = sorted([(int(t - t0), name, int(tag)) for name, a in roles.items() for t, tag in a])
prev = {}
for t, name, tag in ev[lo:hi]:
    d = t - prev.get(name, t)
    prev[name] = t
    print(f"{t:9d} {name} {tag:4d}  (+{d})")
import collections
for name in ("epiX", "epiY"):
    e = roles[name]
    if not len(e):
        continue
    dur, wait, gap = collections.defaultdict(list), collections.defaultdict(list), collections.defaultdict(list)
    tw = td = tdone = None
    for t, tag in e:
        tag = int(tag)
        if 200 <= tag < 220:
            if tdone is not None: gap[tag - 200].append(t - tdone)
            tw = t
        elif 220 <= tag < 240:
            wait[tag - 220].append(t - tw); td = t
        elif tag >= 240:
            dur[tag - 240].append(t - td); tdone = t
    print(f"{name}: layer: median loop gap | D_FULL wait | epilogue duration (cycles)")
    for k in sorted(dur):
        med = lambda a: int(np.median(a)) if len(a) else 0
        print(f"  {k:3d}: gap {med(gap[k]):6d}  wait {med(wait[k]):6d}   epi {med(dur[k]):6d}   n={len(dur[k])}")
for m in (roles["mma"], roles["mmaY"]):
  if len(m):
      w, iss = collections.defaultdict(list), collections.defaultdict(list)
      last = {}
      for t, tag in m:
          tag = int(tag)
          if 110 <= tag < 130: last[tag - 110] = t
          elif 130 <= tag < 150: w[tag - 130].append(t - last.get(tag - 130, t)); last[("r", tag - 130)] = t
          elif tag >= 150: iss[tag - 150].append(t - last.get(("r", tag - 150), t))
      print("MMA issuer (key = 10*stream + layer): median wait for operands (A_READY/A0) | issue time")
      for k in sorted(w):
          print(f"  {k:3d}: wait {int(np.median(w[k])):6d}   issue {int(np.median(iss[k])):6d}")
