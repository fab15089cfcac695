"""Prints the debug timeline written by libcmbpo_b200.so when CMBPO_TRACE=<file> is set (cluster 0:
MMA issuer tags 1xx, epilogue warp 4 tags 2xx; times in SM cycles relative to the first event)."""
import sys
import numpy as np
a = np.fromfile(sys.argv[1], dtype=np.int64).reshape(4, 1024, 2)
ev = []
for role, name in ((0, "mma"), (1, "epi")):
    for t, tag in a[role]:
        if t:
            ev.append((int(t), name, int(tag)))
ev.sort()
t0 = ev[0][0]
limit = int(sys.argv[2]) if len(sys.argv) > 2 else 400
prev = {}
for t, name, tag in ev[:limit]:
    d = t - prev.get(name, t)
    prev[name] = t
    print(f"{t - t0:9d}  {name}  {tag:4d}  (+{d})")
