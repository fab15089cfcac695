"""Counts of the Blackwell-specific SASS mnemonics per kernel (cuobjdump -sass of the sm_100a objects in build/):
UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk (TMA engine), UTCBAR = tcgen05.commit,
SYNCS = mbarrier ops, LDGSTS = cp.async, MUFU.* = special-function unit.   python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections, re, subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pat = re.compile(r"\b(UTC[A-Z]*MMA|UTCBAR|UTCCP|LDTM|STTM|UBLKCP|UTMALDG|UTMASTG|SYNCS|LDGSTS|HMMA|MUFU\.[A-Z0-9]+)")
print("# " + __doc__.strip().replace("\n", "\n# "))
for obj in ("ensemble_fwd", "fused_wide", "wide_fwd", "step_post", "rollout_ops", "gae_ops", "archive_ops", "actor_ops"):
    out = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "build", obj + ".o")], capture_output=True, text=True).stdout
    fn, counts = None, collections.defaultdict(collections.Counter)
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            continue
        m = pat.search(line)
        if m and fn:
            counts[fn][m.group(1)] += 1
    for fn, c in counts.items():
        print(f"{obj:13s} {fn[:70]:70s} " + "  ".join(f"{k}={v}" for k, v in sorted(c.items())))
