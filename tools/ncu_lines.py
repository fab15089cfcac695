"""Executed warp instructions per CUDA source line of one kernel in an ncu --set full --import-source on report.
  python tools/ncu_lines.py gpurun_out/x.ncu-rep <kernel regex> <mangled-name substring> [top N]
The SASS page of the report is matched, instruction by instruction, against nvdisasm -g output of the built library."""
import collections, csv, glob, os, re, subprocess, sys, tempfile
rep, kre, mangled = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# several launches may match: take the first kernel block
hdr_idx = [i for i, r in enumerate(rows) if "Instructions Executed" in r]
h = rows[hdr_idx[0]]
end = hdr_idx[1] - 1 if len(hdr_idx) > 1 else len(rows)
ia, ii = h.index("Source"), h.index("Instructions Executed")
data = [(r[ia].strip(), int(r[ii])) for r in rows[hdr_idx[0] + 1:end] if len(r) > ii and r[ii].isdigit()]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(root, "constrainedmbpo_b200", "libcmbpo_b200.so")], cwd=tmp, capture_output=True)
ins = None
for f in glob.glob(os.path.join(tmp, "*.cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", f], capture_output=True, text=True).stdout
    if mangled not in dis:
        continue
    lines = dis.splitlines()
    start = [i for i, l in enumerate(lines) if ".text." in l and mangled in l and l.startswith("//---")][0]
    cur, ins = None, []
    for l in lines[start + 1:]:
        if l.startswith("//-----") and ".text." in l:
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+.*?;", l):
            ins.append(cur)
    break
assert ins is not None and len(ins) == len(data), (None if ins is None else len(ins), len(data))
agg = collections.Counter()
for (s, n), cur in zip(data, ins):
    agg[cur] += n
tot = sum(agg.values())
print(f"# {kre}: {tot} warp instructions, {len(data)} SASS instructions")
cache = {}
for k, v in agg.most_common(top):
    txt = ""
    if k:
        if k[0] not in cache:
            try:
                cache[k[0]] = open(k[0]).read().splitlines()
            except OSError:
                cache[k[0]] = []
        src = cache[k[0]]
        txt = src[k[1] - 1].strip()[:120] if k[1] - 1 < len(src) else ""
    print(f"{100 * v / tot:5.1f}%  {os.path.basename(k[0]) if k else '?'}:{k[1] if k else 0:<5d} {txt}")
