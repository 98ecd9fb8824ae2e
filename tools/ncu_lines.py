#!/usr/bin/env python
"""Per-source-line instruction and stall-sample totals of one kernel in an .ncu-rep.

  python tools/ncu_lines.py <rep> <kernel regex> <cubin> <mangled function substring> [top]

Joins `ncu --page source --csv` (SASS rows, in address order) with `nvdisasm -g -c` line info
of the same cubin (cuobjdump -xelf all <.so>)."""
import csv, io, re, subprocess, sys, collections

rep, kre, cubin, fn = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
lines = txt.splitlines()
# first kernel only
starts = [i for i, l in enumerate(lines) if l.startswith('"Kernel Name"')]
end = starts[1] if len(starts) > 1 else len(lines)
rows = list(csv.reader(io.StringIO("\n".join(lines[starts[0] + 1:end]))))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
sass = [r for r in rows[1:] if len(r) == len(hdr)]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
cur = None
inside = False
linemap = []
for l in dis:
    if l.startswith(".text."):
        inside = fn in l
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        linemap.append(cur)
n = min(len(linemap), len(sass))
print("sass rows", len(sass), "disasm insts", len(linemap))
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for i in range(n):
    r = sass[i]
    v = (int(r[ix["Instructions Executed"]]), int(r[ix["Thread Instructions Executed"]]), int(r[ix["# Samples"]]))
    for k in range(3):
        agg[linemap[i]][k] += v[k]
        tot[k] += v[k]
print("total inst %d  thread-inst %d (%.1f thr/inst)  samples %d" % (tot[0], tot[1], tot[1] / max(tot[0], 1), tot[2]))
src_cache = {}
def src(f, ln):
    if f not in src_cache:
        import glob
        c = glob.glob("/root/repo/**/" + f, recursive=True)
        src_cache[f] = open(c[0]).read().splitlines() if c else []
    s = src_cache[f]
    return s[ln - 1].strip()[:100] if 0 < ln <= len(s) else ""
for key, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
    f, ln = key if key else ("?", 0)
    print("%5.1f%% smp %5.1f%% inst thr=%4.1f  %s:%d  %s" % (100.0 * v[2] / max(tot[2], 1), 100.0 * v[0] / max(tot[0], 1),
                                                          v[1] / max(v[0], 1), f, ln, src(f, ln)))
