"""Development aid: join ncu's per-SASS-instruction counts (`ncu -i rep --page source --csv`) with nvdisasm line
info of the built library, and print executed warp-instructions per source line.
usage: sass_lines.py <src.csv> <mangled-kernel-substring> [top]"""
import csv, re, subprocess, sys, os, tempfile
from collections import defaultdict
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
srccsv, kname = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "pebblesdr_b200", "libpebblecuda.so")], cwd=tmp,
               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
instr = None
for f in os.listdir(tmp):
    if not f.endswith(".cubin"):
        continue
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                         text=True).stdout.split("\n")
    start = end = None
    for i, l in enumerate(txt):
        if l.startswith(".text.") and kname in l:
            start = i
        elif start is not None and l.startswith("\t.section"):
            end = i
            break
    if start is None:
        continue
    cur, instr = None, []
    for l in txt[start:end]:
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            instr.append((m.group(2).strip(), cur))
    break
assert instr, "kernel not found"
rows = list(csv.reader(open(srccsv)))
hdr, data = rows[1], rows[2:]
ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
assert len(data) == len(instr), (len(data), len(instr))
agg, tot, totsmp = defaultdict(lambda: [0, 0]), 0, 0
for d, (txt, cur) in zip(data, instr):
    n, s = int(d[ie]), int(d[ss])
    tot += n; totsmp += s
    k = cur or ("?", 0)
    agg[k][0] += n; agg[k][1] += s
cache = {}
def line(f, ln):
    if f not in cache:
        p = os.path.join(ROOT, "pebblesdr_b200", "csrc", f)
        cache[f] = open(p).read().split("\n") if os.path.exists(p) else []
    return cache[f][ln - 1].strip()[:95] if 0 < ln <= len(cache[f]) else ""
print("total warp-instr", tot, "samples", totsmp)
for (f, ln), (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% smp%5.1f%% %-13s:%4d %s" % (100 * n / tot, 100 * s / max(totsmp, 1), f, ln, line(f, ln)))
