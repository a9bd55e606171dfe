"""Static SASS opcode histogram of the kernels whose name contains one of the given substrings (cuobjdump -sass of the built
library): how many instructions of each opcode family the compiled kernel CONTAINS (not executed counts — those are in the ncu
reports). usage: python tools/sass_hist.py <substr> [...] > profiles/r2_sass_opcodes.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "pebblesdr_b200", "libpebblecuda.so")], stdout=subprocess.PIPE, text=True).stdout
name, hist = None, None
kernels = []
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = m.group(1); hist = collections.Counter(); kernels.append((name, hist)); continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)(\.[A-Z0-9_.]+)?", line)
    if m and hist is not None:
        hist[m.group(1)] += 1
def demangle(n):
    return subprocess.run(["c++filt", n], stdout=subprocess.PIPE, text=True).stdout.strip()
for want in sys.argv[1:]:
    for n, h in kernels:
        d = demangle(n)
        if want in d and sum(h.values()) > 0:
            tot = sum(h.values())
            print("## %s\n   %d SASS instructions; DFMA+DMUL+DADD = %d (%.0f %%), LDS/STS = %d, LDG/STG = %d, SHFL = %d, BAR = %d, UTMA/UBLKCP (TMA) = %d, tcgen05 (UTC*) = %d" % (
                d[:150], tot, h["DFMA"] + h["DMUL"] + h["DADD"], 100.0 * (h["DFMA"] + h["DMUL"] + h["DADD"]) / tot, h["LDS"] + h["STS"],
                h["LDG"] + h["STG"], h["SHFL"], h["BAR"], h["UTMALDG"] + h["UBLKCP"] + h["UTMASTG"], sum(v for k, v in h.items() if k.startswith("UTC"))))
            print("   " + ", ".join("%s %d" % kv for kv in h.most_common(18)))
            print()
