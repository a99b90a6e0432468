#!/usr/bin/env python
"""Join an ncu SASS source-page CSV with nvdisasm -g line info and report executed warp instructions /
stall samples per CUDA source line.  usage: ncu_by_line.py src_sass.csv func.sass [topN]"""
import csv, re, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
H = rows[1]; si = H.index('Source'); ie = H.index('Instructions Executed'); ss = H.index('# Samples')
data = rows[2:]
lines = []
cur = ("?", 0)
for ln in open(sys.argv[2]):
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s*/\*[0-9a-f]{4,}\*/\s+\S', ln):
        lines.append(cur)
assert len(lines) == len(data), (len(lines), len(data))
tot = sum(int(r[ie]) for r in data); tots = sum(int(r[ss]) for r in data)
agg = collections.Counter(); aggs = collections.Counter(); fp = collections.Counter()
for loc, r in zip(lines, data):
    agg[loc] += int(r[ie]); aggs[loc] += int(r[ss])
    op = r[si].strip().split()
    op = op[1] if op[0].startswith('@') else op[0]
    if op.split('.')[0] in ('DFMA', 'DADD', 'DMUL', 'DSETP', 'DMNMX'):
        fp[loc] += int(r[ie])
src = {}
def srcline(f, n):
    import os
    for base in ('signals_b200/csrc/', ''):
        p = base + f
        if os.path.exists(p):
            if p not in src: src[p] = open(p).read().split('\n')
            return src[p][n - 1].strip()[:90] if 0 < n <= len(src[p]) else ''
    return ''
top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
print("total warp-instr %d, samples %d" % (tot, tots))
byfile = collections.Counter()
for (f, n), c in agg.items(): byfile[f] += c
print({k: round(v / tot * 100, 1) for k, v in byfile.items()})
print("%-16s %5s %7s %7s %7s  %s" % ("file", "line", "exec%", "fp64%", "stall%", "source"))
for (f, n), c in agg.most_common(top):
    print("%-16s %5d %7.2f %7.2f %7.2f  %s" % (f, n, c / tot * 100, fp[(f, n)] / tot * 100, aggs[(f, n)] / tots * 100, srcline(f, n)))
