#!/usr/bin/env python
"""Condensed region view of one kernel in an ncu report: consecutive SASS instructions with the same
execution count, with their share of issued instructions and average active lanes.
  python tools/ncu_regions.py gpurun_out/prof_X.ncu-rep <kernel-substring> [min_pct]"""
import csv, os, subprocess, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ncu_by_line import parse_dis
rep, ksub = sys.argv[1:3]
min_pct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.4
d = tempfile.mkdtemp()
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
subprocess.run(f"cd {d} && cuobjdump -xelf all {root}/terminal-raytracer_b200/libtrt_b200.so >/dev/null && for c in trt_api trt_native; do nvdisasm -g -c $c.sm_100a.cubin; done > all.dis 2>&1; ncu -i {os.path.abspath(rep)} --page source --csv 2>/dev/null > src.csv", shell=True)
rows = list(csv.reader(open(d + '/src.csv')))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address'); hdr = rows[hi]; ci = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) >= len(hdr) and r[0].startswith('0x')]
dis = parse_dis(d + '/all.dis', ksub)
assert len(dis) == len(data), (len(dis), len(data))
tot = sum(float(r[ci['Instructions Executed']]) for r in data)
reg = []
for (off, f, l, t), r in zip(dis, data):
    n = float(r[ci['Instructions Executed']]); th = float(r[ci['Thread Instructions Executed']])
    key = round(n / 1e5)
    tag = (f.replace('trt_', '').replace('.cuh', ''), l)
    if reg and reg[-1]['key'] == key:
        g = reg[-1]; g['n'] += n; g['th'] += th; g['cnt'] += 1; g['lines'].add(tag)
    else:
        reg.append(dict(key=key, n=n, th=th, cnt=1, start=off, lines={tag}))
print(f"total warp-inst {tot:.3e}; rows {len(data)}")
for g in reg:
    if 100 * g['n'] / tot < min_pct: continue
    ls = sorted(g['lines'])
    print(f"{g['start']:6x} exec {g['n']/g['cnt']/1e6:7.2f}M x{g['cnt']:4d} = {100*g['n']/tot:5.2f}%  lanes {g['th']/max(g['n'],1):5.1f}  {ls[0][0]}:{ls[0][1]}..{ls[-1][0]}:{ls[-1][1]}")
