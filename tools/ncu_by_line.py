#!/usr/bin/env python
"""Join an `ncu --page source --csv` dump (per-SASS-instruction counters) with `nvdisasm -g` line
info of the same kernel and aggregate by source line.

  cuobjdump -xelf all terminal-raytracer_b200/libtrt_b200.so      (-> *.cubin)
  nvdisasm -g -c trt_api.sm_100a.cubin > all.dis
  ncu -i prof.ncu-rep --page source --csv > src.csv
  python tools/ncu_by_line.py src.csv all.dis <kernel-name-substring> [top_n]
"""
import csv
import re
import sys
from collections import defaultdict


def parse_dis(path, kernel_sub):
    """-> list of (offset, file, line, text) for the first .text section whose name contains kernel_sub"""
    out, active = [], False
    cur_file, cur_line = "?", 0
    sec = re.compile(r"\.section\s+\.text\.(\S+),")
    loc = re.compile(r'//## File "([^"]+)", line (\d+)')
    ins = re.compile(r"/\*([0-9a-f]{4,})\*/\s+(.*?);")
    for ln in open(path, errors="replace"):
        m = sec.search(ln)
        if m:
            if active:
                break
            active = kernel_sub in m.group(1)
            continue
        if not active:
            continue
        m = loc.search(ln)
        if m:
            cur_file, cur_line = m.group(1).split("/")[-1], int(m.group(2))
            continue
        m = ins.search(ln)
        if m:
            out.append((int(m.group(1), 16), cur_file, cur_line, m.group(2).strip()))
    return out


def main():
    src_csv, dis, ksub = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src_csv)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    ci = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[hdr_i + 1:] if len(r) >= len(hdr) and r[0].startswith("0x")]
    base = int(data[0][0], 16)
    dis_rows = parse_dis(dis, ksub)
    by_off = {o: (f, l, t) for o, f, l, t in dis_rows}
    agg = defaultdict(lambda: defaultdict(float))
    tot = defaultdict(float)
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    missing = 0
    for r in data:
        off = int(r[0], 16) - base
        f, l, _ = by_off.get(off, ("?", 0, ""))
        if f == "?":
            missing += 1
        key = (f, l)
        vals = {"inst": float(r[ci["Instructions Executed"]] or 0), "thr": float(r[ci["Thread Instructions Executed"]] or 0),
                "samples": float(r[ci["# Samples"]] or 0), "n_sass": 1.0}
        for s in stall_cols:
            vals[s] = float(r[ci[s]] or 0)
        for k, v in vals.items():
            agg[key][k] += v
            tot[k] += v
    print(f"kernel rows {len(data)}, disassembly rows {len(dis_rows)}, unmatched {missing}")
    print(f"total warp-inst {tot['inst']:.3e}  thread-inst {tot['thr']:.3e}  avg lanes {tot['thr'] / max(1, tot['inst']):.2f}  samples {tot['samples']:.0f}")
    top_st = sorted(stall_cols, key=lambda s: -tot[s])[:6]
    print("stall totals: " + ", ".join(f"{s[6:]} {100 * tot[s] / max(1, tot['samples']):.1f}%" for s in top_st))
    print(f"{'file:line':34s} {'inst%':>6s} {'smp%':>6s} {'lanes':>6s} {'sass':>5s}  top stalls")
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[:top]:
        st = sorted(stall_cols, key=lambda s: -v[s])[:3]
        sts = ", ".join(f"{s[6:]} {100 * v[s] / max(1, v['samples']):.0f}%" for s in st if v[s] > 0)
        print(f"{key[0] + ':' + str(key[1]):34s} {100 * v['inst'] / tot['inst']:6.2f} {100 * v['samples'] / max(1, tot['samples']):6.2f} "
              f"{v['thr'] / max(1, v['inst']):6.1f} {int(v['n_sass']):5d}  {sts}")


if __name__ == "__main__":
    main()
