#!/usr/bin/env python
"""Development aid: aggregate an `ncu --page source --csv --print-source cuda,sass` dump by CUDA source line.
usage: ncu_lines.py dump.csv [top]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None
agg = collections.OrderedDict()
hdr = None
tot_inst = tot_samp = 0
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    d = dict(zip(hdr, r))
    # each row = one SASS instruction attributed to a source line
    try:
        inst = int(d["Instructions Executed"] or 0); samp = int(d["# Samples"] or 0)
    except ValueError:
        continue
    key = (cur_file, d["Line No"])
    a = agg.setdefault(key, [0, 0, 0, ""])
    a[0] += inst; a[1] += samp; a[2] += 1
    if not a[3]: a[3] = r[1].strip()[:110]
    tot_inst += inst; tot_samp += samp
print("total warp-instr %d, samples %d" % (tot_inst, tot_samp))
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%5.1f%% samp %5.1f%% inst %4d sass  %s:%s  %s" % (100.0 * a[1] / max(1, tot_samp), 100.0 * a[0] / max(1, tot_inst), a[2], f, ln, a[3]))
