#!/usr/bin/env python
"""Development aid: profiles/traffic.json (the `roofline.traffic` figures bench.py quotes) from this round's ncu --set full captures.
usage: make_traffic.py key=report.ncu-rep[:suffix] ...   e.g.
    make_traffic.py "blob1k@512^3 (BASELINE configs[1])=gpurun_out/prof.ncu-rep" \
                    "blob1k@512^3 (BASELINE configs[1])=gpurun_out/prof_native.ncu-rep: native" \
                    "blob10k@1024^3 z-slab sharded (BASELINE configs[3]) rank 0 of 8=gpurun_out/prof_slab.ncu-rep"
Each kernel's dram__bytes_read.sum + dram__bytes_write.sum of its (last) captured launch, in bytes."""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def kernels(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    ri, wi = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    out = {}
    for r in data:
        m = re.search(r"(\w+)(<[^>]*>)?\(", r[ki])
        name = m.group(1) if m else r[ki]
        out[name] = int(float(r[ri]) * UNIT[units[ri]] + float(r[wi]) * UNIT[units[wi]])
    return out


def main():
    res = {}
    for arg in sys.argv[1:]:
        key, rest = arg.split("=", 1)
        rep, _, suffix = rest.partition(":")
        for k, v in kernels(rep).items():
            res.setdefault(key, {})[k + suffix] = v
    commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
    res["_source"] = ("ncu --set full --clock-control none captures of tools/prof_step.py at commit %s (summaries: profiles/r02*_ncu_full_*.csv): "
                      "dram__bytes_read.sum + dram__bytes_write.sum per launch; the N > 1 entry is rank 0's slab captured on one GPU" % commit)
    json.dump(res, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
