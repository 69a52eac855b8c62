"""Development aid: per-SOURCE-LINE totals of an ncu capture. `ncu --page source --csv` prints SASS rows only; this joins them with the line
table of the cubin (nvdisasm -g) by instruction offset.
    python tools/ncu_lines.py <report.ncu-rep> <kernel regex> <cubin> [top N] [regex of the mangled name]"""
import csv, io, re, subprocess, sys, collections

rep, kern, cubin = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
mangled_re = sys.argv[5] if len(sys.argv) > 5 else kern  # the cubin's section names are mangled: give their regex if it differs
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
name = rows[hi - 1][1] if hi else kern
h = rows[hi]
cs, ci = h.index("# Samples"), h.index("Instructions Executed")
body = []
for r in rows[hi + 1:]:
    if r and r[0] == "Address":
        break  # next kernel instance
    if len(r) > ci and r[0].startswith("0x"):
        body.append((int(r[0], 16), r[1].strip(), int(r[cs] or 0), int(r[ci] or 0)))
base = body[0][0]
dis = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.splitlines()
mangled = None
fn = re.search(r"(\w+)\(", name.split("::")[-1] if "::" in name else name)
short = kern.split("|")[0]
start = [i for i, l in enumerate(dis) if l.startswith(".text.") and re.search(mangled_re, l)][0]
line_of, cur = {}, None
for l in dis[start + 1:]:
    if l.startswith("//-----") or l.startswith("\t.section"):
        if line_of:
            break
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", l)
    if m:
        line_of[int(m.group(1), 16)] = cur
samp, inst = collections.Counter(), collections.Counter()
for addr, sass, s, n in body:
    k = line_of.get(addr - base, ("?", 0))
    samp[k] += s
    inst[k] += n
ts, ti = sum(samp.values()), sum(inst.values())
print("%s: %d samples, %d warp instructions" % (kern, ts, ti))
src = {}
for (f, ln), s in samp.most_common(top):
    if f not in src:
        try:
            src[f] = open("parsip_b200/csrc/" + f).read().splitlines()
        except OSError:
            src[f] = []
    text = src[f][ln - 1].strip()[:110] if 0 < ln <= len(src[f]) else ""
    print("%5.1f%% samp %5.1f%% inst  %s:%d  %s" % (100.0 * s / ts, 100.0 * inst[(f, ln)] / ti, f, ln, text))
