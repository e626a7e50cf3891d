#!/usr/bin/env python
"""Where do warps WAIT?  Warp-stall samples of one kernel (ncu --page source --csv of a --set full capture) attributed to
CUDA source lines (nvdisasm --print-line-info of the matching object) and broken down by stall reason.
usage: stall_lines.py <ncu_source.csv> <object.o> <kernel-name-substring> <source.cu> [top]"""
import collections, csv, os, re, subprocess, sys, tempfile
src_csv, obj, kname, cu = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 30
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.split("\n")
i0 = [i for i, l in enumerate(sass) if l.startswith(".text.") and kname in l][0]
cur, seq = None, []
for l in sass[i0 + 1:]:
    if (l.startswith(".text.") or l.startswith(".section")) and seq:
        break
    m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+.*?;", l):
        seq.append(cur)
rows = list(csv.reader(open(src_csv)))
h = rows[1]
si = h.index("# Samples"); ii = h.index("Instructions Executed")
reasons = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
body = [r for r in rows[2:] if len(r) > si and r[si].strip().isdigit()]
n = min(len(body), len(seq))
by_line, by_reason, line_reason = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
inst = collections.Counter()
for r, k in zip(body[:n], seq[:n]):
    s = int(r[si]); by_line[k] += s; inst[k] += int(r[ii]) if r[ii].strip().isdigit() else 0
    for i, c in reasons:
        v = int(r[i]) if r[i].strip().isdigit() else 0
        by_reason[c] += v; line_reason[k][c] += v
tot = sum(by_line.values())
print("total stall samples %d over %d instructions" % (tot, n))
print("by reason: " + ", ".join("%s %.1f%%" % (c.replace("stall_", ""), 100 * v / max(1, sum(by_reason.values()))) for c, v in by_reason.most_common(8)))
text = open(cu).read().split("\n")
for (f, ln), c in by_line.most_common(top):
    t = text[ln - 1].strip()[:80] if f == os.path.basename(cu) and ln - 1 < len(text) else ""
    rs = ", ".join("%s %d%%" % (k.replace("stall_", ""), round(100 * v / max(1, c))) for k, v in line_reason[(f, ln)].most_common(2))
    print("%5.1f%% samples  %5.1f%% instr  %-18s %4d  [%s]  %s" % (100 * c / tot, 100 * inst[(f, ln)] / max(1, sum(inst.values())), f, ln, rs, t))
