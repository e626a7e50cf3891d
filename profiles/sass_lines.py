#!/usr/bin/env python
"""Attribute executed warp instructions of one kernel (ncu --page source --csv) to CUDA source lines
(nvdisasm --print-line-info of the matching cubin).
usage: sass_lines.py <ncu_source.csv> <object.o> <kernel-name-substring> [top]"""
import collections, csv, os, re, subprocess, sys, tempfile
src_csv, obj, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp(dir=os.path.dirname(os.path.abspath(src_csv)))
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
h = rows[1]; ii = h.index("Instructions Executed")
ex = []
for r in rows[2:]:
    try:
        ex.append(int(r[ii]))
    except (ValueError, IndexError):
        pass
n = min(len(ex), len(seq)); ex, seq = ex[:n], seq[:n]   # (ncu may omit a trailing instruction)
agg = collections.Counter()
for k, n in zip(seq, ex):
    agg[k] += n
tot = sum(agg.values())
print("total warp instructions", tot)
cache = {}
for (f, ln), c in sorted(agg.items(), key=lambda x: -x[1])[:top]:
    text = ""
    for d in ("detectron-pytorch_b200/csrc", "include"):
        pth = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), d, f)
        if os.path.exists(pth):
            cache.setdefault(pth, open(pth).read().split("\n"))
            text = cache[pth][ln - 1].strip()[:100] if ln - 1 < len(cache[pth]) else ""
    print("%6.2f%% %-20s %4d  %s" % (100 * c / tot, f, ln, text))
