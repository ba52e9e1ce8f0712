"""Attribute executed warp instructions of one kernel in an .ncu-rep to CUDA source lines.
usage: ncu_lines.py report.ncu-rep kernel-regex lib.so mangled-substring
(SASS order of `ncu --page source` == order of `nvdisasm -g`, whose `//## File ..., line N` markers give the line.)"""
import collections
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile

rep, kre, lib, sub = sys.argv[1:5]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if "Instructions Executed" in r)
ie, so, sm = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
first = rows.index(hdr) + 1
sass = []
for r in rows[first:]:
    if len(r) != len(hdr):
        break   # next kernel instance
    sass.append((r[so].strip(), int(r[ie]), int(r[sm] or 0)))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
lines = None
for f in glob.glob(tmp + "/*.cubin"):
    d = subprocess.run(["nvdisasm", "-g", "-c", f], capture_output=True, text=True).stdout
    if sub in d and len(os.path.basename(f).split("-")) == 1:
        lines = d.splitlines()
        break
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and sub in l and l.rstrip().endswith(":"))
cur, seq = None, []
for l in lines[start + 1:]:
    if l.startswith(".text.") or "//-----" in l:
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);", l)
    if m:
        seq.append(cur)
assert len(seq) == len(sass), (len(seq), len(sass))
agg, smp = collections.Counter(), collections.Counter()
for (txt, n, s), where in zip(sass, seq):
    agg[where] += n
    smp[where] += s
tot, stot = sum(agg.values()), sum(smp.values())
src = {}
print(f"total warp instructions {tot}, samples {stot}")
for where, n in sorted(agg.items(), key=lambda kv: -kv[1])[:int(sys.argv[5]) if len(sys.argv) > 5 else 40]:
    if where and where[0] not in src:
        p = [q for q in glob.glob("**/" + where[0], recursive=True)]
        src[where[0]] = open(p[0]).read().splitlines() if p else []
    text = src[where[0]][where[1] - 1].strip() if where and len(src[where[0]]) >= where[1] else ""
    print(f"{100 * n / tot:5.1f}% inst {100 * smp[where] / max(stot, 1):5.1f}% stall  {where[0] if where else '?'}:{where[1] if where else 0:<4d} {text[:110]}")
