"""Static evidence per kernel of libb200_imagenet.so (no GPU needed): registers / shared memory from `cuobjdump -res-usage`
and counts of the SASS mnemonics that show how data moves (UBLKCP = cp.async.bulk, SYNCS = mbarrier, LDG/STG widths,
packed bf16 ops).  Usage: python tools/sass_evidence.py > profiles/r01_sass_evidence.txt"""
import collections
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(
    os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
    "imagent_distributed_training_pytorch_with_slurm_b200", "libb200_imagenet.so")
res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
usage = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = m.group(1)
        continue
    m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
    if m and cur:
        usage[cur] = (int(m.group(1)), int(m.group(2)))
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ["UBLKCP", "SYNCS", "LDG.E.128", "STG.E.128", "LDS.128", "LDS.64", "HSET2", "HMNMX2", "PRMT", "SHFL", "ATOM", "RED",
        "ACQBULK", "BAR.SYNC"]
hist, total, name = {}, {}, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        hist[name], total[name] = collections.Counter(), 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(.*?)\s*;", line)
    if m and name:
        total[name] += 1
        txt = m.group(1)
        for k in KEYS:
            if re.search(r"(^|\s|@!?U?P\d\s+)" + re.escape(k), txt):
                hist[name][k] += 1
demangle = subprocess.run(["c++filt"] + list(hist), capture_output=True, text=True).stdout.splitlines()
print("kernel | registers | static smem B | SASS instructions | mnemonic counts")
for n, d in sorted(zip(hist, demangle), key=lambda t: t[1]):
    d = re.sub(r"\(anonymous namespace\)::", "", d)
    d = re.sub(r"\(.*", "", d).replace("void ", "")
    r, s = usage.get(n, (0, 0))
    print(f"{d[:70]:70s} | {r:3d} | {s:6d} | {total[n]:5d} | " + ", ".join(f"{k} {c}" for k, c in hist[n].items()))
