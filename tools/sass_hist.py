"""Opcode histogram per kernel of an object file / shared library (cuobjdump -sass), filter by substring."""
import collections
import re
import subprocess
import sys

out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
pat = sys.argv[2] if len(sys.argv) > 2 else ""
name, hist = None, None
res = {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        hist = res.setdefault(name, collections.Counter())
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(.*?)\s*/\*", line)
    if m and hist is not None:
        toks = m.group(1).split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        hist[op.split(".")[0].rstrip(";")] += 1
for k, h in res.items():
    if pat in k:
        print(k[:90], sum(h.values()))
        print("   ", ", ".join(f"{o} {c}" for o, c in h.most_common(18)))
