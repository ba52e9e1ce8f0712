"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of the
last step (steps are delimited by the input-stage kernel).  Usage: python summarize_launches.py file.csv"""
import collections
import csv
import re
import sys


def short(n):
    n = n.replace("(anonymous namespace)::", "").replace("void ", "")
    n = re.sub(r"\(.*", "", n)
    m = re.match(r"([A-Za-z0-9_:]+)(<.*)?", n)
    base = m.group(1) if m else n
    targs = (m.group(2) or "")[:60] if m else ""
    if base.startswith("bn_") or base.startswith("input_") or base.startswith("ce_") or base.startswith("sgd_"):
        return base + targs
    return base[-64:]


def main(path, top=45):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = [(r["Kernel Name"], float(r["Metric Value"])) for r in csv.DictReader(lines)]
    idx = [i for i, (n, _) in enumerate(rows) if "input_resample_kernel" in n]
    last = rows[idx[-1] - 1:]
    tot = sum(t for _, t in last)
    print(f"{path}: {len(rows)} launches captured; last step = {len(last)} launches, {tot/1e6:.3f} ms (sum of kernel times)")
    agg = collections.defaultdict(lambda: [0, 0.0])
    ours = 0.0
    for n, t in last:
        k = short(n)
        agg[k][0] += 1
        agg[k][1] += t
        if "anonymous namespace" in n or "<unnamed>" in n:
            ours += t
    print(f"our kernels: {ours/1e6:.3f} ms = {100*ours/tot:.1f}% of the step")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{t/1e6:8.3f} ms {100*t/tot:5.1f}% x{c:4d}  {k}")


if __name__ == "__main__":
    main(sys.argv[1])
