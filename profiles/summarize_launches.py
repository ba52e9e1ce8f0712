"""Summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
--clock-control none --csv --log-file X.csv python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-e2e --no-graph`).

Writes, next to this script: <round>_launch_summary.txt (per-kernel totals of the LAST step = region B of bench.py),
<round>_launches_last_step.csv (one row per launch) and traffic.json (mean DRAM bytes per BatchNorm call, read by
bench.py).  Steps are delimited by the input-stage coefficient kernel.
Usage: python summarize_launches.py X.csv [round, default r02]"""
import collections
import csv
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
UNIT = {"ns": 1.0, "us": 1e3, "ms": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def short(n):
    n = n.replace("(anonymous namespace)::", "").replace("<unnamed>::", "").replace("void ", "")
    n = re.sub(r"\(.*", "", n)
    m = re.match(r"([A-Za-z0-9_:]+)(<.*)?", n)
    base = m.group(1) if m else n
    targs = (m.group(2) or "")[:40] if m else ""
    if re.match(r"(bn_|input_|ce_|sgd_|maxpool_|pack_)", base):
        return base + targs
    return base[-64:]


def main(path, rnd="r02"):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    launches = collections.OrderedDict()
    for r in csv.DictReader(lines):
        d = launches.setdefault(int(r["ID"]), {"id": int(r["ID"]), "kernel": r["Kernel Name"], "grid": r["Grid Size"],
                                               "block": r["Block Size"]})
        try:
            v = float(r["Metric Value"].replace(",", "")) * UNIT.get(r["Metric Unit"], 1.0)
        except ValueError:
            v = 0.0
        if v != v:
            v = 0.0   # ncu prints nan for a metric it could not collect on a launch
        d[r["Metric Name"]] = v
    rows = list(launches.values())
    starts = [i for i, d in enumerate(rows) if "input_coeff_kernel" in d["kernel"]]
    last = rows[starts[-1]:]
    T, R, W = "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum"
    tot = sum(d[T] for d in last)
    own = [d for d in last if "<unnamed>" in d["kernel"] or "anonymous namespace" in d["kernel"]]
    out = [f"ncu --metrics {T},{R},{W} --clock-control none, command: python bench.py --steps 2 --warmup 2 "
           f"--no-cpu-baseline --no-e2e --no-graph (eager launches: ncu cannot attach to kernels inside a stream capture)",
           f"{len(rows)} launches captured; the last step (region B of bench.py: side stream off) = {len(last)} launches, "
           f"{tot / 1e6:.3f} ms summed kernel time (cold cache, serialised)",
           f"owned kernels: {sum(d[T] for d in own) / 1e6:.3f} ms = {100 * sum(d[T] for d in own) / tot:.1f} % of the "
           f"summed kernel time"]
    fam = {"bn_fwd": ("bn_stats", "bn_fwd_finalize", "bn_apply"), "bn_bwd": ("bn_bwd_",)}   # (calls = finalize launches)
    traffic = {}
    for name, pre in fam.items():
        sel = [d for d in own if any(p in d["kernel"] for p in pre)]
        t, b = sum(d[T] for d in sel), sum(d.get(R, 0) + d.get(W, 0) for d in sel)
        calls = sum(1 for d in sel if "finalize" in d["kernel"])
        out.append(f"family {name}: {t / 1e6:.3f} ms = {100 * t / tot:.1f} % of the step, DRAM traffic {b / 1e9:.2f} GB "
                   f"over {calls} calls")
        traffic[name] = b / max(calls, 1)
    out.append("")
    out.append("   time      share  launches  DRAM traffic   DRAM rate   kernel")
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for d in last:
        a = agg[short(d["kernel"])]
        a[0] += 1
        a[1] += d[T]
        a[2] += d.get(R, 0) + d.get(W, 0)
    for k, (c, t, b) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:48]:
        out.append(f"{t / 1e6:7.3f} ms {100 * t / tot:5.1f}%  x{c:3d}  {b / 1e9:8.2f} GB  {b / t / 1e3 if t else 0:5.2f} TB/s  {k}")
    open(os.path.join(HERE, f"{rnd}_launch_summary.txt"), "w").write("\n".join(out) + "\n")
    with open(os.path.join(HERE, f"{rnd}_launches_last_step.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel", "grid", "block", "time_ns", "dram_read_MB", "dram_write_MB"])
        for d in last:
            w.writerow([d["id"], d["kernel"][:160], d["grid"], d["block"], int(d[T]), round(d.get(R, 0) / 1e6, 3),
                        round(d.get(W, 0) / 1e6, 3)])
    traffic["unit"] = ("DRAM bytes (read+write) per BatchNorm call (its 2-3 launches), mean over the sites of one "
                       "ResNet-50 step, N=256 bf16")
    traffic["source"] = f"profiles/{rnd}_launches_last_step.csv (ncu dram__bytes_read.sum + dram__bytes_write.sum)"
    json.dump(traffic, open(os.path.join(HERE, "traffic.json"), "w"), indent=1)
    print("\n".join(out[:12]))


if __name__ == "__main__":
    main(*sys.argv[1:3])
