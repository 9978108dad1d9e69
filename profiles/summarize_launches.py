"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of ONE training step.  usage: summarize_launches.py launches.csv [title]"""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = []
for x in csv.DictReader(lines):
    if x.get("Metric Name") == "gpu__time_duration.sum":
        v = float(x["Metric Value"].replace(",", ""))
        u = x["Metric Unit"]
        v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
        rows.append((x["Kernel Name"], v))
# one training step = the launches between two consecutive embed_posenc_kernel launches (exactly one per forward pass;
# the staged optimiser launches adam_multi_kernel several times per step, so it no longer marks a boundary)
idx = [i for i, (n, v) in enumerate(rows) if "embed_posenc" in n]
if len(idx) < 2:
    idx = [i for i, (n, v) in enumerate(rows) if "adam_multi" in n]
step = rows[idx[-2]:idx[-1]]
tot = sum(v for _, v in step)
print("#", " ".join(sys.argv[2:]))
print("# per-launch times are cold-cache + serialised (ncu replay): compare SHARES, not absolutes")
print(f"# launches in step: {len(step)}  sum of kernel durations: {tot:.1f} us")
agg = collections.defaultdict(lambda: [0, 0.0])
for n, v in step:
    n = re.sub(r"\(.*", "", n)[:80]
    agg[n][0] += 1
    agg[n][1] += v
print(f"{'total_us':>10} {'share':>6} {'count':>5} {'avg_us':>7}  kernel")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{t:10.1f} {100 * t / tot:5.1f}% {c:5d} {t / c:7.1f}  {n}")
