"""profiles/rNN_gemm_traffic.json from an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv`
launch list of one eager training step: DRAM bytes per GEMM launch (what bench.py reports as roofline.traffic).
usage: gemm_traffic_from_launches.py launches.csv out.json "<the ncu command>" """
import csv
import json
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
per = {}   # launch id -> {metric: value}, name
order = []
for x in csv.DictReader(lines):
    k = x["ID"]
    if k not in per:
        per[k] = {"name": x["Kernel Name"]}
        order.append(k)
    v = float(x["Metric Value"].replace(",", ""))
    u = x["Metric Unit"]
    m = x["Metric Name"]
    if m == "gpu__time_duration.sum":
        v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
    elif u.lower().startswith("k"):
        v *= 1e3
    elif u.lower().startswith("m"):
        v *= 1e6
    elif u.lower().startswith("g"):
        v *= 1e9
    per[k][m] = v
rows = [per[k] for k in order]
idx = [i for i, r in enumerate(rows) if "embed_posenc" in r["name"]]
step = rows[idx[-2]:idx[-1]] if len(idx) >= 2 else rows
g = [r for r in step if "gemm_bf16_tc" in r["name"] or "gemm_tf32_tc" in r["name"]]
out = {"command": sys.argv[3] if len(sys.argv) > 3 else "",
       "gemm_launches_in_step": len(g),
       "avg_dram_bytes_per_launch": sum(r.get("dram__bytes_read.sum", 0) + r.get("dram__bytes_write.sum", 0) for r in g) / max(1, len(g)),
       "avg_us_per_launch_under_ncu": sum(r.get("gpu__time_duration.sum", 0) for r in g) / max(1, len(g)),
       "note": "ncu serialises launches and flushes caches between them: every launch re-reads its operands from DRAM, so this is "
               "an upper bound of the traffic inside the graph-replayed step"}
json.dump(out, open(sys.argv[2], "w"))
print(out)
