"""Average DRAM traffic per GEMM launch of ONE training step from an ncu CSV
(`ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv`).
usage: gemm_traffic.py launches.csv "<command>" > r01_gemm_traffic.json"""
import collections, csv, json, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
per = collections.OrderedDict()
for x in csv.DictReader(lines):
    k = int(x["ID"])
    d = per.setdefault(k, {"name": x["Kernel Name"]})
    v = float(x["Metric Value"].replace(",", ""))
    u = x["Metric Unit"]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3}.get(u, 1)
    d[x["Metric Name"]] = v * scale
rows = list(per.values())
idx = [i for i, r in enumerate(rows) if "adam_multi" in r["name"]]
step = rows[idx[-2] + 1: idx[-1] + 1]
g = [r for r in step if "gemm_tf32_tc" in r["name"]]
tot = sum(r.get("dram__bytes_read.sum", 0) + r.get("dram__bytes_write.sum", 0) for r in g)
print(json.dumps({"command": sys.argv[2] if len(sys.argv) > 2 else "", "gemm_launches_in_step": len(g),
                  "avg_dram_bytes_per_launch": tot / max(1, len(g)),
                  "avg_us_per_launch_under_ncu": sum(r.get("gpu__time_duration.sum", 0) for r in g) / max(1, len(g)),
                  "note": "ncu serialises launches and flushes caches between them: every launch re-reads its operands "
                          "from DRAM, so this is an upper bound of the traffic inside the graph-replayed step"}))
