O=gpurun_out/r2final2; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "default bench exit $?"; tail -c 200 $O/bench_default.err
for w in worker manager; do timeout 600 python bench.py --workload $w --scorer cider --steps 30 --warmup 5 --no-cpu --no-eager > $O/bench_${w}_cider.json 2> $O/bench_${w}_cider.err; echo "$w scored exit $?"; tail -c 200 $O/bench_${w}_cider.err; done
python - <<'PY'
import json
for f in ("bench_default.json", "bench_worker_cider.json", "bench_manager_cider.json"):
    try:
        d = json.loads(open("gpurun_out/r2final2/" + f).read().strip().splitlines()[-1])
        print(f, "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "steps", d["steps"], "host_ms", d.get("host_part_ms_last_step"))
    except Exception as e:
        print(f, "unreadable", e)
PY
