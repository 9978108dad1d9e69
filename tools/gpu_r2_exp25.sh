O=gpurun_out/r2exp25; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 900 python -m pytest tests/test_gpu_agent.py -x -q -m gpu > $O/pytest.log 2>&1; tail -5 $O/pytest.log | cut -c1-250
for w in worker manager; do timeout 600 python bench.py --workload $w --scorer cider --steps 30 --warmup 5 --no-cpu --no-eager > $O/bench_${w}_cider.json 2> $O/bench_${w}_cider.err; echo "$w scored exit $?"; tail -c 200 $O/bench_${w}_cider.err; done
python - <<'PY'
import json
for f in ("bench_worker_cider.json", "bench_manager_cider.json"):
    try:
        d = json.loads(open("gpurun_out/r2exp25/" + f).read().strip().splitlines()[-1])
        print(f, "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "host_ms", d.get("host_part_ms_last_step"))
    except Exception as e:
        print(f, "unreadable", e)
PY
