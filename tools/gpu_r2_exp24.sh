O=gpurun_out/r2exp24; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 600 python -m pytest tests/test_gpu_agent.py -x -q -m gpu -k "pipelined" > $O/pytest.log 2>&1; tail -3 $O/pytest.log | cut -c1-250
timeout 900 python bench.py --steps 30 --warmup 5 --no-eager > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -c 300 $O/bench.err
timeout 900 python bench.py --steps 30 --warmup 5 --no-eager --no-cpu --no-phases --sync-loss > $O/bench_sync.json 2> $O/bench_sync.err
python - <<'PY'
import json
for f in ("bench.json", "bench_sync.json"):
    d = json.loads(open("gpurun_out/r2exp24/" + f).read().strip().splitlines()[-1])
    print(f, "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_step"], 3), d["e2e"]["loss_read"][:40],
          {k: (round(v["ms_per_step"], 3), round(v["e2e"]["ms_per_step"], 3)) for k, v in d.get("phases", {}).items()})
PY
