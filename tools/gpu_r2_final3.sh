O=gpurun_out/r2final3; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest.log | cut -c1-200
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $O/pytest_noalloc_alone.log 2>&1; echo "no_alloc alone exit $?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "default bench exit $?"; tail -c 200 $O/bench_default.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2final3/bench_default.json").read().strip().splitlines()[-1])
print("value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "launches", d["launches_per_step"], "roof", round(d["roofline"]["frac"], 4), "hbm", round(d["step_hbm_roofline"]["frac"], 4), "cpu", round(d["cpu_baseline"]["value"], 1), {k: round(v["ms_per_step"], 3) for k, v in d["phases"].items()}, "eager", round(d["eager_module_swap"]["ms_per_step"], 1), d["clocks"])
PY
