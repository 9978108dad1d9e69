set -x
O=gpurun_out/step; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -4 $O/pytest.log
timeout 400 python bench.py --no-cpu > $O/bench_warmstart.json 2> $O/bench_warmstart.err; echo "bench exit $?"
timeout 300 python tools/microbench.py gemm attn > $O/microbench.txt 2>&1
python - <<'PY'
import json
d=json.loads(open("gpurun_out/step/bench_warmstart.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "roof", d["roofline"]["achieved"], d["roofline"]["avg_launch_us"])
PY
grep -A1 "linear fwd" $O/microbench.txt | grep -v "^--" | paste - - | awk '{print $3,$4,$5,$6,$7,$8,$9,$10}'
grep "mha_core" -A1 $O/microbench.txt | head -12
