O=gpurun_out/attn; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_attn_bf16.py -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -5 $O/pytest.log
timeout 300 python tools/attn_sweep.py > $O/attn_sweep.jsonl 2> $O/attn_sweep.err; echo "sweep exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/attn/attn_sweep.jsonl"):
    d=json.loads(l); print(d["kernel"][:22], d["Sq"], d["Sk"], d["mask"], round(d["us"],1), "us", round(d["tflops"],1), "TF", round(d["frac"],3))
PY
