O=gpurun_out/attn; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_attn_bf16.py -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -25 $O/pytest.log | cut -c1-220
ATTN_SHAPES=800x800,300x300,300x800,800x300 timeout 300 python tools/attn_sweep.py --quick > $O/quick.jsonl 2> $O/quick.err; echo "exit $?"; tail -3 $O/quick.err
python - <<'PY'
import json
for l in open("gpurun_out/attn/quick.jsonl"):
    d=json.loads(l); print(d["kernel"][:22], d["Sq"], d["Sk"], d["mask"], round(d["us"],1), "us", round(d["tflops"],1), "TF", round(d["frac"],3))
PY
