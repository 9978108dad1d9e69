ATTN_SHAPES=${ATTN_SHAPES:-300x300,800x800} timeout 300 python tools/attn_sweep.py --quick > gpurun_out/attn/quick.jsonl 2> gpurun_out/attn/quick.err; echo "exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/attn/quick.jsonl"):
    d=json.loads(l); print(d["kernel"][:22], d["Sq"], d["Sk"], d["mask"], round(d["us"],1), "us", round(d["tflops"],1), "TF", round(d["frac"],3))
PY
