O=gpurun_out/sweep; mkdir -p $O
timeout 500 python tools/attn_sweep.py --bwd > $O/attn_sweep.jsonl 2> $O/sweep.err; echo "sweep exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/sweep/attn_sweep.jsonl"):
    d=json.loads(l); print(d['kernel'][:24], d['Sq'], d['Sk'], d['mask'], d['pass'], round(d['us'],1), 'us', round(d.get('tflops',0),1), 'TF', round(d.get('frac',0),3))
PY
tail -3 $O/sweep.err
