O=gpurun_out/attn_bwd; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_attn_bf16_bwd.py -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -5 $O/pytest.log | cut -c1-250
ATTN_SHAPES=${ATTN_SHAPES:-300x300,800x800} timeout 300 python tools/attn_sweep.py --quick --bwd 2> $O/sweep.err | tee $O/sweep.jsonl | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(d['Sq'], d['Sk'], d['mask'], d['pass'], round(d['us'],1), 'us', round(d.get('tflops',0),1), 'TF', round(d.get('frac',0),3))
"
tail -3 $O/sweep.err
