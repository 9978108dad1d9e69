O=gpurun_out/attn; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_attn_bf16.py -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -3 $O/pytest.log | cut -c1-220
for m in 1 2; do
echo "pair mode $m"
ATTN_PAIR=$m ATTN_SHAPES=800x800,300x300,300x800,800x300 timeout 300 python tools/attn_sweep.py --quick 2> $O/quick.err | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(d['Sq'], d['Sk'], d['mask'], round(d['us'],1), 'us', round(d['tflops'],1), 'TF', round(d['frac'],3))
"
done
