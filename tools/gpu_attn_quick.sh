mkdir -p gpurun_out/attn
for g in 148 111 74 37; do
echo "grid cap $g"
BMHRL_ATTN_GRID=$g ATTN_SHAPES=800x800 timeout 300 python tools/attn_sweep.py --quick 2> gpurun_out/attn/quick.err | head -1 | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(d['Sq'], d['Sk'], d['mask'], round(d['us'],1), 'us', round(d['tflops'],1), 'TF')
"
done
