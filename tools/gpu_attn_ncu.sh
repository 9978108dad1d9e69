set -x
O=gpurun_out/attn; mkdir -p $O
ATTN_SHAPES=800x800,300x300 timeout 300 python tools/attn_sweep.py --quick > $O/sweep_quick.jsonl 2>&1 || exit 1
ATTN_SHAPES=800x800,300x300 timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_fwd_bf16 --launch-skip 2 --launch-count 1 -o $O/attn_bf16_800 python tools/attn_sweep.py --quick > $O/ncu800.log 2>&1
ATTN_SHAPES=300x300 timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_fwd_bf16 --launch-skip 2 --launch-count 1 -o $O/attn_bf16_300 python tools/attn_sweep.py --quick > $O/ncu300.log 2>&1
ls -la $O
