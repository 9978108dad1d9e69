O=gpurun_out/attn; mkdir -p $O
ATTN_SHAPES=${ATTN_SHAPES:-800x800} timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_fwd_bf16 --launch-skip 2 --launch-count 1 -f -o $O/attn_bf16_cap python tools/attn_sweep.py --quick > $O/ncu_cap.log 2>&1
tail -3 $O/ncu_cap.log
