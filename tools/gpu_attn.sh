set -x
O=gpurun_out/attn; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_attn_bf16.py -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -25 $O/pytest.log
timeout 300 python tools/attn_sweep.py > $O/attn_sweep.jsonl 2> $O/attn_sweep.err; echo "sweep exit $?"
cat $O/attn_sweep.jsonl | cut -c1-260
