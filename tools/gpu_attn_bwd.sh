O=gpurun_out/attn_bwd; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_attn_bf16_bwd.py -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -25 $O/pytest.log | cut -c1-250
