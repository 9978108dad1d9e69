O=gpurun_out/attn; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_attn_bf16.py -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -12 $O/pytest.log
