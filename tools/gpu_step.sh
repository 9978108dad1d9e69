O=gpurun_out/step; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_agent.py -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
tail -4 $O/pytest.log
