O=gpurun_out/r2exp21; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 600 python -m pytest tests/test_gpu_dropout_sites.py -q -m gpu -s > $O/pytest.log 2>&1; grep "^fp32\|^bf16" $O/pytest.log | cut -c1-900; tail -6 $O/pytest.log | cut -c1-250
