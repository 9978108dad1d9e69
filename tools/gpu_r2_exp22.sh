O=gpurun_out/r2exp22; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 900 python -m pytest tests/test_gpu_bf16_mode.py tests/test_gpu_gemm_bf16.py tests/test_gpu_dropout_sites.py -x -q -m gpu > $O/pytest.log 2>&1; tail -5 $O/pytest.log | cut -c1-250
timeout 300 python tools/trace_step.py > $O/trace.txt 2> $O/trace.err; head -4 $O/trace.txt; grep "colsum16\|gemm_bf16_tc_kernel<64" $O/trace.txt | head -3; tail -3 $O/trace.err
