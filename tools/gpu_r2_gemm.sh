O=gpurun_out/r2gemm; mkdir -p $O
# rebuild on the box: the snapshot may have been taken while a local build was in progress
(make -C bmhrl_b200/csrc -j16 > gpurun_out/make.log 2>&1 || (tail -20 gpurun_out/make.log; exit 1)) && echo "build ok"
timeout 600 python -m pytest tests/test_gpu_gemm_bf16.py -m gpu -x -q > $O/pytest_gemm.log 2>&1; echo "gemm pytest exit $?"; tail -15 $O/pytest_gemm.log
timeout 600 python -m pytest tests/test_gpu_attn16.py -m gpu -q > $O/pytest_attn16.log 2>&1; echo "attn16 pytest exit $?"; tail -25 $O/pytest_attn16.log
timeout 600 python tools/bench_gemm16.py > $O/bench_gemm16.jsonl 2> $O/bench_gemm16.err; echo "bench_gemm16 exit $?"; tail -3 $O/bench_gemm16.err; cat $O/bench_gemm16.jsonl
bash tools/gpu_r2_base.sh
