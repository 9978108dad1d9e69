O=gpurun_out/r2exp11; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 600 python -m pytest tests/test_gpu_ops.py -x -q -m gpu -k "rnn or critic" > $O/pytest_rnn.log 2>&1; tail -5 $O/pytest_rnn.log
timeout 300 python tools/rnn_bench.py > $O/rnn_bench.txt 2> $O/rnn_bench.err; cat $O/rnn_bench.txt
timeout 300 python tools/gemm_trace.py > $O/gemm_trace.txt 2> $O/gemm_trace.err; head -50 $O/gemm_trace.txt; tail -3 $O/gemm_trace.err
timeout 300 python tools/trace_step.py > $O/trace.txt 2> $O/trace.err; head -3 $O/trace.txt
BMHRL_RNN_EXCHANGE=0 timeout 300 python tools/trace_step.py > $O/trace_barrier.txt 2> $O/trace_barrier.err; head -3 $O/trace_barrier.txt
