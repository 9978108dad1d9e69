O=gpurun_out/r2exp13; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_gemm_bf16.py tests/test_gpu_bf16_mode.py -x -q -m gpu > $O/pytest.log 2>&1; tail -5 $O/pytest.log
timeout 300 python tools/rnn_bench.py > $O/rnn_bench.txt 2> $O/rnn_bench.err; cat $O/rnn_bench.txt; tail -3 $O/rnn_bench.err
timeout 300 python tools/gemm_trace.py > $O/gemm_trace.txt 2> $O/gemm_trace.err; grep -A1 "^video\|^caption\|^audio" $O/gemm_trace.txt | cut -c1-200; tail -3 $O/gemm_trace.err
BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py > $O/trace.txt 2> $O/trace.err; head -3 $O/trace.txt; grep rnn_layer $O/trace.txt | head -3
BMHRL_RNN_EXCHANGE=0 timeout 300 python tools/trace_step.py > $O/trace_barrier.txt 2> $O/trace_barrier.err; head -3 $O/trace_barrier.txt
BMHRL_PDL=0 timeout 600 python tools/bench_gemm16.py > $O/bench_gemm16.jsonl 2> $O/err.log; python - <<'PY'
import json
for l in open('gpurun_out/r2exp13/bench_gemm16.jsonl'):
    d=json.loads(l); print(d['what'], d['M'],d['N'],d['K'], 'fwd', d['bf16_fwd_us'], 'fwd32', d['bf16_fwd32_us'], 'dx', d['bf16_dx_us'], 'dw', d['bf16_dw_us'], 'bn128', d['bf16_fwd_bn128_s1_us'], 'bn64', d['bf16_fwd_bn64_s1_us'])
PY
