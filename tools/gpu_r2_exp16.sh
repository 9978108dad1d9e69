O=gpurun_out/r2exp16; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 900 python -m pytest tests/test_gpu_gemm_bf16.py tests/test_gpu_bf16_mode.py -x -q -m gpu > $O/pytest.log 2>&1; tail -2 $O/pytest.log
timeout 300 python tools/gemm_trace.py > $O/gemm_trace.txt 2> $O/gemm_trace.err; grep -A2 "^video\|^caption\|^audio" $O/gemm_trace.txt | cut -c1-200; tail -3 $O/gemm_trace.err
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu --no-eager > $O/bench.json 2> $O/bench.err; cut -c1-700 $O/bench.json; tail -2 $O/bench.err
BMHRL_PDL=0 timeout 600 python tools/bench_gemm16.py > $O/bench_gemm16.jsonl 2> $O/err.log; python - <<'PY'
import json
for l in open('gpurun_out/r2exp16/bench_gemm16.jsonl'):
    d=json.loads(l); print(d['what'], d['M'],d['N'],d['K'], 'fwd', d['bf16_fwd_us'], 'fwd32', d['bf16_fwd32_us'], 'dx', d['bf16_dx_us'], 'dw', d['bf16_dw_us'], 'bn128', d['bf16_fwd_bn128_s1_us'], 'bn64', d['bf16_fwd_bn64_s1_us'])
PY
