O=gpurun_out/r2exp9; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
BMHRL_PDL=0 timeout 600 python tools/bench_gemm16.py > $O/bench_gemm16.jsonl 2> $O/err.log; python - <<'PY'
import json
for l in open('gpurun_out/r2exp9/bench_gemm16.jsonl'):
    d=json.loads(l); print(d['what'], d['M'],d['N'],d['K'], 'fwd cold', d['bf16_fwd_us'], 'hot', d['bf16_fwd_hotW_us'], 'dx cold', d['bf16_dx_us'], 'hot', d['bf16_dx_hotW_us'], 'dw', d['bf16_dw_us'], 'tf32', d['tf32_fwd_us'])
PY
