O=gpurun_out/r2exp3; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
B="python bench.py --steps 20 --warmup 5 --no-cpu --no-phases --no-eager"
export BMHRL_GB_BN=64 BMHRL_GB_STAGES=3
for cfg in "" "BMHRL_RNN_UPB=8" "BMHRL_RNN_UPB=10" "BMHRL_RNN_UPB=15" "BMHRL_RNN_UPB=15 BMHRL_GB_STAGES=2" "BMHRL_RNN_UPB=15 BMHRL_GB_BN=128 BMHRL_GB_STAGES=2"; do
  echo "== $cfg"; env $cfg timeout 300 $B 2> $O/err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['launches_per_step'], d['e2e']['ms_per_step'])"
done
BMHRL_RNN_UPB=15 BMHRL_MATH=bf16 BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py warmstart > $O/trace_bf16.txt 2> $O/trace.err; echo "trace exit $?"; sed -n 1,3p $O/trace_bf16.txt; grep -n "total_us" -A 14 $O/trace_bf16.txt
timeout 600 python -m pytest tests/test_gpu_agent.py tests/test_gpu_bf16_mode.py -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $O/pytest.log | cut -c1-250
