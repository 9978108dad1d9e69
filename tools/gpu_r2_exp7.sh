O=gpurun_out/r2exp7; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 300 python tools/microbench16.py > $O/microbench16.jsonl 2> $O/mb.err; echo "microbench exit $?"; cat $O/microbench16.jsonl; tail -3 $O/mb.err
timeout 300 python tools/diag_overlap.py > $O/diag_overlap.txt 2>&1; echo "diag exit $?"; grep -v "^     " $O/diag_overlap.txt | tail -12
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -8 $O/pytest.log | cut -c1-250; grep -n "fused vocab\|bf16 errors" $O/pytest.log | cut -c1-400
B="python bench.py --steps 20 --warmup 5 --no-cpu --no-phases --no-eager"
for cfg in "" "--no-pipeline"; do echo "== bench $cfg"; timeout 300 $B $cfg 2> $O/err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['launches_per_step'], d['e2e']['ms_per_step'])"; tail -3 $O/err.log; done
BMHRL_MATH=bf16 BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py warmstart > $O/trace_bf16.txt 2> $O/trace.err; echo "trace exit $?"; sed -n 1,3p $O/trace_bf16.txt; grep -n "total_us" -A 16 $O/trace_bf16.txt
