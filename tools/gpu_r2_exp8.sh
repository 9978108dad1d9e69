O=gpurun_out/r2exp8; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
B="python bench.py --steps 20 --warmup 5 --no-cpu --no-phases --no-eager"
for cfg in "BMHRL_PDL=1" "BMHRL_PDL=0"; do echo "== bench $cfg"; env $cfg timeout 300 $B 2> $O/err.log | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['launches_per_step'], d['e2e']['ms_per_step'], d['loss'])"; tail -3 $O/err.log; done
timeout 300 python tools/diag_overlap.py > $O/diag_overlap.txt 2>&1; echo "diag exit $?"; grep -v "^     " $O/diag_overlap.txt | tail -12
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -8 $O/pytest.log | cut -c1-250
timeout 200 python tools/microbench16.py > $O/microbench16.jsonl 2>&1; head -8 $O/microbench16.jsonl
BMHRL_MATH=bf16 BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py warmstart > $O/trace_bf16.txt 2> $O/trace.err; echo "trace exit $?"; sed -n 1,3p $O/trace_bf16.txt; grep -n "total_us" -A 12 $O/trace_bf16.txt
