O=gpurun_out/r2exp15; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest.log 2>&1; tail -4 $O/pytest.log
timeout 600 python bench.py --steps 30 --warmup 5 > $O/bench.json 2> $O/bench.err; cut -c1-1500 $O/bench.json; tail -2 $O/bench.err
timeout 300 python tools/gemm_trace.py > $O/gemm_trace.txt 2> $O/gemm_trace.err; grep -A2 "^video\|^caption\|^audio" $O/gemm_trace.txt | cut -c1-200; tail -3 $O/gemm_trace.err
for upb in 16 19; do BMHRL_CRITIC_UPB=$upb timeout 300 python tools/trace_step.py > $O/trace_cupb$upb.txt 2> $O/trace.err; head -1 $O/trace_cupb$upb.txt; done
