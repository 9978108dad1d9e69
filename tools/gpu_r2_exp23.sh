O=gpurun_out/r2exp23; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest.log 2>&1; tail -4 $O/pytest.log | cut -c1-250
timeout 300 python tools/trace_step.py > $O/trace.txt 2> $O/trace.err; head -4 $O/trace.txt
