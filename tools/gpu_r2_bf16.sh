O=gpurun_out/r2bf16; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_agent.py::test_scored_step_two_graphs_equals_eager > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -30 $O/pytest.log | cut -c1-250
BMHRL_MATH=bf16 BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py warmstart > $O/trace_bf16.txt 2> $O/trace.err; echo "trace exit $?"; head -60 $O/trace_bf16.txt
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -c 400 $O/bench.err
