# Graphed greedy decoder + library-made side streams: agent tests, config-1 decode timing, then the whole GPU suite.
O=gpurun_out/gdec2; mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_agent.py -q -x > $O/pytest_agent.log 2>&1; echo "pytest agent exit $?" | tee -a $O/pytest_agent.log; tail -4 $O/pytest_agent.log
timeout 100 python tools/bench_decode.py > $O/bench_decode.json 2> $O/bench_decode.err; echo "bench_decode exit $?"; tail -c 700 $O/bench_decode.json; tail -3 $O/bench_decode.err
timeout 240 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest.log; tail -4 $O/pytest.log
