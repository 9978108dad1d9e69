# Last GPU call of round 1 (short budget): new tests first, then the host scorer / decoder timings, a worker-step
# bench line (rl_losses was reordered), the full GPU suite and the default bench line.
O=gpurun_out/r1last; mkdir -p $O
nproc > $O/host.txt; grep -m1 "model name" /proc/cpuinfo >> $O/host.txt
timeout 300 python -m pytest tests/test_gpu_agent.py -q -k "cider_scorer or encode_once" > $O/new_tests.log 2>&1; echo "new tests exit $?" | tee -a $O/new_tests.log; tail -15 $O/new_tests.log
timeout 120 python tools/bench_scorer.py > $O/bench_scorer.json 2> $O/bench_scorer.err; echo "bench_scorer exit $?"
timeout 200 python tools/bench_decode.py > $O/bench_decode.json 2> $O/bench_decode.err; echo "bench_decode exit $?"; tail -c 900 $O/bench_decode.json
timeout 200 python bench.py --workload worker --no-cpu --steps 30 > $O/bench_worker.json 2> $O/bench_worker.err; echo "bench worker exit $?"
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest.log; tail -5 $O/pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 300 python bench.py > $O/bench_warmstart.json 2> $O/bench_warmstart.err; echo "bench exit $?"; cut -c1-300 $O/bench_warmstart.json
