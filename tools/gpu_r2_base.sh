# Round 2 baseline: full GPU suite, bench with phases / eager module swap, then ONE sanitizer tool on the op-level suite
O=gpurun_out/r2base; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest.log; tail -3 $O/pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -c 600 $O/bench.err
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_ops.py tests/test_gpu_tc.py -m gpu -x -q -k "not 4096" > $O/sanitizer_memcheck.log 2>&1; echo "memcheck exit $?"; tail -5 $O/sanitizer_memcheck.log
