O=gpurun_out/r2noalloc; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/pytest_alone.log 2>&1; echo "alone exit $?"; tail -3 $O/pytest_alone.log | cut -c1-200
timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest_all.log 2>&1; echo "all exit $?"; tail -3 $O/pytest_all.log | cut -c1-200
