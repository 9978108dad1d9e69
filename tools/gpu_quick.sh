# Short GPU iteration: the tests named in $1 (pytest -k expression), output under gpurun_out/quick
O=gpurun_out/quick; mkdir -p $O
timeout 400 python -m pytest tests -m gpu -x -q -k "$1" > $O/pytest.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest.log; tail -40 $O/pytest.log
