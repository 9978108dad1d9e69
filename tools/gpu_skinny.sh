# Skinny linear kernel: whole GPU suite, then the graphed decoder with the kernel on (default) and off.
O=gpurun_out/skinny; mkdir -p $O
timeout 70 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest.log; tail -3 $O/pytest.log
timeout 25 python tools/bench_decode.py --only graphed --modes tf32 --reps 10 > $O/bench_on.json 2> $O/on.err; echo "on exit $?"; tail -c 130 $O/bench_on.json
BMHRL_SKINNY_LINEAR=0 timeout 25 python tools/bench_decode.py --only graphed --modes tf32 --reps 10 > $O/bench_off.json 2> $O/off.err; echo "off exit $?"; tail -c 130 $O/bench_off.json
