O=gpurun_out/r2prof3; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest.log | cut -c1-200
timeout 900 python bench.py --steps 30 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -c 300 $O/bench.err; cut -c1-300 $O/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference arm exit $?"
BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py > $O/trace.txt 2> $O/trace.err; head -5 $O/trace.txt
STEPCMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu --no-phases --no-eager --no-pipeline"
timeout 300 $STEPCMD > $O/step_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file $O/launches.csv $STEPCMD > $O/ncu_launch.log 2>&1; echo "ncu launches exit $?"
