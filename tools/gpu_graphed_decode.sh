# Graphed greedy decoder: whole GPU suite, then (only after it exited 0 without ncu) the launch list of ONE graphed
# caption (config 1, tf32) for profiles/.
O=gpurun_out/gdec4; mkdir -p $O
timeout 240 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest.log; tail -4 $O/pytest.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -1 $O/smoke.log
timeout 60 python tools/bench_decode.py --only graphed --modes tf32 --reps 10 > $O/bench_graphed.json 2> $O/bench_graphed.err && \
timeout 150 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_graphed_decode.csv \
  python tools/bench_decode.py --only graphed --modes tf32 --reps 1 --profile > $O/ncu.log 2>&1; echo "ncu exit $?"; tail -c 300 $O/bench_graphed.json
python tools/summarize_region.py $O/launches_graphed_decode.csv "one graphed caption, config 1 (B=1, Sv=14, Sa=36 in 16/40 buffers, 30 tokens), tf32" > $O/launches_graphed_decode.txt 2>&1; head -25 $O/launches_graphed_decode.txt
