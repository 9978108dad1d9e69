O=gpurun_out/final; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log; tail -3 $O/pytest.log
timeout 400 python bench.py > $O/bench_warmstart.json 2> $O/bench_warmstart.err; echo "bench exit $?"
timeout 300 python bench.py --workload worker --no-cpu > $O/bench_worker.json 2> $O/bench_worker.err
timeout 300 python bench.py --workload manager --no-cpu > $O/bench_manager.json 2> $O/bench_manager.err
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-graph --no-cpu > $O/ncu_launch.log 2>&1; echo "ncu launches exit $?"
ATTN_SHAPES=800x800 timeout 400 ncu --set full --clock-control none --import-source on -k regex:attn_fwd_bf16 --launch-skip 2 --launch-count 1 -f -o $O/attn_bf16_800 python tools/attn_sweep.py --quick > $O/ncu_attn800.log 2>&1
ATTN_SHAPES=300x300 timeout 400 ncu --set full --clock-control none --import-source on -k regex:attn_fwd_bf16 --launch-skip 2 --launch-count 1 -f -o $O/attn_bf16_300 python tools/attn_sweep.py --quick > $O/ncu_attn300.log 2>&1
ls -la $O
