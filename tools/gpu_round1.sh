set -x
mkdir -p gpurun_out/r1
O=gpurun_out/r1
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?" >> $O/pytest.log
timeout 400 python bench.py > $O/bench_warmstart.json 2> $O/bench_warmstart.err; echo "exit $?" >> $O/bench_warmstart.err
timeout 300 python bench.py --workload worker --no-cpu > $O/bench_worker.json 2> $O/bench_worker.err
timeout 300 python bench.py --workload manager --no-cpu > $O/bench_manager.json 2> $O/bench_manager.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
timeout 300 python tools/microbench.py gemm attn ln > $O/microbench.txt 2>&1
timeout 300 python tools/trace_step.py warmstart > $O/trace_warmstart.txt 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-graph --no-cpu > $O/ncu_launch.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gemm_tf32_tc_kernel --launch-skip 40 -c 6 -o $O/gemm_full python tools/microbench.py gemm > $O/ncu_full.log 2>&1
ls -la $O
