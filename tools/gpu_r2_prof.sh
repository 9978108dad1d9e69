O=gpurun_out/r2prof; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 900 python -m pytest tests -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest.log | cut -c1-200
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; tail -c 300 $O/bench.err
timeout 200 python tools/microbench16.py > $O/microbench16.jsonl 2>&1
STEPCMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu --no-phases --no-eager --no-pipeline"
timeout 300 $STEPCMD > $O/step_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file $O/launches.csv $STEPCMD > $O/ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 200 python tools/ncu_targets.py all > $O/targets_plain.log 2>&1 && {
for spec in "gemm:gemm_bf16_tc:1:gemm_fwd" "gemm:gemm_bf16_tc:5:gemm_dw" "attn:attn16_fwd:1:attn16_fwd" "attn:attn16_bwd:1:attn16_bwd" "norm:layernorm_fwd16:1:ln_fwd16" "norm:layernorm_bwd_det:1:ln_bwd" "norm:cast16_kernel:1:cast16" "norm:colsum16:1:colsum16" "loss:gemm_bf16_tc:1:vocab_stats" "adam:adam_multi:1:adam"; do
  IFS=: read what pat skip name <<< "$spec"
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 1 -f -o $O/$name python tools/ncu_targets.py $what > $O/ncu_$name.log 2>&1; echo "ncu $name exit $?"
done; }
ls -la $O | head -40
