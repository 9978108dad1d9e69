N=${1:-2}
O=gpurun_out/r2multi3; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
run() {  # name, env...
  name=$1; shift
  env "$@" BMHRL_TIMELINE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/trace_step.py > $O/trace_$name.txt 2> $O/trace_$name.err
  echo "$name: $(grep 'back-to-back' $O/trace_$name.txt | cut -c1-80)"
}
timeout 300 python -m pytest tests/test_gpu_agent.py -x -q -m gpu -k "overlap or pipelined or graph" > $O/pytest.log 2>&1; tail -2 $O/pytest.log
timeout 300 python tools/trace_step.py > $O/trace_n1.txt 2> $O/trace_n1.err; grep 'back-to-back' $O/trace_n1.txt | cut -c1-80
run default X=1
run bucket32 BMHRL_BUCKET_MB=32
run bucket512 BMHRL_BUCKET_MB=512
run nooverlap BMHRL_EARLY_STAGES=0
