# usage: bash tools/gpu_r2_multi2.sh N   (inside gpurun --gpus N): data-parallel step under NCCL settings, rank-0 timelines
N=${1:-2}
O=gpurun_out/r2multi2; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
run() {  # name, env...
  name=$1; shift
  env "$@" BMHRL_TIMELINE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/trace_step.py > $O/trace_$name.txt 2> $O/trace_$name.err
  echo "$name: $(grep 'back-to-back' $O/trace_$name.txt | cut -c1-80)"; grep -i "nccl" $O/trace_$name.txt | grep -v "^ *[0-9.]* *[0-9.]* [0-9]" | head -4
}
timeout 300 python tools/trace_step.py > $O/trace_n1.txt 2> $O/trace_n1.err; grep 'back-to-back' $O/trace_n1.txt | cut -c1-80
run default X=1
run ch4 NCCL_MAX_NCHANNELS=4
run ch8 NCCL_MAX_NCHANNELS=8
run ch16 NCCL_MAX_NCHANNELS=16
run nooverlap BMHRL_EARLY_STAGES=0
run bucket8 BMHRL_BUCKET_MB=8
run bucket128 BMHRL_BUCKET_MB=128
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,GRAPH,TUNING timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 tools/trace_step.py > $O/trace_debug.txt 2> $O/trace_debug.err; grep -i "channels\|NVLS\|algo\|Connected" $O/trace_debug.err $O/trace_debug.txt | head -20 | cut -c1-200
