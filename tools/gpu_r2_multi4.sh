N=${1:-2}
O=gpurun_out/r2multi4_n$N; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
BMHRL_TIMELINE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/trace_step.py > $O/trace_peer_n$N.txt 2> $O/trace_peer_n$N.err
grep 'back-to-back' $O/trace_peer_n$N.txt | cut -c1-80; grep "adam_peer\|peer_wait\|peer_signal" $O/trace_peer_n$N.txt | cut -c1-100
