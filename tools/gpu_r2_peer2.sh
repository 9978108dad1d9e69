N=${1:-2}
O=gpurun_out/r2peer; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 tools/peer_check.py > $O/peer_check_n$N.txt 2> $O/peer_check_n$N.err; echo "exit $?"; cat $O/peer_check_n$N.txt; tail -5 $O/peer_check_n$N.err | cut -c1-300
