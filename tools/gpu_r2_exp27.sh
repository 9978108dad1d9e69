O=gpurun_out/r2exp27; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
for v in 0 1; do echo -n "BMHRL_NODE_PRIORITY=$v: "; BMHRL_NODE_PRIORITY=$v timeout 300 python tools/trace_step.py > $O/trace_$v.txt 2> $O/t$v.err; grep "back-to-back" $O/trace_$v.txt | cut -c1-80; tail -2 $O/t$v.err | cut -c1-300; done
