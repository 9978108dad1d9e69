O=gpurun_out/r2exp20; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
run() { name=$1; shift; env "$@" timeout 300 python tools/trace_step.py 2> $O/$name.err | grep "back-to-back" | sed "s/^/$name: /" | cut -c1-90; }
run base X=1
run deep4 BMHRL_GB_DEEP=4
run deep5 BMHRL_GB_DEEP=5
run deep6 BMHRL_GB_DEEP=6
run bn128 BMHRL_GB_BN=128
run bn128s3 BMHRL_GB_BN=128 BMHRL_GB_STAGES=3
run stages4 BMHRL_GB_STAGES=4
run stages2 BMHRL_GB_STAGES=2
run nosplit BMHRL_GB_SPLIT=1
run base2 X=1
