O=gpurun_out/r2exp28; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
run() { name=$1; shift; echo -n "$name: "; env "$@" timeout 300 python tools/trace_step.py 2> $O/$name.err | grep "back-to-back" | cut -c1-80; }
run base X=1
run wbn128_s2 BMHRL_GB_WGRAD_BN=128 BMHRL_GB_WGRAD_STAGES=2
run wbn128_s3 BMHRL_GB_WGRAD_BN=128 BMHRL_GB_WGRAD_STAGES=3
run wbn128_s4 BMHRL_GB_WGRAD_BN=128 BMHRL_GB_WGRAD_STAGES=4
run wbn64_s2 BMHRL_GB_WGRAD_BN=64 BMHRL_GB_WGRAD_STAGES=2
run wbn64_s4 BMHRL_GB_WGRAD_BN=64 BMHRL_GB_WGRAD_STAGES=4
run base2 X=1
