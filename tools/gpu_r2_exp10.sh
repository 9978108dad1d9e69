O=gpurun_out/r2exp10; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
for mc in default 32 16 4; do
  if [ $mc = default ]; then unset CUDA_DEVICE_MAX_CONNECTIONS; else export CUDA_DEVICE_MAX_CONNECTIONS=$mc; fi
  BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py > $O/trace_mc$mc.txt 2> $O/trace_mc$mc.err
  head -3 $O/trace_mc$mc.txt
done
unset CUDA_DEVICE_MAX_CONNECTIONS
BMHRL_DIAG_NO_CRITIC=1 timeout 300 python tools/trace_step.py > $O/trace_nocritic.txt 2> $O/trace_nocritic.err; head -3 $O/trace_nocritic.txt
BMHRL_PIPELINE=0 timeout 300 python tools/trace_step.py > $O/trace_nopipe.txt 2> $O/trace_nopipe.err; head -3 $O/trace_nopipe.txt
CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu --no-phases --no-eager > $O/bench_mc32.json 2> $O/bench_mc32.err; cat $O/bench_mc32.json | cut -c1-400
