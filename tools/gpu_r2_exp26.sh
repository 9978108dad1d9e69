O=gpurun_out/r2exp26; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
for c in 0 148 296 444 592; do echo -n "BMHRL_ADAM_CTAS=$c: "; BMHRL_ADAM_CTAS=$c timeout 300 python tools/trace_step.py 2> $O/t$c.err | grep "back-to-back" | cut -c1-80; done
BMHRL_TIMELINE=1 timeout 300 python tools/trace_step.py > $O/trace.txt 2> $O/trace.err; grep "adam_multi" $O/trace.txt | head -8 | cut -c1-100
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_peer.py tests/test_gpu_bf16_mode.py -x -q -m gpu > $O/pytest.log 2>&1; tail -3 $O/pytest.log | cut -c1-200
