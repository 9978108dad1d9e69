O=gpurun_out/r2exp18; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 600 python -m pytest tests/test_gpu_peer.py tests/test_gpu_ops.py -x -q -m gpu -k "peer or adam or Adam" > $O/pytest.log 2>&1; tail -5 $O/pytest.log
