O=gpurun_out/r2exp19; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 900 python -m pytest tests/test_gpu_bf16_mode.py -x -q -m gpu -s > $O/pytest.log 2>&1; grep "blocks that started" $O/pytest.log; tail -4 $O/pytest.log
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu --no-eager > $O/bench.json 2> $O/bench.err; cut -c1-330 $O/bench.json; tail -2 $O/bench.err
