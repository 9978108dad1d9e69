O=gpurun_out/r2n4; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --steps 20 --warmup 5 --no-phases > $O/bench_n4.json 2> $O/bench_n4.err; echo "n4 exit $?"; tail -c 300 $O/bench_n4.err; python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2n4/bench_n4.json').read().strip().splitlines()[-1])
print("n", d["n_gpus"], "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), d["config"]["parallelism"][:80], "launches", d["launches_per_step"])
PY
