N=${1:-2}
O=gpurun_out/multi; mkdir -p $O
t0=$(date +%s)
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "exit $? after $(( $(date +%s) - t0 )) s"
python - $O/bench_n$N.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("n", d["n_gpus"], "ms", round(d["ms_per_step"],3), "clips/s", round(d["value"],1), "e2e", round(d["e2e"]["value"],1))
PY
t0=$(date +%s)
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $O/ref_n$N.json 2> $O/ref_n$N.err; echo "ref exit $? after $(( $(date +%s) - t0 )) s"
cut -c1-200 $O/ref_n$N.json
