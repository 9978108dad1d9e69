# N-GPU bench line only (after a change to the stream plumbing): torchrun, one rank per GPU, NCCL.
N=${1:-2}
O=gpurun_out/multi; mkdir -p $O
t0=$(date +%s)
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "exit $? after $(( $(date +%s) - t0 )) s"
tail -3 $O/bench_n$N.err
python - $O/bench_n$N.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("n", d["n_gpus"], "ms", round(d["ms_per_step"],3), "clips/s", round(d["value"],1), "e2e", round(d["e2e"]["value"],1))
PY
