# usage: bash tools/gpu_r2_multi.sh N   (inside gpurun --gpus N)
N=${1:-2}
O=gpurun_out/r2multi; mkdir -p $O
(make -C bmhrl_b200/csrc -j16 > $O/make.log 2>&1 || (tail -20 $O/make.log; exit 1)) && echo "build ok"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-phases --no-eager > $O/bench_n1.json 2> $O/bench_n1.err; echo "n1 exit $?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "n$N exit $?"; tail -c 600 $O/bench_n$N.err
for f in $O/bench_n1.json $O/bench_n$N.json; do python - "$f" <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "n", d["n_gpus"],
          {k: round(v["ms_per_step"], 3) for k, v in d.get("phases", {}).items()})
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
