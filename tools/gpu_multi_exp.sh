N=${1:-2}
O=gpurun_out/multi; mkdir -p $O
run() { name=$1; shift; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 40 --warmup 5 --no-cpu > $O/exp_$name.json 2> $O/exp_$name.err; python - $O/exp_$name.json $name <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], "n", d["n_gpus"], "ms", round(d["ms_per_step"],3), "clips/s", round(d["value"],1))
except Exception as e: print(sys.argv[2], "FAILED", e)
PY
}
run default A=1
run cta4 NCCL_MAX_CTAS=4
run cta8 NCCL_MAX_CTAS=8
run cta16 NCCL_MAX_CTAS=16
run nvls0 NCCL_NVLS_ENABLE=0
