O=gpurun_out/exp; mkdir -p $O
run() { name=$1; shift; env "$@" timeout 300 python bench.py --no-cpu --steps 30 > $O/$name.json 2> $O/$name.err; python - $O/$name.json $name <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], "ms", round(d["ms_per_step"],3), "clips/s", round(d["value"],1), "roof_us", round(d["roofline"]["avg_launch_us"],2))
except Exception as e: print(sys.argv[2], "FAILED", e)
PY
}
run base A=1
run cluster BMHRL_TC_CLUSTER=1
run min60 BMHRL_TC_MIN_TILES=60
run min200 BMHRL_TC_MIN_TILES=200
run min300 BMHRL_TC_MIN_TILES=300
run noconc BMHRL_CONCURRENT=0
