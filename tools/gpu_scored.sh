O=gpurun_out/scored; mkdir -p $O
for w in worker manager; do
  timeout 200 python bench.py --workload $w --scorer cider --no-cpu --steps 40 > $O/bench_${w}_cider.json 2> $O/bench_${w}_cider.err; echo "$w exit $?"; tail -3 $O/bench_${w}_cider.err
  python -c "
import json,sys
d=json.loads(open('$O/bench_${w}_cider.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e'], d.get('scorer_ms_last_step'), d['launches_per_step'])"
done
