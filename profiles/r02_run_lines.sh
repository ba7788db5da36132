#!/bin/bash
# round 2: the bench lines once more with the final traffic.json in place (roofline.frac for every workload)
set -u
O=gpurun_out/r02lines; mkdir -p $O
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_c3.json 2> $O/bench_c3.err
for w in c1 c2; do timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err; done
timeout 600 python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d.get('roofline') or {}
print(sys.argv[1].split('/')[-1], round(d['value'],1), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],1), r.get('kernel'), 'frac', r.get('frac'), 'lanes', r.get('lane_efficiency'), 'dram_frac', r.get('dram_frac'), 'contract', r['contract']['frac'])
PY
done
