#!/bin/bash
# round 2, last call: the final commit once more as the driver will run it - GPU tests, smoke, the default bench line, the reference arm
set -u
O=gpurun_out/r02verify; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_c3.json 2> $O/bench_c3.err; echo "rc=$?" >> $O/bench_c3.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_c3_reference.json 2> $O/bench_c3_reference.err
tail -n 3 $O/pytest_gpu.log; tail -n 2 $O/smoke.log
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02verify/bench_c3.json').read().strip().splitlines()[-1]); r=d['roofline']
print('c3', round(d['value'],1), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],1), 'launches', d['gpu_launches'], r['kernel'], round(r['frac'],4), 'clock samples', d['clocks']['samples'], 'cpu', round(d['cpu_baseline']['value'],4))
d=json.loads(open('gpurun_out/r02verify/bench_c3_reference.json').read().strip().splitlines()[-1]); print('reference', d['value'], d.get('impl'))
PY
