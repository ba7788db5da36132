#!/bin/bash
# round 2, GPU call F: fused-by-size policy (parity suite, bench lines, shares), inner nodes per step for scattered rays
set -u
O=gpurun_out/r02f; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
for w in c3 c2 c1; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err
done
timeout 600 python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err
timeout 300 python profiles/share_time.py --workload c3 > $O/share_c3.log 2>&1
timeout 300 python profiles/share_time.py --workload c2 > $O/share_c2.log 2>&1
for k in 32 16 8 4 2; do
  RTP_STEP_NODES=$k timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > $O/step${k}_c5.json 2> $O/step${k}_c5.err
done
for k in 16 8 4; do
  RTP_STEP_NODES=$k timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > $O/step${k}_c4.json 2> $O/step${k}_c4.err
  RTP_STEP_NODES=$k RTP_DEEP=wavefront timeout 300 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > $O/step${k}_c2wave.json 2> $O/step${k}_c2wave.err
done
RTP_DEEP=wavefront timeout 300 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > $O/step32_c2wave.json 2> $O/step32_c2wave.err
tail -n 4 $O/pytest_gpu.log
for f in $O/bench_*.json $O/step*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), d['roofline'].get('kernel'), d['roofline'].get('frac'), d['roofline'].get('kernels_ms'))
except Exception as e: print(sys.argv[1], 'failed', e)
PY
done
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | cut -c1-150; done
