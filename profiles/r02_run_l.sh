#!/bin/bash
# round 2, GPU call L: k_paths with helper lanes (idle lanes trace shadow rays of their warp's paths) - parity, A/B
set -u
O=gpurun_out/r02l; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
for hlp in 1 0; do
  RTP_PATH_HELP=$hlp timeout 300 python profiles/share_time.py --workload c2 --worlds 8 --all-ranks --steps 20 > $O/share_c2_help$hlp.log 2>&1
  RTP_PATH_HELP=$hlp timeout 300 python profiles/share_time.py --workload c2 --worlds 1,2,4 --steps 20 > $O/share_c2_w124_help$hlp.log 2>&1
  RTP_PATH_HELP=$hlp timeout 600 python profiles/share_time.py --workload c4 --worlds 4,8 --steps 8 > $O/share_c4_help$hlp.log 2>&1
  RTP_PATH_HELP=$hlp RTP_DEEP=paths timeout 300 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > $O/c2_paths_help$hlp.json 2> $O/c2_paths_help$hlp.err
  RTP_PATH_HELP=$hlp RTP_DEEP=paths timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > $O/c4_paths_help$hlp.json 2> $O/c4_paths_help$hlp.err
  RTP_PATH_HELP=$hlp timeout 300 python bench.py --workload c1 --steps 20 --warmup 5 --no-cpu-baseline > $O/c1_help$hlp.json 2> $O/c1_help$hlp.err
done
tail -n 3 $O/pytest_gpu.log
for f in $O/*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), d['roofline'].get('kernel'), d['roofline'].get('kernels_ms'))
except Exception as e: print(sys.argv[1], 'failed', e)
PY
done
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | python -c "
import sys,re
for l in sys.stdin: print(l.split(' ')[0], [round(float(x),4) for x in re.findall(r\"'ms': ([0-9.]+)\",l)])
"; done
