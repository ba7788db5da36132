#!/bin/bash
# round 2, GPU call I: what a 1/8 share waits for (ncu of k_camera on rank 0's share of C3), every rank's share of C2 / C3 at
# world 8, C5 with 9 CTAs per SM for the ray-batch any-hit kernel
set -u
O=gpurun_out/r02i; mkdir -p $O
timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q -k "ray_batch or noalloc" > $O/pytest_rays.log 2>&1
timeout 300 python profiles/share_time.py --workload c3 --worlds 8 --all-ranks --steps 20 > $O/share_c3_all.log 2>&1
timeout 300 python profiles/share_time.py --workload c2 --worlds 8 --all-ranks --steps 20 > $O/share_c2_all.log 2>&1
RTP_TIMELINE=1 timeout 300 python profiles/share_time.py --workload c2 --worlds 8 --steps 3 --timeline > $O/timeline_c2_w8.log 2>&1
RTP_GROUPS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_camera|k_coop_camera" --launch-skip 6 -c 2 -o $O/ncu_c3_share8 -f \
  python profiles/share_time.py --workload c3 --worlds 8 --steps 3 > $O/ncu_c3_share8.log 2>&1
du -sh $O; tail -n 3 $O/pytest_rays.log
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02i/bench_c5.json').read().strip().splitlines()[-1]); print('c5', round(d['value']), round(d['ms_per_step'],3), d['roofline'].get('frac'))
PY
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | python -c "
import sys,re
for l in sys.stdin: print([round(float(x),4) for x in re.findall(r\"'ms': ([0-9.]+)\",l)])
"; done
tail -n 24 $O/timeline_c2_w8.log | cut -c1-120
