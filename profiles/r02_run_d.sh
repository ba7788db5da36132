#!/bin/bash
# round 2, GPU call D: the camera level in one kernel (k_camera) - parity suite, A/B against the separate launches
# (RTP_FUSED=0) on C3 / C2 / C4 and on a rank's share, bounds-checked build, C5 capture.
set -u
O=gpurun_out/r02d; mkdir -p $O
P=ray-tracing-project_b200
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
for f in 1 0; do
  for w in c3 c2 c1; do
    RTP_FUSED=$f timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/fused${f}_$w.json 2> $O/fused${f}_$w.err
  done
  RTP_FUSED=$f timeout 600 python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline > $O/fused${f}_c4.json 2> $O/fused${f}_c4.err
  RTP_FUSED=$f timeout 300 python profiles/share_time.py --workload c3 > $O/share_c3_fused$f.log 2>&1
  RTP_FUSED=$f timeout 300 python profiles/share_time.py --workload c2 --worlds 1,8 > $O/share_c2_fused$f.log 2>&1
done
RTP_FUSED=1 RTP_GROUPS=1 timeout 300 python profiles/share_time.py --workload c3 > $O/share_c3_fused1_g1.log 2>&1
for b in 2 3 4; do RTP_BUDGET=$b timeout 300 python profiles/share_time.py --workload c3 --worlds 1,8 > $O/share_c3_fused1_budget$b.log 2>&1; done
timeout 900 python profiles/checked_run.py > $O/checked_default.log 2>&1
RTP_BUDGET=1 RTP_PARK_LANES=32 RTP_PARK_ANY=1 RTP_PARK_SMALL=1 timeout 900 python profiles/checked_run.py > $O/checked_coop.log 2>&1
T=/tmp/ncu_r02d; mkdir -p $T
timeout 900 ncu --set full --clock-control none --kernel-name-base demangled -k regex:"RaysAny" --launch-skip 3 -c 1 -o $T/ncu_c5 -f \
  python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c5.log 2>&1
ncu -i $T/ncu_c5.ncu-rep --page raw --csv > $O/ncu_c5_raw.csv 2>/dev/null
RTP_GROUPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_camera|k_coop_camera" --launch-skip 8 -c 4 -o $O/ncu_c3_camera -f \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c3_camera.log 2>&1
du -sh $O
tail -n 4 $O/pytest_gpu.log
for f in $O/checked_*.log; do echo "== $f"; tail -n 2 $f | cut -c1-300; done
for f in $O/fused*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), d['roofline'].get('kernel'), d['roofline'].get('kernels_ms'))
except Exception as e: print(sys.argv[1], 'failed', e)
PY
done
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | cut -c1-130; done
