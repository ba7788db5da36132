#!/bin/bash
# round 2, GPU call C: C5 capture, bounds-checked build over the parity suite, design-item variant builds, L2 persistence,
# launch timelines of a 1/8 share.
set -u
O=gpurun_out/r02c; mkdir -p $O
P=ray-tracing-project_b200
T=/tmp/ncu_r02c; mkdir -p $T
timeout 900 ncu --set full --clock-control none -k regex:"RaysAny" --launch-skip 3 -c 1 -o $T/ncu_c5 -f \
  python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c5.log 2>&1
ncu -i $T/ncu_c5.ncu-rep --page raw --csv > $O/ncu_c5_raw.csv 2>/dev/null
# bounds-checked build: default policy, everything through k_coop, k_paths forced, per-level kernels forced
timeout 900 python profiles/checked_run.py > $O/checked_default.log 2>&1
RTP_BUDGET=1 RTP_PARK_LANES=32 RTP_PARK_ANY=1 RTP_PARK_SMALL=1 timeout 900 python profiles/checked_run.py > $O/checked_coop.log 2>&1
RTP_DEEP=paths timeout 900 python profiles/checked_run.py > $O/checked_paths.log 2>&1
RTP_DEEP=wavefront timeout 900 python profiles/checked_run.py > $O/checked_wavefront.log 2>&1
RTP_LIB=$P/libflyscene_b200_checked.so timeout 300 python profiles/sanitize_case.py > $O/checked_sequences.log 2>&1
# design items: each variant on C3 / C2 / C5, same command as the driver's
for v in "" _smem _pf _mb6 _mb10; do
  for w in c3 c2; do
    RTP_LIB=$P/libflyscene_b200$v.so timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/var${v}_$w.json 2> $O/var${v}_$w.err
  done
  RTP_LIB=$P/libflyscene_b200$v.so timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > $O/var${v}_c5.json 2> $O/var${v}_c5.err
done
for mb in 16 32 64; do
  for w in c3 c2; do
    RTP_L2_PERSIST=$mb timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/l2p${mb}_$w.json 2> $O/l2p${mb}_$w.err
  done
done
# launch timeline of rank 0's share of the C3 frame at world 8 (one GPU), 3 bands and 1 band
RTP_TIMELINE=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 8 --steps 4 --timeline > $O/timeline_w8_g3.log 2>&1
RTP_TIMELINE=1 RTP_GROUPS=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 8 --steps 4 --timeline > $O/timeline_w8_g1.log 2>&1
RTP_GROUPS=1 timeout 300 python profiles/share_time.py --workload c3 > $O/share_c3_g1.log 2>&1
RTP_GROUPS=2 timeout 300 python profiles/share_time.py --workload c3 --worlds 4,8 > $O/share_c3_g2.log 2>&1
du -sh $O
for f in $O/checked_*.log; do echo "== $f"; tail -n 3 $f | cut -c1-300; done
for f in $O/var*.json $O/l2p*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], round(d['value']), round(d['ms_per_step'],4), d['roofline'].get('kernels_ms'))
except Exception as e: print(sys.argv[1], 'failed', e)
PY
done
tail -n 3 $O/share_c3_g1.log $O/share_c3_g2.log | cut -c1-400
