#!/bin/bash
# round 2, GPU call R: size of the live parked-pixel consumer
set -u
O=gpurun_out/r02r; mkdir -p $O
RTP_LIVE_COOP=0 timeout 300 python profiles/share_time.py --workload c3 --worlds 2,4,8 --steps 30 > $O/share_c3_live0.log 2>&1
for cfgs in "148 32" "74 32" "296 32" "37 128" "148 64"; do
  set -- $cfgs
  RTP_LIVE_BLOCKS=$1 RTP_LIVE_THREADS=$2 timeout 300 python profiles/share_time.py --workload c3 --worlds 2,4,8 --steps 30 > $O/share_c3_live_b$1_t$2.log 2>&1
done
RTP_TIMELINE=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 8 --steps 3 --timeline > $O/timeline_c3_w8_live.log 2>&1
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | python -c "
import sys,re
for l in sys.stdin: print(l.split(' ')[0], [round(float(x),4) for x in re.findall(r\"'ms': ([0-9.]+)\",l)])
"; done
grep -A10 "rank 0 frame 4" $O/timeline_c3_w8_live.log | head -12
