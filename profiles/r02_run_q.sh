#!/bin/bash
# round 2, GPU call Q: the parked-pixel consumer beside k_camera (k_coop_camera<LIVE>) - parity, shares, timeline
set -u
O=gpurun_out/r02q; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
for l in 1 0; do
  RTP_LIVE_COOP=$l timeout 300 python profiles/share_time.py --workload c3 --worlds 2,4,8 --steps 30 > $O/share_c3_live$l.log 2>&1
done
RTP_LIVE_COOP=1 RTP_GROUPS=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 4,8 --steps 30 > $O/share_c3_live1_g1.log 2>&1
RTP_TIMELINE=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 8 --steps 3 --timeline > $O/timeline_c3_w8_live.log 2>&1
RTP_LIVE_COOP=1 RTP_FUSED=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 1 --steps 30 > $O/share_c3_w1_fused_live.log 2>&1
RTP_LIVE_COOP=1 RTP_FUSED=1 timeout 600 python profiles/share_time.py --workload c4 --worlds 1,8 --steps 8 > $O/share_c4_fused_live.log 2>&1
tail -n 3 $O/pytest_gpu.log
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | python -c "
import sys,re
for l in sys.stdin: print(l.split(' ')[0], [round(float(x),4) for x in re.findall(r\"'ms': ([0-9.]+)\",l)])
"; done
grep -A12 "rank 0 frame 4" $O/timeline_c3_w8_live.log | head -16
