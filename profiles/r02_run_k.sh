#!/bin/bash
# round 2, GPU call K: the parked-pixel tail of a small share - CTAs of k_coop_camera, lane threshold of parking
set -u
O=gpurun_out/r02k; mkdir -p $O
for cb in 2 3 5; do
  RTP_COOP_BLOCKS=$cb timeout 300 python profiles/share_time.py --workload c3 --worlds 1,4,8 --steps 20 > $O/share_c3_coop$cb.log 2>&1
done
for pl in 8 16; do
  RTP_PARK_LANES=$pl timeout 300 python profiles/share_time.py --workload c3 --worlds 4,8 --steps 20 > $O/share_c3_lanes$pl.log 2>&1
done
RTP_COOP_BLOCKS=5 RTP_GROUPS=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 8 --steps 20 > $O/share_c3_coop5_g1.log 2>&1
RTP_COOP_BLOCKS=5 RTP_GROUPS=2 timeout 300 python profiles/share_time.py --workload c3 --worlds 8 --steps 20 > $O/share_c3_coop5_g2.log 2>&1
RTP_COOP_BLOCKS=5 timeout 300 python profiles/share_time.py --workload c2 --worlds 1,8 --steps 20 > $O/share_c2_coop5.log 2>&1
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | python -c "
import sys,re
for l in sys.stdin: print(l.split(' ')[0], [round(float(x),4) for x in re.findall(r\"'ms': ([0-9.]+)\",l)])
"; done
