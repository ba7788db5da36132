#!/bin/bash
set -u
O=gpurun_out/r02p; mkdir -p $O
for g in 1 3; do
  RTP_TIMELINE=1 RTP_GROUPS=$g timeout 300 python profiles/share_time.py --workload c3 --worlds 8 --steps 3 --timeline > $O/timeline_c3_w8_g$g.log 2>&1
  RTP_TIMELINE=1 RTP_GROUPS=$g timeout 300 python profiles/share_time.py --workload c3 --worlds 4 --steps 3 --timeline > $O/timeline_c3_w4_g$g.log 2>&1
done
for f in $O/timeline_*.log; do echo "== $f"; grep -B1 -A8 "rank 0 frame 4" $f | head -12; done
