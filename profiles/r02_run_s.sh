#!/bin/bash
# round 2, GPU call S: did the live-consumer plumbing (publish fence in cam_park, exit counter, sized grid) cost the
# shipped path anything?  previous commit's library against the current one on the same box, interleaved
set -u
O=gpurun_out/r02s; mkdir -p $O
P=ray-tracing-project_b200
for rep in 1 2; do
  RTP_LIB=$P/libflyscene_b200_prev.so timeout 300 python profiles/share_time.py --workload c3 --worlds 2,4,8 --steps 30 > $O/share_c3_prev_$rep.log 2>&1
  timeout 300 python profiles/share_time.py --workload c3 --worlds 2,4,8 --steps 30 > $O/share_c3_cur_$rep.log 2>&1
  RTP_CAM_FULLGRID=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 2,4,8 --steps 30 > $O/share_c3_curfull_$rep.log 2>&1
done
RTP_LIB=$P/libflyscene_b200_prev.so timeout 300 python profiles/share_time.py --workload c2 --worlds 1,8 --steps 30 > $O/share_c2_prev.log 2>&1
timeout 300 python profiles/share_time.py --workload c2 --worlds 1,8 --steps 30 > $O/share_c2_cur.log 2>&1
RTP_CAM_FULLGRID=1 timeout 300 python profiles/share_time.py --workload c2 --worlds 1,8 --steps 30 > $O/share_c2_curfull.log 2>&1
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | python -c "
import sys,re
for l in sys.stdin: print(l.split(' ')[0], [round(float(x),4) for x in re.findall(r\"'ms': ([0-9.]+)\",l)])
"; done
