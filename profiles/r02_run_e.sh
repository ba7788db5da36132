#!/bin/bash
# round 2, GPU call E: parking budget of the fused camera kernel (k_coop_camera is dearer per parked pixel than k_coop per
# parked ray), C4 share fused / unfused
set -u
O=gpurun_out/r02e; mkdir -p $O
for b in 8 10 14 20 0; do
  RTP_FUSED=1 RTP_BUDGET=$b timeout 300 python profiles/share_time.py --workload c3 --worlds 1,2,4,8 --steps 20 > $O/share_c3_fused1_budget$b.log 2>&1
done
for b in 10 20; do
  RTP_FUSED=1 RTP_BUDGET=$b RTP_BUDGET_ANY=6 timeout 300 python profiles/share_time.py --workload c3 --worlds 4,8 --steps 20 > $O/share_c3_fused1_budget${b}_any6.log 2>&1
done
for f in 0 1; do
  RTP_FUSED=$f timeout 600 python profiles/share_time.py --workload c4 --worlds 4,8 --steps 10 > $O/share_c4_fused$f.log 2>&1
done
RTP_FUSED=1 RTP_BUDGET=14 timeout 600 python profiles/share_time.py --workload c4 --worlds 1,8 --steps 10 > $O/share_c4_fused1_budget14.log 2>&1
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | cut -c1-150; done
