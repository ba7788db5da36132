#!/bin/bash
# round 2, GPU call T: after taking the live-consumer plumbing out - parity suite, shares against the previous commit's library
set -u
O=gpurun_out/r02t; mkdir -p $O
P=ray-tracing-project_b200
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
RTP_LIB=$P/libflyscene_b200_prev.so timeout 300 python profiles/share_time.py --workload c3 --worlds 1,2,4,8 --steps 30 > $O/share_c3_prev.log 2>&1
timeout 300 python profiles/share_time.py --workload c3 --worlds 1,2,4,8 --steps 30 > $O/share_c3_cur.log 2>&1
timeout 900 python profiles/checked_run.py > $O/checked_default.log 2>&1
tail -n 3 $O/pytest_gpu.log; tail -n 1 $O/checked_default.log
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | python -c "
import sys,re
for l in sys.stdin: print(l.split(' ')[0], [round(float(x),4) for x in re.findall(r\"'ms': ([0-9.]+)\",l)])
"; done
