#!/bin/bash
# round 2, GPU call J: hashed-rotation tile interleave - parity suite, every rank's share of C2 / C3 / C4 at world 8 (and 4)
set -u
O=gpurun_out/r02j; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
timeout 300 python profiles/share_time.py --workload c3 --worlds 2,4,8 --all-ranks --steps 20 > $O/share_c3_all.log 2>&1
timeout 300 python profiles/share_time.py --workload c2 --worlds 2,4,8 --all-ranks --steps 20 > $O/share_c2_all.log 2>&1
timeout 600 python profiles/share_time.py --workload c4 --worlds 8 --all-ranks --steps 5 > $O/share_c4_all.log 2>&1
tail -n 3 $O/pytest_gpu.log
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | python -c "
import sys,re
for l in sys.stdin: print([round(float(x),4) for x in re.findall(r\"'ms': ([0-9.]+)\",l)])
"; done
