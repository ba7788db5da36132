#!/bin/bash
# round 2, GPU call X: prioritised tile-group streams (band 0 first) so that the band copies overlap the later bands' kernels
set -u
O=gpurun_out/r02x; mkdir -p $O
for sg in 1 0; do
  RTP_STAGGER=$sg timeout 300 python bench.py --workload c3 --steps 40 --warmup 5 --no-cpu-baseline > $O/c3_stagger$sg.json 2> $O/c3_stagger$sg.err
  RTP_STAGGER=$sg timeout 300 python bench.py --workload c2 --steps 40 --warmup 5 --no-cpu-baseline > $O/c2_stagger$sg.json 2> $O/c2_stagger$sg.err
  RTP_STAGGER=$sg timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > $O/c4_stagger$sg.json 2> $O/c4_stagger$sg.err
done
RTP_STAGGER=1 RTP_TIMELINE=1 timeout 300 python profiles/share_time.py --workload c3 --worlds 1 --steps 3 --timeline 2>&1 | grep -A10 "rank 0 frame 4" | head -12
RTP_STAGGER=1 timeout 600 python -m pytest tests/test_gpu_groups.py tests/test_gpu_async.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
for f in $O/*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'pipe', round(d['e2e'].get('pipelined',{}).get('ms_per_step',0),4))
PY
done
