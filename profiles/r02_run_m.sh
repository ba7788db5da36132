#!/bin/bash
# round 2, GPU call M: the camera level's shading folded into its shadow launch (FrameAnyShade) - parity, A/B
set -u
O=gpurun_out/r02m; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
for a in 1 0; do
  RTP_ANY_SHADE=$a timeout 300 python bench.py --workload c3 --steps 40 --warmup 5 --no-cpu-baseline > $O/c3_anyshade$a.json 2> $O/c3_anyshade$a.err
  RTP_ANY_SHADE=$a timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > $O/c4_anyshade$a.json 2> $O/c4_anyshade$a.err
  RTP_ANY_SHADE=$a RTP_FUSED=0 timeout 300 python bench.py --workload c2 --steps 40 --warmup 5 --no-cpu-baseline > $O/c2_unfused_anyshade$a.json 2> $O/c2_unfused_anyshade$a.err
done
timeout 300 python bench.py --workload c3 --steps 40 --warmup 5 --no-cpu-baseline > $O/c3_anyshade1_again.json 2> /dev/null
tail -n 3 $O/pytest_gpu.log
for f in $O/*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'pipe', round(d['e2e'].get('pipelined',{}).get('ms_per_step',0),4), d['roofline'].get('kernel'), d['roofline'].get('kernels_ms'), d.get('gpu_launches'))
except Exception as e: print(sys.argv[1], 'failed', e)
PY
done
