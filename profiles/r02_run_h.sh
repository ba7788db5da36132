#!/bin/bash
# round 2, GPU call H: resident CTAs per SM once more (9 = 56 registers; any-hit kernels only / all trace kernels)
set -u
O=gpurun_out/r02h; mkdir -p $O
P=ray-tracing-project_b200
for v in "" _any9 _mb9; do
  for w in c3 c2; do
    RTP_LIB=$P/libflyscene_b200$v.so timeout 300 python bench.py --workload $w --steps 40 --warmup 5 --no-cpu-baseline > $O/var${v}_$w.json 2> $O/var${v}_$w.err
  done
  RTP_LIB=$P/libflyscene_b200$v.so timeout 300 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > $O/var${v}_c5.json 2> $O/var${v}_c5.err
  RTP_LIB=$P/libflyscene_b200$v.so timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > $O/var${v}_c4.json 2> $O/var${v}_c4.err
  RTP_LIB=$P/libflyscene_b200$v.so timeout 300 python profiles/share_time.py --workload c3 --worlds 2,8 --steps 20 > $O/share_c3$v.log 2>&1
done
for f in $O/var*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), d['roofline'].get('kernel'), d['roofline'].get('kernels_ms'))
except Exception as e: print(sys.argv[1], 'failed', e)
PY
done
for f in $O/share_*.log; do echo "== $f"; grep -E "^[0-9]+ \[" $f | cut -c1-150; done
