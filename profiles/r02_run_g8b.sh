#!/bin/bash
# round 2, 8-GPU box, final code: the driver's scaling series on C3, the one-process path on 8, C5 on 8, multi-device tests
set -u
O=gpurun_out/r02g8b; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/c3_n1.json 2> $O/c3_n1.err
for n in 2 4 8; do
  timeout 300 $TR --nproc-per-node $n --master-port $((29600 + n)) bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline > $O/c3_n$n.json 2> $O/c3_n$n.err
done
timeout 300 python bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu-baseline > $O/c3_oneproc_n8.json 2> $O/c3_oneproc_n8.err
timeout 600 $TR --nproc-per-node 8 --master-port 29632 bench.py --gpus 8 --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > $O/c5_n8.json 2> $O/c5_n8.err
timeout 600 python -m pytest tests/test_gpu_multi_device.py -m gpu -x -q > $O/pytest_multi_device.log 2>&1; echo "rc=$?" >> $O/pytest_multi_device.log
tail -n 3 $O/pytest_multi_device.log
for f in $O/c*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], 'n', d['n_gpus'], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), d['roofline'].get('kernel'), d['roofline'].get('frac'))
except Exception as e: print(sys.argv[1], 'failed', e); print(open(sys.argv[1][:-5]+'.err').read()[-600:])
PY
done
