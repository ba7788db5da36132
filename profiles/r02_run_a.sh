#!/bin/bash
# round 2, GPU call A: parity suite, bench lines of every config, launch list + full ncu capture of C3.
# Run from the repo root on the GPU box: bash profiles/r02_run_a.sh
set -u
O=gpurun_out/r02a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python bench.py > $O/bench_c3.json 2> $O/bench_c3.err; echo "rc=$?" >> $O/bench_c3.err
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_c3_driver.json 2> $O/bench_c3_driver.err
for w in c1 c2; do timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err; done
timeout 600 python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_c3_reference.json 2> $O/bench_c3_reference.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1
RTP_GROUPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace|k_coop|k_shade|k_paths" --launch-skip 8 -c 4 \
  -o $O/ncu_c3 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_full.log 2>&1
ls -la $O
tail -3 $O/pytest_gpu.log; cat $O/bench_c3.json
