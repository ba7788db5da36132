#!/bin/bash
# round 2, GPU call B: no-alloc test alone + whole suite, ncu full captures of C2 / C4 / C5 (brought back as raw-page
# CSV: the .ncu-rep files of 20-50 launches exceed what gpurun copies back), compute-sanitizer passes.
set -u
O=gpurun_out/r02b; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> $O/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_c3.json 2> $O/bench_c3.err
timeout 300 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err
timeout 600 python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err
NCU="ncu --set full --clock-control none"
T=/tmp/ncu_r02b; mkdir -p $T
RTP_GROUPS=1 timeout 900 $NCU -k regex:"k_trace|k_coop" --launch-skip 16 -c 20 -o $T/ncu_c2 -f \
  python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c2.log 2>&1
RTP_GROUPS=1 timeout 1200 $NCU -k regex:"k_trace|k_coop" --launch-skip 40 -c 48 -o $T/ncu_c4 -f \
  python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c4.log 2>&1
timeout 900 $NCU -k regex:"k_trace" --launch-skip 3 -c 2 -o $T/ncu_c5 -f \
  python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c5.log 2>&1
for w in c2 c4 c5; do ncu -i $T/ncu_$w.ncu-rep --page raw --csv > $O/ncu_${w}_raw.csv 2>/dev/null; done
# sanitizers: default policy, then every ray forced through k_coop, then k_paths / per-level kernels forced
S=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck racecheck synccheck initcheck; do
  timeout 900 $S --tool $tool python profiles/sanitize_case.py > $O/san_${tool}_default.log 2>&1; echo "rc=$?" >> $O/san_${tool}_default.log
  RTP_BUDGET=1 RTP_PARK_LANES=32 RTP_PARK_ANY=1 RTP_PARK_SMALL=1 timeout 900 $S --tool $tool python profiles/sanitize_case.py > $O/san_${tool}_coop.log 2>&1; echo "rc=$?" >> $O/san_${tool}_coop.log
done
RTP_DEEP=paths timeout 900 $S --tool memcheck python profiles/sanitize_case.py > $O/san_memcheck_paths.log 2>&1; echo "rc=$?" >> $O/san_memcheck_paths.log
RTP_DEEP=paths timeout 900 $S --tool racecheck python profiles/sanitize_case.py > $O/san_racecheck_paths.log 2>&1; echo "rc=$?" >> $O/san_racecheck_paths.log
RTP_DEEP=wavefront timeout 900 $S --tool memcheck python profiles/sanitize_case.py > $O/san_memcheck_wavefront.log 2>&1; echo "rc=$?" >> $O/san_memcheck_wavefront.log
# a rank's share of the frame alone on one GPU (the floor of the N-GPU frame)
timeout 300 python profiles/share_time.py --workload c3 > $O/share_c3.log 2>&1
timeout 300 python profiles/share_time.py --workload c2 > $O/share_c2.log 2>&1
du -sh $O; ls -la $O
for f in $O/pytest_noalloc_alone.log $O/pytest_gpu.log $O/san_*.log $O/share_*.log; do echo "== $f"; tail -n 6 $f | cut -c1-300; done
