#!/bin/bash
# round 2, final 1-GPU set: parity suite, smoke, bench lines of every config by the driver's command, reference arm,
# ncu launch list + full captures (C3 with source, C2 / C1 / C5 as raw-page CSV), bounds-checked runs, shares
set -u
O=gpurun_out/r02final; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit,driver_version --format=csv > $O/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
timeout 400 python bench.py > $O/bench_c3_default.json 2> $O/bench_c3_default.err
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_c3.json 2> $O/bench_c3.err
for w in c1 c2; do timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_$w.json 2> $O/bench_$w.err; done
timeout 600 python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err
timeout 600 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c5.json 2> $O/bench_c5.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_c3_reference.json 2> $O/bench_c3_reference.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1
NCU="ncu --set full --clock-control none"
T=/tmp/ncu_final; mkdir -p $T
RTP_GROUPS=1 timeout 900 $NCU --import-source on -k regex:"k_trace|k_coop" --launch-skip 9 -c 4 -o $O/ncu_c3 -f \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c3.log 2>&1
RTP_GROUPS=1 timeout 900 $NCU -k regex:"k_camera|k_coop_camera|k_trace|k_coop|k_paths" --launch-skip 14 -c 24 -o $T/ncu_c2 -f \
  python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c2.log 2>&1
RTP_GROUPS=1 timeout 900 $NCU -k regex:"k_camera|k_coop_camera|k_trace|k_coop|k_paths" --launch-skip 6 -c 10 -o $T/ncu_c1 -f \
  python bench.py --workload c1 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c1.log 2>&1
timeout 900 $NCU --kernel-name-base demangled -k regex:"RaysAny" --launch-skip 3 -c 1 -o $T/ncu_c5 -f \
  python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c5.log 2>&1
for w in c2 c1 c5; do ncu -i $T/ncu_$w.ncu-rep --page raw --csv > $O/ncu_${w}_raw.csv 2>/dev/null; done
timeout 900 python profiles/checked_run.py > $O/checked_default.log 2>&1
RTP_BUDGET=1 RTP_PARK_LANES=32 RTP_PARK_ANY=1 RTP_PARK_SMALL=1 timeout 900 python profiles/checked_run.py > $O/checked_coop.log 2>&1
RTP_DEEP=paths timeout 900 python profiles/checked_run.py > $O/checked_paths.log 2>&1
RTP_FUSED=0 RTP_DEEP=wavefront timeout 900 python profiles/checked_run.py > $O/checked_separate_wavefront.log 2>&1
RTP_LIB=ray-tracing-project_b200/libflyscene_b200_checked.so timeout 300 python profiles/sanitize_case.py > $O/checked_sequences.log 2>&1
timeout 300 python profiles/share_time.py --workload c3 > $O/share_c3.log 2>&1
du -sh $O; tail -n 3 $O/pytest_gpu.log $O/smoke.log 2>/dev/null | cut -c1-200
for f in $O/checked_*.log; do echo "== $f"; tail -n 1 $f | cut -c1-300; done
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d.get('roofline') or {}; print(sys.argv[1].split('/')[-1], round(d['value'],3), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],3), r.get('kernel'), r.get('frac'), d.get('clocks',{}).get('samples') if d.get('clocks') else None)
except Exception as e: print(sys.argv[1], 'failed', e)
PY
done
