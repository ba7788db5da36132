P=ray-tracing-project_b200
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
pp() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1',round(d['value'],1),d['roofline']['kernels_ms'])"; }
timeout 100 python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>&1 | pp mb8
for mb in 6 5 4; do RTP_LIB=$PWD/$P/libvar_mb$mb.so timeout 100 python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>&1 | pp mb$mb; done
for r in 1 16; do RTP_REFILL=$r timeout 100 python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>&1 | pp refill$r; done
timeout 100 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --workload c2 2>&1 | pp c2_mb8
RTP_LIB=$PWD/$P/libvar_mb5.so timeout 100 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --workload c2 2>&1 | pp c2_mb5
timeout 400 ncu --set full --import-source on --clock-control none -k regex:k_trace -s 8 -c 2 -o gpurun_out/c3_v2b_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; tail -3 gpurun_out/ncu_full.log
