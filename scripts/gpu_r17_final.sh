#!/bin/bash
# v17 final: parity tests, smoke, bench and launch list on the defaults (emit: early last-mask load; reduce: unrolled + redux; boundary as v16)
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_v17f.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_v17f.log; tail -3 gpurun_out/pytest_gpu_v17f.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke_v17f.log 2>&1; tail -1 gpurun_out/smoke_v17f.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_v17f.json 2> gpurun_out/bench_v17f.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_v17f.json
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v17f.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/ncu_l_v17f.log 2>&1
