#!/bin/bash
# v16 GPU pass, short form (fits the GPU-minutes left in the round): parity tests, smoke, bench with the defaults,
# reference arm, ncu launch list, then the kernel-group sweep of the host-buffer call as time allows.
TAG=${1:-v16}
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -4 gpurun_out/pytest_gpu_$TAG.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke_$TAG.log 2>&1; tail -2 gpurun_out/smoke_$TAG.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cat gpurun_out/bench_$TAG.json
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; cat gpurun_out/bench_ref_$TAG.json
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/ncu_l_$TAG.log 2>&1
: > gpurun_out/e2e_sweep_$TAG.txt
for cfg in "1 0" "1 16" "1 32" "0 0"; do
  set -- $cfg
  QUBER_LABEL_U8=$1 QUBER_E2E_KERNEL_FRAMES=$2 timeout 120 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/sweep.json 2> gpurun_out/sweep.err
  python - "$1" "$2" >> gpurun_out/e2e_sweep_$TAG.txt <<P
import json, sys
try:
    d = json.load(open('gpurun_out/sweep.json'))
    print("label_u8=%s kernel_frames=%s value=%.0f ms=%.4f post_ms=%.4f e2e=%.0f e2e_ms=%.4f d2h=%d" % (sys.argv[1], sys.argv[2], d['value'], d['ms_per_step'], d['stage_ms']['post'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['d2h_bytes_per_step']))
except Exception as e:
    print("label_u8=%s kernel_frames=%s FAILED %r" % (sys.argv[1], sys.argv[2], e))
P
done
cat gpurun_out/e2e_sweep_$TAG.txt
