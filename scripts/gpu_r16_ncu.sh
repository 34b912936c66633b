#!/bin/bash
# v16: one ncu full-set capture of the hot-path kernels of a bench step (after the plain run exited 0 in gpu_r16_short.sh)
TAG=${1:-v16}
mkdir -p gpurun_out
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'msda|boundary|encode|group|emit|nms' -s 36 -c 12 -o gpurun_out/full_$TAG -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/ncu_f_$TAG.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out | tail -5
