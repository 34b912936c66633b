#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (no profiler), then the ncu launch list and one full capture.
set -x
mkdir -p gpurun_out
TAG=${1:-v3}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -3 gpurun_out/pytest_gpu_$TAG.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_$TAG.log 2>&1; tail -2 gpurun_out/smoke_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cat gpurun_out/bench_$TAG.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; cat gpurun_out/bench_ref_$TAG.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/ncu_l_$TAG.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'msda|boundary|encode|group|emit|nms' -s 36 -c 12 -o gpurun_out/full_$TAG -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/ncu_f_$TAG.log 2>&1
ls -la gpurun_out
timeout 300 python scripts/bench_norm.py > gpurun_out/bench_norm_$TAG.txt 2>&1; tail -4 gpurun_out/bench_norm_$TAG.txt
for mb in 4 5 6; do echo "LDS MINB=$mb f32: $(QUBER_MSDA_MINB=$mb timeout 300 python scripts/bench_msda.py 32 f32 2>&1 | head -1)"; done | tee gpurun_out/msda_minb_$TAG.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gn_|upsample_concat' -c 8 -o gpurun_out/full_norm_$TAG -f python scripts/bench_norm.py > gpurun_out/ncu_norm_$TAG.log 2>&1
timeout 300 python scripts/bench_gemm.py > gpurun_out/bench_gemm_$TAG.txt 2>&1; cat gpurun_out/bench_gemm_$TAG.txt
timeout 300 python scripts/bench_encoder.py 2>&1 | grep "encoder x6" | tee gpurun_out/bench_encoder_$TAG.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gemm_bias_act|add_layer_norm|msda_prepare' -c 6 -o gpurun_out/full_gemm_$TAG -f python scripts/bench_encoder.py > gpurun_out/ncu_gemm_$TAG.log 2>&1
