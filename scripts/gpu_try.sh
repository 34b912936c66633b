#!/bin/bash
# quick GPU pass: parity tests, MSDA occupancy variants, bench line (no profiler)
TAG=${1:-try}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -15 gpurun_out/pytest_gpu_$TAG.log
for mb in 5; do for dt in f32 bf16; do
  echo "MINB=$mb $dt: $(QUBER_MSDA_MINB=$mb timeout 300 python scripts/bench_msda.py 32 $dt 2>&1 | tr '\n' ' ')"
done; done | tee gpurun_out/msda_variants_$TAG.log
for dt in f32 bf16; do
  echo "BCAST=lds $dt: $(QUBER_MSDA_BCAST=lds timeout 300 python scripts/bench_msda.py 32 $dt 2>&1 | tr '\n' ' ')"
done | tee -a gpurun_out/msda_variants_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
python - <<P
import json
d=json.load(open('gpurun_out/bench_$TAG.json'))
for k in ['value','ms_per_step','e2e','stage_ms','roofline']: print(k, d.get(k))
P
