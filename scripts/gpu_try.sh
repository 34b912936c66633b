#!/bin/bash
# quick GPU pass: parity tests, MSDA broadcast variants (A/B), encoder aside, bench line (no profiler)
TAG=${1:-try}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -5 gpurun_out/pytest_gpu_$TAG.log
QUBER_MSDA_BCAST=rec4 timeout 600 python -m pytest tests -m gpu -x -q -k "msda or encoder or pixel_decoder" > gpurun_out/pytest_rec4_$TAG.log 2>&1; echo "rec4 pytest rc=$?"; tail -3 gpurun_out/pytest_rec4_$TAG.log
for bc in lds rec4; do for dt in f32 bf16; do for mb in 4 5; do
  echo "BCAST=$bc MINB=$mb $dt: $(QUBER_MSDA_BCAST=$bc QUBER_MSDA_MINB=$mb timeout 300 python scripts/bench_msda.py 32 $dt 2>&1 | tr '\n' ' ')"
done; done; done | tee gpurun_out/msda_variants_$TAG.log
timeout 300 python scripts/bench_encoder.py 2>&1 | grep "encoder x6" | tee gpurun_out/bench_encoder_$TAG.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_$TAG.log 2>&1; tail -2 gpurun_out/smoke_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
python - <<P
import json
d=json.load(open('gpurun_out/bench_$TAG.json'))
for k in ['value','ms_per_step','e2e','stage_ms','roofline','encoder']: print(k, d.get(k))
P
