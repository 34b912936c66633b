#!/bin/bash
# v17: A/B of the encode emit kernel (QUBER_ENCODE_EMIT=0 old, 1 early last-mask load, 2 early + 5 CTAs/SM), parity tests on the default
mkdir -p gpurun_out
: > gpurun_out/emit_ab_v17.txt
for m in 0 1 2 1 0; do
  QUBER_ENCODE_EMIT=$m timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/sweep.json 2> gpurun_out/sweep.err
  python - $m >> gpurun_out/emit_ab_v17.txt <<P
import json, sys
d = json.load(open('gpurun_out/sweep.json'))
print("emit_mode=%s value=%.0f ms=%.4f encode_ms=%.4f e2e_ms=%.4f" % (sys.argv[1], d['value'], d['ms_per_step'], d['stage_ms']['encode'], d['e2e']['ms_per_step']))
P
done
cat gpurun_out/emit_ab_v17.txt
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_v17.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_v17.log; tail -3 gpurun_out/pytest_gpu_v17.log
for m in 0 2; do QUBER_ENCODE_EMIT=$m timeout 200 python -m pytest tests -m gpu -x -q -k "encode or golden or config" > gpurun_out/pytest_emit$m.log 2>&1; tail -1 gpurun_out/pytest_emit$m.log; done
