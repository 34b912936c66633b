#!/bin/bash
# v17: parity on the defaults (emit kernel with early last-mask load, 3-stage boundary ring), then A/B of
# QUBER_ENCODE_EMIT (0 old, 1 early, 2 early + 5 CTAs/SM) x QUBER_BOUNDARY_STAGES (2, 3, 4) x QUBER_ENCODE_REDUCE (0 old, 1 unrolled + redux, 2 + double-buffered), then parity on the alternatives.
# (QUBER_BOUNDARY_STAGES existed only in the experimental boundary kernel of this run: not kept)
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_v17.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_v17.log; tail -3 gpurun_out/pytest_gpu_v17.log
: > gpurun_out/ab_v17.txt
for cfg in "0 2 0" "1 3 0" "1 2 1" "2 3 2" "1 4 1" "1 3 2"; do
  set -- $cfg
  QUBER_ENCODE_EMIT=$1 QUBER_BOUNDARY_STAGES=$2 QUBER_ENCODE_REDUCE=$3 timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/sweep.json 2> gpurun_out/sweep.err
  python - $1 $2 $3 >> gpurun_out/ab_v17.txt <<P
import json, sys
try:
    d = json.load(open('gpurun_out/sweep.json'))
    print("emit=%s stages=%s reduce=%s value=%.0f ms=%.4f encode_ms=%.4f boundary_ms=%.4f e2e_ms=%.4f" % (sys.argv[1], sys.argv[2], sys.argv[3], d['value'], d['ms_per_step'], d['stage_ms']['encode'], d['stage_ms']['boundary'], d['e2e']['ms_per_step']))
except Exception as e:
    print("emit=%s stages=%s FAILED %r" % (sys.argv[1], sys.argv[2], e))
P
done
cat gpurun_out/ab_v17.txt
QUBER_ENCODE_EMIT=0 QUBER_BOUNDARY_STAGES=2 QUBER_ENCODE_REDUCE=1 timeout 200 python -m pytest tests -m gpu -x -q -k "encode or boundary or golden or config or parity" > gpurun_out/pytest_alt02.log 2>&1; tail -1 gpurun_out/pytest_alt02.log
QUBER_ENCODE_EMIT=2 QUBER_BOUNDARY_STAGES=4 QUBER_ENCODE_REDUCE=2 timeout 200 python -m pytest tests -m gpu -x -q -k "encode or boundary or golden or config or parity" > gpurun_out/pytest_alt24.log 2>&1; tail -1 gpurun_out/pytest_alt24.log
timeout 200 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_v17.json 2> gpurun_out/bench_v17.err; echo "bench rc=$?"
