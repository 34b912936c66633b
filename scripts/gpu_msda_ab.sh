#!/bin/bash
# MSDA variant A/B on one B200: parity (pytest -k msda) and time per launch for each broadcast variant / occupancy.
TAG=${1:-ab}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
tail -4 gpurun_out/pytest_gpu_$TAG.log
QUBER_MSDA_BCAST=rec4 QUBER_MSDA_BLK=1 timeout 600 python -m pytest tests -m gpu -x -q -k "msda or encoder or pixel_decoder" > gpurun_out/pytest_rec4blk_$TAG.log 2>&1; echo "rec4 blk pytest rc=$?"; tail -2 gpurun_out/pytest_rec4blk_$TAG.log
run() { echo "BCAST=$1 BLK=$4 MINB=$2 $3: $(QUBER_MSDA_BCAST=$1 QUBER_MSDA_MINB=$2 QUBER_MSDA_BLK=$4 timeout 300 python scripts/bench_msda.py 32 $3 2>&1 | head -1)"; }
{
  for dt in f32 bf16; do
    run lds 5 $dt 0; run rec4 5 $dt 0
    for mb in 3 4 5; do run rec4 $mb $dt 1; done
  done
} | tee gpurun_out/msda_variants_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
python - <<P
import json
d=json.load(open('gpurun_out/bench_$TAG.json'))
for k in ['value','ms_per_step','e2e','stage_ms']: print(k, d.get(k))
P
