#!/bin/bash
# N-GPU pass: the driver's launch line for bench.py at N ranks, ours and the reference arm.
N=${1:-2}; TAG=${2:-n2}
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_${TAG}_1gpu.json 2> gpurun_out/bench_${TAG}_1gpu.err; echo "1gpu rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "N=$N rc=$?"
tail -5 gpurun_out/bench_${TAG}.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err; echo "ref rc=$?"
python - <<P
import json
for f in ['gpurun_out/bench_${TAG}_1gpu.json','gpurun_out/bench_${TAG}.json','gpurun_out/bench_ref_${TAG}.json']:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k:d.get(k) for k in ['n_gpus','value','ms_per_step','e2e','stage_ms','network','clocks']})
    except Exception as e: print(f, 'ERR', e, open(f).read()[-500:])
P
