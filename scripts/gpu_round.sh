#!/bin/bash
# One GPU-box pass, parameterised:  scripts/gpu_round.sh TAG [steps...]
#   steps (default "test smoke bench"): test | testk:<one-word pytest -k expr> | smoke | bench | bench:<c2|c3|c4|c5> | ref | launches | full[:regex] | extra
# Everything lands in gpurun_out/ with the TAG in the name; copy what should be judged into profiles/.
# ncu steps run only after the same bench command has exited 0 without ncu in this call (the `bench` step).
TAG=${1:-r02}; shift
STEPS=${@:-test smoke bench}
mkdir -p gpurun_out
BENCH_OK=0
for s in $STEPS; do
  case $s in
    test)   timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log; tail -5 gpurun_out/pytest_gpu_$TAG.log ;;
    testk:*) timeout 1200 python -m pytest tests -m gpu -x -q -s -k "${s#testk:}" > gpurun_out/pytest_k_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_k_$TAG.log; tail -30 gpurun_out/pytest_k_$TAG.log ;;
    smoke)  timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_$TAG.log 2>&1; tail -2 gpurun_out/smoke_$TAG.log ;;
    bench)  timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err && BENCH_OK=1; echo "bench ok=$BENCH_OK"; cat gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err ;;
    bench:*) CFG=${s#bench:}; timeout 900 python bench.py --config $CFG --steps 20 --warmup 3 --no-network > gpurun_out/bench_${CFG}_$TAG.json 2> gpurun_out/bench_${CFG}_$TAG.err; echo "bench $CFG rc=$?"; cat gpurun_out/bench_${CFG}_$TAG.json; tail -3 gpurun_out/bench_${CFG}_$TAG.err ;;
    ref)    timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; cat gpurun_out/bench_ref_$TAG.json ;;
    launches) timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/plain_l_$TAG.log 2>&1 &&
            timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/ncu_l_$TAG.log 2>&1; tail -2 gpurun_out/ncu_l_$TAG.log ;;
    full*)  RX=${s#full}; RX=${RX#:}; RX=${RX:-msda|boundary|encode|group|emit|nms|rle}
            timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/plain_f_$TAG.log 2>&1 &&
            timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RX" -s 40 -c 14 -o gpurun_out/full_$TAG -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-network > gpurun_out/ncu_f_$TAG.log 2>&1; tail -2 gpurun_out/ncu_f_$TAG.log ;;
    extra)  for f in bench_msda bench_gemm bench_encoder bench_norm; do [ -f scripts/$f.py ] && timeout 300 python scripts/$f.py > gpurun_out/${f}_$TAG.txt 2>&1; grep -v "^-\|^ *Name\|^Self\|cudaLaunch\|^$" gpurun_out/${f}_$TAG.txt | head -30; done
            timeout 300 python scripts/bench_msda.py 32 bf16 > gpurun_out/bench_msda_bf16_$TAG.txt 2>&1; cat gpurun_out/bench_msda_bf16_$TAG.txt ;;
    *) echo "unknown step $s" ;;
  esac
done
ls -la gpurun_out | tail -20
