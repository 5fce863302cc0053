#!/bin/bash
# One GPU-box pass: parity tests, bench (ours + reference), ncu launch list, ncu --set full of the bank kernel.
# usage (from the repo root on the box): bash tools/gpu_check.sh [tag]     env: SKIP_TESTS SKIP_REF SKIP_NCU SKIP_FULL BENCH_ARGS KERNEL (regex of the
# kernel captured with --set full; default the tensor-core table GEMM)
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
if [ "${SKIP_TESTS:-0}" != "1" ]; then
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -15 $OUT/pytest_gpu.log
fi
timeout 600 python bench.py ${BENCH_ARGS:-} > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
tail -c 6000 $OUT/bench.json; tail -5 $OUT/bench.err
if [ "${SKIP_REF:-0}" != "1" ]; then
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "bench ref rc=$?"
fi
if [ "${SKIP_NCU:-0}" != "1" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-worst --no-alt --sustain 0 ${BENCH_ARGS:-} > $OUT/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
fi
if [ "${SKIP_FULL:-0}" != "1" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${KERNEL:-k_tgemm_tc} -s 3 -c 1 -o $OUT/top_full -f \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-worst --no-alt --sustain 0 ${BENCH_ARGS:-} > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
