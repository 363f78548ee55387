#!/bin/bash
# Runs on the GPU box: GPU tests, then the bench plain, then the same command under ncu (launch list + one full capture).
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
CMD="python bench.py --steps 2 --warmup 3 --cpu-budget 0.5 --e2e-requests 1 ${BENCH_ARGS:-}"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KERNEL:-k_copy} -s ${SKIP:-8} -c 3 -f -o gpurun_out/prof_${TAG:-copy} $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"
tail -2 gpurun_out/plain.log
