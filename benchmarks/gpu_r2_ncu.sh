#!/usr/bin/env bash
# One ncu capture per call (B200_PROFILING.md): plain run first, then --set full of the three
# emulated pipeline kernels and the TMA copy.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python benchmarks/ncu_target.py > gpurun_out/ncu_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'reduce_kernel|move_kernel|tma_copy' -s 3 -c 5 \
    -f -o gpurun_out/prof_r2 python benchmarks/ncu_target.py > gpurun_out/ncu_r2.log 2>&1
echo "exit $?"; tail -5 gpurun_out/ncu_plain.log; tail -8 gpurun_out/ncu_r2.log; ls -la gpurun_out/prof_r2.ncu-rep
