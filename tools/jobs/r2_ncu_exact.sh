#!/bin/bash
# GPU job (1 GPU): ncu full-set capture of the first exact-mode tensor-core conv launches of a step (stage 1)
mkdir -p gpurun_out
export B200RN_GRAPH=0
CMD="python tools/profile_step.py --batch 256 --steps 1 --warmup 1 --mode exact"
$CMD > gpurun_out/r2_ncu_exact_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_tf32x3 -s 52 -c 7 -f -o gpurun_out/r2_tf32 $CMD > gpurun_out/r2_ncu_exact.log 2>&1
tail -2 gpurun_out/r2_ncu_exact_plain.log; tail -2 gpurun_out/r2_ncu_exact.log
