#!/bin/bash
# GPU job (1 GPU): full-set ncu capture (with source) of the wide-epilogue kernels' launches of one forward step
mkdir -p gpurun_out
export B200RN_GRAPH=0
CMD="python tools/profile_step.py --batch 256 --steps 1 --warmup 1"
$CMD > gpurun_out/r2_ncu_plain4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"conv_wide|conv_pairwide" -s 14 -c 14 -f -o gpurun_out/r2_wide $CMD > gpurun_out/r2_ncu_wide.log 2>&1
tail -2 gpurun_out/r2_ncu_wide.log; ls -la gpurun_out/*.ncu-rep
