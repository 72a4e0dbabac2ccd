#!/bin/bash
# GPU job (1 GPU): ncu launch list of one exact-mode (3 x TF32) forward step
mkdir -p gpurun_out
export B200RN_GRAPH=0
CMD="python tools/profile_step.py --batch 256 --steps 1 --warmup 1 --mode exact"
MET=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
$CMD > gpurun_out/r2_ncu_exact_plain.log 2>&1 && \
ncu --metrics $MET --clock-control none -k regex:"conv|stem7x7|maxpool|gap_|input_xform|frame" -s 57 -c 57 --csv --log-file gpurun_out/r2_launches_exact.csv $CMD > gpurun_out/r2_ncu_exact_list.log 2>&1
tail -1 gpurun_out/r2_ncu_exact_plain.log; tail -1 gpurun_out/r2_ncu_exact_list.log
