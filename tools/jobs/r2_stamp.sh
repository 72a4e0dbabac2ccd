#!/bin/bash
# GPU job (1 GPU): GPU suite, a few batch sizes, and the ncu launch list of one step (DRAM traffic stamp for bench.py's roofline.traffic)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gputests_d.log 2>&1; tail -2 gpurun_out/r2_gputests_d.log
for b in 8 16 32 256; do python tools/profile_step.py --batch $b --steps 20 --warmup 5 2>&1 | tail -1; done
export B200RN_GRAPH=0
CMD="python tools/profile_step.py --batch 256 --steps 1 --warmup 1"
MET=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 && \
ncu --metrics $MET --clock-control none -k regex:"conv|stem7x7|maxpool|gap_|input_xform" -s 51 -c 51 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
tail -1 gpurun_out/r2_ncu_list.log
