#!/bin/bash
# GPU job (1 GPU): ncu launch list of one forward step + full-set captures of the pair and halo kernels; H2D ceiling at N=1
mkdir -p gpurun_out
export B200RN_GRAPH=0
CMD="python tools/profile_step.py --batch 256 --steps 1 --warmup 1"
MET=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 && \
ncu --metrics $MET --clock-control none -k regex:"conv_|stem7x7|maxpool|gap_|input_xform" -s 51 -c 51 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
$CMD > gpurun_out/r2_ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_pair -s 18 -c 6 -f -o gpurun_out/r2_pair $CMD > gpurun_out/r2_ncu_pair.log 2>&1
$CMD > gpurun_out/r2_ncu_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"halo|conv_tail|stem7x7" -s 9 -c 9 -f -o gpurun_out/r2_halo_tail_stem $CMD > gpurun_out/r2_ncu_halo.log 2>&1
cat gpurun_out/r2_ncu_plain.log | tail -2; tail -2 gpurun_out/r2_ncu_list.log; tail -2 gpurun_out/r2_ncu_pair.log; tail -2 gpurun_out/r2_ncu_halo.log
unset B200RN_GRAPH
python tools/h2d_ceiling.py > gpurun_out/r2_h2d_n1.json 2> gpurun_out/r2_h2d_n1.err; cat gpurun_out/r2_h2d_n1.json; tail -3 gpurun_out/r2_h2d_n1.err
ls -la gpurun_out/*.ncu-rep
