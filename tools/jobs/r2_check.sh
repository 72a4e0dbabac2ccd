#!/bin/bash
# GPU job (1 GPU): whole GPU test suite, smoke, default bench line, ncu launch list of one step
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gputests_c.log 2>&1; tail -4 gpurun_out/r2_gputests_c.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -3 gpurun_out/r2_smoke.log
python bench.py > gpurun_out/r2_bench_c.json 2> gpurun_out/r2_bench_c.err; tail -c 400 gpurun_out/r2_bench_c.json
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r2_bench_refarm.json 2> gpurun_out/r2_bench_refarm.err; head -c 300 gpurun_out/r2_bench_refarm.json
export B200RN_GRAPH=0
CMD="python tools/profile_step.py --batch 256 --steps 1 --warmup 1"
MET=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 && \
ncu --metrics $MET --clock-control none -k regex:"conv|stem7x7|maxpool|gap_|input_xform" -s 51 -c 51 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
tail -2 gpurun_out/r2_ncu_list.log
