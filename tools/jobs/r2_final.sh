#!/bin/bash
# GPU job (1 GPU): whole GPU suite, smoke, the default bench line, exact-mode benches, the batch sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gputests_d.log 2>&1; tail -3 gpurun_out/r2_gputests_d.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -3 gpurun_out/r2_smoke.log
python bench.py --layers-out gpurun_out/r2_layers_b256_fp16.json > gpurun_out/r2_bench_d.json 2> gpurun_out/r2_bench_d.err
python bench.py --mode exact --layers-out gpurun_out/r2_layers_exact_tf32.json > gpurun_out/r2_bench_exact_tf32.json 2> gpurun_out/r2_bench_exact_tf32.err
python bench.py --mode exact --exact-simt --no-cpu-baseline --sustain-s 0 --steps 5 > gpurun_out/r2_bench_exact_simt.json 2> gpurun_out/r2_bench_exact_simt.err
timeout 900 python tools/sweep.py --out gpurun_out/r2_sweep.json > /dev/null 2> gpurun_out/r2_sweep.err
python - <<'PY'
import json
for f in ("r2_bench_d", "r2_bench_exact_tf32", "r2_bench_exact_simt"):
    try:
        j = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        print(f, round(j["ms_per_step"], 3), round(j["value"]), "e2e", round(j["e2e"]["value"]), "parity", j["e2e"]["parity"]["max_abs"],
              "sustained", j["sustained"] and round(j["sustained"]["images_per_s"]), "roof", round(j["roofline"]["achieved"], 1), j["roofline"]["traffic"])
    except Exception as e:
        print(f, "ERR", e)
PY
tail -9 gpurun_out/r2_sweep.err
# ncu launch list of one step (DRAM traffic stamp for bench.py's roofline.traffic)
export B200RN_GRAPH=0
CMD="python tools/profile_step.py --batch 256 --steps 1 --warmup 1"
MET=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 && \
ncu --metrics $MET --clock-control none -k regex:"conv|stem7x7|maxpool|gap_|input_xform" -s 51 -c 51 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
tail -1 gpurun_out/r2_ncu_list.log
