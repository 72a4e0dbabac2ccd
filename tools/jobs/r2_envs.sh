#!/bin/bash
# GPU job (1 GPU): per-layer times under a list of environment settings: usage r2_envs.sh "VAR=a" "VAR=b OTHER=c" ...
mkdir -p gpurun_out
i=0
for setting in "$@"; do
  env $setting python tools/profile_step.py --batch 256 --steps 6 --warmup 3 --layers-out gpurun_out/r2_env$i.json > gpurun_out/r2_env$i.log 2>&1
  echo "[$i] $setting: $(tail -1 gpurun_out/r2_env$i.log)"
  i=$((i+1))
done
python - "$#" <<'PY'
import json, sys
n = int(sys.argv[1])
L = [json.load(open("gpurun_out/r2_env%d.json" % i))["layers"] for i in range(n)]
print("layer sums", [round(sum(l["ms"] for l in x), 4) for x in L])
names = []
for x in L:
    for l in x:
        if l["name"] not in names:
            names.append(l["name"])
for nm in names:
    ms = [next((l["ms"] for l in x if l["name"] == nm), None) for x in L]
    have = [m for m in ms if m is not None]
    if len(have) < n or max(have) - min(have) > 0.002:
        print(nm[:44].ljust(44), " ".join("   -  " if m is None else "%.4f" % m for m in ms))
PY
