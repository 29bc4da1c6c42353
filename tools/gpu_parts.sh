#!/bin/bash
# C5 e2e through zflacb200_batch_decode for several sub-batch counts.
tag=${1:-parts}
mkdir -p gpurun_out
for p in 1 2 3 4 5 6 8; do
  ZFLACB200_BATCH_PARTS=$p python bench.py --workload c5 --steps 4 --warmup 2 --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_p${p}.json 2> gpurun_out/${tag}_p${p}.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_p${p}.json").read().strip().splitlines()[-1]); print("parts $p: step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"])
except Exception as e: print("parts $p: no line", e)
PY
done
