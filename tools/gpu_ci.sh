#!/bin/bash
# One gpurun call: GPU parity tests, then a short C2 bench. Logs land in gpurun_out/<tag>_*.
tag=${1:-ci}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/${tag}_pytest_gpu.log
python bench.py --steps 5 --warmup 3 "$@" > gpurun_out/${tag}_bench_c2.json 2> gpurun_out/${tag}_bench_c2.err
echo "bench rc=$?"; tail -5 gpurun_out/${tag}_bench_c2.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_bench_c2.json").read().strip().splitlines()[-1])
    print("value",d["value"],"ms",d["ms_per_step"],d["kernel_ms"],"e2e",d.get("e2e") and d["e2e"]["ms_per_step"],"md5",d.get("ms_per_step_with_md5"))
except Exception as e: print("no bench line",e)
PY
