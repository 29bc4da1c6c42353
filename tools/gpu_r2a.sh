#!/bin/bash
tag=${1:-r2a}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${tag}_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench_c2.json 2> gpurun_out/${tag}_bench_c2.err; echo "bench c2 rc=$?"; tail -3 gpurun_out/${tag}_bench_c2.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_bench_c2.json").read().strip().splitlines()[-1])
print("value",d["value"],"ms",d["ms_per_step"],d["kernel_ms"]); print("gated",d["gated"]); print("e2e",d["e2e"]); print("cpu",d["cpu_baseline"])
PY
