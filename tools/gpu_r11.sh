#!/bin/bash
# after the i64-container accumulator rule: the GPU suite and the C3 line (the CB = 4 pass B)
tag=${1:-r11}; o=gpurun_out/${tag}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > ${o}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 ${o}_pytest_gpu.log
python bench.py --workload c3 --steps 10 --warmup 3 > ${o}_bench_c3.json 2> ${o}_bench_c3.err; echo "bench c3 rc=$?"
python - <<PY
import json
d=json.loads(open("${o}_bench_c3.json").read().strip().splitlines()[-1])
print("value",d.get("value"),"ms",d.get("ms_per_step"),d.get("kernel_ms"),"e2e",d["e2e"].get("ms_per_step"))
PY
