#!/bin/bash
# the random-encoder parity test first, then the whole GPU suite, then the default line
tag=${1:-r13}; o=gpurun_out/${tag}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_zz_degenerate.py -m gpu -x -q > ${o}_fuzz.log 2>&1; echo "fuzz rc=$?"; tail -25 ${o}_fuzz.log
timeout 900 python -m pytest tests -m gpu -x -q > ${o}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 ${o}_pytest_gpu.log
python bench.py > ${o}_bench_default.json 2> ${o}_bench_default.err; echo "bench default rc=$?"
python - <<PY
import json
for n in ("bench_default",):
    d=json.loads(open("${o}_%s.json"%n).read().strip().splitlines()[-1])
    print(n,"value",d.get("value"),"ms",d.get("ms_per_step"),d.get("kernel_ms"),"e2e",d["e2e"].get("ms_per_step"),d["e2e"].get("value"),"gated",d.get("gated") and (d["gated"].get("value"),d["gated"].get("vs_cpu")))
PY
