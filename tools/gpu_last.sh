#!/bin/bash
# closing check: the GPU suite, smoke and the two bench lines whose numbers moved after the r08 checkpoint
tag=${1:-r09}; o=gpurun_out/${tag}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > ${o}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 ${o}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > ${o}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py > ${o}_bench_default.json 2> ${o}_bench_default.err; echo "bench default rc=$?"
python bench.py --workload c5 --steps 5 --warmup 3 > ${o}_bench_c5.json 2> ${o}_bench_c5.err; echo "bench c5 rc=$?"
python - <<PY
import json
for n in ("bench_default","bench_c5"):
    d=json.loads(open("${o}_%s.json"%n).read().strip().splitlines()[-1])
    print(n,"value",d.get("value"),"ms",d.get("ms_per_step"),d.get("kernel_ms"),"e2e",d["e2e"].get("ms_per_step"),d["e2e"].get("value"),"gated",d.get("gated") and (d["gated"].get("value"),d["gated"].get("vs_cpu")))
    for k in ("strong_c2r","strong_c5"):
        if d.get(k): print("  ",k,d[k]["value"],d[k]["ms_per_step"],d[k]["e2e"]["value"],d[k]["e2e"]["ms_per_step"])
PY
