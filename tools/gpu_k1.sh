#!/bin/bash
# warp-per-tile K1: parity tests, then C2 / C3 / C5 short benches and the A/B with the CTA-per-tile kernel
tag=${1:-k1}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest_gpu.log
for wl in c2 c3 c5; do
  python bench.py --workload $wl --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline --no-strong > gpurun_out/${tag}_bench_${wl}.json 2> gpurun_out/${tag}_bench_${wl}.err; echo "bench $wl rc=$?"; tail -2 gpurun_out/${tag}_bench_${wl}.err
done
true
python - <<PY
import json
for wl in ("c2","c3","c5","c2_cta"):
    try:
        d=json.loads(open("gpurun_out/${tag}_bench_%s.json"%wl).read().strip().splitlines()[-1])
        print(wl,"value",d["value"],"ms",d["ms_per_step"],d["kernel_ms"],"e2e",d.get("e2e") and d["e2e"]["ms_per_step"],"launches",d["gpu_launches"])
    except Exception as e: print(wl,"no bench line",e)
PY
