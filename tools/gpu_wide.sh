#!/bin/bash
# position reader with int32 rows (17..31-bit streams): parity tests, then C3 / C2 / C5 benches and the A/B with the word pipeline
tag=${1:-wide}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${tag}_pytest_gpu.log
for wl in c3 c2 c5; do
  python bench.py --workload $wl --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline --no-strong > gpurun_out/${tag}_bench_${wl}.json 2> gpurun_out/${tag}_bench_${wl}.err; echo "bench $wl rc=$?"; tail -2 gpurun_out/${tag}_bench_${wl}.err
done
ZFLACB200_EXTRA_FLAGS=400 python bench.py --workload c3 --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline --no-e2e --no-strong > gpurun_out/${tag}_bench_c3_wp.json 2> gpurun_out/${tag}_bench_c3_wp.err
python - <<PY
import json
for wl in ("c3","c3_wp","c2","c5"):
    try:
        d=json.loads(open("gpurun_out/${tag}_bench_%s.json"%wl).read().strip().splitlines()[-1])
        print(wl,"value",d["value"],"ms",d["ms_per_step"],d["kernel_ms"],"e2e",d.get("e2e") and d["e2e"]["ms_per_step"],d.get("roofline",{}).get("kernel"))
    except Exception as e: print(wl,"no bench line",e)
PY
