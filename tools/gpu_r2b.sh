#!/bin/bash
# narrow K2a check: parity tests, then C2/C3/C5 short benches (no MD5 / CPU legs), then the A/B with the int32 path
tag=${1:-r2b}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${tag}_pytest_gpu.log
for wl in c2 c3 c5; do
  python bench.py --workload $wl --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_bench_${wl}.json 2> gpurun_out/${tag}_bench_${wl}.err; echo "bench $wl rc=$?"; tail -3 gpurun_out/${tag}_bench_${wl}.err
done
ZFLACB200_NO_NARROW=1 python bench.py --workload c2 --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_bench_c2_wide.json 2> gpurun_out/${tag}_bench_c2_wide.err
python - <<PY
import json
for wl in ("c2","c3","c5","c2_wide"):
    try:
        d=json.loads(open("gpurun_out/${tag}_bench_%s.json"%wl).read().strip().splitlines()[-1])
        print(wl,"value",d["value"],"ms",d["ms_per_step"],d["kernel_ms"],"e2e",d.get("e2e") and d["e2e"]["ms_per_step"],"launches",d["gpu_launches"])
    except Exception as e: print(wl,"no bench line",e)
PY
