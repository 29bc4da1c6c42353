#!/bin/bash
# Quick GPU iteration: parity tests, then short device-resident benches (no MD5 variant, no CPU baseline).
tag=${1:-q}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${tag}_pytest_gpu.log
for wl in c2 c3 c5; do
  python bench.py --workload $wl --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_bench_${wl}.json 2> gpurun_out/${tag}_bench_${wl}.err; echo "bench $wl rc=$?"; tail -3 gpurun_out/${tag}_bench_${wl}.err
done
ZFLACB200_EXTRA_FLAGS=20 python bench.py --workload c2 --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_bench_c2_single.json 2> gpurun_out/${tag}_bench_c2_single.err
python - <<PY
import json
for wl in ("c2","c3","c5","c2_single"):
    try:
        d=json.loads(open("gpurun_out/${tag}_bench_%s.json"%wl).read().strip().splitlines()[-1])
        print(wl,"value",d["value"],"ms",d["ms_per_step"],d["kernel_ms"],"e2e",d.get("e2e") and d["e2e"]["ms_per_step"],"launches",d["gpu_launches"])
    except Exception as e: print(wl,"no bench line",e)
PY
ZFLACB200_BENCH_TWO_CALL=1 ZFLACB200_TRACE=1 python bench.py --workload c5 --steps 3 --warmup 2 --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_bench_c5_twocall.json 2> gpurun_out/${tag}_bench_c5_twocall.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_bench_c5_twocall.json").read().strip().splitlines()[-1]); print("c5 two-call e2e", d["e2e"]["ms_per_step"])
except Exception as e: print("no two-call line", e)
PY
ZFLACB200_TRACE=1 python bench.py --workload c5 --steps 2 --warmup 1 --md5-steps 0 --no-cpu-baseline > /dev/null 2> gpurun_out/${tag}_c5_trace.err; grep -c . gpurun_out/${tag}_c5_trace.err
