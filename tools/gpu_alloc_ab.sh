#!/bin/bash
# A/B of the device-memory cache (devcache.cuh) against the driver's stream-ordered pool, on the e2e figures.
tag=${1:-ab}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest_gpu.log
for alloc in cache pool; do
  for wl in c2 c3 c5; do
    ZFLACB200_ALLOC=$alloc python bench.py --workload $wl --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_${alloc}_${wl}.json 2> gpurun_out/${tag}_${alloc}_${wl}.err; echo "bench $alloc $wl rc=$?"
  done
  ZFLACB200_ALLOC=$alloc ZFLACB200_BENCH_TWO_CALL=1 ZFLACB200_TRACE=1 python bench.py --workload c5 --steps 3 --warmup 2 --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_${alloc}_c5two.json 2> gpurun_out/${tag}_${alloc}_c5two.err
done
python - <<PY
import json
for alloc in ("cache","pool"):
    for wl in ("c2","c3","c5","c5two"):
        try:
            d=json.loads(open("gpurun_out/${tag}_%s_%s.json"%(alloc,wl)).read().strip().splitlines()[-1])
            print(alloc,wl,"value",d["value"],"ms",d["ms_per_step"],"e2e",d.get("e2e") and d["e2e"]["ms_per_step"],"launches",d["gpu_launches"])
        except Exception as e: print(alloc,wl,"no bench line",e)
PY
tail -4 gpurun_out/${tag}_cache_c5two.err | cut -c1-260
tail -4 gpurun_out/${tag}_pool_c5two.err | cut -c1-260
