#!/bin/bash
# What caps end-to-end at N GPUs: the C2 e2e leg alone, default chunking vs fewer chunks, with the library's host trace
tag=${1:-n8e2e}; n=${2:-8}
mkdir -p gpurun_out
{ echo "nproc $(nproc)"; cat /sys/fs/cgroup/cpu.max 2>/dev/null; lscpu | grep -E "^CPU\(s\)|Socket|NUMA|Model name|Thread"; nvidia-smi topo -m; free -g | head -2; } > gpurun_out/${tag}_host.txt 2>&1
run() {  # name, env...
  name=$1; shift
  port=$((29800 + RANDOM % 100))
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n \
    --steps 5 --warmup 2 --no-strong --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_${name}.json 2> gpurun_out/${tag}_${name}.err
  echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_${name}.json").read().strip().splitlines()[-1])
    print("  value",d["value"],"ms",d["ms_per_step"],"e2e",d["e2e"]["value"],"ms",d["e2e"]["ms_per_step"],"floor",d["e2e"].get("pcie_floor_ms"),d["e2e"].get("numa"))
except Exception as e: print("  no line",e)
PY
  grep "zflacb200 trace" gpurun_out/${tag}_${name}.err | tail -4 | cut -c1-400
}
run default ZFLACB200_TRACE=1
run chunk48 ZFLACB200_TRACE=1 ZFLACB200_PIPELINE_CHUNK=50331648
run chunk96 ZFLACB200_PIPELINE_CHUNK=100663296
cat gpurun_out/${tag}_host.txt | head -40
