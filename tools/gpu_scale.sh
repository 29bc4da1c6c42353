#!/bin/bash
# The driver's scaling command at one N: the default bench line (weak C2 + strong_c2r + strong_c5), saved under gpurun_out/
tag=${1:-scale}; n=${2:-2}; port=$((29700 + n))
mkdir -p gpurun_out
t0=$(date +%s)
if [ "$n" = 1 ]; then
  python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/${tag}_default_n${n}.json 2> gpurun_out/${tag}_default_n${n}.err
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n \
    --steps 10 --warmup 3 > gpurun_out/${tag}_default_n${n}.json 2> gpurun_out/${tag}_default_n${n}.err
fi
echo "bench n=$n rc=$? wall=$(( $(date +%s) - t0 )) s"; tail -3 gpurun_out/${tag}_default_n${n}.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_default_n${n}.json").read().strip().splitlines()[-1])
    print("value",d["value"],"ms",d["ms_per_step"],"e2e",d["e2e"]["value"],d["e2e"]["ms_per_step"],"floor",d["e2e"].get("pcie_floor_ms"))
    for k in ("strong_c2r","strong_c5"):
        s=d.get(k)
        if s: print(k,"value",s["value"],"eff",s.get("efficiency_vs_n1"),"e2e",s["e2e"]["value"],"e2e_eff",s.get("e2e_efficiency_vs_n1"),"limit",s.get("limiting_kernel"))
    print("gated",d.get("gated")); print("gen",d["workload_detail"].get("generator"))
except Exception as e: print("no line",e)
PY
