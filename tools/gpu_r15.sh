#!/bin/bash
# Closing checkpoint of round 2 on one GPU: the driver's own commands (pytest -m gpu, smoke, default bench), the C3 and C5 lines,
# the ncu launch list of the bench command and a --set full summary of the C2 kernels (pass B changed after r08).
tag=${1:-r15}; o=gpurun_out/${tag}
mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q > ${o}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 ${o}_pytest_gpu.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > ${o}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 ${o}_smoke.log
timeout 120 python bench.py > ${o}_bench_default.json 2> ${o}_bench_default.err; echo "bench default rc=$?"
timeout 60 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline > ${o}_bench_c3.json 2> ${o}_bench_c3.err; echo "bench c3 rc=$?"
timeout 90 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file ${o}_launches_c2.csv \
  python bench.py --steps 2 --warmup 1 --md5-steps 0 --no-cpu-baseline --no-strong --no-e2e > ${o}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 120 ncu --set full --clock-control none --import-source on -k regex:"k_sync_scan|k_hint_chain|k_parse_frames|k_predict_epilogue|k_verify_chain" -c 6 \
  -o ${o}_c2_kernels -f python tools/profile_run.py c2 3600 1 0 > ${o}_c2_kernels_ncu.log 2>&1; echo "ncu c2 rc=$?"
python tools/ncu_summary.py ${o}_c2_kernels.ncu-rep > ${o}_c2_kernels_ncu_full.txt 2>&1
rm -f ${o}_*.ncu-rep
timeout 90 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > ${o}_bench_c5.json 2> ${o}_bench_c5.err; echo "bench c5 rc=$?"
python - <<PY
import json
for n in ("bench_default","bench_c3","bench_c5"):
    try:
        d=json.loads(open("${o}_%s.json"%n).read().strip().splitlines()[-1])
        print(n,"value",d.get("value"),"ms",d.get("ms_per_step"),d.get("kernel_ms"),"e2e",d.get("e2e") and d["e2e"].get("ms_per_step"),"gated",d.get("gated") and (d["gated"].get("value"),d["gated"].get("vs_cpu")), "roofline", d.get("roofline") and (d["roofline"].get("kernel"), d["roofline"].get("frac")))
        for k in ("strong_c2r","strong_c5"):
            if d.get(k): print("  ",k,d[k]["value"],d[k]["ms_per_step"],d[k]["e2e"]["value"],d[k]["e2e"]["ms_per_step"])
    except Exception as e: print(n,"no line",e)
PY
