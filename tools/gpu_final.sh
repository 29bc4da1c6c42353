#!/bin/bash
# Round checkpoint on one GPU: the driver's own commands (pytest -m gpu, smoke, bench both arms), C3 / C5 lines, the ncu launch
# list of the bench command and --set full summaries of every hot kernel (C2, C3, C5, the device MD5 chain).
tag=${1:-r07}
mkdir -p gpurun_out
o=gpurun_out/${tag}
timeout 1500 python -m pytest tests -m gpu -x -q > ${o}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 ${o}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > ${o}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 ${o}_smoke.log
python bench.py --impl reference > ${o}_bench_reference_default.json 2> ${o}_bench_reference_default.err; echo "ref rc=$?"
python bench.py > ${o}_bench_default.json 2> ${o}_bench_default.err; echo "bench default rc=$?"; tail -2 ${o}_bench_default.err
python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline > ${o}_bench_c3.json 2> ${o}_bench_c3.err; echo "bench c3 rc=$?"
python bench.py --workload c5 --steps 5 --warmup 3 > ${o}_bench_c5.json 2> ${o}_bench_c5.err; echo "bench c5 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file ${o}_launches_c2.csv \
  python bench.py --steps 2 --warmup 1 --md5-steps 0 --no-cpu-baseline --no-strong --no-e2e > ${o}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
full() {  # name, kernel regex, workload, size, count, extra env
  timeout 900 env $6 ncu --set full --clock-control none --import-source on -k regex:"$2" -c $5 -o ${o}_$1 -f \
    python tools/profile_run.py $3 $4 1 ${7:-0} > ${o}_$1_ncu.log 2>&1; echo "ncu $1 rc=$?"
  python tools/ncu_summary.py ${o}_$1.ncu-rep > ${o}_$1_ncu_full.txt 2>&1
}
full c2_kernels "k_sync_scan|k_hint_chain|k_parse_frames|k_predict_epilogue|k_verify_chain" c2 3600 6 ""
full c3_kernels "k_sync_scan|k_parse_frames|k_predict_epilogue" c3 600 4 ""
full c5_kernels "k_parse_frames|k_predict_epilogue|k_md5_lanes" c5 10000 4 "" 1
full md5_chain "k_md5" c2 60 1 "ZFLACB200_EXTRA_FLAGS=100" 1
python - <<PY
import json
for n in ("bench_default","bench_reference_default","bench_c3","bench_c5"):
    try:
        d=json.loads(open("${o}_%s.json"%n).read().strip().splitlines()[-1])
        print(n,"value",d.get("value"),"ms",d.get("ms_per_step"),d.get("kernel_ms"),"e2e",d.get("e2e") and d["e2e"].get("ms_per_step"),"gated",d.get("gated") and (d["gated"].get("value"),d["gated"].get("vs_cpu")), "roofline", d.get("roofline") and (d["roofline"].get("kernel"), d["roofline"].get("frac")))
    except Exception as e: print(n,"no line",e)
PY
rm -f ${o}_*.ncu-rep
