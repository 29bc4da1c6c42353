#!/bin/bash
# One gpurun call for a round checkpoint: GPU parity tests, C2 bench (both arms), C3 + C5 bench, ncu launch list.
tag=${1:-r01f}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench_c2.json 2> gpurun_out/${tag}_bench_c2.err; echo "bench c2 rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference_c2.json 2> gpurun_out/${tag}_bench_reference_c2.err; echo "ref rc=$?"
python bench.py --workload c3 --steps 5 --warmup 3 --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_bench_c3.json 2> gpurun_out/${tag}_bench_c3.err; echo "bench c3 rc=$?"
python bench.py --workload c5 --steps 5 --warmup 3 > gpurun_out/${tag}_bench_c5.json 2> gpurun_out/${tag}_bench_c5.err; echo "bench c5 rc=$?"
python bench.py --workload c2r --steps 5 --warmup 3 > gpurun_out/${tag}_bench_c2r.json 2> gpurun_out/${tag}_bench_c2r.err; echo "bench c2r rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_c2.csv \
  python bench.py --steps 2 --warmup 1 --md5-steps 0 --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_parse_frames|k_predict_epilogue|k_sync_scan" -c 4 -o gpurun_out/${tag}_c2_full \
  python tools/profile_run.py c2 3600 1 0 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -c 1500 gpurun_out/${tag}_bench_c2.json
