#!/bin/bash
# ncu --set full of selected kernels on a device-resident stream.  usage: gpu_ncu.sh tag kernel_regex [workload] [seconds] [count]
tag=$1; kre=$2; wl=${3:-c2}; secs=${4:-3600}; cnt=${5:-3}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$kre" -c $cnt -o gpurun_out/${tag} -f \
  python tools/profile_run.py $wl $secs 1 0 > gpurun_out/${tag}_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/${tag}_ncu.log
python tools/ncu_summary.py gpurun_out/${tag}.ncu-rep > gpurun_out/${tag}_summary.txt 2>&1
cat gpurun_out/${tag}_summary.txt | head -120
