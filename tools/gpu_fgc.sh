#!/bin/bash
# CUDA generator check: bytes against the CPU generator, then timing of the full-size C2 build
tag=${1:-fgc}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_flacgen.py -q --no-header -p no:cacheprovider -s > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"
tail -150 gpurun_out/${tag}_pytest.log
