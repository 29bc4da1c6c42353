#!/bin/bash
# the no-alloc entry point: its test alone in a fresh process, then after every other GPU test (pytest -x, as the round's end runs it)
tag=${1:-noalloc}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${tag}_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/${tag}_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "all rc=$?"; tail -8 gpurun_out/${tag}_pytest_gpu.log
