#!/bin/bash
# Hang-safe GPU test run (used under gpurun): per-test timeout kills the process, log goes to a file.
mkdir -p gpurun_out
timeout -s KILL ${2:-600} python -m pytest tests -m gpu -q --timeout=${1:-90} --timeout-method=thread -p no:cacheprovider ${@:3} > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 60 gpurun_out/pytest_gpu.log
