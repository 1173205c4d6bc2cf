#!/bin/bash
# GPU parity tests only; optional pytest -k expression as $1.
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests -q -m gpu -x --timeout=300 -p no:cacheprovider ${1:+-k "$1"} > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"
tail -40 gpurun_out/pytest_gpu.log
