#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_s14.txt 2>&1
tail -25 gpurun_out/r2_pytest_s14.txt
