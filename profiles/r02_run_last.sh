#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out && rm -f gpurun_out/test_notes.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu.log
tail -3 gpurun_out/r02_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -1
