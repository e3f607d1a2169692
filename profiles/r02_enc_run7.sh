#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_encoder.py tests/test_gpu_retrieval.py -x -q 2>&1 | tail -3
timeout 300 python profiles/enc_probe.py 2048 256
timeout 300 python profiles/enc_probe.py 2048 64
