#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
L=gpurun_out/r02_enc_gemm_debug.log; : > $L
timeout 600 python -m pytest tests/test_gpu_encoder.py -x -q 2>&1 | tail -3 >> $L
timeout 300 python profiles/enc_probe.py 2048 256 >> $L 2>&1
BFK_ENC_GEMM_DEBUG=1 timeout 300 python profiles/enc_probe.py 512 256 2>&1 | grep "enc gemm" | tail -8 >> $L
cat $L
