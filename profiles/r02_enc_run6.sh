#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
BFK_ENC_GEMM_DEBUG=1 timeout 300 python profiles/enc_probe.py 512 256 2>&1 | grep -A6 "CTA 10" | tail -14 | tee gpurun_out/r02_enc_gemm_boundary.log
