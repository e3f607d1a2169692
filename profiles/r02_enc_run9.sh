#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
for a in 1 0; do echo "== BFK_ENC_ARES=$a"; BFK_ENC_ARES=$a BFK_ENC_GEMM_DEBUG=1 timeout 300 python profiles/enc_probe.py 512 256 2>&1 | grep "EPI=[01]" | tail -2 | cut -c1-330; done | tee gpurun_out/r02_enc_ares_ab.log
