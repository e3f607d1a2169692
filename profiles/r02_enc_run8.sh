#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
for t in 65536 131072 262144 524288; do echo "== BFK_ENC_MB_TOKENS=$t"; BFK_ENC_MB_TOKENS=$t timeout 300 python profiles/enc_probe.py 4096 256; done 2>&1 | tee gpurun_out/r02_enc_microbatch.log
