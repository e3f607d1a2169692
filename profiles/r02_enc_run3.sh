#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
L=gpurun_out/r02_enc_exp.log; : > $L
for x in 0 1 2 4; do echo "== BFK_ENC_EXP=$x" >> $L; BFK_ENC_EXP=$x timeout 300 python profiles/enc_probe.py 2048 256 >> $L 2>&1; done
BFK_ENC_EXP=4 timeout 600 python -m pytest tests/test_gpu_encoder.py -x -q 2>&1 | tail -4 >> $L
cat $L
