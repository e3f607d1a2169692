#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
L=gpurun_out/r02_enc_ws.log; : > $L
timeout 600 python -m pytest tests/test_gpu_encoder.py -x -q 2>&1 | tail -8 >> $L
timeout 300 python profiles/enc_probe.py 2048 256 >> $L 2>&1
BFK_ENC_ATTN=ws timeout 300 python profiles/enc_probe.py 2048 256 >> $L 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_enc_ws.csv \
  python profiles/enc_probe.py 512 256 > gpurun_out/ncu_enc.log 2>&1
grep attention gpurun_out/r02_launches_enc_ws.csv | tail -3 | awk -F'","' '{print $5, $(NF)}' | cut -c1-100 >> $L
cat $L
