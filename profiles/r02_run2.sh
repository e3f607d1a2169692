#!/bin/bash
# GPU run 2: decomposition of the step (device-clock stamps), spin wait A/B, pass-size experiments at nq=4096
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
L=gpurun_out/r02_step_probe2.log; : > $L
timeout 300 python profiles/step_probe.py 1 64 1024 4096 >> $L 2>&1
BFK_TC_SPIN=0 timeout 300 python profiles/step_probe.py 1 64 >> $L 2>&1
BFK_TC_PDL=0 timeout 300 python profiles/step_probe.py 1 64 >> $L 2>&1
for c in 8 16 32; do BFK_TC_CHUNK_TILES=$c timeout 300 python profiles/step_probe.py 2048 4096 >> $L 2>&1; done
BFK_TC_PAIR=0 BFK_TC_CHUNK_TILES=16 timeout 300 python profiles/step_probe.py 2048 4096 >> $L 2>&1
cat $L
timeout 600 python -m pytest tests/test_gpu_tensor.py tests/test_gpu_flat.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
