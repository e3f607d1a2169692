#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
L=gpurun_out/r02_f2.log; : > $L
timeout 900 python -m pytest tests/test_gpu_tensor.py tests/test_gpu_flat.py tests/test_gpu_selectors.py -x -q 2>&1 | tail -2 >> $L
for i in 1 2; do
timeout 600 python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c3', j['value'], j['ms_per_step'], j['roofline']['frac'], j['clocks'])" >> $L 2>&1
done
cat $L
