#!/bin/bash
# encoder: parity tests, throughput with the switches on and off, per-kernel launch list
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
L=gpurun_out/r02_enc_gemm.log; : > $L
timeout 600 python -m pytest tests/test_gpu_encoder.py -x -q 2>&1 | tail -5 >> $L
for v in 1 0; do
  echo "== BFK_ENC_ARES=$v" >> $L
  BFK_ENC_ARES=$v timeout 300 python profiles/enc_probe.py 2048 256 >> $L 2>&1
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_enc.csv \
  python profiles/enc_probe.py 512 256 > gpurun_out/ncu_enc.log 2>&1
python - >> $L <<'PY'
import csv, collections, re
rows = [r for r in csv.reader(open("gpurun_out/r02_launches_enc.csv")) if len(r) > 10 and r[0].isdigit()]
agg = collections.defaultdict(list)
for r in rows[len(rows) * 2 // 3:]:   # the last of the three encode_add calls
    name = re.sub(r"\(.*", "", r[4])
    agg[name].append(float(r[-1].replace(",", "")) / 1e3)
for k, v in agg.items():
    print(f"{k:60s} n={len(v):3d} mean {sum(v)/len(v):8.1f} us  total {sum(v):9.1f} us")
PY
cat $L
