#!/bin/bash
# per-kernel times at a micro-batch that fits L2 (64 and 128 chunks) against the 512-chunk one
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
L=gpurun_out/r02_enc_small.log; : > $L
for B in 64 128 512; do
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_enc_b$B.csv \
  python profiles/enc_probe.py $B 256 > gpurun_out/ncu_enc.log 2>&1
python - $B >> $L <<'PY'
import csv, collections, re, sys
B = sys.argv[1]
rows = [r for r in csv.reader(open(f"gpurun_out/r02_launches_enc_b{B}.csv")) if len(r) > 10 and r[0].isdigit()]
ids = sorted({int(r[0]) for r in rows})
keep = set(ids[len(ids) * 2 // 3:])
agg = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows:
    if int(r[0]) not in keep: continue
    name = re.sub(r"\(.*", "", r[4])
    agg[name][r[-3]].append(float(r[-1].replace(",", "")))
print("B =", B)
for k, m in agg.items():
    t = m.get("gpu__time_duration.sum", [0]); rd = m.get("dram__bytes_read.sum", [0]); wr = m.get("dram__bytes_write.sum", [0])
    print(f"{k:50s} n={len(t):3d} mean {sum(t)/len(t)/1e3:8.1f} us  dram rd {sum(rd)/len(rd):10.1f} wr {sum(wr)/len(wr):10.1f}")
PY
done
cat $L
