#!/bin/bash
# ncu captures of the shipped launches (one GPU): full sets for the scan (HBM regime, tensor regime as shipped,
# config-3 shape), K5, K1; DRAM bytes only for the 100M-row launch of the head line
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
cap() { # name kernel-regex skip cmd...
  name=$1; kern=$2; skip=$3; shift 3
  "$@" > gpurun_out/${name}_plain.log 2>&1 || { echo "$name: plain run failed"; tail -3 gpurun_out/${name}_plain.log; return; }
  timeout 900 $NCU -k regex:$kern -s $skip -c 1 -f -o gpurun_out/$name "$@" > gpurun_out/${name}_ncu.log 2>&1 || { echo "$name: ncu failed"; tail -5 gpurun_out/${name}_ncu.log; }
  tail -1 gpurun_out/${name}_plain.log
}
cap r02_ncu_scan_c2_nq64 scan_tc_kernel 2 python profiles/ncu_probe.py 64
cap r02_ncu_scan_c2_nq1 scan_tc_kernel 2 python profiles/ncu_probe.py 1
cap r02_ncu_scan_c2_nq4096 scan_tc_kernel 4 python profiles/ncu_probe.py 4096
cap r02_ncu_scan_c3shape_nq2048 scan_tc_kernel 2 python profiles/ncu_probe.py 2048 2000000 768 100 ip bf16
cap r02_ncu_rerank_c2_nq64 rerank_kernel 2 python profiles/ncu_probe.py 64
cap r02_ncu_pool_stream_bf16_full pool_stream_kernel 3 python profiles/pool_probe.py bf16 full
cap r02_ncu_pool_stream_bf16_ragged pool_stream_kernel 3 python profiles/pool_probe.py bf16 ragged
# head line launch (100M x 384 bf16-stored, nq=64): DRAM bytes only, one pass, no replay of a 12 ms kernel x 40
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:scan_tc_kernel -s 2 -c 1 --csv --log-file gpurun_out/r02_ncu_scan_c5_nq64_dram.csv python profiles/ncu_probe.py 64 100000000 384 10 l2 bf16 > gpurun_out/r02_ncu_scan_c5_plain.log 2>&1
tail -3 gpurun_out/r02_ncu_scan_c5_nq64_dram.csv
# launch list of the default bench command (cold-cache, serialised: shares only)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_bench_c2.csv python bench.py --workload c2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_launches_bench_c2.log 2>&1
ls -la gpurun_out/*.ncu-rep | head -20
