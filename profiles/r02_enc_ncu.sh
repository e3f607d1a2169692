#!/bin/bash
# ncu --set full of the encoder's kernels as launched for a 512-chunk x 256-token micro-batch
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
for k in ${KERNELS:-attn gelu ln qkv}; do
  case $k in
    attn) kern=enc_attention; skip=7;;
    gelu) kern=enc_gemm_kernel; skip=26;;   # gemm launches of a layer: QKV, out+LN, GELU, down+LN
    ln) kern=enc_gemm_kernel; skip=27;;
    qkv) kern=enc_gemm_kernel; skip=24;;
  esac
  timeout 600 $NCU -k "regex:$kern" -s $skip -c 1 -f -o gpurun_out/r02_ncu_enc_$k python profiles/enc_probe.py 512 256 > gpurun_out/r02_ncu_enc_$k.log 2>&1 || tail -5 gpurun_out/r02_ncu_enc_$k.log
done
ls -la gpurun_out/r02_ncu_enc_*.ncu-rep
