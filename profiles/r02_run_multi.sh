#!/bin/bash
# multi-GPU check: sharded parity tests with W > 1 on record, then the bench line at N = $1
N=${1:-2}
cd "$(dirname "$0")/.." && mkdir -p gpurun_out && rm -f gpurun_out/test_notes.log
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > gpurun_out/r02_gpus_multi.txt
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -q > gpurun_out/r02_pytest_sharded_${N}gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_sharded_${N}gpu.log
cat gpurun_out/test_notes.log >> gpurun_out/r02_pytest_sharded_${N}gpu.log
tail -4 gpurun_out/r02_pytest_sharded_${N}gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_n${N}.json 2> gpurun_out/r02_bench_n${N}.err; echo "bench rc=$?"
tail -c 400 gpurun_out/r02_bench_n${N}.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_ref_n${N}.json 2>/dev/null
python - <<PY
import json
for f in ("gpurun_out/r02_bench_n${N}.json", "gpurun_out/r02_bench_ref_n${N}.json"):
    try:
        j = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, j["value"], j["ms_per_step"], j.get("scan_ms"), (j.get("roofline") or {}).get("frac"), j["e2e"]["value"], j.get("parity"), j.get("result_sha"), j["cpu_baseline"])
        for b, r in (j.get("batches") or {}).items():
            print("   ", b, r["value"], r["ms_per_step"], "scan", r["scan_ms"], r["roofline"]["frac"], "e2e", r["e2e"])
    except Exception as e:
        print(f, "parse failed", e)
PY
