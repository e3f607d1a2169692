#!/bin/bash
# final validation of the round: every GPU test, smoke, the default bench line and the reference arm
cd "$(dirname "$0")/.." && mkdir -p gpurun_out && rm -f gpurun_out/test_notes.log
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r02_gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu.log
tail -5 gpurun_out/r02_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02_smoke.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$?"
tail -c 400 gpurun_out/r02_bench_n1.err
timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_ref_n1.json 2>&1
python - <<'PY'
import json
try:
    j = json.load(open("gpurun_out/r02_bench_n1.json"))
    print("c5 head", j["value"], j["ms_per_step"], "scan", j["scan_ms"], "frac", j["roofline"]["frac"], "e2e", j["e2e"]["value"], "parity", j["parity"]["ok"], j["result_sha"][:12])
    for k in ("c2", "c3", "c4", "c4e", "c1"):
        s = j.get(k, {})
        print(k, s.get("value"), s.get("ms_per_step"), (s.get("roofline") or {}).get("frac"), (s.get("e2e") or {}).get("value"), s.get("error"))
        for b, r in (s.get("batches") or {}).items():
            print("   ", b, r["value"], r["ms_per_step"], "scan", r["scan_ms"], r["roofline"]["frac"], "e2e ms", r["e2e_ms_per_step"])
except Exception as e:
    print("bench parse failed", e)
PY
