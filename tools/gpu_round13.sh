#!/bin/bash
# last call of round 2: the full GPU suite on HEAD, T5-large training (BASELINE configs[3]) and the selection workload
TAG=${1:-r5k}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log | cut -c1-300
timeout 600 python bench.py --workload train --model large --steps 10 --warmup 3 --no-library-baseline > gpurun_out/${TAG}_bench_train_large.json 2> gpurun_out/${TAG}_err.log; echo "large rc=$?"
timeout 600 python bench.py --workload select --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_select.json 2>> gpurun_out/${TAG}_err.log; echo "select rc=$?"
python - <<PY
import json
j = json.load(open("gpurun_out/${TAG}_bench_train_large.json")); r = j["roofline"]
print("T5-large train", round(j["value"], 2), j["unit"], round(j["ms_per_step"], 1), "ms", "model_frac", round(r.get("model_frac", 0), 3), "gemm", round(r.get("achieved", 0)), "e2e", round(j["e2e"]["value"], 2), j["clocks"]["sm_mhz"], "MHz")
j = json.load(open("gpurun_out/${TAG}_bench_select.json"))
print("select", round(j["value"]), j["unit"], {k: round(v, 3) for k, v in j["config"].items() if k.endswith("_frac") or k.endswith("_ms")})
PY
