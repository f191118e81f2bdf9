#!/bin/bash
# re-rank chunks alternating between two streams (PN_OVERLAP_CHUNKS=1) against one stream: identical scores, step A B A B
TAG=${1:-r5l}
mkdir -p gpurun_out
timeout 300 python - <<'PY'
import sys, torch
sys.path.insert(0, '.')
from pacednegatives_b200 import scorer as sc
dims = sc.T5Dims(vocab=512, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2)
sd = sc.random_state_dict(dims, seed=3)
m = sc.MonoT5Scorer(dims, sd, rel_id=5, nrel_id=6, chunk_docs=64)
g = torch.Generator().manual_seed(1)
ids = torch.randint(2, 512, (300, 128), generator=g)
m.overlap_chunks = False
a = m.score(ids).clone()
m.overlap_chunks = True
b = m.score(ids).clone()
lens = torch.randint(1, 129, (300,), generator=g)
m.overlap_chunks = False
c = m.score(ids, lengths=lens).clone()
m.overlap_chunks = True
d = m.score(ids, lengths=lens).clone()
torch.cuda.synchronize()
print("two-stream scores identical:", bool((a == b).all()), bool((c == d).all()), float(a.abs().sum()))
PY
for rep in 1 2; do
  for v in 0 1; do
    PN_OVERLAP_CHUNKS=$v python bench.py --workload rerank --steps 5 --warmup 3 --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_rerank_$v$rep.json 2>/dev/null
    python - <<PY
import json
j = json.load(open("gpurun_out/${TAG}_rerank_$v$rep.json")); r = j.get("roofline", {})
print("overlap=$v rep $rep rerank", round(j["ms_per_step"], 2), "ms", round(j["value"], 1), j["unit"], "e2e", round(j["e2e"]["value"], 1), j["clocks"]["sm_mhz"], "MHz")
PY
  done
done | tee gpurun_out/${TAG}_step_ab.txt
