#!/bin/bash
# Build-time variants of the continuous scoring kernel, rebuilt and timed on the GPU box (nvcc is there):
#   tools/scratch/variants.sh "-DSC_MINB=4" "-DSC_GROUP=64" ...
export KB_BENCH_SECONDARY=0 KB_BENCH_LATENCY=0 KB_BENCH_MERGED=0 KB_BENCH_PRUNED=0 KB_BENCH_SUGGEST=0
for flags in "$@"; do
  touch katib_b200/csrc/kb_tpe_score.cu
  make -C katib_b200/csrc EXTRA_kb_tpe_score="$flags" > /dev/null 2>&1 || { echo "$flags: build failed"; continue; }
  timeout 200 python bench.py --cand 24 --steps 3 --warmup 2 --no-e2e > gpurun_out/var.json 2> gpurun_out/var.err
  python - "$flags" <<'PY'
import json, sys
d = json.loads([l for l in open("gpurun_out/var.json") if l.startswith("{")][-1])
k = d["roofline"]["kernels"]
print(sys.argv[1], round(d["value"]), [(round(e["ms_total"], 1), round(e.get("frac", 0), 4)) for e in k[1:3]], d["winners_vs_oracle"]["identical"], d["score_error_max"]["continuous"])
PY
done
