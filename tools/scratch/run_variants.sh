#!/bin/bash
# Times prebuilt variant libraries (tools/scratch/build_variants.sh) on the GPU box; restores the default at the end.
export KB_BENCH_SECONDARY=0 KB_BENCH_LATENCY=0 KB_BENCH_MERGED=0 KB_BENCH_PRUNED=0 KB_BENCH_SUGGEST=0
cp katib_b200/libkatib_b200.so /tmp/kb_default.so
for name in "$@"; do
  if [ "$name" = default ]; then cp /tmp/kb_default.so katib_b200/libkatib_b200.so; else cp tools/scratch/variants/$name.so katib_b200/libkatib_b200.so; fi
  timeout 200 python bench.py --cand ${VAR_CAND:-24} --steps 3 --warmup 2 --no-e2e > gpurun_out/var.json 2> gpurun_out/var.err
  python - "$name" <<'PY'
import json, sys
d = json.loads([l for l in open("gpurun_out/var.json") if l.startswith("{")][-1])
k = d["roofline"]["kernels"]
print(sys.argv[1], round(d["value"]), [(round(e["ms_total"], 1), round(e.get("frac", 0), 4)) for e in k[1:3]], d["winners_vs_oracle"]["identical"], d["score_error_max"]["continuous"], flush=True)
PY
done
cp /tmp/kb_default.so katib_b200/libkatib_b200.so
