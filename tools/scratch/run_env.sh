#!/bin/bash
# Times the default library under environment variants:  tools/scratch/run_env.sh "KB_CONT_LOCAL=0" "KB_CONT_SW_LOCAL=12" ...
export KB_BENCH_SECONDARY=0 KB_BENCH_LATENCY=0 KB_BENCH_MERGED=0 KB_BENCH_PRUNED=0 KB_BENCH_SUGGEST=0
for v in "$@"; do
  env $v timeout 200 python bench.py --cand ${VAR_CAND:-24} --steps 3 --warmup 2 --no-e2e > gpurun_out/var.json 2> gpurun_out/var.err
  python - "$v" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open("gpurun_out/var.json") if l.startswith("{")][-1])
    k = d["roofline"]["kernels"]
    lb = d.get("local_origin_blocks") or {}
    print(sys.argv[1], round(d["value"]), [(round(e["ms_total"], 1), round(e.get("frac", 0), 4)) for e in k[1:3]], d["winners_vs_oracle"]["identical"], d["score_error_max"]["continuous"], lb.get("own_origin"), lb.get("hi_lo_form"), d["winners_index_head"], flush=True)
except Exception as e:
    print(sys.argv[1], "failed", e, open("gpurun_out/var.err").read()[-600:])
PY
done
