#!/bin/bash
# Rebuild the quantised-TPE kernel with different tile / occupancy / fit-degree settings on the GPU box and time each.
set -e
cd "$(dirname "$0")/.."
for v in "8 2 12 7.5" "4 3 12 7.5" "6 2 12 7.5" "8 2 11 6.0" "4 3 11 6.0" "4 4 11 6.0"; do
  set -- $v
  rm -f katib_b200/csrc/build/kb_tpe_score.o katib_b200/csrc/build/kb_tpe_build.o
  make -s -C katib_b200/csrc EXTRA="-DSQ_R=$1 -DSQ_MINB=$2 -DKB_QF_DEG=$3 -DKB_QF_YMAX=$4" > /dev/null
  python bench.py --cand 22 --steps 2 --warmup 1 --no-e2e > /tmp/b.json 2>/tmp/b.err || { tail -3 /tmp/b.err; continue; }
  python - "$v" <<'PY'
import json,sys
d=json.load(open('/tmp/b.json'))
k={e['name']:e for e in d['roofline']['kernels']}
print(sys.argv[1], "| quant ms %.1f frac %.3f | cont ms %.1f frac %.3f | cand/s %.3e"%(k['tpe_score_quant_kernel']['ms_total'],k['tpe_score_quant_kernel']['frac'],k['tpe_score_cont_kernel']['ms_total'],k['tpe_score_cont_kernel']['frac'],d['value']))
PY
  python -m pytest tests/test_gpu_tpe.py -m gpu -q -s -k config5 2>&1 | grep -E "max rel|failed"
done
rm -f katib_b200/csrc/build/kb_tpe_score.o katib_b200/csrc/build/kb_tpe_build.o
make -s -C katib_b200/csrc > /dev/null
