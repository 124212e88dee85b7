#!/bin/bash
# Sweep of B200FEAT_KALDI_TPC (tiles per CTA of the kaldi kernel) on the cfg4 workload.
timeout 600 python -m pytest tests -m gpu -x -q -k "kaldi or abi" 2>&1 | tail -3
for t in ${TPCS:-1 2 4 8 16}; do
  B200FEAT_KALDI_TPC=$t python bench.py --workload cfg4 --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/k_$t.json
  python - "$t" <<'PY'
import sys, json
t = sys.argv[1]
d = json.loads(open(f"gpurun_out/k_{t}.json").read())
print("tpc", t, d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["value"])
PY
done
