#!/bin/bash
# Sweep of tiles per CTA of the round-1 kaldi kernel (B200FEAT_KALDI_KERNEL=1): tools/tpc_sweep.sh kaldi "1 2 4 8"
kind=${1:-kaldi}; tpcs=${2:-"1 2 4 8"}
if [ $kind != kaldi ]; then echo "only the round-1 kaldi kernel reads a tiles-per-CTA variable"; exit 1; fi
var=B200FEAT_KALDI_TPC; wl=cfg4; export B200FEAT_KALDI_KERNEL=1
timeout 600 python -m pytest tests -m gpu -x -q -k "$kind or abi" 2>&1 | tail -3
for t in $tpcs; do
  env $var=$t python bench.py --workload $wl --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/tpc_${kind}_$t.json
  python - "$kind" "$t" <<'PY'
import sys, json
k, t = sys.argv[1:3]
d = json.loads(open(f"gpurun_out/tpc_{k}_{t}.json").read())
print(k, "tpc", t, round(d["ms_per_step"], 4), round(d["roofline"]["frac"], 4), round(d["e2e"]["value"]))
PY
done
