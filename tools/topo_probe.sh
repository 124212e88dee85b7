#!/bin/bash
# host / PCIe topology of the box + cfg2 end-to-end leg with and without NUMA binding
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)" >> gpurun_out/topo.txt
for n in ${NS:-4}; do
  B200FEAT_NO_NUMA_BIND=1 python bench.py --gpus $n --steps 5 --warmup 3 2>/dev/null | tail -1 > gpurun_out/numa_off_n$n.json
  python bench.py --gpus $n --steps 5 --warmup 3 2>/dev/null | tail -1 > gpurun_out/numa_on_n$n.json
  python - $n <<'PY'
import json, sys
n = sys.argv[1]
for k in ("off", "on"):
    d = json.loads(open(f"gpurun_out/numa_{k}_n{n}.json").read())
    print("N", n, "bind", k, "e2e", round(d["e2e"]["value"]), "ms", round(d["e2e"]["ms_per_step"], 2), d["e2e"].get("host_cpus_bound"))
PY
done
cat gpurun_out/topo.txt
