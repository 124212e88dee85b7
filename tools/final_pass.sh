#!/bin/bash
# End-of-round evidence: GPU tests, smoke, the bench lines of every workload, the reference arm, and the ncu launch
# list of the default bench command.  Outputs under gpurun_out/final/ (copied to profiles/ by hand).
out=gpurun_out/final; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.txt 2>&1; tail -2 $out/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.txt 2>&1; tail -1 $out/smoke.txt
for wl in cfg2 cfg3 cfg4 cfg5; do
  python bench.py --workload $wl --steps 20 --warmup 3 2>$out/bench_$wl.err | tail -1 > $out/bench_$wl.json
done
python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 > $out/bench_reference.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches.csv \
  python bench.py --steps 2 --warmup 3 > $out/ncu_bench.log 2>&1
python - <<'PY'
import csv, json, collections
rows = list(csv.reader(l for l in open("gpurun_out/final/launches.csv") if l.startswith('"')))
hdr = rows[0]; k = hdr.index("Kernel Name"); v = hdr.index("Metric Value"); u = hdr.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[1:]:
    val = float(r[v].replace(",", ""))
    val = val / 1e3 if r[u] in ("ns", "nsecond") else val * (1e3 if r[u] in ("ms", "msecond") else 1.0)   # -> us
    n, t = tot.get(r[k], (0, 0.0)); tot[r[k]] = (n + 1, t + val)
with open("gpurun_out/final/launch_summary.txt", "w") as f:
    for name, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{n:5d} launches {t:12.1f} us total {t / n:10.1f} us avg  {name[:110]}\n")
print(open("gpurun_out/final/launch_summary.txt").read()[:1500])
for wl in ("cfg2", "cfg3", "cfg4", "cfg5"):
    d = json.loads(open(f"gpurun_out/final/bench_{wl}.json").read())
    print(wl, round(d["ms_per_step"], 4), round(d["roofline"]["frac"], 4), round(d["e2e"]["value"]), d["gpu_launches"], d["clocks"])
PY
