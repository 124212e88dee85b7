#!/usr/bin/env python
"""Per-instruction stall profile from `ncu --page source --csv` output.
    python tools/ncu_source.py src.csv out.txt   -> writes one line per SASS instruction, prints a 128-instruction summary"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
with open(sys.argv[2], "w") as f:
    for n, r in enumerate(data):
        st = {k: int(r[ix[k]]) for k in ['stall_barrier', 'stall_long_sb', 'stall_math', 'stall_mio', 'stall_no_inst',
                                         'stall_short_sb', 'stall_wait', 'stall_dispatch', 'stall_lg', 'stall_branch_resolving']}
        top = [f"{k[6:]}={v}" for k, v in st.items() if v >= 3]
        f.write(f"{n} {r[ix['Source']][:64].ljust(66)} {r[ix['# Samples']].rjust(5)} {r[ix['Instructions Executed']].rjust(8)} {' '.join(top)}\n")
tot_s = sum(int(r[ix['# Samples']]) for r in data)
tot_i = sum(int(r[ix['Instructions Executed']]) for r in data)
print("total samples", tot_s, "inst", tot_i, "n sass", len(data))
blk = int(sys.argv[3]) if len(sys.argv) > 3 else 128
for b in range(0, len(data), blk):
    seg = data[b:b + blk]
    s = sum(int(r[ix['# Samples']]) for r in seg)
    i = sum(int(r[ix['Instructions Executed']]) for r in seg)
    g = lambda k: sum(int(r[ix[k]]) for r in seg)
    ops = {}
    for r in seg:
        t = r[ix['Source']].split()
        op = t[1] if t[0].startswith('@') else t[0]
        ops[op] = ops.get(op, 0) + 1
    top = sorted(ops.items(), key=lambda x: -x[1])[:3]
    print(f"{b:5d} smp {s:6d} ({100*s/tot_s:4.1f}%) inst {i:9d} ({100*i/tot_i:4.1f}%) noinst {g('stall_no_inst'):5d} bar {g('stall_barrier'):5d} "
          f"short {g('stall_short_sb'):5d} wait {g('stall_wait'):5d} mio {g('stall_mio'):5d} long {g('stall_long_sb'):5d} math {g('stall_math'):4d} {top}")
