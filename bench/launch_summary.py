"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: one training step."""
import csv
import re
import sys

path = sys.argv[1]
rows = [r for r in csv.reader(open(path)) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
data = [(re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("ddpb::", "")[:48], float(r[vi].replace(",", "")) / 1e3)
        for r in rows[1:]]
starts = [i for i, d in enumerate(data) if "pack_weights" in d[0] or "first_layer_operand" in d[0]]
# a step starts at pack_weights (if present right before the operand kernel) or the operand kernel
begin = [i for i, d in enumerate(data) if "first_layer_operand" in d[0]]
s, e = begin[0], begin[1]
if s > 0 and "pack_weights" in data[s - 1][0]:
    s, e = s - 1, e - 1
step = data[s:e]
if "-v" in sys.argv:
    for name, us in step:
        print(f"{name:50s} {us:8.2f}")
agg = {}
for name, us in step:
    agg.setdefault(name, [0.0, 0])
    agg[name][0] += us
    agg[name][1] += 1
tot = sum(us for _, us in step)
print(f"kernels in step: {len(step)}   sum of durations: {tot:.1f} us")
for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{name:50s} {us:9.1f} us  x{n:<3d} {100 * us / tot:5.1f}%")
