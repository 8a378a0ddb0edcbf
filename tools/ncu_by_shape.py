"""Joins an ncu launch list (--metrics gpu__time_duration.sum --csv) of `bench.py --profile-only` with the op tags the
same run wrote (gpurun_out/taglog.json) and prints device time per (kernel, shape).  ncu times are cold-cache and
serialised: compare shares, not absolutes.   python tools/ncu_by_shape.py launches.csv taglog.json"""
import collections
import csv
import json
import sys

KERNELS_PER_OP = {"groupnorm": 2}
rows = [l for l in open(sys.argv[1]) if not l.startswith("==")]
launches = []
for r in csv.DictReader(rows):
    name = r["Kernel Name"]
    if "b200sd::" not in name:
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = v / 1e3 if unit.startswith("n") else (v if unit.startswith("u") else v * 1e3)
    launches.append((name.split("(")[0].replace("void ", "").replace("b200sd::", ""), us))
tags = json.load(open(sys.argv[2]))
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
i = 0
for kind, tag, work in tags:
    n = KERNELS_PER_OP.get(kind, 1)
    us = sum(x[1] for x in launches[i:i + n])
    i += n
    a = agg[f"{kind} {tag}"]
    a[0] += 1; a[1] += us; a[2] += work
assert i == len(launches), (i, len(launches))
tot = sum(a[1] for a in agg.values())
print(f"total {tot / 1e3:.2f} ms over {i} kernels")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 50]:
    print(f"{a[1] / 1e3:8.3f} ms {100 * a[1] / tot:5.1f}%  n={a[0]:4d}  avg {a[1] / a[0]:8.1f} us  {a[2] / (a[1] * 1e-6) / 1e12:8.2f} T(FLOP|B)/s  {k}")
