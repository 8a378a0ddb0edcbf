"""Summarises an ncu launch list of one eager pass (`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,
dram__bytes_write.sum --clock-control none --csv --log-file L python bench.py --profile-only --batch 16`):
per kernel launches / time share / DRAM bytes per launch -> profiles/rNN_launch_summary_b16.json, and per kernel class
the DRAM bytes per launch that bench.py reports as `roofline.traffic` -> profiles/roofline_traffic.json.
ncu times are serialised and cold-cache: compare shares, not absolutes.
python tools/ncu_launch_summary.py launches.csv[.gz] rNN"""
import collections
import csv
import gzip
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
src, tag = sys.argv[1], sys.argv[2]
op = gzip.open if src.endswith(".gz") else open
lines = [l for l in op(src, "rt") if not l.startswith("==")]
per = collections.defaultdict(lambda: collections.defaultdict(float))
count = collections.Counter()
ids = set()
for r in csv.DictReader(lines):
    name = r["Kernel Name"].split("(")[0].replace("void ", "")
    v = float(r["Metric Value"].replace(",", ""))
    m, unit = r["Metric Name"], r["Metric Unit"]
    if m == "gpu__time_duration.sum":
        v = v / 1e6 if unit.startswith("n") else (v / 1e3 if unit.startswith("u") else v)  # -> ms
    else:
        v = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
    per[name][m] += v
    if (r["ID"], name) not in ids:
        ids.add((r["ID"], name))
        count[name] += 1
total = sum(d["gpu__time_duration.sum"] for d in per.values())
kernels = [{"kernel": k, "launches": count[k], "ms": round(d["gpu__time_duration.sum"], 3),
            "share": round(d["gpu__time_duration.sum"] / total, 4),
            "dram_read_MB_per_launch": round(d["dram__bytes_read.sum"] / count[k] / 1e6, 3),
            "dram_write_MB_per_launch": round(d["dram__bytes_write.sum"] / count[k] / 1e6, 3)}
           for k, d in sorted(per.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"])]
cmd = ("ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv "
       "python bench.py --profile-only --batch 16")
(ROOT / "profiles" / f"{tag}_launch_summary_b16.json").write_text(json.dumps(
    {"command": cmd, "note": "one eager pass (25 UNet steps + VAE decode, 16 images); ncu times are serialised and "
     "cold-cache: compare shares, not absolutes", "total_ms": round(total, 2), "kernels": kernels}, indent=1))
classes = {"gemm": "gemm_kernel", "attention": "attention_kernel", "groupnorm_apply": "gn_apply_kernel",
           "groupnorm": "gn_stats_kernel", "layernorm": "layernorm_kernel", "gn_tile_sums": "gn_tile_sums_kernel"}
out = {"source": f"profiles/{tag}_launch_summary_b16.json ({cmd}; averages over ALL launches of the class)"}
for cls, pat in classes.items():
    ks = [k for k in per if pat in k]
    n = sum(count[k] for k in ks)
    if not n:
        continue
    rd = sum(per[k]["dram__bytes_read.sum"] for k in ks)
    wr = sum(per[k]["dram__bytes_write.sum"] for k in ks)
    ms = sum(per[k]["gpu__time_duration.sum"] for k in ks)
    out[cls] = {"launches": n, "dram_read_bytes_per_launch": int(rd / n), "dram_write_bytes_per_launch": int(wr / n),
                "dram_bytes_per_launch": int((rd + wr) / n), "ncu_ms": round(ms, 2), "ncu_share": round(ms / total, 4)}
(ROOT / "profiles" / "roofline_traffic.json").write_text(json.dumps(out, indent=1))
print(f"total {total:.1f} ms over {sum(count.values())} launches")
for k in kernels[:14]:
    print(k)
