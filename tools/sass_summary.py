"""Instruction histogram of every kernel in libb200sd.so (cuobjdump -sass), for profiles/rNN_sass_summary.txt.
Shows which kernels carry the Blackwell instructions: UTCHMMA (tcgen05.mma), UTMALDG / UTMASTG / UTMAREDG (TMA load /
store / reduce), LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit), SYNCS (mbarrier), FFMA2 / FADD2 / FMUL2 (packed fp32).
python tools/sass_summary.py [path/to/lib.so]"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
lib = Path(sys.argv[1]) if len(sys.argv) > 1 else ROOT / "scalable-hw-agnostic-inference_b200" / "b200sd" / "libb200sd.so"
sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True, check=True).stdout
KEY = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "LDTM", "STTM", "SYNCS", "MUFU", "FFMA2", "FADD2",
       "FMUL2", "F2FP", "HMMA", "BAR", "ATOM", "RED", "SHFL", "LDG", "STG", "LDS", "STS"]
kern, hist, total, order = None, collections.defaultdict(collections.Counter), collections.Counter(), []
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(CUtensorMap_st.*|\(void const\*.*|\(b200sd::.*|\(float const\*.*|\(int const\*.*|\(unsigned short const\*.*", "", kern)
        order.append(kern)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Za-z0-9_]+)*)", line)
    if m and kern:
        op, mods = m.group(1), m.group(2)
        total[kern] += 1
        for k in KEY:
            if op == k or (k in ("UTMALDG", "UTMASTG", "UTMAREDG", "UTCHMMA") and op.startswith(k)):
                tag = k
                if k == "UTCHMMA" and ".2CTA" in mods:
                    tag = "UTCHMMA.2CTA"
                if k == "UTMALDG" and ".2CTA" in mods:
                    tag = "UTMALDG.2CTA"
                hist[kern][tag] += 1
print(f"# {lib.name}: {len(order)} kernels, {sum(total.values())} SASS instructions (sm_100a)")
print("# columns: kernel | instructions | counts of the instructions that matter")
grand = collections.Counter()
for k in order:
    grand.update(hist[k])
    body = "  ".join(f"{n}={c}" for n, c in sorted(hist[k].items(), key=lambda kv: -kv[1]))
    print(f"{k} | {total[k]} | {body}")
print("# whole library: " + "  ".join(f"{n}={c}" for n, c in sorted(grand.items(), key=lambda kv: -kv[1])))
