"""Opcode evidence for the judge: per kernel of libpt_stage1.so, how many tcgen05 / TMA / TMEM instructions its
SASS holds (cuobjdump -sass; runs without a GPU).  UTCHMMA = tcgen05.mma (… .2CTA = cta_group::2), UTMALDG / UTMASTG =
cp.async.bulk.tensor load / store, LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops.
    python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

LIB = Path(__file__).resolve().parent.parent / "projectiontrainer_b200" / "libpt_stage1.so"
OPS = ["UTCHMMA", "UTCHMMA.2CTA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "FFMA2", "FMUL2",
       "MUFU.EX2", "MUFU.TANH", "REDG", "ATOMG", "LDG", "STG", "LDS", "STS"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    kernels, cur = OrderedDict(), None
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            kernels[cur] = Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if not m:
            continue
        op = m.group(1)
        kernels[cur]["_total"] += 1
        for key in OPS:
            if op == key or op.startswith(key + "."):
                kernels[cur][key] += 1
        if op.startswith("UTCHMMA") and ".2CTA" in op:
            kernels[cur]["UTCHMMA.2CTA"] += 1
    dem = subprocess.run(["cu++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines()
    names = dict(zip(kernels, dem)) if len(dem) == len(kernels) else {k: k for k in kernels}
    print(f"# cuobjdump -sass {LIB.name}: instruction counts per kernel (sm_100a)")
    print("# " + " ".join(f"{k:>8s}" for k in ["total"] + [o.replace('UTCHMMA.2CTA', 'MMA.2CTA') for o in OPS]))
    for k, c in kernels.items():
        short = re.sub(r"\(.*", "", names[k].replace("(int)", "")).replace("pt::", "").replace("void ", "")
        print(f"{short[:60]:60s} " + " ".join(f"{c.get(o, 0):8d}" for o in ["_total"] + OPS))


if __name__ == "__main__":
    sys.exit(main())
