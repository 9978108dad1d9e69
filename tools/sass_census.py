#!/usr/bin/env python
"""Per-kernel census of the Blackwell-native SASS mnemonics in libbmhrl_b200.so (cuobjdump -sass; no GPU needed).
    python tools/sass_census.py > profiles/r02_sass_census.txt
UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA tensor loads/stores, UTCBAR = tcgen05.commit,
SYNCS = mbarrier ops, HMMA = legacy mma.sync (must be 0)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "bmhrl_b200", "lib", "libbmhrl_b200.so")
PATS = ["UTCHMMA", "UTCQMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "HMMA", "LDGSTS",
        "LDG.E.128", "STG.E.128", "RED", "ATOM"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", cur)
            kernels[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        kernels[cur]["_total"] += 1
        for p in PATS:
            if op.startswith(p) or (p in ("LDG.E.128", "STG.E.128") and op.startswith(p.split(".")[0]) and ".128" in op):
                kernels[cur][p] += 1
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)}: instruction counts per kernel (static, not executed counts)")
    print("# " + " ".join(f"{p:>9}" for p in ["instrs"] + PATS) + "  kernel")
    for k, c in kernels.items():
        print("  " + " ".join(f"{c.get(p, 0):>9}" for p in ["_total"] + PATS) + "  " + k)
    tc = [k for k, c in kernels.items() if c.get("UTCHMMA", 0) + c.get("UTCQMMA", 0) + c.get("UTCMMA", 0) > 0]
    print(f"# {len(kernels)} kernels; tcgen05.mma in {len(tc)}; legacy HMMA in "
          f"{sum(1 for c in kernels.values() if c.get('HMMA', 0) > 0)}")


if __name__ == "__main__":
    sys.exit(main())
