#!/usr/bin/env python
"""Opcode counts per kernel of the built library: `python scripts/sass_summary.py > profiles/rNN_sass_opcodes.md`.

`cuobjdump -sass` of stereo_matching_b200/libstereo_b200.so; per kernel the number of SASS instructions and of the
opcodes that show which hardware path a kernel takes: UTCHMMA (tcgen05.mma; `.2CTA` = cta_group::2), LDTM / STTM
(tcgen05.ld / st), UBLKCP (cp.async.bulk), UTMALDG / UTMASTG (tensor-map TMA), SYNCS (mbarrier), REDUX / CREDUX
(redux.sync), HMMA / IMMA (legacy mma.sync - none expected)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "stereo_matching_b200", "libstereo_b200.so")
KEYS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "REDUX", "HMMA", "FFMA", "DFMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = {}
    counts = collections.OrderedDict()
    name = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            counts[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            op = m.group(1)
            c = counts[name]
            c["total"] += 1
            base = op.split(".")[0]
            if base in ("UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "HMMA", "IMMA", "FFMA", "DFMA"):
                c[base] += 1
            if base in ("REDUX", "CREDUX"):
                c["REDUX"] += 1
            if base == "UTCHMMA" and ".2CTA" in op:
                c["UTCHMMA.2CTA"] += 1
    names = list(counts)
    dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    demangle = dict(zip(names, dem)) if len(dem) == len(names) else {n: n for n in names}
    print("| kernel | SASS instructions | " + " | ".join(KEYS) + " |")
    print("|---|---|" + "---|" * len(KEYS))
    total = collections.Counter()
    for n in sorted(names, key=lambda k: demangle[k]):
        c = counts[n]
        short = re.sub(r"\(.*", "", demangle[n]).replace("stereo::", "").replace("(anonymous namespace)::", "")
        print(f"| `{short[:90]}` | {c['total']} | " + " | ".join(str(c[k]) if c[k] else "" for k in KEYS) + " |")
        total.update(c)
    print(f"| **all {len(names)} kernels** | {total['total']} | " + " | ".join(str(total[k]) for k in KEYS) + " |")


if __name__ == "__main__":
    main()
