#!/usr/bin/env python
"""Per-kernel SASS opcode evidence (B200_PROFILING.md "What proves a Blackwell-native kernel"): for every kernel in
ufm_b200/libufm_b200.so the counts of tcgen05 MMAs (UTC*MMA), tensor-memory loads / stores (LDTM / STTM), TMA loads /
stores / reduces / prefetches (UTMALDG / UTMASTG / UTMAREDG / UTMAPF), MUFU, plus legacy tensor opcodes (HMMA: none).

    python scripts/sass_opcodes.py > profiles/r02_sass_opcodes.txt        (CPU only: cuobjdump -sass)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ufm_b200", "libufm_b200.so")
KEYS = ["UTCHMMA", "UTCHMMA.2CTA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "UBLKCP", "SYNCS", "MUFU",
        "HMMA", "HGMMA", "LDGSTS"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kern = None
    counts = collections.OrderedDict()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            kern = re.sub(r"ufm::\(anonymous namespace\)::", "", kern)
            counts[kern] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_.]+)", line)
        if m and kern:
            op = m.group(1)
            c = counts[kern]
            c["total"] += 1
            base = op.split(".")[0]
            if base in KEYS:
                c[base] += 1
            if op.startswith("UTCHMMA") and ".2CTA" in op:
                c["UTCHMMA.2CTA"] += 1
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} (sm_100a): opcode counts per kernel; UTCHMMA = tcgen05.mma kind::f16, LDTM / STTM = "
          "tcgen05.ld / st, UTMA* = cp.async.bulk.tensor (TMA), HMMA / HGMMA = legacy mma.sync / wgmma (expected: none)")
    print(f"{'kernel':90s} {'instrs':>7s} " + " ".join(f"{k:>8s}" for k in KEYS))
    tot = collections.Counter()
    for k, c in counts.items():
        print(f"{k[:90]:90s} {c['total']:7d} " + " ".join(f"{c[x]:8d}" for x in KEYS))
        tot.update(c)
    print(f"{'TOTAL':90s} {tot['total']:7d} " + " ".join(f"{tot[x]:8d}" for x in KEYS))


if __name__ == "__main__":
    main()
