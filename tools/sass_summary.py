#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of the in-tree libsrb200.so (cuobjdump -sass; no GPU needed):
    python tools/sass_summary.py > profiles/rNN_sass_summary.txt
The mnemonics that prove the Blackwell paths: UTCHMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit),
UBLKCP (cp.async.bulk: the TMA engine's bulk copy; UTMALDG would be its tensor-map form), SYNCS (mbarrier), FFMA2
(fma.rn.f32x2), LDGSTS (cp.async), DFMA / DADD (fp64)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "speech-recognition_b200", "srb200", "libsrb200.so")
WANT = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "FFMA2", "FFMA", "LDGSTS", "DFMA", "DADD",
        "MUFU", "ATOM", "RED", "REDG"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    counts, name = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)", "anon").split("(")[0]
            counts[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and name:
            counts[name][m.group(1)] += 1
    print(f"# SASS mnemonic counts per kernel, {os.path.relpath(SO, ROOT)} (sm_100a)")
    print(f"{'kernel':70s} " + " ".join(f"{w:>7s}" for w in WANT) + "   total")
    for k, c in counts.items():
        print(f"{k[:70]:70s} " + " ".join(f"{c.get(w, 0):7d}" for w in WANT) + f" {sum(c.values()):7d}")
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    print(f"{'ALL':70s} " + " ".join(f"{tot.get(w, 0):7d}" for w in WANT) + f" {sum(tot.values()):7d}")


if __name__ == "__main__":
    sys.exit(main())
