#!/usr/bin/env python
"""cuobjdump -sass of libb200shard.so -> per-kernel counts of the Blackwell tensor-core / TMA mnemonics
(UTCHMMA = tcgen05.mma, UTMALDG = TMA tensor load, UTMASTG = TMA tensor store, LDTM / STTM = tcgen05.ld / st,
UTCBAR = tcgen05.commit, FFMA2 = packed fp32 FMA, HMMA = mma.sync).  Writes profiles/sass_summary.txt."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "llm-sharding_b200", "lib", "libb200shard.so")
PAT = ["UTCHMMA.2CTA", "UTCHMMA", "UTMALDG.2D.2CTA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UTCBAR", "HMMA", "FFMA2", "SYNCS", "UCGABAR", "ACQBULK"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", cur)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        for p in PAT:
            if op.startswith(p):
                counts[cur][p] += 1
                break
    lines = ["# SASS mnemonic counts per kernel (cuobjdump -sass llm-sharding_b200/lib/libb200shard.so, sm_100a)", ""]
    total = collections.Counter()
    for k, c in counts.items():
        if not c:
            continue
        total.update(c)
        lines.append(f"{k}\n    " + "  ".join(f"{p}={c[p]}" for p in PAT if c[p]))
    lines += ["", "TOTAL  " + "  ".join(f"{p}={total[p]}" for p in PAT if total[p])]
    path = os.path.join(ROOT, "profiles", "sass_summary.txt")
    open(path, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[-3:]))


if __name__ == "__main__":
    sys.exit(main())
