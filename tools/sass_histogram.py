#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of the in-tree libb200ddpm.so (cuobjdump -sass): the mnemonics that prove the
tcgen05 / TMEM / TMA path (UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM / STTM = tcgen05.ld / st, UTMALDG /
UTMASTG = cp.async.bulk.tensor loads / stores, UTMAPF = prefetch.tensormap, SYNCS = mbarrier ops, LDGSTS = cp.async).

    python tools/sass_histogram.py > profiles/r02_sass_opcodes.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "generative_models_b200", "libb200ddpm.so")
KEY = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTMACCTL", "SYNCS", "LDGSTS",
       "HMMA", "MUFU", "ATOMS", "ATOMG", "REDG", "RED", "BAR", "SHFL", "LDS", "STS", "LDG", "STG", "FFMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur is not None:
            op = m.group(1)
            cur[op] += 1
            if op in ("UTMALDG", "UTMASTG"):
                cur[op + m.group(2)] += 1
    dem = subprocess.run(["cu++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines()
    print("# SASS opcode histogram of generative_models_b200/libb200ddpm.so (sm_100a), per kernel\n")
    print("`cuobjdump -sass` of the in-tree library; counts are static instruction counts.  Columns: total instructions, then "
          "the mnemonics that matter here.\n")
    print("| kernel | total | " + " | ".join(KEY) + " | TMA variants |")
    print("|---|---|" + "---|" * (len(KEY) + 1))
    tot = collections.Counter()
    for (name, c), d in zip(kernels.items(), dem):
        mm = re.match(r"(?:void )?([\w:]+(?:<.*?>)?)\(", d)
        short = (mm.group(1) if mm else d).replace("b200::", "")
        variants = ", ".join(f"{k}×{v}" for k, v in sorted(c.items()) if k.startswith(("UTMALDG.", "UTMASTG.")))
        print(f"| `{short}` | {sum(v for k, v in c.items() if '.' not in k)} | " + " | ".join(str(c.get(k, 0)) for k in KEY) + f" | {variants} |")
        tot.update(c)
    print("\nLibrary totals: " + ", ".join(f"{k} {tot.get(k, 0)}" for k in KEY if tot.get(k, 0)))


if __name__ == "__main__":
    main()
