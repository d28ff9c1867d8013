#!/usr/bin/env python3
"""Per-kernel SASS opcode counts of librslqr.so (cuobjdump -sass): the evidence for which units a kernel uses
(DMMA = fp64 tensor path, UBLKCP = TMA bulk copies, LDGSTS = cp.async, SYNCS = mbarrier, SHFL / LDS / STS = LSU data pipe)."""
import collections
import re
import subprocess
import sys

WANT = ["DMMA", "DFMA", "DMUL", "DADD", "MUFU", "SHFL", "LDS", "STS", "LDG", "STG", "LDGSTS", "UBLKCP", "SYNCS", "BAR", "UTMALDG", "UTCHMMA", "LDTM"]


def main():
    so = sys.argv[1] if len(sys.argv) > 1 else "rslqr_b200/librslqr.so"
    filt = sys.argv[2] if len(sys.argv) > 2 else ""
    txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    fn = None
    counts = collections.OrderedDict()
    wide = collections.defaultdict(collections.Counter)
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            fn = fn.replace("void rslqr::", "")
            counts[fn] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
        if m and fn:
            counts[fn][m.group(1)] += 1
            counts[fn]["_total"] += 1
            if m.group(1) in ("LDS", "STS", "LDG", "STG"):
                w = "128" if ".128" in m.group(2) else ("64" if ".64" in m.group(2) else "32")
                wide[fn][m.group(1) + "." + w] += 1
    print("| kernel | total | " + " | ".join(WANT) + " | LDS/STS widths |")
    print("|---|---|" + "---|" * (len(WANT) + 1))
    for fn, c in counts.items():
        if filt and filt not in fn:
            continue
        w = ", ".join(f"{k}:{v}" for k, v in sorted(wide[fn].items()) if k.startswith(("LDS", "STS")))
        print(f"| `{fn}` | {c['_total']} | " + " | ".join(str(c[k]) for k in WANT) + f" | {w} |")


if __name__ == "__main__":
    main()
