#!/usr/bin/env python3
"""Opcode histogram of the kernels in a cubin / shared object (cuobjdump -sass), one block per kernel.

    python tools/sass_hist.py duckdb-cudasp-extension_b200/libcudasp_b200.so [name-substring ...] > profiles/rNN_sass_hist.txt

Static counts (instructions in the binary, not executed).  For k_ladder the loop body is the unrolled doubling and mixed
addition, so the static mix of the body is also the dynamic mix up to the op-list proportions.
"""
import collections
import re
import subprocess
import sys


def main():
    path, pats = sys.argv[1], sys.argv[2:]
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    cur, hist = None, collections.OrderedDict()
    ins = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)")
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            hist[cur] = collections.Counter()
            continue
        m = ins.match(line)
        if m and cur:
            hist[cur][m.group(1)] += 1
    for name, h in hist.items():
        if pats and not any(p in name for p in pats):
            continue
        total = sum(h.values())
        wide = sum(v for k, v in h.items() if k.startswith("IMAD.WIDE"))
        print("== %s" % name)
        print("   %d instructions, %d IMAD.WIDE* (%.1f %%), %.2f other instructions per IMAD.WIDE" %
              (total, wide, 100.0 * wide / max(total, 1), (total - wide) / max(wide, 1)))
        groups = collections.Counter()
        for k, v in h.items():
            groups[k.split(".")[0]] += v
        print("   by base opcode: " + ", ".join("%s %d" % kv for kv in groups.most_common(14)))
        for k, v in h.most_common(40):
            print("   %6d  %s" % (v, k))
        print()


if __name__ == "__main__":
    main()
