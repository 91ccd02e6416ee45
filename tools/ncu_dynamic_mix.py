#!/usr/bin/env python3
"""Dynamic instruction mix of one kernel from an ncu report's source page (per-SASS-instruction execution counts):

    ncu -i prof.ncu-rep --page source --csv > k.csv ; python tools/ncu_dynamic_mix.py k.csv [rows]

Prints executed warp instructions by opcode, per row-warp when `rows` (rows processed by the launch) is given, the share that
runs on the fmaheavy pipe (IMAD*) and that pipe's cycles under the measured cost model (IMAD.WIDE 4 cycles, other IMAD 2)."""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    r = list(csv.reader(open(path)))
    hdr = next(i for i, x in enumerate(r) if x and x[0] == "Address")
    h = r[hdr]
    si, ei = h.index("Source"), h.index("Instructions Executed")
    stall_cols = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
    mix, stalls = collections.Counter(), collections.Counter()
    for x in r[hdr + 1:]:
        if len(x) <= ei:
            continue
        m = re.match(r"\s*(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)", x[si])
        if not m:
            continue
        mix[m.group(1)] += int(x[ei] or 0)
        for i in stall_cols:
            stalls[h[i]] += int(x[i] or 0)
    total = sum(mix.values())
    warps = rows / 32.0 if rows else 1.0
    wide = sum(v for k, v in mix.items() if k.startswith("IMAD.WIDE"))
    imad_other = sum(v for k, v in mix.items() if k.startswith("IMAD") and not k.startswith("IMAD.WIDE"))
    print("%s: %d warp instructions executed%s" % (r[0][1][:80] if r[0] else "?", total, ", %.0f per row" % (total / warps) if rows else ""))
    print("IMAD.WIDE*: %d (%.1f %%)%s;  other IMAD* (fmaheavy too): %d (%.1f %%);  everything else: %d (%.1f %%)" % (
        wide, 100.0 * wide / total, ", %.0f per row" % (wide / warps) if rows else "", imad_other, 100.0 * imad_other / total,
        total - wide - imad_other, 100.0 * (total - wide - imad_other) / total))
    fm = 4 * wide + 2 * imad_other
    print("fmaheavy pipe cycles by the cost model (4 per IMAD.WIDE, 2 per other IMAD): %d, of which other IMAD %.1f %%" % (fm, 200.0 * imad_other / fm))
    for k, v in mix.most_common(25):
        print("  %12d  %5.1f %%  %s%s" % (v, 100.0 * v / total, k, "   (%.0f per row)" % (v / warps) if rows else ""))
    st = sum(stalls.values())
    print("stall samples (all): " + ", ".join("%s %.1f %%" % (k[6:], 100.0 * v / st) for k, v in stalls.most_common(8)))


if __name__ == "__main__":
    main()
