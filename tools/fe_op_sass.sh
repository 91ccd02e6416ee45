#!/bin/bash
# per-field-operation SASS instruction mix (see tools/fe_op_sass.cu); prints total / IMAD.WIDE / other-IMAD / rest per op, net of the load-store frame
cd "$(dirname "$0")/.."
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -cubin -o /tmp/fe_op_sass.cubin tools/fe_op_sass.cu "$@" || exit 1
python - <<'PY'
import subprocess, re, collections
txt = subprocess.run(["cuobjdump", "-sass", "/tmp/fe_op_sass.cubin"], capture_output=True, text=True).stdout
cur, h = None, {}
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = re.sub(r"_Z\d+(op_\w+?)PK.*", r"\1", m.group(1)); h[cur] = collections.Counter(); continue
    m = re.match(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)\s*(.*?);", line)
    if m and cur:
        op = m.group(1)
        if op == "IMAD" and "0x3d1" in m.group(2): op = "IMAD*977"
        if op in ("IMAD.MOV.U32", "MOV", "HFMA2") and re.search(r"RZ, RZ, RZ|R\d+, RZ\s*$|-RZ, RZ, 0, 0", m.group(2)): op = "zero(" + op.split(".")[0] + ")"
        h[cur][op] += 1
base = h.get("op_nop", collections.Counter())
for name in sorted(h):
    if name == "op_nop": continue
    c = h[name] - base
    tot = sum(c.values()); wide = sum(v for k, v in c.items() if k.startswith("IMAD.WIDE")); oi = sum(v for k, v in c.items() if k.startswith("IMAD") and not k.startswith("IMAD.WIDE"))
    print("%-8s total %3d  IMAD.WIDE %2d  other IMAD %2d  rest %3d   fmaheavy cycles %3d | %s" % (name, tot, wide, oi, tot - wide - oi, 4 * wide + 2 * oi,
          ", ".join("%s %d" % kv for kv in c.most_common() if not kv[0].startswith("IMAD.WIDE"))))
PY
