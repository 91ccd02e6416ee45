#!/usr/bin/env python3
"""Generate tests/golden/golden.json by running the REFERENCE's own Python.

Run in the build container only (needs /root/reference, which does not exist on the GPU
box):   python tests/golden/make_golden.py

What is imported / executed from the reference (read-only, nothing copied into the repo):
  * gECC/scripts/{constants,field,ec}.py      -- Jacobian / affine secp256k1 arithmetic
  * scripts/verify_bip352_test_vector.py      -- point_add/point_double/point_multiply/
                                                 compress_point/tagged_hash + its printed values
  * test/sql/cudasp.test                      -- parsed for the BLOB literals and expected counts
  * gECC/scripts/verify_gpu_output.py         -- the three (s, P, s*P in Montgomery form) GPU dumps

The JSON holds only inputs and the reference's outputs, so the oracle (oracle/) and the
CUDA path can both be checked against it with no access to the reference.
"""
import contextlib
import io
import json
import os
import random
import re
import runpy
import sys

REF = os.environ.get("CUDASP_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))

sys.path.insert(0, os.path.join(REF, "gECC", "scripts"))
import constants  # noqa: E402  (reference)
import ec  # noqa: E402  (reference)
import field  # noqa: E402  (reference)

curve = ec.G1_SECP256K1
fq = field.Fq_SECP256K1
p = constants.SECP256K1_q
n = constants.SECP256K1_n
G = constants.SECP256K1_g1_generator

# Execute the reference verifier once; keep its functions and its printed intermediates.
_buf = io.StringIO()
with contextlib.redirect_stdout(_buf):
    ver = runpy.run_path(os.path.join(REF, "scripts", "verify_bip352_test_vector.py"))
ver_stdout = _buf.getvalue()
point_add, point_multiply = ver["point_add"], ver["point_multiply"]
compress_point, tagged_hash = ver["compress_point"], ver["tagged_hash"]


def h(x, nbytes=32):
    return format(x, "0%dx" % (2 * nbytes))


def upper64_signed(x):
    v = (x >> 192) & 0xFFFFFFFFFFFFFFFF
    return v - (1 << 64) if v >> 63 else v


def ref_pipeline(scan_sk, tweak, spend, labels):
    """The verifier script's step sequence (scripts/verify_bip352_test_vector.py:149-210),
    extended with the label rule of src/cudasp_gpu.cu:153-172 (final_point + label)."""
    S = point_multiply(scan_sk, tweak)
    ser = compress_point(S[0], S[1]) + b"\x00\x00\x00\x00"
    t = int.from_bytes(tagged_hash(b"BIP0352/SharedSecret", ser), "big")
    T = point_multiply(t, G)
    P0 = point_add(T, spend)
    cands = [upper64_signed(P0[0])]
    for L in labels:
        cands.append(upper64_signed(point_add(P0, L)[0]))
    return dict(S=[h(S[0]), h(S[1])], msg=ser.hex(), t=h(t), T=[h(T[0]), h(T[1])],
                P0=[h(P0[0]), h(P0[1])], cands=cands)


out = {"generator": "tests/golden/make_golden.py", "reference_commit": "98601999"}

# ---- 1. field KATs through the reference's Fp class (gECC/scripts/field.py:55-70)
rng = random.Random(233)
fk = []
edge = [0, 1, 2, p - 1, p - 2, (1 << 255), (1 << 256) - (1 << 32) - 978, 0x1000003D1, 0xFFFFFFFF, (1 << 128) - 1]
pairs = [(a, b) for a in edge for b in edge][:60] + [(rng.randrange(p), rng.randrange(p)) for _ in range(64)]
for a, b in pairs:
    a %= p
    b %= p
    fk.append(dict(a=h(a), b=h(b), add=h(fq.add(a, b)), sub=h(fq.sub(a, b)), mul=h(fq.mul(a, b)),
                   inv=h(fq.inv(a)) if a else h(0)))
out["field"] = fk

# ---- 2. scalar-mult KATs: the seed-42 generator of constants_generator.py:158-186, replayed
random.seed(42)
N_GEN = 3972
for _ in range(1):
    random_r = [random.randint(0, fq.p - 1) for _ in range(N_GEN)]
    random_s = [random.randint(0, fq.p - 1) for _ in range(N_GEN)]
    random_e = [random.randint(0, fq.p - 1) for _ in range(N_GEN)]
    random_priv = [random.randint(0, fq.p - 1) for _ in range(N_GEN)]
    random_k = [random.randint(0, fq.p - 1) for _ in range(N_GEN)]
keys = [curve.random_element() for _ in range(24)]
pm = []
for i in range(24):
    s, P = random_s[i], keys[i]
    R = curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(P), s))
    pm.append(dict(s=h(s), x=h(P[0]), y=h(P[1]), rx=h(R[0]), ry=h(R[1])))
out["pmul_seed42"] = pm

# the three GPU dumps recorded in gECC/scripts/verify_gpu_output.py:20-45 (Montgomery form)
src = open(os.path.join(REF, "gECC", "scripts", "verify_gpu_output.py")).read()
dumps = []
for m in re.finditer(r"'s': (0x[0-9a-f]+),\s*'x': (0x[0-9a-f]+),\s*'y': (0x[0-9a-f]+),\s*"
                     r"'result_x_mont': (0x[0-9a-f]+),\s*'result_y_mont': (0x[0-9a-f]+)", src):
    s, x, y, rxm, rym = (int(g, 16) for g in m.groups())
    dumps.append(dict(s=h(s), x=h(x), y=h(y), rx=h(fq.from_mont(rxm)), ry=h(fq.from_mont(rym))))
out["pmul_gpu_dumps"] = dumps
assert len(dumps) == 3
# the first dump is also row h=100 of test/sql/cudasp.test (scalar and point literals there)

# generate_batch_pmul_test replayed (constants_generator.py:378-413): seed 42, scalars <= 2^128
random.seed(42)
bp = []
for i in range(10):
    P = curve.random_element()
    s = random.randint(1, 2 ** 128)
    R = curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(P), s))
    bp.append(dict(s=h(s), x=h(P[0]), y=h(P[1]), rx=h(R[0]), ry=h(R[1])))
out["pmul_batch_seed42"] = bp

# small multiples of G (constants_generator.py:505-512 EC_G_TIMES_{2,3,5,10}) + edge scalars
gm = []
for k in [1, 2, 3, 5, 10, n - 1, n - 2, n + 1, n + 5, (1 << 256) - 1, 1 << 255, (1 << 128) - 1, 1 << 128]:
    R = curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(G), k))
    gm.append(dict(k=h(k), rx=h(R[0]), ry=h(R[1])))
out["g_multiples"] = gm

# ---- 3. end-to-end pipeline KATs via the verifier's own functions
rng = random.Random(352)
rows = []
for i in range(40):
    sk = rng.randrange(1, 1 << 256) if i % 5 == 0 else rng.randrange(1, n)   # some scalars >= n
    tweak = curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(G), rng.randrange(1, n)))
    spend = curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(G), rng.randrange(1, n)))
    L = [0, 0, 1, 2, 5][i % 5]
    labels = [curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(G), rng.randrange(1, n))) for _ in range(L)]
    r = ref_pipeline(sk, tweak, spend, labels)
    r.update(sk=h(sk), tweak=[h(tweak[0]), h(tweak[1])], spend=[h(spend[0]), h(spend[1])],
             labels=[[h(l[0]), h(l[1])] for l in labels])
    rows.append(r)
out["pipeline"] = rows

# the intermediates the verifier prints for the in-tree BIP-352 vector
bip = {}
for key, pat in [("S_x", r"shared_secret\.x = ([0-9a-f]{64})"), ("S_y", r"shared_secret\.y = ([0-9a-f]{64})"),
                 ("msg", r"serialized \(37 bytes\) = ([0-9a-f]{74})"), ("t", r"hash = ([0-9a-f]{64})"),
                 ("T_x", r"output_point\.x = ([0-9a-f]{64})"), ("P0_x", r"final_point\.x = ([0-9a-f]{64})"),
                 ("upper64", r"Upper 64 bits \(unsigned\): (\d+)")]:
    bip[key] = re.search(pat, ver_stdout).group(1)
bip.update(scan_sk_be=ver["scan_private_key_hex"], spend_le=ver["spend_public_key_hex"], tweak_le=ver["tweak_key_hex"],
           expected=ver["expected_output"])
out["bip352_vector"] = bip

# ---- 4. test/sql/cudasp.test, parsed: every INSERT row and every cudasp_scan query with its expectation
sql = open(os.path.join(REF, "test", "sql", "cudasp.test")).read()


def blob(lit):
    return bytes(int(x, 16) for x in re.findall(r"\\x([0-9a-fA-F]{2})", lit)).hex()


ins = []
for m in re.finditer(r"INSERT INTO test_data VALUES\s*\(BLOB '([^']*)', (\d+), BLOB '([^']*)', \[([^\]]*)\]\);", sql):
    ins.append(dict(txid=blob(m.group(1)), height=int(m.group(2)), tweak=blob(m.group(3)),
                    outputs=[int(x) for x in m.group(4).split(",")]))
assert [r["height"] for r in ins] == [100, 200, 300, 400]
queries = []
qre = re.compile(r"query I\nSELECT (COUNT\(\*\)|height) FROM cudasp_scan\(\(SELECT txid, height, tweak_key, outputs FROM (\w+)"
                 r"(?: WHERE height = (\d+))?\), BLOB '([^']*)', BLOB '([^']*)', (CAST\(\[\] AS BLOB\[\]\)|\[[^\]]*\])"
                 r"(?:, batch_size := (\d+))?\);\n----\n(-?\d+)")
for m in qre.finditer(sql):
    labels = [blob(x) for x in re.findall(r"BLOB '([^']*)'", m.group(6))]
    queries.append(dict(select=m.group(1), table=m.group(2), where_height=int(m.group(3)) if m.group(3) else None,
                        scan_sk=blob(m.group(4)), spend=blob(m.group(5)), labels=labels,
                        batch_size=int(m.group(7)) if m.group(7) else None, expect=int(m.group(8))))
assert len(queries) == 13, len(queries)
# rows visible to each query: test_data grows as the file proceeds
pos_ins = [m.start() for m in re.finditer(r"INSERT INTO test_data", sql)]
for q, m in zip(queries, qre.finditer(sql)):
    q["rows_visible"] = sum(1 for pi in pos_ins if pi < m.start())
rep = re.search(r"CREATE TABLE large_test AS\s*SELECT\s*BLOB '([^']*)' AS txid,\s*(\d+) AS height,\s*BLOB '([^']*)' AS tweak_key,"
                r"\s*\[([^\]]*)\] AS outputs\s*FROM range\((\d+)\)", sql)
out["sql"] = dict(test_data=ins, queries=queries,
                  replicated=dict(txid=blob(rep.group(1)), height=int(rep.group(2)), tweak=blob(rep.group(3)),
                                  outputs=[int(x.split("::")[0]) for x in rep.group(4).split(",")],
                                  large_test=20000, very_large_test=200000))

# ---- 5. int64 extraction sign cases (gECC/test/ec_util_test.cu:25-49)
out["upper64_cases"] = [["0000000000000000", 0], ["ffffffffffffffff", -1], ["0102030405060708", 0x0102030405060708],
                        ["7fffffffffffffff", 0x7FFFFFFFFFFFFFFF], ["8000000000000000", -(1 << 63)]]

# ---- 6. tagged-hash KATs through the verifier's tagged_hash (hashlib)
rng = random.Random(55)
th = []
for i in range(12):
    msg = bytes([2 + (i & 1)]) + rng.randbytes(32) + b"\x00\x00\x00\x00"
    th.append(dict(msg=msg.hex(), hash=tagged_hash(b"BIP0352/SharedSecret", msg).hex()))
out["tagged_hash"] = th

# ---- 7. the BIP-352 receive loop for k = 0..4 (match details), through the verifier's own primitives
# (scripts/verify_bip352_test_vector.py:25-90 point_multiply / point_add / compress_point / tagged_hash): per transaction
# the chain t_k = hash(ser(S) || ser32(k)), P_k = t_k*G + B_spend, the label sums P_k + L_i, their 64-bit prefixes, and an
# outputs list a sender paying this recipient n_pay times (some through labels) would produce, shuffled among foreign
# outputs, with the records the receiver must find: (k, output_index, label_index | -1, t_k).
rng = random.Random(3520)
recv = []
for i in range(20):
    sk = rng.randrange(1, n)
    tweak = curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(G), rng.randrange(1, n)))
    spend = curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(G), rng.randrange(1, n)))
    L = [0, 1, 2, 3, 5][i % 5]
    labels = [curve.get_xy(curve.multiply_jacobian(curve.to_jacobian(G), rng.randrange(1, n))) for _ in range(L)]
    S = point_multiply(sk, tweak)
    chain = []
    for k in range(5):
        ser = compress_point(S[0], S[1]) + k.to_bytes(4, "big")
        t = int.from_bytes(tagged_hash(b"BIP0352/SharedSecret", ser), "big")
        assert 0 < t < n
        Pk = point_add(point_multiply(t, G), spend)
        sums = [point_add(Pk, Lb) for Lb in labels]
        chain.append(dict(k=k, msg=ser.hex(), t=h(t), P=[h(Pk[0]), h(Pk[1])], label_sums=[[h(q[0]), h(q[1])] for q in sums],
                          cands=[upper64_signed(Pk[0])] + [upper64_signed(q[0]) for q in sums]))
    n_pay = 1 + i % 5                                     # the sender pays k = 0 .. n_pay-1
    pays = []                                             # (k, label index or -1)
    for k in range(n_pay):
        pays.append((k, rng.randrange(L) if L and rng.random() < 0.5 else -1))
    outs = [chain[k]["cands"][1 + li] for k, li in pays] + [rng.randrange(-(1 << 63), 1 << 63) for _ in range(rng.randrange(0, 4))]
    order = list(range(len(outs)))
    rng.shuffle(order)
    outputs = [outs[j] for j in order]
    where = {j: pos for pos, j in enumerate(order)}
    records = [dict(k=k, output_index=where[idx], label_index=li, t=chain[k]["t"]) for idx, (k, li) in enumerate(pays)]
    recv.append(dict(sk=h(sk), tweak=[h(tweak[0]), h(tweak[1])], spend=[h(spend[0]), h(spend[1])],
                     labels=[[h(l[0]), h(l[1])] for l in labels], S=[h(S[0]), h(S[1])], chain=chain, outputs=outputs,
                     records=records))
out["receive_loop"] = recv

with open(os.path.join(HERE, "golden.json"), "w") as f:
    json.dump(out, f, indent=1)
print("wrote", os.path.join(HERE, "golden.json"), os.path.getsize(os.path.join(HERE, "golden.json")), "bytes")
