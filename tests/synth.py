"""Seeded synthetic tweak tables with planted matches (SURVEY.md section 8d / BASELINE.md section 3).
TEST / BENCH INFRASTRUCTURE: uses the CPU oracle to produce valid curve points and the planted
candidates; never imported by the product."""
from __future__ import annotations

import random

import numpy as np

from oracle import c_oracle as CO
from oracle import ref_scan as R


def make_keys(seed: int, n_labels: int):
    rng = random.Random(seed)
    sk = rng.randrange(1, R.N)
    spend = CO.pt_mul(rng.randrange(1, R.N).to_bytes(32, "little"), R.encode_point(R.G))[0]
    labels = [CO.pt_mul(rng.randrange(1, R.N).to_bytes(32, "little"), R.encode_point(R.G))[0] for _ in range(n_labels)]
    return sk.to_bytes(32, "little"), spend, labels


def make_table(rows: int, seed: int, n_labels: int = 0, n_base: int = 100, n_per_label: int = 0, geom_p: float = 0.5,
               max_extra: int = 15, batch_rows: int = 300000, specials: bool = True):
    """-> dict(sk, spend, labels, tweaks[rows,64], outputs int64, offsets uint32[rows+1], expected sorted ids)"""
    rng = random.Random(seed)
    nrng = np.random.default_rng(seed)
    sk, spend, labels = make_keys(seed, n_labels)
    g = R.encode_point(R.G)
    p0 = CO.pt_mul(rng.randrange(1, R.N).to_bytes(32, "little"), g)[0]
    q = CO.pt_mul(rng.randrange(1, R.N).to_bytes(32, "little"), g)[0]
    tweaks = CO.gen_chain(p0, q, rows)
    tweaks = tweaks[nrng.permutation(rows)]
    n_out = 1 + np.minimum(nrng.geometric(geom_p, rows) - 1, max_extra)
    empty = set()
    if specials and rows >= 200:
        empty = set(rng.sample(range(rows), 10))
        for i in empty:
            n_out[i] = 0
    offsets = np.zeros(rows + 1, np.uint32)
    offsets[1:] = np.cumsum(n_out)
    outputs = nrng.integers(-(1 << 63), (1 << 63) - 1, int(offsets[-1]), dtype=np.int64)
    # planted rows: first/last row of every batch range + random ones
    edges = [r for b in range(0, rows, batch_rows) for r in (b, min(b + batch_rows, rows) - 1) if r not in empty]
    want = n_base + n_per_label * n_labels
    if rows <= 100000:      # historical draw (keeps the small tables of the test-suite unchanged)
        cand_rows = [r for r in range(rows) if r not in empty]
        drawn = rng.sample(cand_rows, min(len(cand_rows), want + 40))
    else:                   # large tables: sample from the range, then drop the few empty-output rows
        drawn = [r for r in rng.sample(range(rows), min(rows, want + 60)) if r not in empty]
    pool = list(dict.fromkeys(edges + drawn))
    planted = pool[:want]
    near = pool[want:want + 20] if specials else []
    sel = np.array(planted + near, dtype=np.int64)
    expected = []
    if len(sel):
        _, cands = CO.scan(sk, spend, labels, tweaks[sel], None, None, threads=8, want_cands=True)
        for j, r in enumerate(planted):
            which = 0 if j < n_base else 1 + (j - n_base) // max(n_per_label, 1)
            slot = int(offsets[r]) + rng.randrange(int(n_out[r]))
            outputs[slot] = cands[j, which]
            expected.append(r)
        for j, r in enumerate(near):
            slot = int(offsets[r]) + rng.randrange(int(n_out[r]))
            outputs[slot] = cands[len(planted) + j, 0] ^ 1     # low bit flipped: must NOT match
    return dict(sk=sk, spend=spend, labels=labels, tweaks=tweaks, outputs=outputs, offsets=offsets,
                expected=np.array(sorted(expected), dtype=np.uint32))


def sender_outputs(sk: bytes, spend: bytes, labels, tweak: bytes, recipe, rng=None):
    """Sender side of BIP-352 for ONE transaction (test infrastructure, uses the oracle's EC arithmetic): the list of
    64-bit output prefixes a sender would create paying this recipient len(recipe) times, recipe[k] = -1 (B_spend) or a
    label index, with k counting up from 0 -- plus the expected receiver records [(k, label, t_k)].
    P_k = B_m + t_k*G with t_k = hash(ser(S) || ser32(k)), S = scan_sk * tweak (= a_sum*input_hash * B_scan on the
    sender's side; the same point)."""
    from oracle import ref_scan as R
    S = R.pt_mul(R.parse_scalar(sk), R.parse_point(tweak))
    B = R.parse_point(spend)
    outs, recs = [], []
    for k, m in enumerate(recipe):
        msg = bytes([0x02 + (S[1] & 1)]) + S[0].to_bytes(32, "big") + k.to_bytes(4, "big")
        t = int.from_bytes(R.tagged_hash(msg), "big")
        Bm = B if m < 0 else R.pt_add(B, R.parse_point(labels[m]))
        Pk = R.pt_add(R.pt_mul(t, R.G), Bm)
        outs.append(R.upper64_signed(Pk[0]))
        recs.append((k, m, t))
    return outs, recs
