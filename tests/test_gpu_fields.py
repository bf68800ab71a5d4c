"""GPU parity tests of the field engine (general path kernels) through the C ABI."""
import random

import pytest

import helpers
from oracle import fxoracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fx(built):
    import fxtract_b200
    return fxtract_b200


def run_fields(fx, mode, patterns, fields, want_id=False):
    ctx = fx.Context(mode, want_pattern_id=want_id, batch_bytes=max(1 << 20, sum(map(len, fields)) + 64))
    ctx.set_patterns(patterns)
    ctx.submit_fields(fields)
    r = ctx.collect()
    flags = [r.flags[i] for i in range(len(fields))]
    ids = [r.pattern_id[i] for i in range(len(fields))] if want_id else None
    nm = r.n_matched
    ctx.close()
    return flags, ids, nm


def rand_fields(rng, n, alphabet, lo=0, hi=200):
    return [bytes(rng.choice(alphabet) for _ in range(rng.randint(lo, hi))) for _ in range(n)]


@pytest.mark.parametrize("seed", range(6))
def test_fields_literal(fx, seed):
    rng = random.Random(seed)
    alpha = b"ACGT" if seed % 2 == 0 else b"ACGTN:_ 0123"
    fields = rand_fields(rng, rng.choice([1, 33, 1000, 5000]), alpha)
    for plen in (0, 1, 2, 5, 8, 20, 64):
        f = rng.choice([x for x in fields if len(x) >= plen] or [b"A" * plen])
        s = rng.randint(0, len(f) - plen)
        pat = f[s:s + plen]
        pats = [pat] if rng.random() < 0.5 else [pat, fxoracle.revcomp(pat)]
        flags, _, nm = run_fields(fx, fx.MODE_LITERAL, pats, fields)
        want = [int(any(p in x for p in pats)) for x in fields]
        assert flags == want
        assert nm == sum(want)


@pytest.mark.parametrize("seed", range(6))
def test_fields_multi_first_hit(fx, seed):
    rng = random.Random(50 + seed)
    alpha = b"ACGT" if seed % 2 == 0 else b"abcAC:1"
    fields = rand_fields(rng, rng.choice([5, 200, 3000]), alpha, 0, 120)
    pats = set()
    for _ in range(rng.randint(1, 60)):
        f = rng.choice(fields)
        if len(f) < 2:
            continue
        n = rng.randint(1, min(12, len(f)))
        s = rng.randint(0, len(f) - n)
        pats.add(f[s:s + n])
    pats = sorted(pats) or [b"A"]
    ac = fxoracle.AC(pats)
    want_ids = [ac.first(x) for x in fields]
    flags, ids, nm = run_fields(fx, fx.MODE_MULTI, pats, fields, want_id=True)
    assert ids == want_ids
    assert flags == [int(i >= 0) for i in want_ids]


def test_fields_multi_reference_pattern_files(fx):
    import os
    rng = random.Random(9)
    pdir = os.path.join(helpers.ROOT, "tests", "golden", "patterns")
    for fn in ("random_100_30bp_patterns.txt", "random_1000_patterns_minus_small.txt"):
        pats = sorted(set(l for l in open(os.path.join(pdir, fn), "rb").read().split(b"\n") if l))
        fields = rand_fields(rng, 300, b"ACGT", 80, 160)
        for k in range(0, 300, 7):            # plant
            p = rng.choice(pats)
            pos = rng.randint(0, max(0, len(fields[k]) - len(p)))
            fields[k] = (fields[k][:pos] + p + fields[k][pos + len(p):])
        ac = fxoracle.AC(pats)
        want = [ac.first(x) for x in fields]
        flags, ids, _ = run_fields(fx, fx.MODE_MULTI, pats, fields, want_id=True)
        assert ids == want and sum(flags) >= 40


@pytest.mark.parametrize("seed", range(4))
def test_fields_exact(fx, seed):
    rng = random.Random(80 + seed)
    fields = rand_fields(rng, rng.choice([3, 500, 4000]), b"ACGT:0123456789", 0, 60)
    keys = sorted(set(rng.sample(fields, min(len(fields), 50)) + [b"absent", b""]))
    flags, ids, nm = run_fields(fx, fx.MODE_EXACT, keys, fields, want_id=True)
    lut = {k: i for i, k in enumerate(keys)}
    want = [lut.get(x, -1) for x in fields]
    assert ids == want
    assert nm == sum(1 for w in want if w >= 0)


def test_fields_exact_word_boundaries(fx):
    """Keys of every length 0..33 (the look-up hashes and verifies 8 bytes per step with a masked tail word),
    each present at every byte misalignment of the packed buffer, next to near misses: one byte changed at each
    position, one byte shorter, one byte longer."""
    rng = random.Random(4242)
    keys = [bytes(rng.choice(b"ACGT:01 ") for _ in range(n)) for n in range(34)]
    keys = sorted(set(keys))
    fields = []
    for k in keys:
        for pad in range(9):
            fields.append(bytes(rng.choice(b"xyz") for _ in range(pad)))      # shifts the next field's alignment
            fields.append(k)
        for i in range(len(k)):
            fields.append(k[:i] + bytes([k[i] ^ 1]) + k[i + 1:])
        fields.append(k[:-1])
        fields.append(k + b"A")
        fields.append(k + b"\0")
    lut = {k: i for i, k in enumerate(keys)}
    want = [lut.get(x, -1) for x in fields]
    flags, ids, nm = run_fields(fx, fx.MODE_EXACT, keys, fields, want_id=True)
    assert ids == want
    assert nm == sum(1 for w in want if w >= 0) and nm >= 9 * len(keys)


@pytest.mark.parametrize("seed", range(8))
def test_fields_multi_kmer_index(fx, seed):
    """ACGT pattern sets of length >= 12 take the packed 2-bit k-mer index (any-hit, no pattern ids):
    flags must equal plain substring search, also for hits at field edges, across non-ACGT bytes and for
    back-to-back packed fields (no k-mer may straddle two fields)."""
    rng = random.Random(300 + seed)
    minlen = rng.choice([12, 13, 16, 17, 20, 24, 31])
    pats = sorted(set(bytes(rng.choice(b"ACGT") for _ in range(rng.randint(minlen, minlen + rng.choice([0, 3, 20]))))
                      for _ in range(rng.choice([1, 7, 300, 5000]))))
    alpha = b"ACGT" if seed % 2 == 0 else b"ACGTACGTACGTNacgt"
    fields = rand_fields(rng, rng.choice([40, 2000]), alpha, 0, 260)
    for i in range(0, len(fields), 3):                   # plant: start, end, middle, split across two fields
        p = rng.choice(pats)
        f = fields[i]
        mode = rng.randint(0, 3)
        if mode == 0:
            fields[i] = p + f
        elif mode == 1:
            fields[i] = f + p
        elif mode == 2:
            c = rng.randint(0, len(f))
            fields[i] = f[:c] + p + f[c:]
        elif i + 1 < len(fields):
            c = rng.randint(1, len(p) - 1)
            fields[i] = f + p[:c]
            fields[i + 1] = p[c:] + fields[i + 1]
    want = [int(any(p in x for p in pats)) for x in fields]
    flags, _, nm = run_fields(fx, fx.MODE_MULTI, pats, fields)
    assert flags == want
    assert nm == sum(want) and 0 < nm
