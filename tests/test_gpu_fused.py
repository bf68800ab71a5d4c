"""GPU parity tests of the matchers that run INSIDE the scan kernel (round 2): the packed k-mer window matcher
(multi-pattern, ACGT sets) and the whole-field hash look-up (exact), both against the oracle through the C ABI,
plus the record-capacity regime the round-1 advisor flagged (short records)."""
import os
import random

import pytest

import helpers
from oracle import fxoracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fx(built):
    import fxtract_b200
    return fxtract_b200


def run_raw(fx, data, mode, patterns, target=0, invert=False, count_only=False):
    from fxtract_b200 import lib
    ctx = fx.Context(mode, target=target, invert=invert, count_only=count_only, batch_bytes=max(len(data), 1 << 20))
    ctx.set_patterns(patterns)
    ctx.submit_raw_bytes(data, True)
    r = ctx.collect()
    out = None
    if not count_only and r.canonical_ok:
        out = b"".join(data[o:o + l] for o, l in lib.result_runs(r))
    res = dict(n_records=r.n_records, n_matched=r.n_matched, n_selected=r.n_selected, ok=r.canonical_ok,
               violation=r.violation, out=out, n_lines=r.n_lines)
    ctx.close()
    return res


def fastq_with_plants(rng, nrec, pats, every=5, seqlen=(40, 200), letters=b"ACGT", name_len=(5, 40)):
    """Canonical FASTQ whose sequences carry planted patterns at the start, the end and the middle of the read."""
    recs = []
    for i in range(nrec):
        n = rng.randint(*seqlen)
        seq = bytearray(letters[rng.randrange(len(letters))] for _ in range(n))
        if i % every == 0:
            p = pats[rng.randrange(len(pats))]
            if len(p) <= n:
                o = rng.choice([0, n - len(p), rng.randint(0, n - len(p))])
                seq[o:o + len(p)] = p
        qual = bytes(rng.randint(33, 73) for _ in range(n))
        name = b"r%d_" % i + b"x" * rng.randint(*name_len)
        recs.append(b"@" + name + b" c:%d\n" % i + bytes(seq) + b"\n+\n" + qual + b"\n")
    return b"".join(recs)


# shortest pattern -> probe stride of the window index: 12, 13 -> none (index pass + field matcher); 14..17 -> 4;
# 18..25 -> 8; >= 26 -> 16
@pytest.mark.parametrize("minlen,npat", [(12, 30), (13, 400), (14, 5), (17, 700), (18, 60), (20, 1000), (25, 3),
                                         (26, 200), (31, 900), (40, 50), (20, 3000), (16, 6000)])
def test_kmer_fused_strides_and_filters(fx, minlen, npat):
    """Every probe stride, both filter sizes (more than 32768 index entries take the 128 KiB filter), patterns of
    mixed length, hits at read edges and across tile edges; flags -f / -vf / -cf against the oracle."""
    rng = random.Random(minlen * 1000 + npat)
    pats = sorted(set(helpers.rand_dna(rng, rng.randint(minlen, minlen + rng.choice([0, 2, 9]))).encode() for _ in range(npat)) |
                  {helpers.rand_dna(rng, minlen).encode()})
    data = fastq_with_plants(rng, 4000, pats, every=4)
    assert len(data) > 8 * 32768
    pf = b"\n".join(pats) + b"\n"
    want = fxoracle.run(data, "", pattern_file=pf)
    got = run_raw(fx, data, fx.MODE_MULTI, pats)
    assert got["ok"] == 1, got["violation"]
    assert got["out"] == want, (minlen, npat, len(got["out"]), len(want))
    assert want.count(b"\n") // 4 >= 900
    want_v = fxoracle.run(data, "v", pattern_file=pf)
    got_v = run_raw(fx, data, fx.MODE_MULTI, pats, invert=True)
    assert got_v["out"] == want_v
    cnt = run_raw(fx, data, fx.MODE_MULTI, pats, count_only=True)
    assert cnt["n_selected"] == want.count(b"\n") // 4


def test_kmer_fused_candidate_list_overflow(fx):
    """Pattern-dense text: nearly every probed window passes the filter, far more than a tile's candidate list holds
    (kCandOverflow) -- the resolve kernel then walks every probe position of such a tile itself."""
    rng = random.Random(99)
    pats = sorted({b"A" * 20, b"ACGT" * 5, b"TTTTTTTTTTTTTTTTTTTTTTG"})
    recs = []
    for i in range(5000):
        kind = i % 7
        n = rng.randint(60, 160)
        if kind < 3:
            seq = b"A" * n
        elif kind == 3:
            seq = (b"ACGT" * 50)[rng.randint(0, 3):][:n]
        elif kind == 4:
            seq = b"A" * 19 + b"C" + b"A" * 19 + b"G" * (n - 39)          # 19 is not enough
        elif kind == 5:
            seq = b"T" * (n - 1) + b"G"
        else:
            seq = helpers.rand_dna(rng, n).encode()
        qual = b"A" * n if i % 2 else bytes(rng.randint(33, 73) for _ in range(n))
        recs.append(b"@o%d\n" % i + seq + b"\n+\n" + qual + b"\n")
    data = b"".join(recs)
    pf = b"\n".join(pats) + b"\n"
    for flags, kw in (("", {}), ("v", {"invert": True})):
        want = fxoracle.run(data, flags, pattern_file=pf)
        got = run_raw(fx, data, fx.MODE_MULTI, pats, **kw)
        assert got["ok"] == 1, got["violation"]
        assert got["out"] == want, (flags, len(got["out"]), len(want))
    want_q = fxoracle.run(data, "Q", pattern_file=pf)
    got_q = run_raw(fx, data, fx.MODE_MULTI, pats, target=3)
    assert got_q["out"] == want_q and want_q.count(b"\n") // 4 > 1000


def test_kmer_fused_revcomp_and_quality_lookalikes(fx):
    """-rf (reverse complements added), quality lines made of base letters (a pattern spelled in a quality line
    is not a hit), N / lower case / IUPAC letters beside the occurrences."""
    rng = random.Random(4711)
    base = sorted(set(helpers.rand_dna(rng, 20).encode() for _ in range(300)))
    both = sorted(set(base) | set(fxoracle.revcomp(p) for p in base))
    letters = b"ACGTACGTACGTNEQSUWBDFPRVacgtnMKY"
    recs = []
    for i in range(6000):
        n = rng.randint(30, 180)
        seq = bytearray(letters[rng.randrange(len(letters))] for _ in range(n))
        qual = bytearray(rng.choice(b"ACGTIF@+") for _ in range(n))
        p = both[rng.randrange(len(both))]
        if len(p) <= n:
            o = rng.randint(0, n - len(p))
            if i % 3 == 0:
                seq[o:o + len(p)] = p
            elif i % 3 == 1:
                qual[o:o + len(p)] = p            # must NOT select the record
        if qual[:1] == b"+" and n == 1:
            qual = bytearray(b"I")
        recs.append(b"@q%d\n" % i + bytes(seq) + b"\n+\n" + bytes(qual) + b"\n")
    data = b"".join(recs)
    pf = b"\n".join(base) + b"\n"
    want = fxoracle.run(data, "r", pattern_file=pf)
    got = run_raw(fx, data, fx.MODE_MULTI, both)
    assert got["ok"] == 1, got["violation"]
    assert got["out"] == want, (len(got["out"]), len(want))
    n_want = want.count(b"\n") // 4
    assert 1500 < n_want < 3000
    # quality target: now only the quality-line plants count
    want_q = fxoracle.run(data, "Q", pattern_file=b"\n".join(both) + b"\n")
    got_q = run_raw(fx, data, fx.MODE_MULTI, both, target=3)
    assert got_q["ok"] == 1 and got_q["out"] == want_q


def test_kmer_fused_patterns_with_non_base_bytes(fx):
    """The reference's own speed-test sets hold a pattern with an N (fxtract_test/speed_test/random_10000_*): such a
    pattern is indexed by the windows of its longest run of bases and compared in full when a window hits, so the set
    stays on the fused matcher; the N only matches an N in the text."""
    rng = random.Random(31)
    pats = set(helpers.rand_dna(rng, rng.randint(20, 40)).encode() for _ in range(400))
    odd = []
    for _ in range(40):
        p = bytearray(helpers.rand_dna(rng, rng.randint(28, 60)).encode())
        pos = rng.choice([0, len(p) - 1, rng.randint(0, len(p) - 1)])
        p[pos] = rng.choice(b"NRYn")
        odd.append(bytes(p))
    allp = sorted(pats | set(odd))
    recs = []
    for i in range(5000):
        n = rng.randint(70, 200)
        seq = bytearray(helpers.rand_dna(rng, n).encode())
        if i % 3 == 0:
            p = rng.choice(odd if i % 2 else allp)
            o = rng.randint(0, n - len(p))
            seq[o:o + len(p)] = p
            if i % 12 == 3:                                   # the odd byte replaced by a base: must NOT match
                for k, ch in enumerate(p):
                    if ch not in b"ACGT":
                        seq[o + k] = ord("A")
        recs.append(b"@o%d\n" % i + bytes(seq) + b"\n+\n" + b"I" * n + b"\n")
    data = b"".join(recs)
    want = fxoracle.run(data, "", pattern_file=b"\n".join(allp) + b"\n")
    got = run_raw(fx, data, fx.MODE_MULTI, allp)
    assert got["ok"] == 1 and got["out"] == want, (len(got["out"]), len(want))
    assert 1000 < want.count(b"\n") // 4 < 1700


def test_kmer_fused_equals_two_stage(fx, monkeypatch):
    """FXG_TWO_STAGE=1 keeps the round-1 pipeline (index pass + field matcher); both must select the same records."""
    rng = random.Random(5)
    pats = sorted(set(helpers.rand_dna(rng, 20).encode() for _ in range(500)))
    data = fastq_with_plants(rng, 5000, pats, every=3)
    fused = run_raw(fx, data, fx.MODE_MULTI, pats)
    names = sorted(set(l[1:].split(b" ")[0] for l in data.split(b"\n")[0::4] if l))
    keys = sorted(set(rng.sample(names, 200) + [b"absent"]))
    fused_x = run_raw(fx, data, fx.MODE_EXACT, keys, target=1)
    monkeypatch.setenv("FXG_TWO_STAGE", "1")
    two = run_raw(fx, data, fx.MODE_MULTI, pats)
    two_x = run_raw(fx, data, fx.MODE_EXACT, keys, target=1)
    assert fused["ok"] == 1 and two["ok"] == 1 and fused["out"] == two["out"] and len(fused["out"]) > 0
    assert fused_x["out"] == two_x["out"] and fused_x["n_selected"] == 200


@pytest.mark.parametrize("seed", range(4))
def test_exact_fused_all_targets(fx, seed):
    """-HXf / -Xf / -CXf / -QXf with keys of every length and records placed so that header, sequence and
    quality lines straddle tile edges (24 KiB tiles; the straddling line is looked up by the finalize kernel)."""
    rng = random.Random(600 + seed)
    recs = []
    for i in range(7000):
        n = rng.randint(1, 170)
        name = b"n%d" % i + bytes(rng.choice(b"abcXYZ:_0123") for _ in range(rng.randint(0, 45)))
        comment = b"" if rng.random() < 0.3 else b" " + bytes(rng.choice(b"abc 123:N") for _ in range(rng.randint(1, 25))).rstrip(b" ") + b"e"
        seq = helpers.rand_dna(rng, n).encode()
        qual = bytes(rng.randint(33, 73) for _ in range(n))
        if n == 1 and qual == b"+":
            qual = b"I"
        recs.append(b"@" + name + comment + b"\n" + seq + b"\n+\n" + qual + b"\n")
    data = b"".join(recs)
    lines = data.split(b"\n")
    names = [l[1:].split(b" ")[0] for l in lines[0::4] if l]
    comments = [l.split(b" ", 1)[1] for l in lines[0::4] if b" " in l]
    for target, flag, pool in ((1, "H", names), (0, "", lines[1::4]), (2, "C", comments), (3, "Q", lines[3::4])):
        keys = sorted(set(rng.sample(pool, 300) + [b"absent_key", pool[0], pool[-1]]))
        keys = [k for k in keys if k and b"\t" not in k]
        pf = b"\n".join(keys) + b"\n"
        for inv in (False, True):
            want = fxoracle.run(data, flag + "X" + ("v" if inv else ""), pattern_file=pf)
            got = run_raw(fx, data, fx.MODE_EXACT, keys, target=target, invert=inv)
            assert got["ok"] == 1, got["violation"]
            assert got["out"] == want, (target, inv, len(got["out"]), len(want))
        cnt = run_raw(fx, data, fx.MODE_EXACT, keys, target=target, count_only=True)
        assert cnt["n_selected"] == fxoracle.run(data, flag + "X", pattern_file=pf).count(b"\n") // 4


def test_exact_fused_dense_hits(fx):
    """Every read name is in the key list and the reads are short: ~200 look-up hits per 24 KiB tile, several rounds
    of the resolve kernel's four-lines-per-lane walk, every lane on the verification path.  Same for sequence keys."""
    rng = random.Random(4242)
    recs, names, seqs = [], [], []
    for i in range(12000):
        n = rng.randint(20, 40)
        name = b"k%d" % i
        seq = helpers.rand_dna(rng, n).encode()
        names.append(name); seqs.append(seq)
        recs.append(b"@" + name + b" x\n" + seq + b"\n+\n" + bytes(rng.randint(33, 73) for _ in range(n)) + b"\n")
    data = b"".join(recs)
    for target, flag, pool in ((1, "H", names), (0, "", seqs)):
        keys = sorted(set(pool[::1 if target == 1 else 2]) | {b"absent"})
        pf = b"\n".join(keys) + b"\n"
        for inv in (False, True):
            want = fxoracle.run(data, flag + "X" + ("v" if inv else ""), pattern_file=pf)
            got = run_raw(fx, data, fx.MODE_EXACT, keys, target=target, invert=inv)
            assert got["ok"] == 1, got["violation"]
            assert got["out"] == want, (target, inv, len(got["out"]), len(want))
    assert len(fxoracle.run(data, "HX", pattern_file=b"\n".join(names) + b"\n")) == len(data)


def test_exact_fused_long_reads_and_redo(fx):
    """Tiles without a "+" anchor (reads longer than a tile) and tiles whose guessed phase is wrong (1-base reads
    with "+" quality) take the REDO launch, which repeats the look-ups under the true phase."""
    rng = random.Random(77)
    recs = []
    for i in range(400):
        kind = rng.random()
        name = b"L%d_%s" % (i, b"y" * rng.randint(36, 60))      # long names keep the line density under 1 per 12 bytes
        if kind < 0.2:
            n = rng.randint(30000, 90000)                      # longer than one or several tiles
            recs.append(b"@" + name + b"\n" + helpers.rand_dna(rng, n).encode() + b"\n+\n" + b"I" * n + b"\n")
        elif kind < 0.6:
            for k in range(60):                                # runs of 1-base reads whose quality line is "+"
                recs.append(b"@" + name + b".%d\n" % k + rng.choice([b"A", b"C", b"G", b"T"]) + b"\n+\n+\n")
        else:
            n = rng.randint(20, 150)
            recs.append(b"@" + name + b" cm\n" + helpers.rand_dna(rng, n).encode() + b"\n+\n" + b"F" * n + b"\n")
    data = b"".join(recs)
    lines = data.split(b"\n")
    names = [l[1:].split(b" ")[0] for l in lines[0::4] if l]
    keys = sorted(set(rng.sample(names, 500) + [b"nope"]))
    pf = b"\n".join(keys) + b"\n"
    want = fxoracle.run(data, "HX", pattern_file=pf)
    got = run_raw(fx, data, fx.MODE_EXACT, keys, target=1)
    assert got["ok"] == 1, got["violation"]
    assert got["out"] == want, (len(got["out"]), len(want))
    assert got["n_selected"] == 500
    seqs = sorted(set(rng.sample(lines[1::4], 300)))
    want = fxoracle.run(data, "X", pattern_file=b"\n".join(seqs) + b"\n")
    got = run_raw(fx, data, fx.MODE_EXACT, seqs, target=0)
    assert got["ok"] == 1 and got["out"] == want
    # the k-mer matcher over the same input (long reads, occurrences across several tile edges)
    pats = sorted(set(helpers.rand_dna(rng, 24).encode() for _ in range(50)) |
                  set(s[o:o + 24] for s in lines[1::4] if len(s) > 30000 for o in (0, 24576 - 10, 32768 - 7, len(s) - 24)
                      if len(s[o:o + 24]) == 24))
    want = fxoracle.run(data, "", pattern_file=b"\n".join(pats) + b"\n")
    got = run_raw(fx, data, fx.MODE_MULTI, pats)
    assert got["ok"] == 1 and got["out"] == want and len(want) > 0


@pytest.mark.parametrize("rec_bytes", [33, 40, 47, 48, 49, 56, 64])
def test_short_records_are_never_dropped(fx, rec_bytes):
    """Round-1 advisor finding: canonical FASTQ with 32..47 bytes per record passed the density check but
    overflowed the record capacity (records silently dropped).  The line table now holds one line per 12 bytes and
    the per-record buffers are sized with that, so such a batch is either handled completely or reported
    (FXG_V_DENSE_LINES) and re-routed -- never truncated."""
    rng = random.Random(rec_bytes)
    nrec = 60000                                        # > 15 KB of input per the finding, many tiles
    recs = []
    for i in range(nrec):
        # "@" name "\n" seq "\n+\n" qual "\n": 6 + |name| + 2 |seq| bytes
        slen = rng.randint(3, max(3, (rec_bytes - 9) // 2))
        nlen = rec_bytes - 6 - 2 * slen
        if nlen < 1:
            slen = (rec_bytes - 7) // 2
            nlen = rec_bytes - 6 - 2 * slen
        name = (b"%x" % i).rjust(nlen, b"n")[-nlen:]
        seq = helpers.rand_dna(rng, slen).encode()
        recs.append(b"@" + name + b"\n" + seq + b"\n+\n" + b"I" * slen + b"\n")
    data = b"".join(recs)
    assert abs(len(data) / nrec - rec_bytes) < 1
    pat = b"ACG"
    for inv, cnt_only in ((False, False), (True, False), (False, True)):
        got = run_raw(fx, data, fx.MODE_LITERAL, [pat], invert=inv, count_only=cnt_only)
        if not got["ok"]:
            assert got["violation"] & 256, got          # reported, to be re-routed by the host
            assert rec_bytes <= 49                      # 48-byte records sit exactly on the table's capacity
            continue
        assert got["n_records"] == nrec
        want = fxoracle.run(data, "v" if inv else "", pattern=pat)
        if cnt_only:
            assert got["n_selected"] == want.count(b"\n") // 4
        else:
            assert got["out"] == want, (rec_bytes, inv, len(got["out"]), len(want))
    if rec_bytes >= 56:
        assert run_raw(fx, data, fx.MODE_LITERAL, [pat], count_only=True)["ok"] == 1


def test_d2h_bytes_follow_the_run_count(fx):
    """Only the runs a batch selected travel back: a 1024-run prefix rides with the counters, the rest is fetched
    once the count is known."""
    from fxtract_b200 import lib
    rng = random.Random(12)
    data = helpers.make_fastq(rng, 20000)
    for pat, inv in ((b"ACGTACGTACGTACGTACGTAAAA", False), (b"AC", False), (b"ACGTAC", True)):
        ctx = fx.Context(fx.MODE_LITERAL, invert=inv, batch_bytes=max(len(data), 1 << 20))
        ctx.set_patterns([pat])
        ctx.submit_raw_bytes(data, True)
        r = ctx.collect()
        runs = lib.result_runs(r)
        out = b"".join(data[o:o + l] for o, l in runs)
        assert out == fxoracle.run(data, "v" if inv else "", pattern=pat)
        d2h = ctx.last_d2h_bytes()
        assert d2h <= 64 + 16 * max(1024, len(runs)), (d2h, len(runs))
        ctx.close()
