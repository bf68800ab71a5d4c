"""GPU parity tests of the raw FASTQ path (fused scan kernel + compaction) against the oracle.
Everything goes through the C ABI (fxtract_b200.lib is a ctypes shim)."""
import ctypes as C
import hashlib
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


TARGET_FLAG = {0: "", 1: "H", 2: "C", 3: "Q"}


@pytest.mark.parametrize("seed", range(12))
def test_literal_seq_matches_oracle(fx, seed):
    rng = random.Random(seed)
    nrec = rng.choice([1, 3, 50, 400, 3000, 20000])
    data = helpers.make_fastq(rng, nrec, seqlen=(30, 200) if seed % 3 else (1, 40))
    for trial in range(6):
        plen = rng.choice([1, 2, 3, 5, 7, 9, 14, 15, 16, 20, 31])
        pat = helpers.pick_substring(rng, data, lo=plen, hi=plen) if rng.random() < 0.8 else helpers.rand_dna(rng, plen).encode()
        use_rc = rng.random() < 0.5
        pats = [pat] + ([fxoracle.revcomp(pat)] if use_rc else [])
        for inv in (False, True):
            flags = ("r" if use_rc else "") + ("v" if inv else "")
            want = fxoracle.run(data, flags, pattern=pat)
            got = run_raw(fx, data, fx.MODE_LITERAL, pats, invert=inv)
            assert got["ok"] == 1, got
            assert got["n_records"] == data.count(b"\n") // 4
            assert got["out"] == want, (seed, trial, pat, inv, len(got["out"]), len(want))
            cnt = int(fxoracle.run(data, flags + "c", pattern=pat))
            assert got["n_selected"] == cnt
            c2 = run_raw(fx, data, fx.MODE_LITERAL, pats, invert=inv, count_only=True)
            assert c2["n_selected"] == cnt and c2["ok"] == 1


@pytest.mark.parametrize("target", [1, 2, 3])
@pytest.mark.parametrize("seed", range(4))
def test_literal_other_targets(fx, target, seed):
    rng = random.Random(100 + seed)
    data = helpers.make_fastq(rng, rng.choice([5, 300, 5000]))
    lines = data.split(b"\n")
    for trial in range(8):
        if target == 1:
            hdr = rng.choice(lines[0::4][:-1] or lines[0:1])
            name = hdr[1:].split(b" ")[0]
            pat = name[rng.randint(0, len(name) // 2):][:rng.randint(1, 12)] or b"r"
            if rng.random() < 0.2:
                pat = b"r1"
        elif target == 2:
            pat = rng.choice([b"N:0", b"1:N", b":N:0:A", b"2", b"r1", b"zzz"])
        else:
            q = rng.choice(lines[3::4])
            s = rng.randint(0, max(0, len(q) - 4))
            pat = q[s:s + rng.randint(1, 4)] or b"I"
        # the reference's XOR rule (main.cpp:559): with -H/-C/-Q and no -r the reverse complement is searched too
        pats = [pat, fxoracle.revcomp(pat)]
        want = fxoracle.run(data, TARGET_FLAG[target], pattern=pat)
        got = run_raw(fx, data, fx.MODE_LITERAL, pats, target=target)
        assert got["ok"] == 1
        assert got["out"] == want, (target, pat)
        want1 = fxoracle.run(data, TARGET_FLAG[target] + "r", pattern=pat)   # -r + target: pattern only
        got1 = run_raw(fx, data, fx.MODE_LITERAL, [pat], target=target)
        assert got1["out"] == want1, (target, pat, "single")


def test_empty_pattern_and_newline_pattern(fx):
    rng = random.Random(5)
    data = helpers.make_fastq(rng, 100)
    got = run_raw(fx, data, fx.MODE_LITERAL, [b""])
    assert got["out"] == data and got["n_selected"] == 100
    got = run_raw(fx, data, fx.MODE_LITERAL, [b""], invert=True)
    assert got["out"] == b"" and got["n_selected"] == 0
    got = run_raw(fx, data, fx.MODE_LITERAL, [b"A\nC"])
    assert got["n_selected"] == 0


@pytest.mark.parametrize("seed", range(6))
def test_multi_and_exact_raw(fx, seed):
    rng = random.Random(200 + seed)
    data = helpers.make_fastq(rng, rng.choice([10, 700, 9000]))
    lines = data.split(b"\n")
    # multi-pattern on sequence
    pats = sorted(set(helpers.pick_substring(rng, data, lo=6, hi=25) for _ in range(rng.randint(1, 30))) |
                  set(helpers.rand_dna(rng, 12).encode() for _ in range(5)))
    pf = b"\n".join(pats) + b"\n"
    for inv in (False, True):
        want = fxoracle.run(data, "v" if inv else "", pattern_file=pf)
        got = run_raw(fx, data, fx.MODE_MULTI, pats, invert=inv)
        assert got["ok"] == 1 and got["out"] == want
    # multi on names
    names = sorted(set(l[1:].split(b" ")[0] for l in lines[0::4] if l))
    sub = sorted(set(rng.sample(names, min(len(names), 7)) + [b"nope"]))
    want = fxoracle.run(data, "H", pattern_file=b"\n".join(sub) + b"\n")
    got = run_raw(fx, data, fx.MODE_MULTI, sub, target=1)
    assert got["out"] == want
    # exact on names (-HXf) and on sequences (-Xf)
    want = fxoracle.run(data, "HX", pattern_file=b"\n".join(sub) + b"\n")
    got = run_raw(fx, data, fx.MODE_EXACT, sub, target=1)
    assert got["ok"] == 1 and got["out"] == want
    seqs = sorted(set(rng.sample(lines[1::4], min(len(lines[1::4]), 5)) + [b"ACGT"]))
    for inv in (False, True):
        want = fxoracle.run(data, "X" + ("v" if inv else ""), pattern_file=b"\n".join(seqs) + b"\n")
        got = run_raw(fx, data, fx.MODE_EXACT, seqs, invert=inv)
        assert got["out"] == want
    # comment / quality field tables
    want = fxoracle.run(data, "C", pattern_file=b"N:0\n")
    got = run_raw(fx, data, fx.MODE_MULTI, [b"N:0"], target=2)
    assert got["out"] == want
    q = rng.choice(lines[3::4])
    want = fxoracle.run(data, "QX", pattern_file=q + b"\n")
    got = run_raw(fx, data, fx.MODE_EXACT, [q], target=3)
    assert got["out"] == want


DIRTY = [
    (b"@r1 c\nACGT\n+\nIIII\n@r2\nAC\nGT\n+\nIIII\n", "wrapped"),
    (b"@r1\tc\nACGT\n+\nIIII\n", "tab"),
    (b"@r1\nACGT\n+r1\nIIII\n", "plusname"),
    (b"@r1\r\nACGT\r\n+\r\nIIII\r\n", "crlf"),
    (b"@r1\nACGT\n+\nIII\n", "shortqual"),
    (b"@r1\nACGT\n+\nIIIII\n", "longqual"),
    (b"@r1\n\n+\n\n", "emptyseq"),
    (b">r1\nACGT\n>r2\nACGT\n", "fasta"),
    (b"@r1\nACGT\n+\nIIII", "nofinalnewline"),
    (b"@r1\nACGT\n+\nIIII\n@r2\nACGT\n+\n", "truncated"),
    (b"@r1\nAC@T\n+\nIIII\n", "at_in_seq"),
    (b"@r1\nAC+T\n+\nIIII\n", "plus_in_seq"),
    (b"@r1\nAC T\n+\nIIII\n", "space_in_seq"),
    (b"@r1\nACGT\n+\nII I\n", "space_in_qual"),
    (b"@r1 \nACGT\n+\nIIII\n", "trailing_space_header"),
    (b"\n@r1\nACGT\n+\nIIII\n", "leading_blank"),
    (b"@r1\nACGT\n+\nIIII\n\n", "trailing_blank"),
    (b"@r1\nAC\x00T\n+\nIIII\n", "nul"),
    (b"@r1\nAC\xc3T\n+\nIIII\n", "highbit"),
    (b"r1\nACGT\n+\nIIII\n", "no_at"),
    (b"@r1\nAC9T\n+\nIIII\n", "digit_in_seq"),
]


@pytest.mark.parametrize("data,name", DIRTY, ids=[d[1] for d in DIRTY])
def test_non_canonical_is_rejected(fx, data, name):
    # pad with clean records on both sides so the defect sits mid-stream and across tile sizes
    rng = random.Random(1)
    clean = helpers.make_fastq(rng, 200)
    for blob in (data, clean + data, clean * 120 + data + (clean if data.endswith(b"\n") else b"")):
        got = run_raw(fx, blob, fx.MODE_LITERAL, [b"ACGTACGTACGTACGTAC"], count_only=True)
        assert got["ok"] == 0, (name, got)


def test_stray_bytes_are_rejected_wherever_they_sit(fx):
    """Bytes kseq drops, stops at or reads as EOF (0x7f in a sequence, 0x80..0xff anywhere: kseq.hpp:138-144, :167-181)
    must reject the batch on every line role and at every position inside a 4-byte word / 16-byte chunk: the range
    check of the main pass is a SWAR sum whose carries cross byte lanes (0xff is caught through the line-end check)."""
    rng = random.Random(11)
    clean = helpers.make_fastq(rng, 400, seqlen=(100, 160))
    lines = clean.split(b"\n")
    starts = []
    o = 0
    for ln in lines[:-1]:
        starts.append(o)
        o += len(ln) + 1
    ok = run_raw(fx, clean, fx.MODE_LITERAL, [b"ACGTACGTACGTACGTAC"], count_only=True)
    assert ok["ok"] == 1
    blobs = []
    for bad in (0x7f, 0x80, 0x9f, 0xa0, 0xa1, 0xc0, 0xfe, 0xff, 0x01, 0x1f):
        for role in (0, 1, 3):
            for k in range(17):
                li = 4 * rng.randrange(5, 390) + role
                pos = starts[li] + 2 + ((k + rng.randrange(0, 4) * 16) % (len(lines[li]) - 3))
                b = bytearray(clean)
                b[pos] = bad
                if k % 5 == 0 and pos + 1 < starts[li] + len(lines[li]) - 1:
                    b[pos + 1] = 0x7e                     # the neighbour a carry would land on
                blobs.append((bad, role, pos, bytes(b)))
    from fxtract_b200 import lib
    ctx = fx.Context(fx.MODE_LITERAL, count_only=True, batch_bytes=max(len(clean), 1 << 20))
    ctx.set_patterns([b"ACGTACGTACGTACGTAC"])
    for bad, role, pos, blob in blobs:
        ctx.submit_raw_bytes(blob, True)
        r = ctx.collect()
        assert r.canonical_ok == 0, (hex(bad), role, pos, pos & 15, r.violation)
    ctx.close()


def test_canonical_edge_shapes(fx):
    # shortest records, '@'/'+' opening quality lines, names without comments, long lines across tiles
    recs = [b"@\nA\n+\nI\n", b"@a\nAC\n+\n@+\n", b"@x y z\nACGTN\n+\n+@+@+\n"]
    pad = b"@pad 1:N:0\n" + b"ACGT" * 30 + b"\n+\n" + b"I" * 120 + b"\n"     # keeps the line density below 1 per 8 bytes
    big = b"@long\n" + b"ACGT" * 30000 + b"\n+\n" + b"I" * 120000 + b"\n"
    grp = b"".join(recs) + pad
    data = grp * 50 + big + grp * 7 + big * 2
    got = run_raw(fx, data, fx.MODE_LITERAL, [b"GTNXX"], invert=True)
    assert got["ok"] == 1, got
    assert got["out"] == data
    assert got["n_records"] == 50 * 4 + 1 + 28 + 2
    want = fxoracle.run(data, "", pattern=b"CGTACGTACGTACGTA")
    got = run_raw(fx, data, fx.MODE_LITERAL, [b"CGTACGTACGTACGTA"])
    assert got["out"] == want and got["n_records"] == 231


def test_dense_lines_are_rerouted_not_misparsed(fx):
    """More than one line per 8 bytes over a whole tile is outside the raw path's table capacity:
    the batch must be reported (FXG_V_DENSE_LINES) so the host re-routes it, never mis-parsed."""
    recs = [b"@\nA\n+\nI\n", b"@a\nAC\n+\n@+\n"]
    data = b"".join(recs) * 4000
    got = run_raw(fx, data, fx.MODE_LITERAL, [b"AC"], count_only=True)
    assert got["ok"] == 0 and got["violation"] & 256, got


def test_synth_device_equals_host_and_planting(fx):
    import torch
    from fxtract_b200 import lib
    spec = fx.SynthSpec(seed=3, first_record=99990, n_records=5000, read_len=150, plant_mode=1, pattern_seed=2,
                        n_patterns=100, pattern_len=20, plant_per_million=50000, plant_revcomp=1)
    host = lib.synth_host(spec)
    n = len(host)
    ctx = fx.Context(fx.MODE_MULTI, count_only=True, batch_bytes=1 << 20)
    d = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    ctx.synth_device(spec, d.data_ptr())
    dev = bytes(d[:n].cpu().numpy().tobytes())
    assert hashlib.sha256(dev).hexdigest() == hashlib.sha256(host).hexdigest()
    pats = sorted(set(lib.synth_pattern(spec, p) for p in range(100)))
    both = sorted(set(pats) | set(fxoracle.revcomp(p) for p in pats))
    ctx.set_patterns(both)
    ctx.scan_device_async(d.data_ptr(), n)
    r = ctx.scan_device_finish()
    want = int(fxoracle.run(host, "rc", pattern_file=b"\n".join(pats) + b"\n"))
    assert r.canonical_ok == 1 and r.n_records == 5000
    assert r.n_selected == want and 150 < want < 400
    ctx.close()


def test_phase_guess_redo_path(fx):
    """1-base reads whose quality line is "+" make the tile-local phase anchor (a line that is
    exactly "+") ambiguous: tiles that anchor on the quality line are re-evaluated by the REDO
    launch.  Output must still equal the oracle's, and real defects must still be caught."""
    rng = random.Random(9)
    recs = []
    for i in range(30000):
        name = "n%d_%s" % (i, "x" * rng.randint(20, 40))
        if rng.random() < 0.7:
            recs.append("@%s\n%s\n+\n+\n" % (name, rng.choice("ACGT")))
        else:
            n = rng.randint(2, 60)
            recs.append("@%s\n%s\n+\n%s\n" % (name, helpers.rand_dna(rng, n), "+" * n))
    data = "".join(recs).encode()
    assert len(data) > 20 * 32768
    for pat in (b"GATT", b"A", b"n77_"):
        tgt = 1 if pat.startswith(b"n") else 0
        flags = "H" if tgt else ""
        pats = [pat, fxoracle.revcomp(pat)] if tgt else [pat]
        want = fxoracle.run(data, flags, pattern=pat)
        got = run_raw(fx, data, fx.MODE_LITERAL, pats, target=tgt)
        assert got["ok"] == 1, got["violation"]
        assert got["out"] == want
    bad = data.replace(b"\n+\n+\n@n15000_", b"\n+\n++\n@n15000_", 1)
    if bad != data:
        got = run_raw(fx, bad, fx.MODE_LITERAL, [b"GATT"], count_only=True)
        assert got["ok"] == 0
    bad = data[:500000] + data[500000:].replace(b"\nA\n+\n+\n", b"\n9\n+\n+\n", 1)
    got = run_raw(fx, bad, fx.MODE_LITERAL, [b"GATT"], count_only=True)
    assert got["ok"] == 0


@pytest.mark.parametrize("seed", range(3))
def test_multi_kmer_raw(fx, seed):
    """Raw path + k-mer index second stage against the oracle (-f, -rf, -vf on sequences)."""
    rng = random.Random(400 + seed)
    data = helpers.make_fastq(rng, rng.choice([300, 6000]), seqlen=(30, 200))
    pats = sorted(set(helpers.pick_substring(rng, data, lo=14, hi=30) for _ in range(rng.randint(2, 40))) |
                  set(helpers.rand_dna(rng, 20).encode() for _ in range(50)))
    pats = [p for p in pats if set(p) <= set(b"ACGT") and len(p) >= 12]
    both = sorted(set(pats) | set(fxoracle.revcomp(p) for p in pats))
    pf = b"\n".join(pats) + b"\n"
    for flags, plist, inv in (("", pats, False), ("r", both, False), ("v", pats, True)):
        want = fxoracle.run(data, flags, pattern_file=pf)
        got = run_raw(fx, data, fx.MODE_MULTI, plist, invert=inv)
        assert got["ok"] == 1 and got["out"] == want, (flags, len(got["out"]), len(want))


def test_literal_fingerprint_table_forms(fx):
    """The 8-byte-unit filter picks its perfect hash cheapest form first (fxg_api.cu, build_literal_table):
    fingerprint lo + hi * mul2 in a 32-slot table, then 64 slots, then two multipliers.  Two keys that differ
    only in byte 0 ("CAAAAAAA" / "AAAAAAAA") cannot be separated with mul1 == 1 and force the general form;
    low-complexity and planted patterns must give the oracle's records in every form."""
    rng = random.Random(77)
    recs = []
    plant = [b"CAAAAAAAAAAAAAAAAAAA", b"ACGTACGTACGTACGTACGT", b"TTTTTTTTTTTTTTTTTTTTTTTTT", b"GATTACAGATTACAGATTACAG"]
    for i in range(6000):
        seq = bytearray(helpers.rand_dna(rng, rng.randint(60, 180)).encode())
        if i % 7 == 0:
            p = plant[(i // 7) % len(plant)]
            if rng.random() < 0.5:
                p = fxoracle.revcomp(p)
            o = rng.randint(0, len(seq) - len(p))
            seq[o:o + len(p)] = p
        qual = bytes(rng.randint(33, 73) for _ in range(len(seq)))
        recs.append(b"@r%d x:%d\n" % (i, i) + bytes(seq) + b"\n+\n" + qual + b"\n")
    data = b"".join(recs)
    for pat in plant:
        for use_rc in (False, True):
            pats = [pat] + ([fxoracle.revcomp(pat)] if use_rc else [])
            want = fxoracle.run(data, "r" if use_rc else "", pattern=pat)
            got = run_raw(fx, data, fx.MODE_LITERAL, pats)
            assert got["ok"] == 1
            assert got["out"] == want, (pat, use_rc, len(got["out"]), len(want))
            assert len(want) > 0


def test_multi_kmer_non_base_letters(fx):
    """The k-mer walk takes words of 0x40..0x47 / 0x50..0x57 letters through its fast path (superset of A, C, G, T:
    fxg_fields.cu, kmer_word).  Occurrences directly beside such letters (and beside N, lower case, IUPAC codes)
    must still be found and nothing else may be reported."""
    rng = random.Random(91)
    pats = sorted(set(helpers.rand_dna(rng, rng.choice([16, 20, 24, 31])).encode() for _ in range(40)))
    letters = b"ACGTACGTACGTNEQSUWBDFPRVacgtnMKY"
    recs = []
    for i in range(5000):
        n = rng.randint(40, 190)
        seq = bytearray(letters[rng.randrange(len(letters))] for _ in range(n))
        if i % 3 == 0:
            p = pats[rng.randrange(len(pats))]
            if len(p) <= n:
                o = rng.randint(0, n - len(p))
                seq[o:o + len(p)] = p
        qual = bytes(rng.randint(33, 73) for _ in range(n))
        recs.append(b"@q%d\n" % i + bytes(seq) + b"\n+\n" + qual + b"\n")
    data = b"".join(recs)
    pf = b"\n".join(pats) + b"\n"
    want = fxoracle.run(data, "", pattern_file=pf)
    got = run_raw(fx, data, fx.MODE_MULTI, pats)
    assert got["ok"] == 1 and got["out"] == want, (len(got["out"]), len(want))
    assert want.count(b"\n") // 4 > 1000
