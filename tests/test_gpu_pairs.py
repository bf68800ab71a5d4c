"""GPU parity tests of mates on the raw path (two files / -I): the device applies each strategy's pair rule
(pair_select_kernel / pair_spread_kernel, fxg_scan.cu) and the host cuts both streams at equal record counts
(run_raw_pairs, strategies.cpp).  Checked through the C ABI against the oracle and through the CLI against the
reference binary (FXTRACT_PATH=raw forbids the fields fallback)."""
import ctypes as C
import os
import random
import subprocess

import pytest

import helpers
from oracle import fxoracle

pytestmark = pytest.mark.gpu

V_PAIRING = 1024


@pytest.fixture(scope="module")
def fx(built):
    import fxtract_b200
    return fxtract_b200


def run_pairs(fx, d1, d2, mode, patterns, target=0, invert=False, count_only=False):
    """d2 None: interleaved.  Returns the output the host would write (pairs in input order)."""
    from fxtract_b200 import lib
    total = len(d1) + (len(d2) if d2 else 0)
    ctx = fx.Context(mode, target=target, invert=invert, count_only=count_only, batch_bytes=max(total, 1 << 20))
    ctx.set_patterns(patterns)
    b1 = C.create_string_buffer(d1, len(d1))
    if d2 is None:
        ctx.submit_raw_interleaved_from(C.addressof(b1), len(d1), True)
    else:
        b2 = C.create_string_buffer(d2, len(d2))
        ctx.submit_raw_pair_from(C.addressof(b1), len(d1), C.addressof(b2), len(d2), True)
    r = ctx.collect()
    out = None
    if not count_only and r.canonical_ok:
        runs = lib.result_runs(r)
        if d2 is None:
            out = b"".join(d1[o:o + l] for o, l in runs)
        else:
            np_ = len(runs) // 2
            assert len(runs) == 2 * np_
            parts = []
            for k in range(np_):
                o, l = runs[k]
                parts.append(d1[o:o + l])
                o, l = runs[np_ + k]
                assert o >= len(d1)
                parts.append(d2[o - len(d1):o - len(d1) + l])
            out = b"".join(parts)
    res = dict(n_records=r.n_records, n_selected=r.n_selected, ok=r.canonical_ok, violation=r.violation, out=out)
    ctx.close()
    return res


def mates(rng, npairs, pats, every=6, seqlen=(40, 180)):
    """Two canonical FASTQ files of npairs records; patterns planted in read 1, read 2, both or neither."""
    f1, f2 = [], []
    for i in range(npairs):
        recs = []
        for which in (1, 2):
            n = rng.randint(*seqlen)
            seq = bytearray(helpers.rand_dna(rng, n).encode())
            kind = (i // every) % 4 if i % every == 0 else 0          # 0 none, 1 read 1, 2 read 2, 3 both
            if kind in (which, 3):
                q = rng.choice(pats)
                if len(q) <= n:
                    o = rng.randint(0, n - len(q))
                    seq[o:o + len(q)] = q
            qual = bytes(rng.randint(33, 73) for _ in range(n))
            recs.append(b"@p%d/%d x%d\n" % (i, which, i % 5) + bytes(seq) + b"\n+\n" + qual + b"\n")
        f1.append(recs[0])
        f2.append(recs[1])
    return b"".join(f1), b"".join(f2), f1, f2


@pytest.mark.parametrize("seed", range(6))
def test_literal_pairs_match_oracle(fx, seed):
    rng = random.Random(400 + seed)
    npairs = rng.choice([1, 2, 33, 500, 4000, 20000])
    pat = helpers.rand_dna(rng, rng.choice([6, 15, 20, 31])).encode()
    d1, d2, f1, f2 = mates(rng, npairs, [pat], every=rng.choice([1, 3, 7]))
    inter = b"".join(a + b for a, b in zip(f1, f2))
    for inv in (False, True):
        for use_rc in (False, True):
            flags = ("v" if inv else "") + ("r" if use_rc else "")
            pats = [pat] + ([fxoracle.revcomp(pat)] if use_rc else [])
            want = fxoracle.run(d1, flags, pattern=pat, data2=d2)
            got = run_pairs(fx, d1, d2, fx.MODE_LITERAL, pats, invert=inv)
            assert got["ok"] == 1 and got["n_records"] == npairs, got
            assert got["out"] == want, (seed, flags, len(got["out"]), len(want))
            assert got["n_selected"] == int(fxoracle.run(d1, flags + "c", pattern=pat, data2=d2))
            cnt = run_pairs(fx, d1, d2, fx.MODE_LITERAL, pats, invert=inv, count_only=True)
            assert cnt["ok"] == 1 and cnt["n_selected"] == got["n_selected"]
            want_i = fxoracle.run(inter, flags + "I", pattern=pat)
            got_i = run_pairs(fx, inter, None, fx.MODE_LITERAL, pats, invert=inv)
            assert got_i["ok"] == 1 and got_i["n_records"] == npairs
            assert got_i["out"] == want_i, (seed, flags, "interleaved")
            assert want_i == want                      # same pairs either way


@pytest.mark.parametrize("seed", range(4))
def test_multi_and_exact_pairs_match_oracle(fx, seed):
    rng = random.Random(500 + seed)
    npairs = rng.choice([50, 3000, 15000])
    pats = sorted(set(helpers.rand_dna(rng, 20).encode() for _ in range(rng.choice([3, 40, 300]))))
    d1, d2, f1, f2 = mates(rng, npairs, pats, every=rng.choice([2, 5]))
    inter = b"".join(a + b for a, b in zip(f1, f2))
    pf = b"\n".join(pats) + b"\n"
    # multipattern_search: -v selects the pairs where NEITHER mate has a hit (main.cpp:427-438)
    for flags in ("", "v"):
        want = fxoracle.run(d1, flags, pattern_file=pf, data2=d2)
        got = run_pairs(fx, d1, d2, fx.MODE_MULTI, pats, invert="v" in flags)
        assert got["ok"] == 1 and got["out"] == want, (seed, flags, len(got["out"]), len(want))
        assert got["n_selected"] == int(fxoracle.run(d1, flags + "c", pattern_file=pf, data2=d2))
        got_i = run_pairs(fx, inter, None, fx.MODE_MULTI, pats, invert="v" in flags)
        assert got_i["ok"] == 1 and got_i["out"] == fxoracle.run(inter, flags + "I", pattern_file=pf)
    # hash_search on names: ids of some first mates, some second mates; -v selects on read 1 only (main.cpp:331-343)
    ids = [f1[i].split(b" ")[0][1:] for i in range(0, npairs, 7)] + [f2[i].split(b" ")[0][1:] for i in range(3, npairs, 11)] + [b"none"]
    idf = b"\n".join(ids) + b"\n"
    for flags in ("HX", "HXv"):
        want = fxoracle.run(d1, flags, pattern_file=idf, data2=d2)
        got = run_pairs(fx, d1, d2, fx.MODE_EXACT, sorted(set(ids)), target=1, invert="v" in flags)
        assert got["ok"] == 1 and got["out"] == want, (seed, flags, len(got["out"]), len(want))
        assert got["n_selected"] == int(fxoracle.run(d1, flags + "c", pattern_file=idf, data2=d2))
        got_i = run_pairs(fx, inter, None, fx.MODE_EXACT, sorted(set(ids)), target=1, invert="v" in flags)
        assert got_i["ok"] == 1 and got_i["out"] == fxoracle.run(inter, flags + "I", pattern_file=idf)


def test_pairing_violations(fx):
    rng = random.Random(7)
    pat = b"ACGTACGTACGTACGTACGT"
    d1, d2, f1, f2 = mates(rng, 300, [pat])
    # unequal parts, an odd interleaved batch, a part that ends inside a record
    bad = [(d1, b"".join(f2[:299])), (b"".join(f1[:299]), d2), (d1 + f2[0], b"".join(f2[1:]))]
    for a, b in bad:
        got = run_pairs(fx, a, b, fx.MODE_LITERAL, [pat])
        assert got["ok"] == 0 and got["violation"] & V_PAIRING, got
    got = run_pairs(fx, d1 + d2[:len(f2[0])], None, fx.MODE_LITERAL, [pat])
    assert got["ok"] == 0 and got["violation"] & V_PAIRING, got
    cut = len(f1[0]) // 2
    got = run_pairs(fx, d1[:-cut], d2, fx.MODE_LITERAL, [pat])
    assert got["ok"] == 0, got
    ok = run_pairs(fx, d1, d2, fx.MODE_LITERAL, [pat])
    assert ok["ok"] == 1 and ok["n_records"] == 300


@pytest.fixture(scope="module")
def cli(built):
    assert os.path.exists(helpers.PRODUCT_CLI), "host CLI not built"
    if not os.path.exists(helpers.REF_CLI):
        pytest.skip("oracle/_ref/fxtract not built")
    return dict(os.environ)


def test_cli_mates_on_the_raw_path(cli, tmp_path):
    """The reference's default mode (files two at a time, main.cpp:694-697,763-764) and -I on canonical FASTQ: every
    strategy, -v, -c, per-pattern output files; small staging buffers force many batches and carried tails; mates of
    different lengths make the two streams drift; three contexts exercise the ordering."""
    rng = random.Random(21)
    wd = str(tmp_path)
    pats = sorted(set(helpers.rand_dna(rng, rng.choice([20, 22, 25])).encode() for _ in range(50)))
    d1, d2, f1, f2 = mates(rng, 40000, pats, every=3, seqlen=(50, 150))
    # read 2 shorter than read 1 on average: the streams advance at different byte rates
    d2s = b"".join(r.split(b"\n")[0] + b"\n" + r.split(b"\n")[1][:60] + b"\n+\n" + r.split(b"\n")[3][:60] + b"\n" for r in f2)
    open(os.path.join(wd, "r1.fq"), "wb").write(d1)
    open(os.path.join(wd, "r2.fq"), "wb").write(d2)
    open(os.path.join(wd, "r2s.fq"), "wb").write(d2s)
    open(os.path.join(wd, "il.fq"), "wb").write(b"".join(a + b for a, b in zip(f1, f2)))
    open(os.path.join(wd, "r2short.fq"), "wb").write(b"".join(f2[:31000]))
    open(os.path.join(wd, "p.txt"), "wb").write(b"\n".join(pats) + b"\n")
    files = ["o_a.fq", "o_b.fq"]
    with open(os.path.join(wd, "pf.txt"), "wb") as f:
        for k, q in enumerate(pats):
            f.write(q + (b"\t" + files[k % 3].encode() if k % 3 < 2 else b"") + b"\n")
    ids = [f1[i].split(b" ")[0][1:] for i in range(0, 40000, 9)] + [f2[i].split(b" ")[0][1:] for i in range(4, 40000, 13)]
    open(os.path.join(wd, "ids.txt"), "wb").write(b"\n".join(ids) + b"\n")
    with open(os.path.join(wd, "idf.txt"), "wb") as f:
        for k, q in enumerate(ids):
            f.write(q + (b"\t" + files[k % 2].encode() if k % 4 else b"") + b"\n")
    pat = pats[5].decode()
    env = dict(cli)
    env.update(FXTRACT_PATH="raw", FXTRACT_BATCH_MIB="1", FXTRACT_DEVICES="0,0,0", FXTRACT_STATS="1")
    argvs = ([pat, "r1.fq", "r2.fq"], ["-v", pat, "r1.fq", "r2.fq"], ["-c", pat, "r1.fq", "r2.fq"], ["-r", pat, "r1.fq", "r2s.fq"],
             ["-f", "p.txt", "r1.fq", "r2.fq"], ["-vf", "p.txt", "r1.fq", "r2s.fq"], ["-cf", "p.txt", "r1.fq", "r2.fq"],
             ["-vcf", "p.txt", "r2.fq", "r1.fq"], ["-rf", "p.txt", "r1.fq", "r2.fq"],
             ["-HXf", "ids.txt", "r1.fq", "r2.fq"], ["-HXvf", "ids.txt", "r1.fq", "r2.fq"], ["-HXcf", "ids.txt", "r1.fq", "r2s.fq"],
             ["-I", pat, "il.fq"], ["-Iv", pat, "il.fq"], ["-Icf", "p.txt", "il.fq"], ["-Ivf", "p.txt", "il.fq"],
             ["-IHXf", "ids.txt", "il.fq"], ["-Q", "II", "r1.fq", "r2.fq"], ["-Hc", "p1", "r1.fq", "r2.fq"],
             ["-f", "pf.txt", "r1.fq", "r2.fq"], ["-If", "pf.txt", "il.fq"], ["-HXf", "idf.txt", "r1.fq", "r2.fq"],
             ["-IHXf", "idf.txt", "il.fq"], ["-c", pat, "r1.fq", "r2.fq", "r2.fq", "r1.fq"])
    for argv in argvs:
        for f in files:
            if os.path.exists(os.path.join(wd, f)):
                os.remove(os.path.join(wd, f))
        r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        want = {f: helpers.read_or_none(wd, f) for f in files}
        for f in files:
            if os.path.exists(os.path.join(wd, f)):
                os.remove(os.path.join(wd, f))
        r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, timeout=600)
        assert r2.returncode == 0, (argv, r2.stderr[-400:])
        assert b"fields 0 B" in r2.stderr and b"raw 0 B" not in r2.stderr, (argv, r2.stderr[-300:])
        assert r1.stdout == r2.stdout, (argv, len(r1.stdout), len(r2.stdout))
        got = {f: helpers.read_or_none(wd, f) for f in files}
        assert got == want, (argv, {f: (len(want[f] or b""), len(got[f] or b"")) for f in files})
        assert len(r1.stdout) > 0 or any(want.values()), argv
    # a shorter second file ends the pair stream (fx.cpp:70-75); what is left of the longer one is never looked at
    env.pop("FXTRACT_PATH")
    for argv in ([pat, "r1.fq", "r2short.fq"], ["-vf", "p.txt", "r2short.fq", "r1.fq"]):
        r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, timeout=600)
        assert r2.returncode == 0 and r1.stdout == r2.stdout, (argv, len(r1.stdout), len(r2.stdout), r2.stderr[-300:])
        assert b"raw 0 B" not in r2.stderr
