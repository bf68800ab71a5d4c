"""CPU tests of the host CLI's own logic (reader with kseq semantics, pair rules per strategy,
output routing, option surface, raw->fields re-routing) against the reference binary and the
reference's golden vectors.

There is no GPU in the build container, and the product has no CPU matcher, so these tests run
the real `fxtract` host binary against tests/cpu_shim (a CPU stand-in for the C ABI's fields
path, TEST INFRASTRUCTURE ONLY, reached through an explicit LD_LIBRARY_PATH).  The same cases
run against the real libfxgpu.so on the GPU box in test_cli_golden.py (-m gpu)."""
import os
import random
import subprocess

import pytest

import cases
import helpers

SHIM_DIR = os.path.join(helpers.ROOT, "tests", "cpu_shim")
SHIM_BUILD = os.path.join(SHIM_DIR, "_build")


@pytest.fixture(scope="module")
def shim_env(built):
    os.makedirs(SHIM_BUILD, exist_ok=True)
    subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-o", os.path.join(SHIM_BUILD, "libfxgpu.so"),
                    os.path.join(SHIM_DIR, "shim_fxgpu.c"), os.path.join(helpers.ROOT, "oracle", "fxo_ac.c")], check=True)
    assert os.path.exists(helpers.PRODUCT_CLI), "host CLI not built"
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = SHIM_BUILD
    return env


ALL = cases.LITERAL_CASES + cases.REGEX_CASES


@pytest.mark.parametrize("case", ALL, ids=[c[0] for c in ALL])
def test_host_cli_golden(case, shim_env, tmp_path):
    wd = helpers.stage_golden(tmp_path)
    ok, detail = helpers.run_case(helpers.PRODUCT_CLI, case, wd, env=shim_env)
    assert ok, detail


@pytest.mark.parametrize("seed", range(40))
def test_host_cli_matches_reference_binary(seed, shim_env, tmp_path):
    if not os.path.exists(helpers.REF_CLI):
        pytest.skip("oracle/_ref/fxtract not built")
    rng = random.Random(1000 + seed)
    wd = str(tmp_path)
    for argv, files in helpers.random_cli_cases(rng, wd, trials=12):
        r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        for f in files:
            helpers.stash(wd, f, ".ref")
        r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                            env=shim_env, timeout=120)
        assert r2.returncode == r1.returncode, (argv, r2.stderr[-300:])
        assert r1.stdout == r2.stdout, (argv, r1.stdout[:400], r2.stdout[:400], r2.stderr[-300:])
        for f in files:
            assert helpers.read_or_none(wd, f) == helpers.read_or_none(wd, f + ".ref"), (argv, f)


def test_cumulative_count_and_multi_input(shim_env, tmp_path):
    wd = helpers.stage_golden(tmp_path)
    for argv in (["-ScH", "HISEQ", "1.fa", "11.fa", "13.fa"], ["-Sc", "A", "1.fa", "empty.fa", "5.fa"],
                 ["-c", "A", "4_1.fa", "4_2.fa", "1.fa"], ["-V"], ["-SlH", "HISEQ", "11.fa", "1.fa"],
                 ["-S", "A", "does_not_exist.fa"], ["-I", "A", "1.fa", "5.fa"]):
        r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=shim_env)
        assert (r1.returncode, r1.stdout) == (r2.returncode, r2.stdout), (argv, r1.stdout, r2.stdout, r2.stderr)


def test_large_streaming_inputs(shim_env, tmp_path):
    """Batches, window refills and oversized records: multi-MB inputs, tiny staging buffers."""
    rng = random.Random(77)
    wd = str(tmp_path)
    big = b">chr1 big\n" + b"\n".join(helpers.rand_dna(rng, 70).encode() for _ in range(60000)) + b"\n"
    fq = helpers.make_fastq(rng, 30000)
    fa = helpers.make_fasta(rng, 20000)
    open(os.path.join(wd, "big.fa"), "wb").write(big + fa)
    open(os.path.join(wd, "r.fq"), "wb").write(fq)
    env = dict(shim_env)
    env["FXTRACT_BATCH_MIB"] = "1"
    pat = helpers.pick_substring(rng, fq, 8, 12).decode()
    for argv in (["-S", pat, "r.fq", "big.fa"], ["-Sv", pat, "r.fq"], ["-Sc", "ACGTAC", "big.fa", "r.fq"],
                 ["-SH", "r12", "r.fq"], ["-SQc", "II", "r.fq"], ["-I", pat, "r.fq"]):
        r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env)
        assert r1.stdout == r2.stdout, (argv, len(r1.stdout), len(r2.stdout), r2.stderr[-300:])


def test_raw_ingest_windows_and_carry(shim_env, tmp_path):
    """Threaded window ingest of the raw path (strategies.cpp, RawIngest): windows smaller than the file, partial
    records carried from window to window, several reader threads, two contexts dealt round-robin -- output must be
    the reference's.  The shim walks the batches on the CPU (FXSHIM_RAW=1); the kernels are not under test here."""
    if not os.path.exists(helpers.REF_CLI):
        pytest.skip("oracle/_ref/fxtract not built")
    rng = random.Random(5)
    wd = str(tmp_path)
    fq = helpers.make_fastq(rng, 40000, seqlen=(30, 300))
    open(os.path.join(wd, "w.fq"), "wb").write(fq)
    pat = helpers.pick_substring(rng, fq, 5, 7).decode()
    for mib, readers, devs in (("1", "1", "0"), ("1", "5", "0,0,0"), ("2", "3", "0,0"), ("64", "4", "0")):
        env = dict(shim_env, FXSHIM_RAW="1", FXTRACT_PATH="raw", FXTRACT_BATCH_MIB=mib, FXTRACT_READERS=readers,
                   FXTRACT_DEVICES=devs, FXTRACT_STATS="1")
        for argv in (["-S", pat, "w.fq"], ["-Sv", pat, "w.fq"], ["-Sc", pat, "w.fq"], ["-Sc", pat, "w.fq", "w.fq"]):
            r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, timeout=120)
            assert r2.returncode == 0, (argv, r2.stderr[-300:])
            assert r1.stdout == r2.stdout, (mib, readers, devs, argv, len(r1.stdout), len(r2.stdout), r2.stderr[-300:])
            assert b"raw 0 B" not in r2.stderr and b"0 re-routed" in r2.stderr, r2.stderr


def test_raw_ingest_mates(shim_env, tmp_path):
    """Mates on the raw path (strategies.cpp, run_raw_pairs): two files cut at equal record counts from the reader
    threads' newline counts, interleaved input cut at whole pairs, carried tails, unequal files, an odd record, a
    truncated last record (re-routed to the fields path).  The shim applies the literal pair rule on the CPU
    (FXSHIM_RAW=1); the kernels are not under test here."""
    if not os.path.exists(helpers.REF_CLI):
        pytest.skip("oracle/_ref/fxtract not built")
    rng = random.Random(9)
    wd = str(tmp_path)
    a = helpers.make_fastq(rng, 30000, seqlen=(30, 300), name_prefix="a")
    b = helpers.make_fastq(rng, 30000, seqlen=(20, 120), name_prefix="b")
    recs_b = b.split(b"\n@b")
    short_b = b"\n@b".join(recs_b[:25000]) + b"\n"
    files = {"a.fq": a, "b.fq": b, "bs.fq": short_b, "odd.fq": b"\n@a".join(a.split(b"\n@a")[:20001]) + b"\n",
             "cut.fq": a[:len(a) - 7]}
    for name, data in files.items():
        open(os.path.join(wd, name), "wb").write(data)
    pat = helpers.pick_substring(rng, a, 5, 6).decode()
    cases_ = ([pat, "a.fq", "b.fq"], ["-v", pat, "a.fq", "b.fq"], ["-c", pat, "a.fq", "b.fq"], ["-cv", pat, "b.fq", "a.fq"],
              [pat, "a.fq", "bs.fq"], [pat, "bs.fq", "a.fq"], ["-I", pat, "a.fq"], ["-Iv", pat, "a.fq"], ["-Ic", pat, "odd.fq"],
              ["-I", pat, "odd.fq"], [pat, "cut.fq", "b.fq"], ["-I", pat, "cut.fq"], ["-c", pat, "a.fq", "b.fq", "b.fq", "a.fq"])
    for mib, readers, devs in (("1", "2", "0"), ("2", "6", "0,0,0"), ("64", "4", "0,0")):
        env = dict(shim_env, FXSHIM_RAW="1", FXTRACT_BATCH_MIB=mib, FXTRACT_READERS=readers, FXTRACT_DEVICES=devs, FXTRACT_STATS="1")
        for argv in cases_:
            r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, timeout=120)
            assert r2.returncode == 0, (argv, r2.stderr[-300:])
            assert r1.stdout == r2.stdout, (mib, readers, devs, argv, len(r1.stdout), len(r2.stdout), r2.stderr[-400:])
            if "cut.fq" not in argv or mib != "64":        # one window: the truncated record sends all of it to the fields path
                assert b"raw 0 B" not in r2.stderr, (argv, r2.stderr)
            if "cut.fq" not in argv and "bs.fq" not in argv:
                assert b" 0 re-routed" in r2.stderr, (argv, r2.stderr)
