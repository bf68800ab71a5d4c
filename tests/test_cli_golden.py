"""GPU tests of the drop-in `fxtract` CLI (host C++ + libfxgpu.so): the reference's own
acceptance suite (fxtract_test/run.sh restated in cases.py) plus randomised and synthetic
workloads, byte-compared with the reference binary built from the unmodified sources."""
import ctypes as C
import hashlib
import os
import random
import subprocess

import pytest

import cases
import helpers

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cli(built):
    assert os.path.exists(helpers.PRODUCT_CLI), "host CLI not built"
    env = dict(os.environ)
    env.pop("LD_LIBRARY_PATH", None)          # the real library, through the binary's rpath
    return env


ALL = cases.LITERAL_CASES + cases.REGEX_CASES


@pytest.mark.parametrize("case", ALL, ids=[c[0] for c in ALL])
def test_cli_golden(case, cli, tmp_path):
    wd = helpers.stage_golden(tmp_path)
    ok, detail = helpers.run_case(helpers.PRODUCT_CLI, case, wd, env=cli)
    assert ok, detail


@pytest.mark.parametrize("seed", range(12))
def test_cli_matches_reference_binary(seed, cli, tmp_path):
    if not os.path.exists(helpers.REF_CLI):
        pytest.skip("oracle/_ref/fxtract not built")
    rng = random.Random(5000 + seed)
    wd = str(tmp_path)
    for argv, files in helpers.random_cli_cases(rng, wd, trials=8):
        r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        for f in files:
            helpers.stash(wd, f, ".ref")
        r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                            env=cli, timeout=300)
        assert r2.returncode == r1.returncode, (argv, r2.stderr[-300:])
        assert r1.stdout == r2.stdout, (argv, r1.stdout[:400], r2.stdout[:400], r2.stderr[-300:])
        for f in files:
            assert helpers.read_or_none(wd, f) == helpers.read_or_none(wd, f + ".ref"), (argv, f)


def _synth_file(path, n, **kw):
    import fxtract_b200 as fx
    from fxtract_b200 import lib
    spec = fx.SynthSpec(seed=kw.get("seed", 11), first_record=kw.get("first", 0), n_records=n, read_len=150,
                        plant_mode=kw.get("plant", 0), pattern_seed=2, n_patterns=kw.get("npat", 0), pattern_len=20,
                        plant_per_million=kw.get("ppm", 0), plant_revcomp=kw.get("rc", 0))
    data = lib.synth_host(spec)
    open(path, "wb").write(data)
    return spec, data


def _both(argv, wd, env):
    r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    r2 = subprocess.run([helpers.PRODUCT_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env,
                        timeout=600)
    return r1, r2


def test_cli_raw_path_synthetic_all_strategies(cli, tmp_path):
    """Canonical FASTQ takes the raw GPU path (checked through FXTRACT_PATH=raw, which forbids the
    fields fallback); small staging buffers force many batches and carried partial records; two
    contexts on the device exercise the round-robin ordering."""
    from fxtract_b200 import lib
    wd = str(tmp_path)
    spec, data = _synth_file(os.path.join(wd, "s.fq"), 60000, plant=1, npat=200, ppm=20000, rc=1)
    recs = data.split(b"\n")
    pat = recs[4 * 1234 + 1][37:57].decode()
    pats = sorted(set(lib.synth_pattern(spec, p) for p in range(200)))
    open(os.path.join(wd, "p.txt"), "wb").write(b"\n".join(pats) + b"\n")
    names = [recs[4 * i][1:].split(b" ")[0] for i in range(0, 60000, 37)] + [b"absent:1"]
    open(os.path.join(wd, "ids.txt"), "wb").write(b"\n".join(names) + b"\n")
    env = dict(cli)
    env.update(FXTRACT_PATH="raw", FXTRACT_BATCH_MIB="1", FXTRACT_DEVICES="0,0")
    for argv in (["-S", pat, "s.fq"], ["-Sr", pat, "s.fq"], ["-Sc", pat, "s.fq"], ["-Sv", pat, "s.fq"],
                 ["-Sf", "p.txt", "s.fq"], ["-Srf", "p.txt", "s.fq"], ["-Scf", "p.txt", "s.fq"], ["-Svcf", "p.txt", "s.fq"],
                 ["-SHXf", "ids.txt", "s.fq"], ["-SHXvcf", "ids.txt", "s.fq"], ["-SH", "FC:3:", "s.fq"],
                 ["-SQc", "II", "s.fq"], ["-SCc", "1:N", "s.fq"], [pat, "s.fq"],
                 # -l on the raw path: the first batch with a match prints the file name and ends the input
                 ["-Sl", pat, "s.fq"], ["-Slf", "p.txt", "s.fq"], ["-SlHXf", "ids.txt", "s.fq"], ["-Slvf", "p.txt", "s.fq"],
                 ["-Slc", pat, "s.fq"]):
        r1, r2 = _both(argv, wd, env)
        assert r2.returncode == 0, (argv, r2.stderr[-400:])
        assert hashlib.sha256(r1.stdout).hexdigest() == hashlib.sha256(r2.stdout).hexdigest(), (argv, len(r1.stdout), len(r2.stdout))
        assert len(r1.stdout) > 0


def test_cli_raw_path_per_pattern_output_files(cli, tmp_path):
    """Pattern files with a second column (main.cpp:222-256) on canonical FASTQ: the raw path returns every selected
    record as its own run with the pattern that selected it (exact: the key; multi: the first reported hit = smallest
    end offset, then longest pattern) and the host routes it like FileManager::find (fileManager.cpp:171-202).
    stdout and every output file must equal the reference binary's; FXTRACT_PATH=raw forbids the fields fallback."""
    rng = random.Random(77)
    wd = str(tmp_path)
    pats = sorted(set(helpers.rand_dna(rng, rng.choice([20, 20, 24, 31])).encode() for _ in range(60)))
    # nested / overlapping patterns: the first-hit rule decides which file a record goes to
    pats += [pats[0][2:] + b"ACGTACGT", pats[1][:18] + b"GG", pats[2] + b"T"]
    pats = sorted(set(pats))
    recs = []
    for i in range(30000):
        n = rng.randint(60, 170)
        seq = bytearray(helpers.rand_dna(rng, n).encode())
        if i % 4 == 0:
            for _ in range(rng.choice([1, 1, 2, 3])):
                q = rng.choice(pats)
                if len(q) <= n:
                    o = rng.randint(0, n - len(q))
                    seq[o:o + len(q)] = q
        recs.append(b"@r%d c%d\n" % (i, i % 7) + bytes(seq) + b"\n+\n" + bytes(rng.randint(33, 73) for _ in range(n)) + b"\n")
    open(os.path.join(wd, "in.fq"), "wb").write(b"".join(recs))
    files = ["o_a.fq", "o_b.fq", "o_c.fq"]
    with open(os.path.join(wd, "pf.txt"), "wb") as f:
        for k, q in enumerate(pats):
            f.write(q + (b"\t" + files[k % 4].encode() if k % 4 < 3 else b"") + b"\n")       # every 4th pattern: stdout
    ids = [b"r%d" % i for i in range(0, 30000, 11)] + [b"nope"]
    with open(os.path.join(wd, "ids.txt"), "wb") as f:
        for k, q in enumerate(ids):
            f.write(q + (b"\t" + files[k % 3].encode() if k % 5 else b"") + b"\n")
    env = dict(cli)
    env.update(FXTRACT_PATH="raw", FXTRACT_BATCH_MIB="1", FXTRACT_STATS="1")
    for argv in (["-Sf", "pf.txt", "in.fq"], ["-Srf", "pf.txt", "in.fq"], ["-SHXf", "ids.txt", "in.fq"],
                 ["-Svf", "pf.txt", "in.fq"], ["-Scf", "pf.txt", "in.fq"]):
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
        assert b"fields 0 B" in r2.stderr, r2.stderr[-300:]                  # everything went through the raw path
        assert r1.stdout == r2.stdout, (argv, len(r1.stdout), len(r2.stdout))
        got = {f: helpers.read_or_none(wd, f) for f in files}
        assert got == want, (argv, {f: (len(want[f] or b""), len(got[f] or b"")) for f in files})
        if argv[0] in ("-Sf", "-SHXf"):
            assert all(want[f] for f in files) and len(r1.stdout) > 0


def test_cli_reroutes_non_canonical_batches(cli, tmp_path):
    """A defect in the middle of an otherwise canonical file: the batches before it go through the
    raw path, the rest through the host reader; the output is still byte-identical."""
    rng = random.Random(3)
    wd = str(tmp_path)
    _, clean = _synth_file(os.path.join(wd, "c.fq"), 20000)
    dirty = helpers.make_fastq(rng, 300, dirty=True)
    if not dirty.endswith(b"\n"):
        dirty += b"\n"
    open(os.path.join(wd, "m.fq"), "wb").write(clean + dirty + clean[:100000].rsplit(b"\n@SYN", 1)[0] + b"\n")
    env = dict(cli)
    env.update(FXTRACT_BATCH_MIB="1", FXTRACT_STATS="1")
    for argv in (["-Sc", "ACGTAC", "m.fq"], ["-Sv", "ACGTACGT", "m.fq"], ["-SHc", "r1", "m.fq"]):
        r1, r2 = _both(argv, wd, env)
        assert r1.stdout == r2.stdout, (argv, len(r1.stdout), len(r2.stdout), r2.stderr[-300:])
        assert b"1 re-routed" in r2.stderr and b"raw 0 B" not in r2.stderr, r2.stderr
