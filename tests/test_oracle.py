"""CPU tests: pins the plain-C restatement (oracle/fx_oracle.c) against
(1) every literal-path golden vector of the reference's own suite and
(2) the reference binary built from the unmodified sources (oracle/_ref/fxtract),
    on randomised clean/dirty FASTA/FASTQ inputs and flag combinations."""
import os
import random
import subprocess

import pytest

import cases
import helpers


@pytest.fixture(scope="module")
def oracle_cli(built):
    assert os.path.exists(helpers.ORACLE_CLI)
    return helpers.ORACLE_CLI


@pytest.mark.parametrize("case", cases.LITERAL_CASES, ids=[c[0] for c in cases.LITERAL_CASES])
def test_oracle_port_golden(case, oracle_cli, tmp_path):
    wd = helpers.stage_golden(tmp_path)
    ok, detail = helpers.run_case(oracle_cli, case, wd)
    assert ok, detail


@pytest.mark.parametrize("case", cases.LITERAL_CASES + cases.REGEX_CASES,
                         ids=[c[0] for c in cases.LITERAL_CASES + cases.REGEX_CASES])
def test_reference_binary_golden(case, built, tmp_path):
    if not os.path.exists(helpers.REF_CLI):
        pytest.skip("oracle/_ref/fxtract not built (no /root/reference here)")
    wd = helpers.stage_golden(tmp_path)
    ok, detail = helpers.run_case(helpers.REF_CLI, case, wd)
    assert ok, detail


FLAGSETS = ["", "-v", "-c", "-vc", "-H", "-Hv", "-C", "-Q", "-r", "-rH", "-X", "-HX", "-HXv", "-rX",
            "-f", "-Hf", "-vf", "-cf", "-rf", "-HXf", "-Cf", "-HXcf", "-l", "-lf", "-I", "-Iv", "-IHf"]


@pytest.mark.parametrize("seed", range(40))
def test_port_matches_reference_binary(seed, oracle_cli, tmp_path):
    if not os.path.exists(helpers.REF_CLI):
        pytest.skip("oracle/_ref/fxtract not built")
    rng = random.Random(seed)
    wd = str(tmp_path)
    kind = rng.choice(["fq", "fq_dirty", "fa"])
    if kind == "fq":
        d1 = helpers.make_fastq(rng, rng.randint(1, 40)); d2 = helpers.make_fastq(rng, rng.randint(1, 40))
    elif kind == "fq_dirty":
        d1 = helpers.make_fastq(rng, rng.randint(1, 40), dirty=True); d2 = helpers.make_fastq(rng, rng.randint(1, 40), dirty=True)
    else:
        d1 = helpers.make_fasta(rng, rng.randint(1, 40)); d2 = helpers.make_fasta(rng, rng.randint(1, 40))
    open(os.path.join(wd, "a.fx"), "wb").write(d1)
    open(os.path.join(wd, "b.fx"), "wb").write(d2)
    for trial in range(12):
        flags = rng.choice(FLAGSETS)
        target = "H" if "H" in flags else ("C" if "C" in flags else ("Q" if "Q" in flags else "S"))
        if "Q" in flags and "f" in flags and kind == "fa":
            continue  # reference UB (main.cpp:394-398)
        pats = []
        for _ in range(rng.randint(1, 4)):
            if target == "H":
                names = [l[1:].split()[0] for l in d1.split(b"\n") if l[:1] in b">@" and len(l) > 1 and l[1:].split()]
                nm = rng.choice(names) if names else b"zz"
                pats.append(nm if ("X" in flags or rng.random() < 0.5) else nm[: max(1, len(nm) // 2)])
            elif target == "C":
                pats.append(rng.choice([b"N:0", b"Accumuli", b"x", b"sample=", b"two"]))
            elif target == "Q":
                pats.append(bytes([rng.randint(40, 70)]) * rng.randint(1, 2))
            else:
                sub = helpers.pick_substring(rng, d1)
                if "X" in flags:
                    lines = [l for l in d1.split(b"\n") if l and l[:1] not in b">@+"]
                    sub = rng.choice(lines) if lines else sub
                pats.append(sub)
        pats = [p for p in pats if p and b"\n" not in p and b"\t" not in p and b"\r" not in p]
        if not pats:
            continue
        argv = []
        if flags:
            argv.append(flags if "f" not in flags else flags.replace("f", "") + "f")
        if "f" in flags:
            pf = os.path.join(wd, "p%d.txt" % trial)
            open(pf, "wb").write(b"\n".join(pats) + b"\n")
            argv.append(pf)
        else:
            argv.append(pats[0].decode("latin1"))
        single = ("I" in flags) or rng.random() < 0.5
        if single and "I" not in flags:
            argv.insert(0, "-S")
        argv += ["a.fx"] if single else ["a.fx", "b.fx"]
        argv = [a for a in argv if a != "-"]
        r1 = subprocess.run([helpers.REF_CLI] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        r2 = subprocess.run([oracle_cli] + argv, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert r1.stdout == r2.stdout, (argv, kind, r1.stdout[:400], r2.stdout[:400])
