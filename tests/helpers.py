"""Shared helpers for the test-suite: golden-case runner and dirty-input generators."""
import gzip
import os
import random
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "fxtract_test")
ORACLE_CLI = os.path.join(ROOT, "oracle", "_build", "fxoracle")
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "fxtract")
PRODUCT_CLI = os.path.join(ROOT, "fxtract_b200", "_build", "fxtract")


def stage_golden(tmp):
    d = os.path.join(str(tmp), "fxtract_test")
    shutil.copytree(GOLDEN, d)
    return d


def run_case(binary, case, workdir, env=None):
    """Runs one cases.py entry with cwd=workdir; returns (ok, detail)."""
    cid, argv, kind, expected = case
    args = []
    for a in argv:
        if isinstance(a, tuple) and a[0] == "PATFILE":
            p = os.path.join(workdir, "_pat_%s.txt" % cid.replace("*", "s"))
            with open(p, "w") as f:
                f.write(a[1])
            args.append(p)
        elif isinstance(a, str) and a.startswith("GZ:"):
            src = os.path.join(workdir, a[3:])
            dst = src + ".gz"
            with open(src, "rb") as fi, gzip.open(dst, "wb") as fo:
                fo.write(fi.read())
            args.append(a[3:] + ".gz")
        else:
            args.append(a)
    if kind == "files":
        for out, _ in expected:
            try:
                os.remove(os.path.join(workdir, out))
            except FileNotFoundError:
                pass
    r = subprocess.run([binary] + args, cwd=workdir, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                       env=env, timeout=300)
    if kind == "stdout":
        want = open(os.path.join(workdir, expected), "rb").read()
        return r.stdout == want, (r.stdout[:300], want[:300], r.stderr[-500:])
    for out, exp in expected:
        got = open(os.path.join(workdir, out), "rb").read() if os.path.exists(os.path.join(workdir, out)) else None
        want = open(os.path.join(workdir, exp), "rb").read()
        if got != want:
            return False, (out, got, want, r.stderr[-500:])
    return True, None


# ------------------------------------------------------------------ generators
DNA = "ACGT"


def rand_dna(rng, n, alphabet=DNA):
    return "".join(rng.choice(alphabet) for _ in range(n))


def make_fastq(rng, nrec, dirty=False, seqlen=(20, 160), name_prefix="r"):
    """Canonical 4-line FASTQ, or (dirty) the variants kseq accepts: TAB separators,
    '+name' lines, CRLF, blank lines, wrapped sequence, lower-case/N, '@' first quality char."""
    out = []
    for i in range(nrec):
        n = rng.randint(*seqlen)
        alpha = DNA if not dirty or rng.random() < 0.7 else "ACGTNacgtn"
        seq = rand_dna(rng, n, alpha)
        qual = "".join(chr(rng.randint(33, 73)) for _ in range(n))
        if rng.random() < 0.2:
            qual = "@" + qual[1:]
        name = "%s%d:%d" % (name_prefix, i, rng.randint(0, 99999))
        comment = ""
        r = rng.random()
        if r < 0.6:
            comment = " %d:N:0:%s" % (rng.randint(1, 2), rand_dna(rng, 6))
        if dirty:
            r = rng.random()
            if r < 0.1:
                comment = "\tx y"
            elif r < 0.2:
                comment = "  two spaces"
            elif r < 0.25:
                comment = " "
        nl = "\r\n" if dirty and rng.random() < 0.05 else "\n"
        plus = "+" if not dirty or rng.random() < 0.8 else "+" + name
        if dirty and rng.random() < 0.15 and n > 10:
            k = rng.randint(1, n - 1)
            seqtxt = seq[:k] + nl + seq[k:]
            qualtxt = qual[:k] + nl + qual[k:]
        else:
            seqtxt, qualtxt = seq, qual
        out.append("@%s%s%s%s%s%s%s%s%s" % (name, comment, nl, seqtxt, nl, plus, nl, qualtxt, nl))
        if dirty and rng.random() < 0.05:
            out.append(nl)
    s = "".join(out)
    if dirty and rng.random() < 0.3 and s.endswith("\n"):
        s = s[:-1]
    return s.encode()


def make_fasta(rng, nrec, wrap=True, seqlen=(20, 200), name_prefix="s"):
    out = []
    for i in range(nrec):
        n = rng.randint(*seqlen)
        seq = rand_dna(rng, n, DNA if rng.random() < 0.8 else "ACGTNRYacgt")
        name = "%s%d_%d" % (name_prefix, i, rng.randint(0, 9999))
        comment = ""
        if rng.random() < 0.5:
            comment = " " + rng.choice(["Candidatus Accumulibacter", "sample=%d" % i, "x", "len=%d" % n])
        if wrap and rng.random() < 0.5:
            w = rng.randint(10, 70)
            body = "\n".join(seq[k:k + w] for k in range(0, n, w))
        else:
            body = seq
        out.append(">%s%s\n%s\n" % (name, comment, body))
        if rng.random() < 0.05:
            out.append("\n")
    return "".join(out).encode()


def pick_substring(rng, data, lo=4, hi=24):
    """A substring of some sequence-looking line of `data` (so the pattern hits)."""
    lines = [l for l in data.split(b"\n") if l and l[:1] not in b">@+" and len(l) > lo]
    if not lines:
        return b"ACGT"
    l = rng.choice(lines)
    n = rng.randint(lo, min(hi, len(l)))
    s = rng.randint(0, len(l) - n)
    return l[s:s + n].strip(b"\r")


# ------------------------------------------------------------------ randomised CLI cases
FLAGSETS = ["", "-v", "-c", "-vc", "-H", "-Hv", "-C", "-Q", "-r", "-rH", "-X", "-HX", "-HXv", "-rX",
            "-f", "-Hf", "-vf", "-cf", "-rf", "-HXf", "-Cf", "-HXcf", "-l", "-lf", "-lX", "-I", "-Iv", "-IHf", "-IX",
            "-Qf", "-vHXf", "-lv"]


def stash(wd, name, suffix):
    p = os.path.join(wd, name)
    if os.path.exists(p):
        os.replace(p, p + suffix)


def read_or_none(wd, name):
    p = os.path.join(wd, name)
    return open(p, "rb").read() if os.path.exists(p) else None


def random_cli_cases(rng, wd, trials=12):
    """Yields (argv, [output files]) over freshly written random inputs a.fx / b.fx in wd.
    Covers the three literal strategies, all targets, -v/-c/-l/-r, pairs, -I, -S, gz input and
    pattern files with per-pattern output files."""
    kind = rng.choice(["fq", "fq_dirty", "fa"])
    if kind == "fq":
        d1 = make_fastq(rng, rng.randint(1, 40)); d2 = make_fastq(rng, rng.randint(1, 40))
    elif kind == "fq_dirty":
        d1 = make_fastq(rng, rng.randint(1, 40), dirty=True); d2 = make_fastq(rng, rng.randint(1, 40), dirty=True)
    else:
        d1 = make_fasta(rng, rng.randint(1, 40)); d2 = make_fasta(rng, rng.randint(1, 40))
    open(os.path.join(wd, "a.fx"), "wb").write(d1)
    open(os.path.join(wd, "b.fx"), "wb").write(d2)
    with gzip.open(os.path.join(wd, "a.fx.gz"), "wb") as f:
        f.write(d1)
    for trial in range(trials):
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
                sub = pick_substring(rng, d1)
                if "X" in flags:
                    lines = [l for l in d1.split(b"\n") if l and l[:1] not in b">@+"]
                    sub = rng.choice(lines) if lines else sub
                pats.append(sub)
        pats = [p for p in pats if p and b"\n" not in p and b"\t" not in p and b"\r" not in p]
        if not pats:
            continue
        argv, outs = [], []
        if flags:
            argv.append(flags if "f" not in flags else flags.replace("f", "") + "f")
        if "f" in flags:
            pf = os.path.join(wd, "p%d.txt" % trial)
            with_files = rng.random() < 0.4
            lines = []
            for k, p in enumerate(pats):
                if with_files and rng.random() < 0.7:
                    out = "out_%d_%d.fx" % (trial, k % 2)
                    outs.append(out)
                    lines.append(p + b"\t" + out.encode())
                else:
                    lines.append(p)
            open(pf, "wb").write(b"\n".join(lines) + b"\n")
            argv.append(pf)
        else:
            argv.append(pats[0].decode("latin1"))
        single = ("I" in flags) or rng.random() < 0.5
        if single and "I" not in flags:
            argv.insert(0, "-S")
        first = "a.fx.gz" if rng.random() < 0.2 else "a.fx"
        argv += [first] if single else [first, "b.fx"]
        argv = [a for a in argv if a != "-"]
        yield argv, sorted(set(outs))
