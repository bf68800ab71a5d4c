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
