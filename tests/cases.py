"""Command lines of the reference's acceptance suite (fxtract_test/run.sh:36-155),
restated as data.  Each case: (id, argv after the program name, kind, expected).

kind "stdout": stdout must equal the golden file `expected`.
kind "files" : the two per-pattern output files must equal the golden pair.
Process substitutions `<(echo X)` (run.sh:89,148,154) become a FIFO-less temp
pattern file holding "X\n" (marker ("PATFILE", text)).
Regex cases (-G/-E/-P) are outside the graft (BASELINE.json north_star) and are
listed separately so the host CLI's POSIX paths can still be checked.
"""

H1 = "HISEQ2000:55:C0JRTACXX:2:1101:11128:12710_1:N:0:CTTGTAAT"
SEQ150 = ("TCGAACGTCGCAAAGACTCGCACCCTCGCTGCGAACGACACGTCTCAATCCCTTTGAATTCAGGGCATCAGTTCGAACTGGAGCAGTACGAC"
          "CACGTTGATCTGAAGTGGCGTCTCAATCCCTTTGAATTCAGGGCATCAGTTCGAACGG")


def PAT(text):
    return ("PATFILE", text)


LITERAL_CASES = [
    ("1a", ["CAAAGGGATTGAGACGCCACTT", "1.fa"], "stdout", "1.expected.fa"),
    ("2a", ["-H", H1, "1.fa"], "stdout", "2.expected.fa"),
    ("2e", ["-HX", H1, "1.fa"], "stdout", "2.expected.fa"),
    ("2f", ["-HX", "1660334", "12.fa"], "stdout", "12.expected.fa"),
    ("2g", ["-rHX", "1660334", "12.fa"], "stdout", "2g.expected.fa"),
    ("3", ["-Hf", "headers.txt", "1.fa"], "stdout", "3.expected.fa"),
    ("4", ["-H", "HSQ868392H08B7ADXX:2:1112:8977:35114", "4_1.fa", "4_2.fa"], "stdout", "4.expected.fa"),
    ("5", ["-C", "Accumulibacter", "5.fa"], "stdout", "5.expected.fa"),
    ("6a", ["-Hv", "647449011", "11.fa"], "stdout", "11.expected.fa"),
    ("6b", ["-HvX", "647449011", "11.fa"], "stdout", "11.expected.fa"),
    ("6f", ["-Hvf", PAT("647449011\n"), "11.fa"], "stdout", "11.expected.fa"),
    ("7", ["-Hc", "HISEQ2000", "1.fa"], "stdout", "7.expected.fa"),
    ("8", ["-H", "1101:11128:12710", "8.fa"], "stdout", "8.expected.fa"),
    ("9", ["-H", "1101:11128:12710", "9.fa"], "stdout", "8.expected.fa"),
    ("10a", ["-zH", "1101:11128:12710", "GZ:8.fa"], "stdout", "8.expected.fa"),
    ("10c", ["-Hz", "HSQ868392H08B7ADXX:2:1112:8977:35114", "GZ:4_1.fa", "GZ:4_2.fa"], "stdout", "4.expected.fa"),
    ("11", ["-CS", "Accumulibacter", "1.fa", "11.fa", "5.fa"], "stdout", "5.expected.fa"),
    ("12a", ["-HXf", "headers2.txt", "1.fa"], "files",
     [("14_out_1.fasta", "14_1.expected"), ("14_out_2.fasta", "14_2.expected")]),
    ("12b", ["-Hf", "headers2.txt", "1.fa"], "files",
     [("14_out_1.fasta", "14_1.expected"), ("14_out_2.fasta", "14_2.expected")]),
    ("13e", ["-HX", "HWI-ST1243:175:C29BRACXX:4:1101:15034:30425", "15.fq"], "stdout", "15.expected.fq"),
    ("14a", ["-r", "AAGTGGCGTCTCAATCCCTTTG", "1.fa"], "stdout", "1.expected.fa"),
    ("14b", ["-rX", SEQ150, "1.fa"], "stdout", "13b.expected.fa"),
    ("14c", ["-rf", PAT("AAGTGGCGTCTCAATCCCTTTG\n"), "1.fa"], "stdout", "1.expected.fa"),
    ("14d", ["-X", SEQ150, "1.fa"], "stdout", "empty.fa"),
    ("14e", ["-f", PAT("AAGTGGCGTCTCAATCCCTTTG\n"), "1.fa"], "stdout", "empty.fa"),
    # run.sh "13a" diffs a stale file; the command itself is literal search on 13.fa
    ("13a*", ["CAAAGGGATTGAGACGCCACTT", "13.fa"], "stdout", "13.expected.fa"),
]

REGEX_CASES = [
    ("1b", ["-G", "CAAAGGGATTGAGACGCCACTT", "1.fa"], "stdout", "1.expected.fa"),
    ("1c", ["-E", "CAAAGGGATTGAGACGCCACTT", "1.fa"], "stdout", "1.expected.fa"),
    ("2b", ["-HG", H1, "1.fa"], "stdout", "2.expected.fa"),
    ("2c", ["-HE", H1, "1.fa"], "stdout", "2.expected.fa"),
    ("6d", ["-HvG", "647449011", "11.fa"], "stdout", "11.expected.fa"),
    ("6e", ["-HvE", "647449011", "11.fa"], "stdout", "11.expected.fa"),
    ("13b", ["-G", "CAAAGGGATTGAGACGCCACTT", "13.fa"], "stdout", "13.expected.fa"),
    ("13c", ["-E", "CAAAGGGATTGAGACGCCACTT", "13.fa"], "stdout", "13.expected.fa"),
]
