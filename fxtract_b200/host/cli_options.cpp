// cli_options.cpp -- the reference's command-line surface (main.cpp:83-198): same getopt
// string, same flag meanings, same messages, so argv stays a drop-in.  Device selection and
// batch sizing never appear in argv; they come from environment variables.
#include "fxhost.h"
#include <cstdlib>
#include <cstring>
#include <unistd.h>

#ifndef FXTRACT_VERSION
#define FXTRACT_VERSION "2.3"       // PACKAGE_VERSION of the reference Makefile (Makefile:15)
#endif

namespace fxh {

static const char kSynopsis[] =
    "[options] {-f pattern_file | pattern} {<read1.fx> <read2.fx>}...\n"
    "       fxtract [options] -I {-f pattern_file | pattern} <read12.fx>...\n"
    "       fxtract [options] -S {-f pattern_file | pattern} <read.fx>...\n";

// one row per documented switch, in the reference's order (main.cpp:87-107; -P is listed only by
// builds that have PCRE, which this one -- like a NO_PCRE=1 reference build -- has not)
static const char *const kSwitchHelp[][2] = {
    {"-H", "Evaluate patterns in the context of headers (default: sequences)"},
    {"-Q", "Evaluate patterns in the context of quality scores (default: sequences)"},
    {"-C", "Evaluate patters in the context of comment strings - everything after\n"
           "\t             the first space on the header line of the record (default: sequences)"},
    {"-G", "pattern is a posix basic regular expression (default: literal substring)"},
    {"-E", "pattern is a posix extended regular expression (default: literal substring)"},
    {"-X", "pattern exactly matches the whole string (default: literal substring)"},
    {"-r", "Match the reverse complement of a literal pattern. Not compatible with regular expressions"},
    {"-I", "The read file is interleaved (both pairs in a single file)"},
    {"-S", "The files do not contain pairs. Allows for multiple files to be given on the command line"},
    {"-v", "Inverse the match criteria. Print pairs that do not contain matches"},
    {"-c", "Print only the count of reads (or pairs) that were found"},
    {"-f <file>", "File containing patterns, one per line"},
    {"-l", "Print file names that contain matches"},
    {"-h", "Print this help"},
    {"-V", "Print version"},
};

void print_usage_and_exit()
{
    // util's usage(msg): "Usage: <prog> <msg>" on stderr, non-zero exit (stream and code are not
    // pinned by any reference test; this matches the stand-in the oracle binary is built with)
    fprintf(stderr, "Usage: fxtract %s", kSynopsis);
    for (const auto &row : kSwitchHelp) fprintf(stderr, "\t%-12s %s\n", row[0], row[1]);
    fputc('\n', stderr);
    exit(2);
}

CliOptions parse_cli(int argc, char **argv)
{
    CliOptions o;
    int ch;
    while ((ch = getopt(argc, argv, "HhCcIf:zjqVvrQGEPX12SlL")) != -1) {
        switch (ch) {
        case 'H': o.H = true; break;
        case 'Q': o.Q = true; break;
        case 'C': o.C = true; break;
        case 'G': o.G = true; break;
        case 'E': o.E = true; break;
        case 'X': o.X = true; break;
        case 'r': o.r = true; break;
        case 'I': o.I = true; break;
        case 'S': o.S = true; break;
        case 'v': o.v = true; break;
        case 'c': o.c = true; break;
        case 'l': o.l = true; break;
        case 'L': o.L = true; break;
        case '1': o.one = true; break;
        case '2': o.two = true; break;
        case 'f': o.f = optarg; break;
        case 'z': break;                                    // accepted, no effect (main.cpp:174)
        case 'V': puts(FXTRACT_VERSION); exit(1);           // main.cpp:143-146: exits 1
        case 'j':
            fprintf(stderr, "\n\nERROR: The -j option has been removed, unfortunately parsing bzip2 files is no longer supported\n\n");
            print_usage_and_exit();
        default:                                            // -h, -q, -P (no PCRE in this build), unknown
            print_usage_and_exit();
        }
    }
    if ((o.G || o.E) && o.r) {
        fprintf(stderr,
                "\n\nERROR: Searching for the reverse complement of a regular\n"
                "expression does not work. If your pattern is a litteral string just\n"
                "dropG or E flags. If your pattern is a regular expression\n"
                "you will have to manually input it as either a separate command or\n"
                "using alternation in the regular expression\n\n");
        print_usage_and_exit();
    }
    o.first_operand = optind;
    return o;
}

// util.h:11-20 as rules: letters -> IUPAC complement (U -> A), case kept; '`' -> '@' like the
// reference table's entry 96; everything else unchanged.  Bytes >= 0x80 index the reference's
// table out of bounds (UB); they are left unchanged here.
static inline unsigned char complement_byte(unsigned char c)
{
    static const char upper[] = "TVGHEFCDIJMLKNOPQYSAABWXRZ";
    if (c >= 'A' && c <= 'Z') return (unsigned char)upper[c - 'A'];
    if (c >= 'a' && c <= 'z') return (unsigned char)(upper[c - 'a'] + ('a' - 'A'));
    if (c == 0x60) return 0x40;
    return c;
}

std::string reverse_complement(const std::string &s)
{
    std::string r(s.size(), '\0');
    const size_t n = s.size();
    for (size_t i = 0; i < n; ++i) r[n - 1 - i] = (char)complement_byte((unsigned char)s[i]);
    return r;
}

std::vector<int> devices_from_env()
{
    std::vector<int> d;
    const char *e = getenv("FXTRACT_DEVICES");
    if (e && *e) {
        const char *p = e;
        while (*p) {
            char *end = nullptr;
            long v = strtol(p, &end, 10);
            if (end == p) break;
            d.push_back((int)v);
            p = (*end == ',') ? end + 1 : end;
            if (*end && *end != ',') break;
        }
    }
    if (d.empty()) d.push_back(0);
    return d;
}

size_t batch_bytes_from_env()
{
    const char *e = getenv("FXTRACT_BATCH_MIB");
    long mib = e ? atol(e) : 64;
    if (mib < 1) mib = 1;
    if (mib > 2048) mib = 2048;
    return (size_t)mib << 20;
}

}  // namespace fxh
