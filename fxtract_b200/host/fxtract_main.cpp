// fxtract_main.cpp -- entry point of the drop-in `fxtract` CLI: pattern loading and the per-input
// loop of the reference's main() (main.cpp:654-770), every search strategy served by the GPU.
#include "fxhost.h"
#include <cstdlib>
#include <cstring>
#include <fstream>

using namespace fxh;

// main.cpp:222-256 (+ split :200-220): one pattern per line, empty lines skipped, TAB-separated
// optional output file, extra columns ignored, no CR stripping.  Streams in one pass (the pattern
// file may be a FIFO, run.sh:89).
static void load_pattern_file(Job &job, std::istream &in)
{
    std::string line;
    while (in.good()) {
        std::getline(in, line);
        if (line.empty()) continue;
        const size_t t1 = line.find('\t');
        const std::string pat = line.substr(0, t1);
        if (t1 == std::string::npos) {
            job.router.add(pat);
            if (job.opt.r) job.router.add(reverse_complement(pat));
        } else {
            const size_t t2 = line.find('\t', t1 + 1);
            const std::string file = line.substr(t1 + 1, t2 == std::string::npos ? std::string::npos : t2 - t1 - 1);
            job.router.add(pat, file);
            if (job.opt.r) job.router.add(reverse_complement(pat), file);
        }
    }
}

int main(int argc, char **argv)
{
    static char outbuf[1 << 20];
    setvbuf(stdout, outbuf, _IOFBF, sizeof outbuf);
    Job job;
    job.opt = parse_cli(argc, argv);
    const char *st = getenv("FXTRACT_STATS");
    job.stats = st && *st && strcmp(st, "0") != 0;
    int idx = job.opt.first_operand;
    if (job.opt.f == nullptr) {
        if (idx >= argc) {
            puts("Please provide a pattern (or pattern file) and at least one input file");
            fflush(stdout);
            print_usage_and_exit();
        } else if (idx >= argc - 1) {
            puts("Please provide an input file");
            fflush(stdout);
            print_usage_and_exit();
        }
        job.pattern = argv[idx++];
        job.router.add(job.pattern);
        if (job.opt.r) job.router.add(reverse_complement(job.pattern));
    } else {
        if (idx > argc - 1) {
            puts("Please provide an input file");
            fflush(stdout);
            print_usage_and_exit();
        }
        std::ifstream in(job.opt.f);
        if (!in.good()) {
            fprintf(stderr, "problem opening pattern file\n");
            return 1;
        }
        load_pattern_file(job, in);
    }
    const bool two_at_a_time = !job.opt.S && !job.opt.I;          // main.cpp:694-697, :763-767
    while (idx < argc) {
        const char *f1 = argv[idx];
        const char *f2 = two_at_a_time ? argv[idx + 1] : nullptr; // argv[argc] == NULL covers a lone trailing file
        int rc = run_input(job, f1, f2);
        if (rc != 0) { fflush(stdout); return rc; }
        if (job.opt.c) printf("%d\n", (int)job.matched_records);  // cumulative, main.cpp:759-761
        idx += two_at_a_time ? 2 : 1;
    }
    fflush(stdout);
    return 0;
}
