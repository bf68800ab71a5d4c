// fxsynth -- stand-alone generator of the synthetic benchmark workload (no CUDA, no libfxgpu): the same bytes as
// fxg_synth_host / the device generator, from fxg_synth_core.h.  bench.py's reference arm and the verification
// legs use it to write FASTQ shards, pattern files and read-id lists.
//
//   fxsynth fastq    <spec> <out>                 synthetic FASTQ of the spec's record range
//   fxsynth patterns <spec> <out>                 the spec's n_patterns planted patterns, one per line
//   fxsynth names    <spec> <stride> <count> <out>  read names of records first, first+stride, ... (one per line)
//   fxsynth bytes    <spec>                       prints the byte size of the range
// <spec> = seed:first_record:n_records:read_len:plant_mode:pattern_seed:n_patterns:pattern_len:plant_per_million:plant_revcomp
#include "../csrc/fxg_synth_core.h"
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

using namespace fxg;

static bool parse_spec(const char *s, fxg_synth_spec *o)
{
    unsigned long long v[10];
    if (sscanf(s, "%llu:%llu:%llu:%llu:%llu:%llu:%llu:%llu:%llu:%llu", &v[0], &v[1], &v[2], &v[3], &v[4], &v[5], &v[6], &v[7], &v[8], &v[9]) != 10)
        return false;
    o->seed = v[0]; o->first_record = v[1]; o->n_records = v[2]; o->read_len = (uint32_t)v[3]; o->plant_mode = (uint32_t)v[4];
    o->pattern_seed = v[5]; o->n_patterns = (uint32_t)v[6]; o->pattern_len = (uint32_t)v[7]; o->plant_per_million = (uint32_t)v[8];
    o->plant_revcomp = (uint32_t)v[9];
    return true;
}

static int usage()
{
    fprintf(stderr, "usage: fxsynth fastq|patterns|names|bytes <spec> [...]  (see fxsynth.cpp)\n");
    return 2;
}

int main(int argc, char **argv)
{
    if (argc < 3) return usage();
    const std::string cmd = argv[1];
    fxg_synth_spec s;
    if (!parse_spec(argv[2], &s)) return usage();
    if (cmd == "bytes") { printf("%llu\n", (unsigned long long)rec_offset(s, s.n_records)); return 0; }
    if (cmd == "fastq") {
        if (argc < 4) return usage();
        FILE *f = fopen(argv[3], "wb");
        if (!f) { perror(argv[3]); return 1; }
        // chunks of records generated on all host threads, written in order
        const unsigned long long chunk = 1ull << 18;
        unsigned nt = std::thread::hardware_concurrency();
        if (nt == 0) nt = 1;
        if (nt > 64) nt = 64;
        std::vector<uint8_t> buf;
        for (unsigned long long r0 = 0; r0 < s.n_records; r0 += chunk) {
            const unsigned long long r1 = r0 + chunk < s.n_records ? r0 + chunk : s.n_records;
            const unsigned long long b0 = rec_offset(s, r0), b1 = rec_offset(s, r1);
            buf.resize((size_t)(b1 - b0));
            auto work = [&](unsigned long long lo, unsigned long long hi) {
                for (unsigned long long i = lo; i < hi; ++i) {
                    const unsigned long long g = s.first_record + i;
                    synth_host_record(s, g, dec_digits(g), buf.data() + (rec_offset(s, i) - b0));
                }
            };
            std::vector<std::thread> th;
            const unsigned long long n = r1 - r0;
            for (unsigned t = 0; t < nt; ++t) th.emplace_back(work, r0 + n * t / nt, r0 + n * (t + 1) / nt);
            for (auto &x : th) x.join();
            if (fwrite(buf.data(), 1, buf.size(), f) != buf.size()) { perror("write"); return 1; }
        }
        fclose(f);
        return 0;
    }
    if (cmd == "patterns") {
        if (argc < 4) return usage();
        FILE *f = fopen(argv[3], "wb");
        if (!f) { perror(argv[3]); return 1; }
        std::string line(s.pattern_len + 1, '\n');
        for (uint32_t p = 0; p < s.n_patterns; ++p) {
            for (uint32_t k = 0; k < s.pattern_len; ++k) line[k] = (char)base_char(pattern_base(s, p, k));
            fwrite(line.data(), 1, line.size(), f);
        }
        fclose(f);
        return 0;
    }
    if (cmd == "names") {
        if (argc < 6) return usage();
        const unsigned long long stride = strtoull(argv[3], nullptr, 10), count = strtoull(argv[4], nullptr, 10);
        FILE *f = fopen(argv[5], "wb");
        if (!f) { perror(argv[5]); return 1; }
        char name[80];
        for (unsigned long long i = 0; i < count; ++i) {
            const unsigned long long g = s.first_record + i * stride;
            const int d = dec_digits(g);
            const unsigned int n = 22u + (unsigned int)d;
            for (unsigned int b = 0; b < n; ++b) name[b] = (char)rec_byte(s, g, d, b + 1);
            name[n] = '\n';
            fwrite(name, 1, n + 1, f);
        }
        fclose(f);
        return 0;
    }
    return usage();
}
