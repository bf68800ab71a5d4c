// fxg_synth_core.h -- deterministic synthetic FASTQ: the counter-based generator shared
// by the CUDA kernel (fxg_synth.cu), the library's host generator and the stand-alone `fxsynth` tool that
// bench.py's reference arm uses (so that arm never loads libfxgpu.so).  Keyed by (seed, global record index,
// word): any shard of the stream can be regenerated anywhere.
//
// Record i (global index g):
//   @SYN:1:FC:<lane>:<tile>:<x>:<g> 1:N:0:ACGT\n     lane 1 digit, tile 4, x 5, g unpadded decimal
//   <read_len bases, uniform ACGT, optionally with a planted pattern>\n
//   +\n
//   <read_len quality bytes, uniform in [33,73], so '@' can open a quality line>\n
#pragma once
#include <stdint.h>
#include <string.h>
#include "../../include/fxgpu.h"
#ifdef __CUDACC__
#define FXG_HD __host__ __device__
#else
#define FXG_HD
#endif

namespace fxg {

FXG_HD inline unsigned long long mix64(unsigned long long z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
FXG_HD inline unsigned long long rnd(unsigned long long seed, unsigned long long g, unsigned int j)
{
    return mix64(seed + g * 0x9E3779B97F4A7C15ull + (unsigned long long)(j + 1) * 0xD1B54A32D192ED03ull);
}
FXG_HD inline int dec_digits(unsigned long long g)
{
    int d = 1;
    while (g >= 10) { g /= 10; ++d; }
    return d;
}
// sum of dec_digits(x) for x in [0, n)
FXG_HD inline unsigned long long digit_sum_below(unsigned long long n)
{
    unsigned long long s = n, p = 10;
    for (int k = 0; k < 19; ++k) {
        if (n > p) s += n - p; else break;
        p *= 10;
    }
    return s;
}
FXG_HD inline unsigned long long rec_fixed_bytes(const fxg_synth_spec &s)
{
    return 39ull + 2ull * s.read_len;   // everything except the decimal index
}
FXG_HD inline unsigned long long rec_offset(const fxg_synth_spec &s, unsigned long long i)
{
    return i * rec_fixed_bytes(s) + digit_sum_below(s.first_record + i) - digit_sum_below(s.first_record);
}
FXG_HD inline unsigned int base_char(unsigned int code)
{
    return code == 0 ? 'A' : (code == 1 ? 'C' : (code == 2 ? 'G' : 'T'));
}
FXG_HD inline unsigned int pattern_base(const fxg_synth_spec &s, unsigned int p, unsigned int k)
{
    unsigned long long w = rnd(s.pattern_seed, p, k >> 5);
    return (unsigned int)(w >> ((k & 31) * 2)) & 3u;
}
// sequence base k of global record g (2-bit code)
FXG_HD inline unsigned int seq_code(const fxg_synth_spec &s, unsigned long long g, unsigned int k)
{
    if (s.plant_mode && s.n_patterns && s.pattern_len && s.pattern_len <= s.read_len) {
        if ((rnd(s.seed, g, 200) % 1000000ull) < s.plant_per_million) {
            unsigned int p = (unsigned int)(rnd(s.seed, g, 201) % s.n_patterns);
            unsigned int o = (unsigned int)(rnd(s.seed, g, 202) % (s.read_len - s.pattern_len + 1));
            if (k >= o && k < o + s.pattern_len) {
                bool rc = s.plant_revcomp && (rnd(s.seed, g, 203) & 1ull);
                unsigned int q = k - o;
                if (!rc) return pattern_base(s, p, q);
                return 3u - pattern_base(s, p, s.pattern_len - 1 - q);    // A<->T, C<->G with codes A0 C1 G2 T3
            }
        }
    }
    unsigned long long w = rnd(s.seed, g, k >> 5);
    return (unsigned int)(w >> ((k & 31) * 2)) & 3u;
}
// byte b of the record with global index g whose index has d decimal digits
FXG_HD inline unsigned int rec_byte(const fxg_synth_spec &s, unsigned long long g, int d, unsigned int b)
{
    const unsigned int R = s.read_len;
    const unsigned int hl = 35u + (unsigned int)d;      // header line incl. '\n'
    if (b < hl) {
        unsigned long long h = rnd(s.seed, g, 100);
        if (b < 10) { const char *pre = "@SYN:1:FC:"; return (unsigned int)pre[b]; }
        if (b == 10) return '1' + (unsigned int)(h % 8);
        if (b == 11) return ':';
        if (b < 16) {
            unsigned int tile = 1101u + (unsigned int)((h >> 8) % 1000);
            unsigned int div = b == 12 ? 1000u : (b == 13 ? 100u : (b == 14 ? 10u : 1u));
            return '0' + (tile / div) % 10;
        }
        if (b == 16) return ':';
        if (b < 22) {
            unsigned int x = 10000u + (unsigned int)((h >> 24) % 90000);
            unsigned int div = 1;
            for (unsigned int q = b; q < 21; ++q) div *= 10;
            return '0' + (x / div) % 10;
        }
        if (b == 22) return ':';
        if (b < 23u + (unsigned int)d) {
            unsigned long long div = 1;
            for (unsigned int q = b; q < 22u + (unsigned int)d; ++q) div *= 10;
            return '0' + (unsigned int)((g / div) % 10);
        }
        unsigned int t = b - (23u + (unsigned int)d);
        const char *suf = " 1:N:0:ACGT\n";
        return (unsigned int)suf[t];
    }
    b -= hl;
    if (b < R) return base_char(seq_code(s, g, b));
    if (b == R) return '\n';
    if (b == R + 1) return '+';
    if (b == R + 2) return '\n';
    b -= R + 3;
    if (b < R) {
        unsigned long long w = rnd(s.seed, g, 8 + (b >> 3));
        unsigned int by = (unsigned int)(w >> ((b & 7) * 8)) & 0xffu;
        return 33u + ((by * 41u) >> 8);
    }
    return '\n';
}

// Host generator: same bytes as rec_byte(), produced a record at a time (the
// per-record random words are drawn once, not once per byte) and on all host threads.
inline void synth_host_record(const fxg_synth_spec &s, unsigned long long g, int d, uint8_t *o)
{
    const unsigned int R = s.read_len;
    unsigned long long h = rnd(s.seed, g, 100);
    memcpy(o, "@SYN:1:FC:", 10); o += 10;
    *o++ = (uint8_t)('1' + (unsigned int)(h % 8)); *o++ = ':';
    unsigned int tile = 1101u + (unsigned int)((h >> 8) % 1000);
    o[0] = '0' + tile / 1000 % 10; o[1] = '0' + tile / 100 % 10; o[2] = '0' + tile / 10 % 10; o[3] = '0' + tile % 10; o += 4;
    *o++ = ':';
    unsigned int x = 10000u + (unsigned int)((h >> 24) % 90000);
    for (int k = 4; k >= 0; --k) { o[k] = '0' + x % 10; x /= 10; }
    o += 5; *o++ = ':';
    { unsigned long long v = g; for (int k = d - 1; k >= 0; --k) { o[k] = (uint8_t)('0' + v % 10); v /= 10; } o += d; }
    memcpy(o, " 1:N:0:ACGT\n", 12); o += 12;
    for (unsigned int k = 0; k < R; k += 32) {
        unsigned long long w = rnd(s.seed, g, k >> 5);
        unsigned int n = R - k < 32 ? R - k : 32;
        for (unsigned int q = 0; q < n; ++q) o[k + q] = (uint8_t)base_char((unsigned int)(w >> (q * 2)) & 3u);
    }
    if (s.plant_mode && s.n_patterns && s.pattern_len && s.pattern_len <= R &&
        (rnd(s.seed, g, 200) % 1000000ull) < s.plant_per_million) {
        unsigned int p = (unsigned int)(rnd(s.seed, g, 201) % s.n_patterns);
        unsigned int off = (unsigned int)(rnd(s.seed, g, 202) % (R - s.pattern_len + 1));
        bool rc = s.plant_revcomp && (rnd(s.seed, g, 203) & 1ull);
        for (unsigned int q = 0; q < s.pattern_len; ++q)
            o[off + q] = (uint8_t)base_char(rc ? 3u - pattern_base(s, p, s.pattern_len - 1 - q) : pattern_base(s, p, q));
    }
    o += R; *o++ = '\n'; *o++ = '+'; *o++ = '\n';
    for (unsigned int b = 0; b < R; b += 8) {
        unsigned long long w = rnd(s.seed, g, 8 + (b >> 3));
        unsigned int n = R - b < 8 ? R - b : 8;
        for (unsigned int q = 0; q < n; ++q) o[b + q] = (uint8_t)(33u + ((((unsigned int)(w >> (q * 8)) & 0xffu) * 41u) >> 8));
    }
    o[R] = '\n';
}


}  // namespace fxg
