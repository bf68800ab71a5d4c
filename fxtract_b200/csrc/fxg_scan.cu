// fxg_scan.cu -- raw FASTQ path (Path A): one fused pass over the batch.
//
// Replaces, for canonical 4-line FASTQ, the reference's per-record loop
//   kstream::read (kseq.hpp:120-164)  ->  record boundaries
//   prepare_data (main.cpp:291-304)   ->  target field
//   strstr x1..2 (main.cpp:568-588)   ->  literal hit
//   on_match (main.cpp:258-271)       ->  count / select
// with a position-centric kernel: every CTA takes a 32 KiB tile, finds '\n' with
// SWAR on 16-byte loads, turns the per-tile counts into global line numbers with a
// single-pass decoupled look-back prefix sum (line phase = line# mod 4 decides
// header / sequence / plus / quality), validates the canonical-FASTQ predicate
// under which kseq-parse-then-print is the identity, and runs
// the literal matcher over aligned 8-byte units.  Hits set one bit per record;
// K6 (compact_*) turns bits + record starts into ordered byte runs.
//
// HBM traffic: each input byte is read once; outputs are 1 bit/record (+ 8 B per
// record of record starts only when records are emitted).
#include "fxg_common.cuh"

namespace fxg {

struct __align__(16) ScanSmem {
    uint4 tile[kChunks];                      // staged tile bytes
    unsigned char cnt[kChunks];               // low-byte ('\n') count per chunk
    unsigned int pre8[kThreads];              // exclusive '\n' prefix per group of 8 chunks
    unsigned short nlpos[kMaxTileLines];      // tile-relative '\n' positions, in order
    unsigned short nlchunk[kChunks];          // chunks that contain a low byte
    int cand_off[kCandCap];                   // tile-relative offset of a candidate unit
    unsigned short cand_mask[kCandCap];       // (pattern,shift) bits to verify
    unsigned long long lit_keys[64];
    unsigned short lit_masks[64];
    unsigned int warp_sums[kThreads / 32];
    unsigned int n_nlchunk, n_cand, tile_hits, viol;
    unsigned int total;                       // low bytes in this tile
    unsigned int tile_id;
    unsigned long long lbase;                 // '\n' count before this tile
    unsigned long long first_bad;
    long long e[3];                           // last three '\n' positions before the tile
};

__device__ __forceinline__ uint4 ld_stream(const uint8_t *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ unsigned long long ld_vol64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ long long ld_vol64s(const long long *p)
{
    long long v;
    asm volatile("ld.volatile.global.s64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}

// bit7 of each byte set iff the byte is < 0x20 (valid for bytes < 0x80)
__device__ __forceinline__ unsigned int low_mask(unsigned int x)
{
    return ~(x + 0x60606060u) & ~x & 0x80808080u;
}
// nonzero bit7 somewhere iff some byte >= 0x7f
__device__ __forceinline__ unsigned int high_bits(unsigned int x)
{
    return (x + 0x01010101u) | x;
}
__device__ __forceinline__ unsigned int has_zero_byte(unsigned int v)
{
    return (v - 0x01010101u) & ~v & 0x80808080u;
}

__device__ __forceinline__ void note_violation(ScanSmem &S, unsigned int bit, unsigned long long off)
{
    atomicOr(&S.viol, bit);
    atomicMin(&S.first_bad, off);
}

// byte of the batch at global offset g: from the staged tile when inside it
__device__ __forceinline__ unsigned int text_byte(const ScanSmem &S, const ScanParams &P,
                                                  unsigned long long base, long long g)
{
    long long rel = g - (long long)base;
    if (rel >= 0 && rel < kTileBytes) return reinterpret_cast<const unsigned char *>(S.tile)[rel];
    return P.buf[g];
}

// '\n' count in tile bytes [0, rel)
__device__ __forceinline__ unsigned int lines_before(const ScanSmem &S, int rel)
{
    int c = rel >> 4;
    unsigned int n = S.pre8[c >> 3];
    for (int q = (c & ~7); q < c; ++q) n += S.cnt[q];
    const unsigned char *tb = reinterpret_cast<const unsigned char *>(S.tile);
    for (int b = c << 4; b < rel; ++b) n += (tb[b] < 0x20);
    return n;
}

// global position of the '\n' with tile-local index i (i >= -3)
__device__ __forceinline__ long long nl_global(const ScanSmem &S, unsigned long long base, int i)
{
    if (i >= kMaxTileLines) return kNoPos;
    return i >= 0 ? (long long)(base + S.nlpos[i]) : S.e[-i - 1];
}

// ---------------------------------------------------------------- verification of one literal candidate
__device__ void verify_candidate(ScanSmem &S, const ScanParams &P, const LiteralTable *lit,
                                 unsigned long long base, int n_tab, int unit_rel, unsigned int bits)
{
    while (bits) {
        int b = __ffs(bits) - 1;
        bits &= bits - 1;
        int pat = b >> 3, s = b & 7;
        int len = lit->pat_len[pat];
        long long p = (long long)base + unit_rel - s;
        if (p < 0 || (unsigned long long)(p + len) > P.nbytes) continue;
        const uint8_t *pb = P.pat_bytes + lit->pat_off[pat];
        bool ok = true;
        for (int k = 0; k < len; ++k) {
            if (text_byte(S, P, base, p + k) != pb[k]) { ok = false; break; }
        }
        if (!ok) continue;
        // line attribution: number of '\n' before p
        int rel = (int)(p - (long long)base);
        unsigned int loc = rel > 0 ? lines_before(S, rel) : 0u;
        unsigned long long L = S.lbase + loc;
        int ph = (int)(L & 3);
        bool hit = false;
        if (P.target == FXG_TARGET_SEQ) hit = (ph == 1);
        else if (P.target == FXG_TARGET_QUAL) hit = (ph == 3);
        else if (ph == 0) {
            // header line: name = (ls, first ' ' or end), comment = after that ' '
            long long ls = nl_global(S, base, (int)loc - 1) + 1;      // offset of '@'
            long long q = ls + 1, sp = -1;
            for (;; ++q) {
                if ((unsigned long long)q >= P.nbytes) break;
                unsigned int ch = text_byte(S, P, base, q);
                if (ch == '\n') break;
                if (ch == ' ') { sp = q; break; }
            }
            if (P.target == FXG_TARGET_NAME) {
                long long name_end = sp >= 0 ? sp : q;
                hit = (p >= ls + 1) && (p + len <= name_end);
            } else {
                hit = (sp >= 0) && (p >= sp + 1);
            }
        }
        if (hit) {
            unsigned long long rec = L >> 2;
            if (rec < P.rec_cap) {
                unsigned int bit = 1u << (rec & 31);
                unsigned int old = atomicOr(&P.hitbits[rec >> 5], bit);
                if (!(old & bit)) atomicAdd(&S.tile_hits, 1u);
            } else {
                note_violation(S, FXG_V_DENSE_LINES, (unsigned long long)p);
            }
        }
    }
}

__device__ __forceinline__ void push_candidate(ScanSmem &S, const ScanParams &P, const LiteralTable *lit,
                                               unsigned long long base, int n_tab, int unit_rel, unsigned int bits)
{
    unsigned int slot = atomicAdd(&S.n_cand, 1u);
    if (slot < kCandCap) { S.cand_off[slot] = unit_rel; S.cand_mask[slot] = (unsigned short)bits; }
    else verify_candidate(S, P, lit, base, n_tab, unit_rel, bits);   // queue full: verify in place
}

// ---------------------------------------------------------------- the kernel
__global__ void __launch_bounds__(kThreads, 4) scan_raw_kernel(const ScanParams P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ScanSmem &S = *reinterpret_cast<ScanSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const LiteralTable *lit = P.lit;
    int lit_unit = 0, lit_nkeys = 0;
    unsigned int lit_mul1 = 0, lit_mul2 = 0;
    if (P.match_enabled) { lit_unit = lit->unit; lit_nkeys = lit->n_keys; lit_mul1 = lit->mul1; lit_mul2 = lit->mul2; }

    if (tid == 0) {
        S.tile_id = atomicAdd(&P.ctr->ticket, 1u);
        S.n_nlchunk = 0; S.n_cand = 0; S.tile_hits = 0; S.viol = 0;
        S.first_bad = ~0ull;
    }
    if (P.match_enabled && tid < 64) { S.lit_keys[tid] = lit->keys[tid]; S.lit_masks[tid] = lit->masks[tid]; }
    __syncthreads();
    const unsigned long long t = S.tile_id;
    const unsigned long long base = t * (unsigned long long)kTileBytes;
    const unsigned long long remain = P.nbytes - base;
    const int tile_len = remain < (unsigned long long)kTileBytes ? (int)remain : kTileBytes;
    unsigned char *tb = reinterpret_cast<unsigned char *>(S.tile);

    // ---- A. coalesced 16-byte loads, stage in smem, SWAR '\n' / range classification
    {
        uint4 v[kChunksPerThread];
#pragma unroll
        for (int j = 0; j < kChunksPerThread; ++j) {
            int c = warp * (kChunks / 8) + j * 32 + lane;
            int off = c << 4;
            if (off + 16 <= tile_len) v[j] = ld_stream(P.buf + base + off);
            else {
                unsigned int w[4] = {0x78787878u, 0x78787878u, 0x78787878u, 0x78787878u};
                for (int k = 0; k < 16 && off + k < tile_len; ++k) {
                    unsigned int by = P.buf[base + off + k];
                    w[k >> 2] = (w[k >> 2] & ~(0xffu << ((k & 3) * 8))) | (by << ((k & 3) * 8));
                }
                v[j] = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
        unsigned int hi = 0;
#pragma unroll
        for (int j = 0; j < kChunksPerThread; ++j) {
            int c = warp * (kChunks / 8) + j * 32 + lane;
            S.tile[c] = v[j];
            unsigned int m = low_mask(v[j].x) | (low_mask(v[j].y) >> 1) | (low_mask(v[j].z) >> 2) | (low_mask(v[j].w) >> 3);
            hi |= high_bits(v[j].x) | high_bits(v[j].y) | high_bits(v[j].z) | high_bits(v[j].w);
            unsigned int n = __popc(m);
            S.cnt[c] = (unsigned char)n;
            if (n) { unsigned int slot = atomicAdd(&S.n_nlchunk, 1u); S.nlchunk[slot] = (unsigned short)c; }
        }
        if (hi & 0x80808080u) note_violation(S, FXG_V_BYTE_RANGE, base);
    }
    __syncthreads();

    // ---- B. block scan of per-chunk counts (8 consecutive chunks per thread)
    {
        uint2 w = *reinterpret_cast<const uint2 *>(&S.cnt[tid * 8]);
        unsigned int s = __dp4a(w.x, 0x01010101u, 0u) + __dp4a(w.y, 0x01010101u, 0u);
        unsigned int inc = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned int o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) S.warp_sums[warp] = inc;
        __syncthreads();
        unsigned int woff = 0, tot = 0;
#pragma unroll
        for (int w2 = 0; w2 < kThreads / 32; ++w2) {
            unsigned int ws = S.warp_sums[w2];
            if (w2 < warp) woff += ws;
            tot += ws;
        }
        S.pre8[tid] = woff + inc - s;
        if (tid == 0) S.total = tot;
    }
    __syncthreads();
    const unsigned int total = S.total;
    const int n_tab = total < (unsigned int)kMaxTileLines ? (int)total : kMaxTileLines;
    if (total > (unsigned int)kMaxTileLines && tid == 0) note_violation(S, FXG_V_DENSE_LINES, base);

    // ---- D1. ordered '\n' position table from the chunks that have low bytes
    {
        const unsigned int nn = S.n_nlchunk;
        for (unsigned int q = tid; q < nn; q += kThreads) {
            int c = S.nlchunk[q];
            unsigned int idx = S.pre8[c >> 3];
            for (int k = (c & ~7); k < c; ++k) idx += S.cnt[k];
            for (int k = 0; k < 16; ++k) {
                unsigned int by = tb[(c << 4) + k];
                if (by < 0x20) {
                    if (by != '\n') note_violation(S, FXG_V_BYTE_RANGE, base + (c << 4) + k);
                    if (idx < (unsigned int)kMaxTileLines) S.nlpos[idx] = (unsigned short)((c << 4) + k);
                    ++idx;
                }
            }
        }
    }
    __syncthreads();

    // ---- look-back: publish aggregate (+ own last three '\n'), fetch exclusive prefix
    if (warp == 0) {
        TileDesc *D = P.desc;
        unsigned long long own3 = n_tab >= 3 ? 3 : n_tab;
        if (lane == 0) {
            D[t].p1 = n_tab >= 1 ? (long long)(base + S.nlpos[n_tab - 1]) : kNoPos;
            D[t].p2 = n_tab >= 2 ? (long long)(base + S.nlpos[n_tab - 2]) : kNoPos;
            D[t].p3 = n_tab >= 3 ? (long long)(base + S.nlpos[n_tab - 3]) : kNoPos;
            __threadfence();
            unsigned long long st = (t == 0 ? kFlagInc : kFlagAgg) | (own3 << 60) | (unsigned long long)total;
            atomicExch(&D[t].status, st);
        }
        unsigned long long excl = 0;
        long long e0 = kNoPos, e1 = kNoPos, e2 = kNoPos;
        int have = 0;
        bool count_done = false;
        long long look = (long long)t - 1;
        while (!count_done || have < 3) {
            long long idx = look - lane;
            unsigned long long st;
            long long q1 = kNoPos, q2 = kNoPos, q3 = kNoPos;
            if (idx >= 0) {
                unsigned int spins = 0;
                do {
                    st = ld_vol64(&D[idx].status);
                    if (++spins > (1u << 25)) { atomicOr(&P.ctr->violation, 1u << 30); st = kFlagInc; break; }   // never hang the GPU
                } while ((st >> 62) == 0);
                __threadfence();
                q1 = ld_vol64s(&D[idx].p1); q2 = ld_vol64s(&D[idx].p2); q3 = ld_vol64s(&D[idx].p3);
            } else if (idx == -1) {
                st = kFlagInc | (1ull << 60);      // virtual '\n' at offset -1 opens line 0
                q1 = -1;
            } else {
                st = kFlagInc;
            }
            if (!count_done) {
                unsigned int incmask = __ballot_sync(0xffffffffu, (st >> 62) == 2);
                int first_inc = incmask ? (__ffs(incmask) - 1) : 32;
                unsigned long long val = (lane <= first_inc) ? (st & kValMask) : 0ull;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
                excl += val;
                if (first_inc < 32) count_done = true;
            }
            int c3 = (int)((st >> 60) & 3);
            for (int l = 0; l < 32 && have < 3; ++l) {
                int cc = __shfl_sync(0xffffffffu, c3, l);
                long long r1 = __shfl_sync(0xffffffffu, q1, l);
                long long r2 = __shfl_sync(0xffffffffu, q2, l);
                long long r3 = __shfl_sync(0xffffffffu, q3, l);
                if (cc >= 1 && have < 3) { if (have == 0) e0 = r1; else if (have == 1) e1 = r1; else e2 = r1; ++have; }
                if (cc >= 2 && have < 3) { if (have == 0) e0 = r2; else if (have == 1) e1 = r2; else e2 = r2; ++have; }
                if (cc >= 3 && have < 3) { if (have == 0) e0 = r3; else if (have == 1) e1 = r3; else e2 = r3; ++have; }
            }
            look -= 32;
            if (look < -1) { have = 3; if (!count_done) count_done = true; }
        }
        if (lane == 0) {
            S.lbase = excl; S.e[0] = e0; S.e[1] = e1; S.e[2] = e2;
            if (t != 0) {
                unsigned long long st = kFlagInc | (own3 << 60) | (excl + (unsigned long long)total);
                atomicExch(&D[t].status, st);
            }
        }
    }
    __syncthreads();
    const unsigned long long lbase = S.lbase;

    // ---- C. phase-aware class checks on chunks without '\n' + literal filter on every chunk
#pragma unroll 2
    for (int j = 0; j < kChunksPerThread; ++j) {
        int c = warp * (kChunks / 8) + j * 32 + lane;
        int off = c << 4;
        if (off >= tile_len) continue;
        uint4 x = S.tile[c];
        if (S.cnt[c] == 0) {
            uint2 w = *reinterpret_cast<const uint2 *>(&S.cnt[c & ~7]);
            int sh = (c & 7) * 8;
            unsigned long long m = sh ? ((1ull << sh) - 1ull) : 0ull;
            unsigned int lo = w.x & (unsigned int)m, hi2 = w.y & (unsigned int)(m >> 32);
            unsigned int pc = S.pre8[c >> 3] + __dp4a(lo, 0x01010101u, 0u) + __dp4a(hi2, 0x01010101u, 0u);
            int ph = (int)((lbase + pc) & 3);
            if (ph == 1) {
                // sequence: every byte must have bit 6 set (0x40..0x7e) and not be '@'
                unsigned int a = x.x & x.y & x.z & x.w;
                unsigned int z = has_zero_byte(x.x ^ 0x40404040u) | has_zero_byte(x.y ^ 0x40404040u) |
                                 has_zero_byte(x.z ^ 0x40404040u) | has_zero_byte(x.w ^ 0x40404040u);
                if ((a & 0x40404040u) != 0x40404040u || z) note_violation(S, FXG_V_SEQ_CLASS, base + off);
            } else if (ph == 3) {
                // quality: no space (all other bytes already range-checked)
                unsigned int z = has_zero_byte(x.x & 0x5f5f5f5fu) | has_zero_byte(x.y & 0x5f5f5f5fu) |
                                 has_zero_byte(x.z & 0x5f5f5f5fu) | has_zero_byte(x.w & 0x5f5f5f5fu);
                if (z) note_violation(S, FXG_V_QUAL_CLASS, base + off);
            }
        }
        if (P.match_enabled) {
            if (lit_unit == 8) {
                unsigned int h0 = ((x.x * lit_mul1) ^ (x.y * lit_mul2)) >> 26;
                unsigned int h1 = ((x.z * lit_mul1) ^ (x.w * lit_mul2)) >> 26;
                unsigned long long k0 = S.lit_keys[h0], k1 = S.lit_keys[h1];
                unsigned long long u0 = ((unsigned long long)x.y << 32) | x.x;
                unsigned long long u1 = ((unsigned long long)x.w << 32) | x.z;
                if (k0 == u0) push_candidate(S, P, lit, base, n_tab, off, S.lit_masks[h0]);
                if (k1 == u1) push_candidate(S, P, lit, base, n_tab, off + 8, S.lit_masks[h1]);
            } else {
                const int U = lit_unit;
                const unsigned int wv[4] = {x.x, x.y, x.z, x.w};
                for (int u = 0; u < 16; u += U) {
                    unsigned int word = wv[u >> 2];
                    unsigned int val = U == 4 ? word : (U == 2 ? ((word >> ((u & 2) * 8)) & 0xffffu)
                                                                : ((word >> ((u & 3) * 8)) & 0xffu));
                    for (int k = 0; k < lit_nkeys; ++k)
                        if ((unsigned int)S.lit_keys[k] == val) push_candidate(S, P, lit, base, n_tab, off + u, S.lit_masks[k]);
                }
            }
        }
    }

    // ---- E. one thread per '\n': record structure + partial-chunk class checks + record table
    for (int i = tid; i < n_tab; i += kThreads) {
        const unsigned long long j = lbase + (unsigned long long)i;
        const int ph = (int)(j & 3);
        const int rel = S.nlpos[i];
        const long long pos = (long long)base + rel;
        const long long prev = nl_global(S, base, i - 1);
        const long long llen = pos - prev - 1;                   // bytes in the line that ends here
        if (llen > 0 && text_byte(S, P, base, pos - 1) == ' ') note_violation(S, FXG_V_HEADER_SPACE, (unsigned long long)pos);
        if (ph == 0) {
            if (llen < 1 || text_byte(S, P, base, prev + 1) != '@') note_violation(S, FXG_V_HEADER_AT, (unsigned long long)(prev + 1));
            if (P.fld_off && (P.target == FXG_TARGET_NAME || P.target == FXG_TARGET_COMMENT)) {
                unsigned long long rec = j >> 2;
                if (rec < P.rec_cap) {
                    long long q = prev + 2, sp = -1;
                    for (; q < pos; ++q) if (text_byte(S, P, base, q) == ' ') { sp = q; break; }
                    if (P.target == FXG_TARGET_NAME) {
                        long long ne = sp >= 0 ? sp : pos;
                        P.fld_off[rec] = (unsigned long long)(prev + 2);
                        P.fld_len[rec] = (unsigned int)(ne > prev + 2 ? ne - (prev + 2) : 0);
                    } else {
                        P.fld_off[rec] = (unsigned long long)(sp >= 0 ? sp + 1 : pos);
                        P.fld_len[rec] = (unsigned int)(sp >= 0 ? pos - (sp + 1) : 0);
                    }
                }
            }
        } else if (ph == 1) {
            if (llen < 1) note_violation(S, FXG_V_EMPTY_SEQ, (unsigned long long)pos);
            if (P.fld_off && P.target == FXG_TARGET_SEQ) {
                unsigned long long rec = j >> 2;
                if (rec < P.rec_cap) { P.fld_off[rec] = (unsigned long long)(prev + 1); P.fld_len[rec] = (unsigned int)llen; }
            }
        } else if (ph == 2) {
            if (llen != 1 || text_byte(S, P, base, pos - 1) != '+') note_violation(S, FXG_V_PLUS_LINE, (unsigned long long)pos);
        } else {
            long long slen = nl_global(S, base, i - 2) - nl_global(S, base, i - 3) - 1;
            if (llen != slen) note_violation(S, FXG_V_QUAL_LEN, (unsigned long long)pos);
            unsigned long long rec = j >> 2;
            if (P.rec_start && rec + 1 <= P.rec_cap) P.rec_start[rec + 1] = (unsigned long long)(pos + 1);
            if (P.fld_off && P.target == FXG_TARGET_QUAL && rec < P.rec_cap) {
                P.fld_off[rec] = (unsigned long long)(prev + 1); P.fld_len[rec] = (unsigned int)llen;
            }
        }
        // partial-chunk class checks around this '\n' (full chunks were checked in C)
        {
            int cb = rel & ~15;
            int lo = (prev + 1 > (long long)base + cb) ? (int)(prev + 1 - (long long)base) : cb;
            if (ph == 1) {
                for (int b = lo; b < rel; ++b) { unsigned int ch = tb[b]; if (!(ch & 0x40) || ch == 0x40) note_violation(S, FXG_V_SEQ_CLASS, base + b); }
            } else if (ph == 3) {
                for (int b = lo; b < rel; ++b) if (tb[b] == ' ') note_violation(S, FXG_V_QUAL_CLASS, base + b);
            }
            int ce = cb + 16 < tile_len ? cb + 16 : tile_len;
            int hi = (i + 1 < n_tab && S.nlpos[i + 1] < ce) ? S.nlpos[i + 1] : ce;
            int ph2 = (ph + 1) & 3;
            if (ph2 == 1) {
                for (int b = rel + 1; b < hi; ++b) { unsigned int ch = tb[b]; if (!(ch & 0x40) || ch == 0x40) note_violation(S, FXG_V_SEQ_CLASS, base + b); }
            } else if (ph2 == 3) {
                for (int b = rel + 1; b < hi; ++b) if (tb[b] == ' ') note_violation(S, FXG_V_QUAL_CLASS, base + b);
            }
        }
    }
    if (t == 0 && tid == 0 && P.rec_start) P.rec_start[0] = 0;
    __syncthreads();

    // ---- V. verify queued literal candidates
    if (P.match_enabled) {
        unsigned int nc = S.n_cand < (unsigned int)kCandCap ? S.n_cand : (unsigned int)kCandCap;
        for (unsigned int q = tid; q < nc; q += kThreads)
            verify_candidate(S, P, lit, base, n_tab, S.cand_off[q], S.cand_mask[q]);
        __syncthreads();
    }

    // ---- epilogue
    if (tid == 0) {
        if (S.tile_hits) atomicAdd(&P.ctr->n_hits, (unsigned long long)S.tile_hits);
        if (S.viol) { atomicOr(&P.ctr->violation, S.viol); atomicMin(&P.ctr->first_bad, S.first_bad); }
        if (base + (unsigned long long)tile_len == P.nbytes) {
            P.ctr->n_lines = lbase + total;
            P.ctr->last_byte = tb[tile_len - 1];
        }
    }
}

size_t scan_smem_bytes() { return sizeof(ScanSmem); }

cudaError_t launch_scan_raw(const ScanParams &P, cudaStream_t stream)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(scan_raw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScanSmem));
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    unsigned long long ntiles = (P.nbytes + kTileBytes - 1) / kTileBytes;
    if (ntiles == 0) return cudaSuccess;
    scan_raw_kernel<<<(unsigned int)ntiles, kThreads, sizeof(ScanSmem), stream>>>(P);
    return cudaGetLastError();
}

// ================================================================= K6: flags -> ordered runs
// sel(k) = hit(k) XOR invert for k < n_records; a run is a maximal block of
// consecutive selected records; runs[r] = {rec_start[first], rec_start[last+1]-rec_start[first]}.
// Replaces the per-record `pair.print(out)` of on_match (main.cpp:258-271, kseq.cpp:32-48)
// for canonical FASTQ, where print(parse(bytes)) == bytes.
struct CompactParams {
    const unsigned int *hitbits;
    const unsigned long long *rec_start;
    const ScanCounters *ctr;        // n_lines -> n_records
    unsigned long long *block_sums; // [nblocks] run starts per block -> exclusive prefix
    unsigned long long *out_counts; // [0] = n_runs, [1] = n_selected
    fxg_run *runs;
    unsigned long long run_cap;
    int invert;
};
constexpr int kCompactThreads = 1024;

__device__ __forceinline__ unsigned int sel_word(const CompactParams &C, unsigned long long w, unsigned long long nrec)
{
    unsigned long long first = w * 32ull;
    if (first >= nrec) return 0u;
    unsigned int v = C.hitbits[w];
    if (C.invert) v = ~v;
    unsigned long long left = nrec - first;
    if (left < 32) v &= (1u << left) - 1u;
    return v;
}

__global__ void __launch_bounds__(kCompactThreads) compact_count_kernel(const CompactParams C)
{
    __shared__ unsigned long long red[2][kCompactThreads / 32];
    const unsigned long long nrec = C.ctr->n_lines >> 2;
    const unsigned long long w = (unsigned long long)blockIdx.x * kCompactThreads + threadIdx.x;
    unsigned int s = sel_word(C, w, nrec);
    unsigned int prev_msb = (w > 0) ? (sel_word(C, w - 1, nrec) >> 31) : 0u;
    unsigned int starts = s & ~((s << 1) | prev_msb);
    unsigned long long a = __popc(starts), b = __popc(s);
    for (int d = 16; d > 0; d >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, d); b += __shfl_xor_sync(0xffffffffu, b, d); }
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = a; red[1][threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long ta = 0, tb2 = 0;
        for (int i = 0; i < kCompactThreads / 32; ++i) { ta += red[0][i]; tb2 += red[1][i]; }
        C.block_sums[blockIdx.x] = ta;
        if (tb2) atomicAdd(&C.out_counts[1], tb2);
    }
}

__global__ void __launch_bounds__(1024) compact_scan_kernel(const CompactParams C, unsigned int nblocks)
{
    // single CTA: exclusive scan of block_sums in place; total -> out_counts[0]
    __shared__ unsigned long long carry;
    __shared__ unsigned long long wsum[32];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (unsigned int b0 = 0; b0 < nblocks; b0 += 1024) {
        unsigned int i = b0 + threadIdx.x;
        unsigned long long v = i < nblocks ? C.block_sums[i] : 0ull;
        unsigned long long inc = v;
        for (int d = 1; d < 32; d <<= 1) { unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d); if ((threadIdx.x & 31) >= d) inc += o; }
        if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
        __syncthreads();
        unsigned long long woff = 0;
        for (int k = 0; k < (int)(threadIdx.x >> 5); ++k) woff += wsum[k];
        unsigned long long c0 = carry;
        if (i < nblocks) C.block_sums[i] = c0 + woff + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = c0 + woff + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) C.out_counts[0] = carry;
}

__global__ void __launch_bounds__(kCompactThreads) compact_write_kernel(const CompactParams C)
{
    __shared__ unsigned long long wsum[kCompactThreads / 32];
    const unsigned long long nrec = C.ctr->n_lines >> 2;
    const unsigned long long w = (unsigned long long)blockIdx.x * kCompactThreads + threadIdx.x;
    unsigned int s = sel_word(C, w, nrec);
    unsigned int prev_msb = (w > 0) ? (sel_word(C, w - 1, nrec) >> 31) : 0u;
    unsigned int next_lsb = sel_word(C, w + 1, nrec) & 1u;
    unsigned int starts = s & ~((s << 1) | prev_msb);
    unsigned int ends = s & ~((s >> 1) | (next_lsb << 31));
    unsigned long long n = __popc(starts), inc = n;
    for (int d = 1; d < 32; d <<= 1) { unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d); if ((threadIdx.x & 31) >= d) inc += o; }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
    __syncthreads();
    unsigned long long woff = 0;
    for (int k = 0; k < (int)(threadIdx.x >> 5); ++k) woff += wsum[k];
    unsigned long long excl = C.block_sums[blockIdx.x] + woff + inc - n;   // run starts before this word
    // a run that is open at the start of this word has index excl-1
    unsigned int m = starts | ends;
    unsigned long long opened = excl;
    while (m) {
        int b = __ffs(m) - 1;
        m &= m - 1;
        unsigned long long rec = w * 32ull + b;
        if (starts & (1u << b)) {
            if (opened < C.run_cap) C.runs[opened].off = C.rec_start[rec];
            ++opened;
        }
        if (ends & (1u << b)) {
            unsigned long long r = opened - 1;
            if (r < C.run_cap) C.runs[r].len = C.rec_start[rec + 1];   // end offset; fixed up below
        }
    }
}

__global__ void compact_fixup_kernel(const CompactParams C)
{
    unsigned long long n = C.out_counts[0] < C.run_cap ? C.out_counts[0] : C.run_cap;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x)
        C.runs[i].len -= C.runs[i].off;
}

cudaError_t launch_compact(const unsigned int *hitbits, const unsigned long long *rec_start, const ScanCounters *ctr,
                           unsigned long long rec_cap, unsigned long long *block_sums, unsigned long long *out_counts,
                           fxg_run *runs, unsigned long long run_cap, int invert, cudaStream_t stream, int *launches)
{
    CompactParams C;
    C.hitbits = hitbits; C.rec_start = rec_start; C.ctr = ctr; C.block_sums = block_sums;
    C.out_counts = out_counts; C.runs = runs; C.run_cap = run_cap; C.invert = invert;
    unsigned long long words = (rec_cap + 31) / 32;
    unsigned int nblocks = (unsigned int)((words + kCompactThreads - 1) / kCompactThreads);
    if (nblocks == 0) nblocks = 1;
    compact_count_kernel<<<nblocks, kCompactThreads, 0, stream>>>(C);
    compact_scan_kernel<<<1, 1024, 0, stream>>>(C, nblocks);
    compact_write_kernel<<<nblocks, kCompactThreads, 0, stream>>>(C);
    compact_fixup_kernel<<<148, 256, 0, stream>>>(C);
    if (launches) *launches += 4;
    return cudaGetLastError();
}

}  // namespace fxg
