// fxg_common.cuh -- shared device/host declarations of libfxgpu (sm_100a only).
#pragma once
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>
#include "../../include/fxgpu.h"

struct fxg_ctx;

namespace fxg {

int set_cuda_error(fxg_ctx *ctx, cudaError_t e, const char *what);

// ---------------------------------------------------------------- geometry
// The scan kernel is compiled for two tile geometries (template parameter of scan_tiles_kernel):
//   G24  24 KiB tiles, 12 warps, 3 CTAs / SM   literal, exact and index-only passes
//   G32  32 KiB tiles, 16 warps, 2 CTAs / SM   fused window matcher with a 32 KiB filter in shared memory
//   G16D 16 KiB tiles, 2 x 16 warps, 1 CTA / SM ... with the 128 KiB filter (large pattern sets): the filter leaves room
//                                              for one CTA only, and a single CTA idles the SM at each of its block
//                                              barriers -- so its 32 warps form two groups that work on different
//                                              tiles and synchronise separately (named barriers), like two CTAs would
//   G16  16 KiB tiles, 16 warps                the REDO launch of G16D batches
// Every warp owns a contiguous 2 KiB segment of its tile in both.  Kernels downstream of the scan (prefix,
// finalize, compaction) read the geometry of the batch from ScanParams at run time.
template <int TILE_BYTES, int THREADS, int MIN_CTAS, int GROUPS = 1> struct Geom {
    static constexpr int kTileBytes = TILE_BYTES;           // bytes of input one group of warps scans per step
    static constexpr int kThreads = THREADS;                 // threads of one group
    static constexpr int kGroups = GROUPS;                   // independent groups per CTA: own tiles, own named barrier,
                                                             // own shared-memory block; they share the matcher's filter
    static constexpr int kCtaThreads = THREADS * GROUPS;
    static constexpr int kMinCtas = MIN_CTAS;                // resident CTAs per SM the kernel is compiled for
    static constexpr int kChunks = TILE_BYTES / 16;          // 16-byte chunks per tile
    static constexpr int kWarps = THREADS / 32;
    static constexpr int kWarpChunks = kChunks / kWarps;     // chunks one warp covers (contiguous)
    static constexpr int kChunksPerThread = kChunks / THREADS;
    // newline-table capacity per tile: one line per 12 bytes.  A denser tile raises FXG_V_DENSE_LINES (the batch is
    // re-routed), so an accepted batch holds at most ntiles * kMaxTileLines / 4 records: the record capacity every
    // per-record buffer is sized with (48 bytes per record).
    static constexpr int kMaxTileLines = ((TILE_BYTES / 12) + 7) & ~7;
    static constexpr int kHitWords = kMaxTileLines / 32 + 1; // line-hit words per tile (local line indices 0..kMaxTileLines)
    static_assert(TILE_BYTES % (THREADS * 16) == 0 && TILE_BYTES <= 65536, "tile = whole 16-byte chunks per thread; u16 positions");
    static_assert(kWarps % 4 == 0, "every line role has its own warps");
};
#if defined(FXG_VAR_G24_TILE)          // geometry experiments (scripts/build_variant.sh): -DFXG_VAR_G24_TILE=16384 -DFXG_VAR_G24_THREADS=256 -DFXG_VAR_G24_CTAS=4
typedef Geom<FXG_VAR_G24_TILE, FXG_VAR_G24_THREADS, FXG_VAR_G24_CTAS> G24;
#else
typedef Geom<24576, 384, 3> G24;
#endif
typedef Geom<32768, 512, 2> G32;
typedef Geom<16384, 512, 1, 2> G16D;
typedef Geom<16384, 512, 2> G16;
constexpr int kCandCap = 512;                  // literal candidates queued per tile pass
constexpr int kKmerCandDiv = 64;               // window matcher: a tile lists up to tile bytes / 64 filter hits for kmer_resolve_kernel

// ---------------------------------------------------------------- per-tile results
// The scan kernel has no cross-tile dependency: every tile publishes its '\n' count, its
// ordered '\n' position table, a 4-bit mask saying for which line phases (global line
// number of the tile's first line, mod 4) the tile is illegal, and one "line has a match"
// bit per local line.  A tiny second kernel turns the counts into global line numbers
// (exclusive scan), picks the phase, checks the few lines that straddle tile edges and
// converts line hits into record hits / record starts / field tables.
struct __align__(16) TileInfo {
    unsigned int total;          // bytes < 0x20 in the tile (== '\n' count unless FXG_V_BYTE_RANGE is raised)
    unsigned int viol;           // bits 0..15: FXG_V_* bits (phase-independent ones, plus -- when bits 16..18 are
                                 // non-zero -- everything evaluated under the guessed phase (bits 16..18) - 1)
    unsigned short cls[4];       // OR of the line flags of the lines with local index == k (mod 4): low byte = the
                                 // line itself, high byte = class flags of the head of the following line
};
constexpr long long kNoPos = -(1ll << 62);

// ---------------------------------------------------------------- counters
struct ScanCounters {
    unsigned long long n_lines;
    unsigned long long n_hits;        // records with >= 1 hit (deduplicated through the hit bitset)
    unsigned long long first_bad;     // min offset of a violation (tile granularity)
    unsigned int violation;           // FXG_V_* bitmask
    unsigned int last_byte;           // last byte of the batch (must be '\n')
    unsigned long long n_pair_sel;    // pair submissions: selected pairs (pair_select_kernel)
};

// ---------------------------------------------------------------- literal matcher tables
// Aligned-unit filter: an occurrence of a pattern of length >= 2U-1 at offset p
// fully contains the U-byte unit at ceil_U(p); the table holds, for every pattern
// and every shift s in [0,U), the U bytes pat[s..s+U).  One lookup per aligned
// unit of text replaces one comparison per text position.  For U == 8 the lookup
// is a 32-bit multiplicative fingerprint (two IMADs) whose top 6 bits index a
// collision-free 64-slot table; every candidate is then verified byte by byte.
struct LiteralTable {
    int n_patterns;              // 1 or 2
    int unit;                    // 8, 4, 2 or 1 bytes
    int n_keys;                  // unique keys
    int all_match;               // empty pattern: every record matches (strstr(hay,"") != NULL)
    int never;                   // pattern contains '\n': can never match a field
    unsigned int mul1, mul2;     // perfect-hash multipliers (unit == 8)
    int shift;                   // unit == 8: slot = fingerprint >> shift (27: 32 slots, 26: 64 slots)
    int pat_len[2];
    int pat_off[2];              // offsets into pat_bytes
    unsigned long long keys[64]; // unit==8: 64-slot perfect hash; else first n_keys entries
    unsigned int fp[64];         // unit==8: fp[h] = lo*mul1 + hi*mul2 of the key in slot h = fp >> shift (else unmatchable)
    unsigned short masks[64];    // bit (pattern*8 + shift) set if keys[i] == pat[shift..shift+U)
};

// Scan modes (template parameter of scan_tiles_kernel)
constexpr int kModeIndex = 0;        // record index + canonical predicate only (a second-stage field matcher follows)
constexpr int kModeLitSmall = 1;     // literal, patterns shorter than 15 bytes: aligned units of 4 / 2 / 1 bytes
constexpr int kModeLit8 = 8;         // literal, aligned 8-byte-unit fingerprints (two multipliers)
constexpr int kModeLit9 = 9;         // ... first multiplier 1
constexpr int kModeKmer = 16;        // + log2(probe stride): 18 (S=4), 19 (S=8), 20 (S=16): fused window matcher (multi-pattern)
constexpr int kModeExact = 30;       // fused whole-field hash look-up (FileManager::find)
__host__ __device__ constexpr bool mode_is_kmer(int m) { return m >= 18 && m <= 20; }
__host__ __device__ constexpr bool mode_is_literal(int m) { return m == 1 || m == 8 || m == 9; }

// Fused multi-pattern matcher (replaces ssearch_scan, main.cpp:406,426, on the raw path; north_star: "warp-cooperative
// scan over coalesced 16-byte loads with overlap handling at chunk edges").  Position-centric: every lane looks at
// the windows of k <= 16 BYTES that start at byte 0, S, 2S, ... of its 16-byte chunk of the staged tile (S = 16, 8 or
// 4; a window that runs past the chunk takes its tail from the head of the next chunk, the last chunk of a tile from
// the 16 bytes staged behind it).  Every pattern contributes the S windows at its first S offsets and k + S - 1 <=
// the shortest pattern, so every occurrence holds a window at a probed text position.  A window is hashed as four
// little-endian words, one multiply-add each (window_hash); the hash selects one word of a blocked Bloom filter in
// shared memory and four bits in it.  Filter hits go to a per-tile list in global memory as {hash, tile-relative
// offset}; kmer_resolve_kernel looks them up afterwards (bucketed table in L2 -> pattern record -> byte-wise
// comparison with the text), one thread per candidate.  No false positives, no false negatives, any alphabet.
struct KmerTile {
    const unsigned int *filter;            // [1 << filter_log2_words] device copy, staged into shared memory per CTA
    // bucketed table: bucket b = 32 bytes (one L2 sector) = 4 window hashes (uint4 at ht[2b]) + 4 values (uint4 at
    // ht[2b+1]); value 0 = empty, else ((record offset / 16 << 7) | window offset in the pattern) + 1; slots fill in
    // order, a full bucket spills into the next one.
    const uint4 *ht;                       // [2 * nbuckets]
    // pattern records, 16-byte aligned: {length u32, pattern id u32, bytes ...}
    const uint4 *rec;
    unsigned int ht_mask;                  // nbuckets - 1
    unsigned int mask2, mask3;             // bytes of window words 2 and 3 that belong to the window (k = 9 .. 16)
    int k, stride;                         // stride in {4, 8, 16}
    int filter_log2_words;                 // 13 (32 KiB, 2 CTAs / SM) or 15 (128 KiB, 1 CTA of 32 warps / SM)
};
// hash of a window given as four little-endian words (bytes past the window masked off); identical on host and device.
// (Measured and dropped: folding the masks into left-shifted multipliers saves two LOP3 per probe, but the last
// bytes of the window then only reach the top bits of the hash -- more filter hits, config 3 resolve + 10 %.)
constexpr unsigned int kWinMul0 = 0x9E3779B1u, kWinMul1 = 0x85EBCA77u, kWinMul2 = 0xC2B2AE3Du, kWinMul3 = 0x27D4EB2Fu;
__host__ __device__ inline unsigned int window_hash(unsigned int w0, unsigned int w1, unsigned int w2, unsigned int w3)
{
    return w0 * kWinMul0 + w1 * kWinMul1 + w2 * kWinMul2 + w3 * kWinMul3;
}
__host__ __device__ inline unsigned int kmer_tile_filter_word(unsigned int h, int log2_words) { return h >> (32 - log2_words); }
// four bits of one filter word per window: {r, r + 16} and {s, s + 8} (mod 32), r = h & 31, s = (h >> 5) & 31
// (Which hash bits give the two rotation counts was measured: 0 / 5 as here; 5 / 10, 7 / 12 cost one more shift per
// probe and gain nothing; 10 / 15 and 12 / 17 overlap the bits that pick the filter WORD, so all entries of a word
// share their rotations and 13 % of the probes pass -- config 3 at a third of its speed.)
#ifndef FXG_FILT_SHIFT_R
#define FXG_FILT_SHIFT_R 0
#define FXG_FILT_SHIFT_S 5
#endif
constexpr int kFiltShiftR = FXG_FILT_SHIFT_R, kFiltShiftS = FXG_FILT_SHIFT_S;
__host__ __device__ inline unsigned int kmer_tile_filter_mask(unsigned int h)
{
    const unsigned int r = (h >> kFiltShiftR) & 31u, s = (h >> kFiltShiftS) & 31u;
    return ((0x00010001u << r) | (0x00010001u >> ((32u - r) & 31u))) | ((0x00000101u << s) | (0x00000101u >> ((32u - s) & 31u)));
}
__host__ __device__ inline unsigned int kmer_tile_bucket(unsigned int h, unsigned int mask) { return ((h * 0x9E3779B1u) >> 9) & mask; }
constexpr unsigned int kCandOverflow = 0x80000000u;   // tile_ncand: the tile had more filter hits than its list holds

// Bucketed table over whole-field keys.  Bucket b = 8 entries {fingerprint, key offset} = 64 bytes (two L2 sectors,
// fetched with four independent 16-byte loads); fingerprint 0 = empty, slots fill in order, a full bucket spills into
// the next one (0.1 % of the buckets at <= 2 keys per bucket).  The key offset points into `kw`: one meta word
// (pattern id << 32 | key length) followed by the key as zero-padded little-endian 8-byte words, so a fingerprint
// match is verified after ONE more round trip.  The tile that issues a look-up waits for its slowest lane: what
// counts is the length of the dependent chain -- one round trip for a miss, two for a hit -- not the average.
struct ExactDevice {
    const uint4 *bkt;                      // [4 * nbuckets] {fp0, off0, fp1, off1} x 4; fingerprint = (hash >> 32) | 1
    const unsigned long long *kw;          // meta + key words per key (+ 8 words of slack at the end)
    unsigned int mask;                     // nbuckets - 1; bucket = (unsigned)hash & mask
    // Pre-filter in front of the table: word-blocked Bloom filter, >= 16 bits per key, four bits of one 32-bit word
    // per key (exact_pre_word / exact_pre_mask).  1 M keys = 2 MiB: resident in L2 while the batch streams past it
    // (the scan's bulk copies are evict-first), so the ~99 % of the records that are not in the list cost one 4-byte
    // L2 hit instead of a 64-byte bucket read from HBM.
    const unsigned int *pre;               // [pre_mask + 1] or nullptr
    unsigned int pre_mask;
};
__host__ __device__ inline unsigned int exact_pre_word(unsigned long long h, unsigned int pre_mask) { return (unsigned int)(h >> 20) & pre_mask; }
__host__ __device__ inline unsigned int exact_pre_mask(unsigned long long h)
{
    return (1u << ((h >> 44) & 31u)) | (1u << ((h >> 49) & 31u)) | (1u << ((h >> 54) & 31u)) | (1u << ((h >> 59) & 31u));
}

struct ScanParams {
    const uint8_t *buf;
    unsigned long long nbytes;
    unsigned long long ntiles;
    // geometry of this batch (run-time copy of the Geom the scan kernel was launched with)
    int tile_bytes;                   // Geom::kTileBytes
    int nl_stride;                    // Geom::kMaxTileLines: u16 entries per tile in tile_nl
    int hit_words;                    // Geom::kHitWords: words per tile in tile_hit
    TileInfo *info;                   // [ntiles]
    unsigned short *tile_nl;          // [ntiles * nl_stride] tile-relative '\n' positions
    unsigned int *tile_hit;           // [ntiles * hit_words] line-hit bits (literal / k-mer / exact modes)
    unsigned long long *lbase;        // [ntiles] '\n' count before each tile, relative to its 4096-tile span
    unsigned long long *bsum;         // [ntiles / 4096 + 1] '\n' count before each span
    ScanCounters *ctr;
    unsigned int *hitbits;            // 1 bit per record
    unsigned long long rec_cap;
    const LiteralTable *lit;          // device copy
    const uint8_t *pat_bytes;         // device copy of pattern bytes
    int target;
    int mode;                         // kMode*: selects the scan kernel variant
    int match_enabled;                // 1: the scan kernel leaves line hits in tile_hit (literal modes);
                                      // 0: index/validate only (a second-stage field matcher uses the field table)
    int lit_unit;                     // host copy of lit->unit
    unsigned int one;                 // == 1 (lets the kernel issue adds as IMAD; see fma_add)
    int l2_evict_first;               // 1: the bulk copies of the batch carry an L2 evict-first hint (FXG_L2_HINT=0 turns it off)
    unsigned int *redo_list;          // [ntiles] tiles whose locally guessed line phase was wrong (normally none)
    unsigned int *redo_count;         // [1]
    int redo_unanchored;              // exact mode: tiles without a phase anchor go to the REDO launch too
    KmerTile kmer;                    // mode_is_kmer(mode)
    uint2 *cand;                      // [ntiles * cand_cap] window matcher: {window hash, tile-relative offset} per filter hit
    unsigned int *tile_ncand;         // [ntiles] entries of the tile's candidate list (| kCandOverflow: the list was too
                                      // short -- kmer_resolve_kernel then walks every probe position of the tile itself)
    int cand_cap;                     // list capacity per tile
    ExactDevice exact;                // mode == kModeExact
    unsigned long long *keyhash;      // [ntiles * nl_stride / 4] exact mode: key hash of the target line with local index 4j + first
    // per-pattern output routing (cfg.route_runs): first reported hit of every record, min-reduced; all ones = none.
    // EXACT: the pattern id.  MULTI: (end offset of the occurrence in its line << 40) | ((0xffff - pattern length)
    // << 24) | pattern id -- smallest end offset first, then the longest pattern (oracle/fxo_ac.h).
    unsigned long long *rec_first;    // [rec_cap] or nullptr
    // field table for the second-stage matchers (multi without k-mer index): target field of record k
    unsigned long long *fld_off;      // [rec_cap] or nullptr
    unsigned int *fld_len;            // [rec_cap]
    int header_whole;                 // name target, exact strategy: the field table holds the whole header line (after '@');
                                      // the look-up kernel ends the name at the first ' ' while it hashes (FieldSrc::stop_at_space)
};

// ---------------------------------------------------------------- field engine
// Byte-class compressed Aho-Corasick DFA.  One transition entry = next state | (high bit: the next state reports a
// pattern), 16 bits wide when the automaton has < 32768 states, else 32; the whole table is staged in shared
// memory when it fits (north_star: "transition table held in shared memory"), else it is walked in L2.  Per text
// byte the walk is ONE dependent look-up (the class of the next bytes does not depend on the state).
struct AcDevice {
    const uint8_t *cls;        // [256] byte -> class
    const void *next;          // [nstates * ncls] u16 or u32 entries
    const int *out;            // [nstates] longest pattern ending here or -1 (read when an entry's high bit is set)
    int ncls;
    int nstates;
    int wide;                  // 1: 32-bit entries
};
// Packed 2-bit k-mer index for pattern sets over {A,C,G,T} (north_star: "packed 2-bit k-mer hash for the ACGT
// alphabet").  Every pattern contributes its first `stride` k-mers (offsets 0..stride-1, stride = min pattern
// length - k + 1), so an occurrence is found by probing every stride-th text position only.  Probe = two bits of a
// 2^20-bit filter (128 KB) held in shared memory, then an open-addressing table in L2, then byte verification.
#ifndef FXG_KMER_FILTER_BITS
#define FXG_KMER_FILTER_BITS 20
#endif
constexpr int kKmerFilterBits = FXG_KMER_FILTER_BITS;
constexpr int kKmerFilterWords = 1 << (kKmerFilterBits - 5);
struct KmerDevice {
    const unsigned int *filter;            // [kKmerFilterWords]
    const unsigned int *ht_key;            // [nslots] packed k-mer
    const unsigned int *ht_val;            // [nslots] 0 = empty, else ((pattern id << 5) | offset) + 1
    const uint8_t *blob;                   // pattern bytes
    const unsigned int *off;               // [n+1]
    unsigned int ht_mask;                  // nslots - 1
    unsigned int kmask;                    // (1 << 2k) - 1
    int k, stride;
};
// code of a base: A 0, C 1, T 2, G 3 ((byte >> 1) & 3); identical on host and device
__host__ __device__ inline unsigned int fxg_base_code(unsigned int b) { return (b >> 1) & 3u; }
__host__ __device__ inline bool fxg_is_acgt(unsigned int b) { return b == 'A' || b == 'C' || b == 'G' || b == 'T'; }
__host__ __device__ inline unsigned int fxg_kmer_hash(unsigned int key) { return key * 0x9E3779B1u; }
__host__ __device__ inline unsigned int fxg_kmer_hash2(unsigned int key) { return key * 0xC2B2AE35u; }   // top bits used

// 64-bit hash of a whole-field key, identical on host and device: the key is read as little-endian 8-byte words
// (the last 1..7 bytes zero padded into one more word), each word is folded in with one multiply, the length goes
// into the final mix.  Word-wise so that the look-up kernel hashes 8 bytes per step (fields_exact_kernel).
constexpr unsigned long long kHashSeed = 0xcbf29ce484222325ull;
__host__ __device__ inline unsigned long long fxg_hash_word(unsigned long long h, unsigned long long w)
{
    h = (h ^ w) * 0x9e3779b97f4a7c15ull;
    return h ^ (h >> 32);
}
__host__ __device__ inline unsigned long long fxg_hash_finish(unsigned long long h, unsigned int n)
{
    h = (h ^ n) * 0xff51afd7ed558ccdull;
    h ^= h >> 33;
    return h ? h : 1ull;
}
__host__ __device__ inline unsigned long long fxg_hash_bytes(const uint8_t *p, unsigned int n)
{
    unsigned long long h = kHashSeed;
    for (unsigned int i = 0; i < n; i += 8) {
        unsigned long long w = 0;
        for (unsigned int b = 0; b < 8 && i + b < n; ++b) w |= (unsigned long long)p[i + b] << (8 * b);
        h = fxg_hash_word(h, w);
    }
    return fxg_hash_finish(h, n);
}

struct FieldSrc {
    const uint8_t *base;
    const unsigned int *off32;           // packed layout: [n+1] offsets, or
    const unsigned long long *off64;     // table layout: start offsets ...
    const unsigned int *len32;           // ... and lengths
    const unsigned long long *n_lines_dev; // raw path: field count = *n_lines_dev / 4 (capped by the launch's n)
    int stop_at_space;                   // fields are header lines: the key is the name, up to the first ' ' (exact kernel)
    __device__ __forceinline__ unsigned long long count(unsigned long long cap) const
    {
        if (!n_lines_dev) return cap;
        unsigned long long v = *n_lines_dev >> 2;
        return v < cap ? v : cap;
    }
    __device__ __forceinline__ void get(unsigned long long k, const uint8_t *&p, unsigned int &len) const
    {
        if (off32) { unsigned int a = off32[k]; p = base + a; len = off32[k + 1] - a; }
        else { p = base + off64[k]; len = len32[k]; }
    }
};

struct FieldOut {
    uint8_t *flags;            // 1 byte per field or nullptr
    int *pattern_id;           // first-hit pattern id per field or nullptr
    unsigned int *hitbits;     // 1 bit per field (raw path) or nullptr
    unsigned long long *n_hits;
};

struct LiteralFieldParams {
    const uint8_t *pat[2];
    int len[2];
    int n_patterns;
};


// launchers (fxg_scan.cu / fxg_fields.cu / fxg_synth.cu)
void scan_geometry(int mode, int filter_log2_words, int *tile_bytes, int *nl_stride, int *hit_words);
cudaError_t launch_scan_raw(const ScanParams &P, cudaStream_t stream, int *launches, cudaEvent_t after_scan);
cudaError_t launch_scan_kmer(const ScanParams &P, cudaStream_t stream, int dev);
cudaError_t launch_scan_kmer_redo(const ScanParams &P, cudaStream_t stream, int dev);
cudaError_t launch_compact(const unsigned int *hitbits, const ScanParams &P, const ScanCounters *ctr,
                           unsigned long long rec_cap, unsigned long long *block_sums, unsigned long long *out_counts,
                           fxg_run *runs, unsigned long long run_cap, int invert, int *run_pid, int per_record,
                           cudaStream_t stream, int *launches);
// mates (fx.cpp:58-77): per-record hit bits -> per-pair selection under the strategy's rule -> selected bits of both
// mates, ready for launch_compact with invert = 0.  split = 0: interleaved; else byte offset of the second file's part.
struct PairRule { int invert; int need_neither; int all_match; };
cudaError_t launch_pair_select(unsigned int *hitbits, unsigned int *pairsel, const ScanParams &P, unsigned long long rec_cap,
                               unsigned long long split, PairRule rule, cudaStream_t stream, int *launches);
cudaError_t launch_fields_literal(const FieldSrc &src, unsigned long long n, const LiteralFieldParams &L,
                                  const FieldOut &O, cudaStream_t st);
cudaError_t launch_fields_ac(const FieldSrc &src, unsigned long long n, const AcDevice &A, const FieldOut &O, cudaStream_t st);
cudaError_t launch_fields_kmer(const FieldSrc &src, unsigned long long n, const KmerDevice &K, const FieldOut &O, cudaStream_t st);
cudaError_t launch_fields_exact(const FieldSrc &src, unsigned long long n, const ExactDevice &E, const FieldOut &O, cudaStream_t st);
cudaError_t launch_synth(const fxg_synth_spec &s, uint8_t *d_out, cudaStream_t st);
cudaError_t scan_launch_config(const void *func, int threads, size_t smem, int dev, int *grid_cap);

#define FXG_CUDA_OK(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return fxg::set_cuda_error(ctx, e__, #expr); } while (0)

}  // namespace fxg
