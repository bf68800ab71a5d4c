// fxg_scan_kernel.cuh -- the fused per-tile scan kernel of the raw FASTQ path (Path A), as a template over the
// tile geometry and the matcher mode.  Instantiated in fxg_scan.cu (G24: literal / exact / index-only) and
// fxg_scan_kmer.cu (G32: fused packed k-mer matcher).
//
// Replaces, for canonical 4-line FASTQ, the reference's per-record loop
//   kstream::read (kseq.hpp:120-164)        ->  record boundaries
//   prepare_data (main.cpp:291-304)         ->  target field
//   strstr x1..2 (main.cpp:568-588)         ->  literal hit            (modes 1 / 8 / 9)
//   ssearch_scan (main.cpp:406,426)         ->  multi-pattern hit      (modes 17..20, ACGT pattern sets)
//   manager.find (main.cpp:315)             ->  whole-field look-up    (mode 30)
//   on_match (main.cpp:258-271)             ->  count / select
// with a position-centric kernel that has NO cross-tile dependency.  Every CTA takes one tile, pulls it into
// shared memory with one bulk async copy (TMA engine, mbarrier completion) and then
//   1. main pass, 16 bytes per lane per step, SWAR only: byte range check, per-chunk class summaries ("has a
//      byte < 0x41", "has a byte < 0x21") as ballot bit rows, the list of chunks that can hold a '\n', and the
//      matcher's filter (literal: fingerprint of aligned 8-byte units; k-mer: packed 2-bit windows against a
//      hashed bit filter in shared memory);
//   2. ordered table of the tile's '\n' positions (warp-local masks, ballot prefix);
//   3. phase anchor: the first "+" line among local lines 1..31 gives the tile's guessed line phase;
//   4. one thread per line, one line role per warp: record structure and character classes under the guessed
//      phase (tiles without anchor: flags for all four phases at once); exact mode: the target-role warps hash
//      their line's key from shared memory and probe the table;
//   5. filter candidates are verified byte by byte and recorded as "local line i has a match";
//   6. the tile publishes {count, violations + guess, '\n' table, line-hit bits}.
// The epilogue kernels (fxg_scan.cu) turn counts into global line numbers, check every guess against the true
// phase (wrong guess -> REDO launch), check the few lines that straddle tile edges and map line hits to record
// hits.  The canonical-FASTQ predicate is what makes kseq-parse-then-print the identity on
// the record bytes; a batch that violates it is reported (never silently mis-parsed) and the host re-routes it.
#pragma once
#include "fxg_common.cuh"
#include "fxg_exact.cuh"

namespace fxg {

constexpr int kPrefixThreads = 1024;
constexpr int kPrefixSpan = 4 * kPrefixThreads;

struct TileState {
    unsigned int cls[4];                  // generic path: line flags OR-ed per (local line index mod 4)
    unsigned int n_cand, viol;
};

template <class G> struct __align__(128) ScanSmem {
    // staged tile bytes, double buffered (bulk-copy destinations).  One more chunk behind each tile: the first 16
    // bytes of the next tile, where the k-mer windows that start in the tile's last chunk end.  The pad keeps the
    // second buffer 128-byte aligned.
    uint4 tile[2][G::kChunks + 8];
    unsigned short nlpos[G::kMaxTileLines + 16];      // tile-relative '\n' positions, in order (+ slack, see step 3)
    unsigned short wl[G::kWarps][G::kWarpChunks];     // per warp: its chunks that contain a byte < 0x21 ('\n' candidates)
    unsigned short wm[G::kWarps][G::kWarpChunks];     // per entry of wl: 16-bit mask of the bytes < 0x20 of that chunk
    unsigned int wtot[2][G::kWarps];                  // '\n' count of each warp's 2 KiB segment (by tile parity)
    unsigned int below[17];                           // interleaved-layout mask of the bytes < p of a chunk
    __align__(16) unsigned int sbits[G::kChunks / 32];              // bit c: chunk c has a byte < 0x41 (not a sequence letter)
    __align__(16) unsigned int qbits[G::kChunks / 32];              // bit c: chunk c has a byte < 0x21 (not a quality char)
    unsigned int linehit[G::kHitWords];               // bit i: local line i has a (phase-conditional) match
    unsigned long long mbar[2];                       // bulk-copy completion barriers
    // per-tile scalars, double buffered by tile parity: a tile's copy is reset by the thread that publishes it, behind
    // the tile's last barrier, and next used two tiles on (no extra barrier between "last read" and "first write")
    TileState st[2];
};

// literal modes: fingerprint table and candidate queue, placed behind ScanSmem in the dynamic allocation (the window
// matcher has its filter there instead)
struct __align__(16) LitSmem {
    unsigned long long keys[64];
    unsigned int fp[64];
    unsigned short masks[64];
    int cand_off[kCandCap];                           // tile-relative offset of a candidate unit
    unsigned short cand_mask[kCandCap];               // (pattern,shift) bits to verify
};

__device__ __forceinline__ void reset_state(TileState &T)
{
    T.n_cand = 0; T.viol = 0;
    T.cls[0] = T.cls[1] = T.cls[2] = T.cls[3] = 0;
}

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ unsigned int smem_u32(const void *p) { return (unsigned int)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *mbar, unsigned int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(mbar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *mbar, unsigned int bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mbar)), "r"(bytes) : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine), completion counted in bytes on the mbarrier
// The batch streams through L2 exactly once, so its lines are marked evict-first: the look-up tables of the matchers
// (window index, key table, DFA) and the per-tile tables the epilogue kernels read back stay resident in the 126 MB L2.
__device__ __forceinline__ unsigned long long l2_stream_policy(int evict_first)
{
    unsigned long long pol;
    if (evict_first) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned int bytes, unsigned long long *mbar, unsigned long long pol)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(mbar)), "l"(pol) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *mbar, unsigned int parity)
{
    asm volatile("{\n"
                 ".reg .pred P1;\n"
                 "LAB_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
                 "@P1 bra DONE;\n"
                 "bra LAB_WAIT;\n"
                 "DONE:\n"
                 "}" ::"r"(smem_u32(mbar)), "r"(parity) : "memory");
}

// ---------------------------------------------------------------- SWAR helpers (bytes < 0x80)
constexpr unsigned int kHi = 0x80808080u;
constexpr unsigned int kAdd20 = 0x60606060u;   // (w + kAdd20) has bit7 set per byte iff byte >= 0x20
constexpr unsigned int kAdd21 = 0x5f5f5f5fu;   // ... iff byte >= 0x21
constexpr unsigned int kAdd41 = 0x3f3f3f3fu;   // ... iff byte >= 0x41

// x + k issued as IMAD (fma pipe) instead of IADD3 (alu pipe): the main pass is bound by the
// alu pipe (LOP3/SHF/IADD3 share it at half rate), so the SWAR threshold adds go to the other
// pipe.  `one` is a kernel parameter equal to 1 that ptxas cannot fold away.
__device__ __forceinline__ unsigned int fma_add(unsigned int x, unsigned int one, unsigned int k)
{
    unsigned int r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(one), "r"(k));
    return r;
}

// Interleaved-layout mask: bit (8k + 7 - i) set iff byte k of word i (byte 4i+k of the chunk) is
// below the threshold encoded by `add`.  4 adds, 3 shifts, 4 LOP3 (bitwise selects).
__device__ __forceinline__ unsigned int il_mask(const uint4 &x, unsigned int add)
{
    unsigned int t = x.x + add;
    t = (t & 0x80808080u) | (((x.y + add) >> 1) & 0x40404040u);
    t = (t & 0xc0c0c0c0u) | (((x.z + add) >> 2) & 0x20202020u);
    t = (t & 0xe0e0e0e0u) | (((x.w + add) >> 3) & 0x10101010u);
    return ~t & 0xf0f0f0f0u;
}

__constant__ unsigned int kBelowTab[17] = {
    0x00000000u, 0x00000080u, 0x00008080u, 0x00808080u, 0x80808080u, 0x808080c0u, 0x8080c0c0u, 0x80c0c0c0u, 0xc0c0c0c0u,
    0xc0c0c0e0u, 0xc0c0e0e0u, 0xc0e0e0e0u, 0xe0e0e0e0u, 0xe0e0e0f0u, 0xe0e0f0f0u, 0xe0f0f0f0u, 0xf0f0f0f0u};

// 16-bit mask, bit b set iff byte b of the chunk is below the threshold encoded by `add`
__device__ __forceinline__ unsigned int lin_mask(const uint4 &x, unsigned int add)
{
    unsigned int n0 = ((~(x.x + add) & kHi) * 0x00204081u) >> 28;
    unsigned int n1 = ((~(x.y + add) & kHi) * 0x00204081u) >> 28;
    unsigned int n2 = ((~(x.z + add) & kHi) * 0x00204081u) >> 28;
    unsigned int n3 = ((~(x.w + add) & kHi) * 0x00204081u) >> 28;
    return n0 | (n1 << 4) | (n2 << 8) | (n3 << 12);
}

__device__ __forceinline__ void note_violation(TileState &T, unsigned int bit) { atomicOr(&T.viol, bit); }
// TileInfo::viol: bits 0..15 violations, 16..18 guessed phase + 1, bit 31 "some line of this tile has a hit"
// (tile_finalize_kernel reads the tile's hit words only then)
constexpr unsigned int kTileHasHit = 0x80000000u;
constexpr unsigned int kGuessMask = 0x7u;

// byte of the batch at global offset g: from the staged tile when inside it
template <class G>
__device__ __forceinline__ unsigned int text_byte(const unsigned char *tb, const ScanParams &P, unsigned long long base, long long g)
{
    long long rel = g - (long long)base;
    if (rel >= 0 && rel < G::kTileBytes) return tb[rel];
    return P.buf[g];
}

// '\n' count in tile bytes [0, rel): lower bound in the finished position table
template <class G> __device__ __forceinline__ unsigned int lines_before(const ScanSmem<G> &S, int n_tab, int rel)
{
    int lo = 0, hi = n_tab;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if ((int)S.nlpos[mid] < rel) lo = mid + 1; else hi = mid;
    }
    return (unsigned int)lo;
}

// any bit set in rows[lo, hi) (chunk indices) of both class bit rows
template <class G>
__device__ __forceinline__ void range_any2(const ScanSmem<G> &S, int lo, int hi, unsigned int &anyS, unsigned int &anyQ)
{
    anyS = 0; anyQ = 0;
    if (lo >= hi) return;
    const int wl = lo >> 5, wh = (hi - 1) >> 5;
    const unsigned int ml = 0xffffffffu << (lo & 31), mh = 0xffffffffu >> (31 - ((hi - 1) & 31));
    if (wl == wh) { anyS = S.sbits[wl] & ml & mh; anyQ = S.qbits[wl] & ml & mh; return; }
    anyS = (S.sbits[wl] & ml) | (S.sbits[wh] & mh);
    anyQ = (S.qbits[wl] & ml) | (S.qbits[wh] & mh);
    for (int w = wl + 1; w < wh; ++w) { anyS |= S.sbits[w]; anyQ |= S.qbits[w]; }
}

// ---------------------------------------------------------------- per-line checks (phase-free)
// Flags of local line i (the line that ends at nlpos[i]) and class flags of the head of
// line i+1 (the bytes after that '\n' inside the same 16-byte chunk).
constexpr unsigned int kFA = 1, kFB = 2, kFC = 4, kFD = 8, kFS = 16, kFQ = 32, kFSP = 64, kFCTL = 128;
//  kFA  not a header line: empty or first byte != '@'          (FXG_V_HEADER_AT)
//  kFSP ends with ' ' (separator would not round-trip)          (FXG_V_HEADER_SPACE)
//  kFB  empty line (as a sequence: FXG_V_EMPTY_SEQ)
//  kFC  not exactly "+"                                         (FXG_V_PLUS_LINE)
//  kFD  length differs from the line two above                  (FXG_V_QUAL_LEN)
//  kFS  has a byte outside [0x41,0x7e]                          (FXG_V_SEQ_CLASS)
//  kFQ  has a byte outside [0x21,0x7e]                          (FXG_V_QUAL_CLASS)
//  kFCTL ends with a control byte other than '\n' (any role)    (FXG_V_BYTE_RANGE)
template <class G>
__device__ __forceinline__ void line_flags(const ScanSmem<G> &S, const uint4 *tile, int i, int n_tab, unsigned int &own,
                                           unsigned int &nxt)
{
    const unsigned char *tb = reinterpret_cast<const unsigned char *>(tile);
    const int pos = S.nlpos[i];
    const bool ctl = tb[pos] != '\n';             // a control byte posing as a line end
    const int prev = i > 0 ? (int)S.nlpos[i - 1] : -1;
    const int next = i + 1 < n_tab ? (int)S.nlpos[i + 1] : G::kTileBytes;
    const int c = pos >> 4, cb = c << 4, cp = prev >> 4;
    const uint4 x = tile[c];
    const unsigned int s16 = il_mask(x, kAdd41), q16 = il_mask(x, kAdd21);
    const int pb = pos - cb;
    const int lo = prev + 1 > cb ? prev + 1 - cb : 0;
    const unsigned int r_tail = S.below[pb] & ~S.below[lo];
    const int hi = next - cb < 16 ? next - cb : 16;
    const unsigned int r_head = S.below[hi] & ~S.below[pb + 1];
    unsigned int ms, mq;
    range_any2(S, cp + 1, c, ms, mq);
    own = (((s16 & r_tail) | ms) ? kFS : 0u) | (((q16 & r_tail) | mq) ? kFQ : 0u);
    nxt = ((s16 & r_head) ? kFS : 0u) | ((q16 & r_head) ? kFQ : 0u);
    if (ctl) own |= kFCTL;
    if (i >= 1) {                              // line starts inside the tile: structure is local
        const int llen = pos - prev - 1;
        const unsigned int fb = tb[prev + 1], lb = tb[pos - 1];
        if (llen < 1) own |= kFA | kFB | kFC;
        else {
            if (fb != '@') own |= kFA;
            if (lb == ' ') own |= kFSP;
            if (llen != 1 || fb != '+') own |= kFC;
        }
        if (i >= 3 && llen != (int)S.nlpos[i - 2] - (int)S.nlpos[i - 3] - 1) own |= kFD;
    }
}

// FXG_V_* bits of a line with flags f if its role is `role` (0 header, 1 sequence, 2 plus, 3 quality)
__host__ __device__ inline unsigned int role_kinds(unsigned int f, int role)
{
    if (role == 0) return ((f & kFA) ? FXG_V_HEADER_AT : 0u) | ((f & kFSP) ? FXG_V_HEADER_SPACE : 0u);
    if (role == 1) return ((f & kFB) ? FXG_V_EMPTY_SEQ : 0u) | ((f & kFS) ? FXG_V_SEQ_CLASS : 0u);
    if (role == 2) return (f & kFC) ? FXG_V_PLUS_LINE : 0u;
    return ((f & kFD) ? FXG_V_QUAL_LEN : 0u) | ((f & kFQ) ? FXG_V_QUAL_CLASS : 0u);
}
// line role the target field lives on
__host__ __device__ inline int target_role(int target) { return target == FXG_TARGET_SEQ ? 1 : (target == FXG_TARGET_QUAL ? 3 : 0); }

// ---------------------------------------------------------------- verification of one literal candidate
// A verified occurrence is recorded on its local line; whether that line has the target's
// role (sequence / quality / header) is decided by tile_finalize_kernel.  Header targets add
// the geometric condition (inside the name / inside the comment), which is phase-free.
template <class G>
__device__ __noinline__ void verify_candidate(ScanSmem<G> &S, TileState &T, const unsigned char *tb, const ScanParams &P,
                                              unsigned long long base, int n_tab, int unit_rel, unsigned int bits)
{
    const LiteralTable *lit = P.lit;
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
            if (text_byte<G>(tb, P, base, p + k) != pb[k]) { ok = false; break; }
        }
        if (!ok) continue;
        // local line: number of '\n' before p inside the tile (an occurrence that starts in the
        // previous tile lies on this tile's line 0)
        int rel = (int)(p - (long long)base);
        unsigned int loc = rel > 0 ? lines_before(S, n_tab, rel) : 0u;
        bool hit = true;
        if (P.target == FXG_TARGET_NAME || P.target == FXG_TARGET_COMMENT) {
            // header line: name = (ls, first ' ' or end), comment = after that ' '
            long long ls;
            if (loc >= 1 && loc <= (unsigned int)n_tab) ls = (long long)base + S.nlpos[loc - 1] + 1;
            else {
                long long q = (p < (long long)base ? p : (long long)base) - 1;
                int guard = 1 << 20;
                while (q >= 0 && P.buf[q] != '\n' && --guard) --q;
                if (!guard) { note_violation(T, FXG_V_DENSE_LINES); continue; }
                ls = q + 1;
            }
            long long q = ls + 1, sp = -1;
            for (;; ++q) {
                if ((unsigned long long)q >= P.nbytes) break;
                unsigned int ch = text_byte<G>(tb, P, base, q);
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
        if (hit) { atomicOr(&S.linehit[loc >> 5], 1u << (loc & 31)); atomicOr(&T.viol, kTileHasHit); }
    }
}

__device__ __forceinline__ void push_candidate(LitSmem &LS, TileState &T, int unit_rel, unsigned int bits)
{
    unsigned int slot = atomicAdd(&T.n_cand, 1u);
    if (slot < kCandCap) { LS.cand_off[slot] = unit_rel; LS.cand_mask[slot] = (unsigned short)bits; }
}

// generic (short-pattern) literal filter: units of 4, 2 or 1 bytes compared against every key
static __device__ __noinline__ void filter_small_units(LitSmem &LS, TileState &T, unsigned int x0, unsigned int x1, unsigned int x2,
                                                unsigned int x3, int off, int U, int nkeys)
{
    for (int u = 0; u < 16; u += U) {
        const int wi = u >> 2;
        unsigned int word = wi == 0 ? x0 : (wi == 1 ? x1 : (wi == 2 ? x2 : x3));
        unsigned int val = U == 4 ? word : (U == 2 ? ((word >> ((u & 2) * 8)) & 0xffffu) : ((word >> ((u & 3) * 8)) & 0xffu));
        for (int k = 0; k < nkeys; ++k)
            if ((unsigned int)LS.keys[k] == val) push_candidate(LS, T, off + u, LS.masks[k]);
    }
}

// ---------------------------------------------------------------- exact mode: key of one whole line of the tile
// Line = tile bytes [a, e).  The scan kernel only HASHES the target field of its lines (shared-memory reads, no
// dependent global load inside the lock-step tile loop) and leaves the 64-bit hashes in a per-tile list; the table
// look-ups are done afterwards by exact_resolve_kernel with one thread per record, where their latency is hidden by
// parallelism instead of stalling a tile.
__device__ __forceinline__ unsigned long long exact_line_hash_tile(const ScanParams &P, const unsigned char *tb, int a, int e)
{
    int ks = a, ke = e;
    bool stop = false;
    if (P.target == FXG_TARGET_NAME) { ks = a + 1; stop = true; }            // name: after '@' up to the first ' '
    else if (P.target == FXG_TARGET_COMMENT) {                               // comment: after the first ' '
        int sp = e;
        for (int q = a + 1; q < e; ++q) if (tb[q] == ' ') { sp = q + 1; break; }
        ks = sp;
    }
    if (ks > ke) ks = ke;                                                    // empty key (no match unless "" is a key)
    TileWords ld;
    ld.w = reinterpret_cast<const unsigned int *>(tb);
    unsigned int keylen;
    return exact_hash(ld, (long long)ks, (unsigned int)(ke - ks), stop, keylen);
}
// the same over global memory: line = batch bytes [a, e)
__device__ __forceinline__ bool exact_line_global(const ScanParams &P, long long a, long long e, int &pid)
{
    long long ks = a, ke = e;
    bool stop = false;
    if (P.target == FXG_TARGET_NAME) { ks = a + 1; stop = true; }
    else if (P.target == FXG_TARGET_COMMENT) {
        long long sp = e;
        for (long long q = a + 1; q < e; ++q) if (P.buf[q] == ' ') { sp = q + 1; break; }
        ks = sp;
    }
    if (ks > ke) ks = ke;
    return exact_lookup_global(P.exact, P.buf + ks, (unsigned int)(ke - ks), stop, pid);
}

// ---------------------------------------------------------------- class check of a byte range
// any bit set in row[lo, hi) (chunk indices)
__device__ __forceinline__ unsigned int row_any(const unsigned int *row, int lo, int hi)
{
    if (lo >= hi) return 0u;
    const int wl = lo >> 5, wh = (hi - 1) >> 5;
    const unsigned int ml = 0xffffffffu << (lo & 31), mh = 0xffffffffu >> (31 - ((hi - 1) & 31));
    if (wl == wh) return row[wl] & ml & mh;
    unsigned int any = (row[wl] & ml) | (row[wh] & mh);
    for (int w = wl + 1; w < wh; ++w) any |= row[w];
    return any;
}
// non-zero iff tile bytes [a, b) hold a byte below 0x41 (seq) / 0x21 (!seq)
template <class G>
__device__ __forceinline__ unsigned int class_bad(const ScanSmem<G> &S, const uint4 *tile, int a, int b, bool seq)
{
    if (b <= a) return 0u;
    const unsigned int add = seq ? kAdd41 : kAdd21;
    const unsigned int *row = seq ? S.sbits : S.qbits;
    const int ca = a >> 4, cb = (b - 1) >> 4;
    const unsigned int ma = il_mask(tile[ca], add) & ~S.below[a - (ca << 4)];
    if (ca == cb) return ma & S.below[b - (ca << 4)];
    const unsigned int mb = il_mask(tile[cb], add) & S.below[b - (cb << 4)];
    return ma | mb | row_any(row, ca + 1, cb);
}

// ---------------------------------------------------------------- kernel 1: per-tile scan
// Persistent CTAs: tile t = blockIdx.x, += gridDim.x.  The next tile is prefetched into the other shared-memory
// buffer while the current one is processed.
// MODE = kMode* (fxg_common.cuh).  REDO = second launch over the (normally empty) list of tiles whose locally
// guessed line phase turned out wrong (exact mode: also the tiles without anchor): those are re-evaluated with the
// phase-free per-line path, exact-mode hashes under the true phase.
template <class G, int MODE, bool REDO>
__global__ void __launch_bounds__(G::kCtaThreads, G::kMinCtas) scan_tiles_kernel(const ScanParams P)
{
    constexpr int kTileBytes = G::kTileBytes, kThreads = G::kThreads, kChunks = G::kChunks, kWarps = G::kWarps;
    constexpr int kWarpChunks = G::kWarpChunks, kChunksPerThread = G::kChunksPerThread, kMaxTileLines = G::kMaxTileLines;
    constexpr int kHitWords = G::kHitWords;
    constexpr bool LIT8 = MODE == kModeLit8 || MODE == kModeLit9;
    constexpr bool LIT = mode_is_literal(MODE);
    constexpr bool KMER = mode_is_kmer(MODE);
    constexpr bool EXACT = MODE == kModeExact;
    constexpr int SB = KMER ? (1 << (MODE - kModeKmer)) : 16;          // probe stride in bytes
    constexpr int NPROBE = 16 / SB;
    constexpr bool HITS = LIT && !REDO;                                // this launch produces line-hit words
    constexpr int kGroups = G::kGroups;
    constexpr bool ROLEMIX = EXACT || (KMER && G::kGroups == 1);        // see step 4
    constexpr bool KDEFER = KMER && G::kGroups == 1;                   // window matcher: filter test deferred by one step
    typedef ScanSmem<G> Smem;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // a CTA is kGroups independent groups of kThreads threads: everything below -- tid, warp, S, the tile loop, the
    // barriers -- is per group; only the matcher's filter behind the groups' blocks is shared
    const int gid = kGroups > 1 ? (int)threadIdx.x / kThreads : 0;
    Smem &S = reinterpret_cast<Smem *>(smem_raw)[gid];
    unsigned int *filt = reinterpret_cast<unsigned int *>(smem_raw + kGroups * sizeof(Smem));   // window filter (dynamic tail), or
    LitSmem &LS = *reinterpret_cast<LitSmem *>(smem_raw + kGroups * sizeof(Smem));              // the literal tables
    const int tid = (int)threadIdx.x - gid * kThreads, lane = tid & 31, warp = tid >> 5;
    const LiteralTable *lit = P.lit;
    const unsigned long long n_work = REDO ? (unsigned long long)*P.redo_count : P.ntiles;
    if ((unsigned long long)blockIdx.x * kGroups >= n_work) return;
    const unsigned long long k_first = (unsigned long long)blockIdx.x * kGroups + gid, k_step = (unsigned long long)gridDim.x * kGroups;
#define FXG_GROUP_SYNC()                                                                                                 \
    do {                                                                                                                \
        if (kGroups == 1) __syncthreads();                                                                              \
        else asm volatile("bar.sync %0, %1;" ::"r"(gid + 1), "n"(kThreads) : "memory");                                 \
    } while (0)
#define FXG_TILE_OF(k) (REDO ? (unsigned long long)P.redo_list[(k)] : (unsigned long long)(k))
    // issue the bulk copy of tile `tt` into buffer `bf` (thread 0); in k-mer mode with the 16 bytes that follow it
    constexpr unsigned int kFetch = (unsigned int)kTileBytes + ((KMER && SB < 16) ? 16u : 0u);
#define FXG_ISSUE_TILE(tt, bf)                                                                                          \
    do {                                                                                                                \
        const unsigned long long b__ = (tt) * (unsigned long long)kTileBytes;                                           \
        const unsigned long long rem__ = P.nbytes - b__;                                                                \
        const unsigned int bulk__ = (unsigned int)(rem__ < (unsigned long long)kFetch ? rem__ : (unsigned long long)kFetch) & ~15u; \
        if (bulk__) {                                                                                                   \
            mbar_expect_tx(&S.mbar[bf], bulk__);                                                                        \
            bulk_g2s(&S.tile[bf][0], P.buf + b__, bulk__, &S.mbar[bf], l2pol);                                                 \
        }                                                                                                               \
    } while (0)

    // ---- once per CTA: tables, barriers, first tile in flight
    const unsigned long long l2pol = l2_stream_policy(P.l2_evict_first);
    if (tid == 0) {
        mbar_init(&S.mbar[0], 1);
        mbar_init(&S.mbar[1], 1);
        if (k_first < n_work) {
            const unsigned long long t0 = FXG_TILE_OF(k_first);
            FXG_ISSUE_TILE(t0, 0);
        }
        reset_state(S.st[0]);
        reset_state(S.st[1]);
    }
    if (tid >= 128 && tid < 145) S.below[tid - 128] = kBelowTab[tid - 128];
    if (LIT) for (int w = tid; w < kHitWords; w += kThreads) S.linehit[w] = 0;
    int lit_unit = 0, lit_nkeys = 0;
    unsigned int lit_mul1 = 0, lit_mul2 = 0, lit_shift = 26;
    if (LIT) {
        lit_unit = lit->unit; lit_nkeys = lit->n_keys; lit_mul1 = lit->mul1; lit_mul2 = lit->mul2; lit_shift = (unsigned int)lit->shift;
        if (tid >= 64 && tid < 128) {
            const int k = tid - 64;
            LS.keys[k] = lit->keys[k]; LS.fp[k] = lit->fp[k]; LS.masks[k] = lit->masks[k];
        }
    }
    unsigned int wmask2 = 0, wmask3 = 0, fshift = 0;
    if (KMER && !REDO) {
        wmask2 = P.kmer.mask2; wmask3 = P.kmer.mask3; fshift = 32u - (unsigned int)P.kmer.filter_log2_words;
        const int nw = 1 << P.kmer.filter_log2_words;
        const uint4 *src = reinterpret_cast<const uint4 *>(P.kmer.filter);
        uint4 *dst = reinterpret_cast<uint4 *>(filt);
        for (int w = (int)threadIdx.x; w < nw / 4; w += G::kCtaThreads) dst[w] = __ldg(src + w);
    }
    const unsigned int one = P.one;
    const unsigned int lt_mask = (1u << lane) - 1u;
    unsigned short *my_list = S.wl[warp];
    unsigned short *my_mask = S.wm[warp];
    const int trole = target_role(P.target);
    __syncthreads();                     // the whole CTA: the filter is staged by every thread

    unsigned int it = 0;
    for (unsigned long long k = k_first; k < n_work; k += k_step, ++it) {
        const unsigned long long t = FXG_TILE_OF(k);
        const int cur = (int)(it & 1u);
        const uint4 *tile = S.tile[cur];
        TileState &T = S.st[cur];
        const unsigned char *tb = reinterpret_cast<const unsigned char *>(tile);
        const unsigned long long base = t * (unsigned long long)kTileBytes;
        const unsigned long long remain = P.nbytes - base;
        const int tile_len = remain < (unsigned long long)kTileBytes ? (int)remain : kTileBytes;
        const int bulk = tile_len & ~15;

        // ---- prefetch the next tile of this CTA into the other buffer (free since the end of the
        //      previous iteration)
        if (tid == 0) {
            const unsigned long long kn = k + k_step;
            if (kn < n_work) {
                const unsigned long long tn = FXG_TILE_OF(kn);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                FXG_ISSUE_TILE(tn, cur ^ 1);
            }
        }
        if (tile_len < kTileBytes) {             // ragged end of the batch: patch by hand, neutral filler
            for (int b = bulk + tid * 16; b < kTileBytes; b += kThreads * 16) {
                unsigned int w[4] = {0x78787878u, 0x78787878u, 0x78787878u, 0x78787878u};
                for (int q = 0; q < 16 && b + q < tile_len; ++q) {
                    unsigned int by = P.buf[base + b + q];
                    w[q >> 2] = (w[q >> 2] & ~(0xffu << ((q & 3) * 8))) | (by << ((q & 3) * 8));
                }
                S.tile[cur][b >> 4] = make_uint4(w[0], w[1], w[2], w[3]);
            }
            FXG_GROUP_SYNC();
        }
        if (bulk) mbar_wait(&S.mbar[cur], (it >> 1) & 1u);

        // ---- 1. main pass: 16 bytes per lane per step, SWAR only.  Per chunk: byte range, "has a
        //         byte < 0x41" / "< 0x21" ballot rows, list of the chunks with a byte < 0x21 (the only
        //         ones that can hold a '\n'), matcher filter.
        unsigned int wcount = 0;
        {
            unsigned int hi_acc = 0;
            unsigned int pr0 = 0, pr1 = 0, pf0 = 1, pf1 = 1;
            unsigned int khs[NPROBE], kfw[NPROBE];
            auto kmer_eval = [&](const unsigned int (&eh)[NPROBE], const unsigned int (&ef)[NPROBE], int ec) {
                unsigned int hits = 0;
#pragma unroll
                for (int q = 0; q < NPROBE; ++q) {
                    const unsigned int h = eh[q];
                    const unsigned int m = __funnelshift_l(0x00010001u, 0x00010001u, h >> kFiltShiftR) | __funnelshift_l(0x00000101u, 0x00000101u, h >> kFiltShiftS);
                    if ((m & ~ef[q]) == 0u) hits |= 1u << q;
                }
                if (__any_sync(0xffffffffu, hits != 0u)) {
                    unsigned int bq[NPROBE], tot = 0;
#pragma unroll
                    for (int q = 0; q < NPROBE; ++q) { bq[q] = __ballot_sync(0xffffffffu, (hits >> q) & 1u); tot += __popc(bq[q]); }
                    unsigned int slot = 0;
                    if (lane == 0) slot = atomicAdd(&T.n_cand, tot);
                    slot = __shfl_sync(0xffffffffu, slot, 0);
                    uint2 *list = P.cand + t * (unsigned long long)P.cand_cap;
#pragma unroll
                    for (int q = 0; q < NPROBE; ++q) {
                        const unsigned int mine = slot + __popc(bq[q] & lt_mask);
                        if (((hits >> q) & 1u) && mine < (unsigned int)P.cand_cap)
                            list[mine] = make_uint2(eh[q], (unsigned int)((ec << 4) + q * SB));
                        slot += __popc(bq[q]);
                    }
                }
            };
#pragma unroll
            for (int j = 0; j < kChunksPerThread; ++j) {
                const int c = warp * kWarpChunks + j * 32 + lane;
                const uint4 x = tile[c];
                // bytes >= 0x7f anywhere: bit 7 of (byte + 1).  Only a byte 0xff escapes this sum (it wraps to 0 and
                // carries into its neighbour) -- and every 0xff is caught on another way: + 0x5f and + 0x60 wrap as
                // well, so it is listed as a '\n' candidate, enters the position table as a byte < 0x20, and the
                // per-line check "the byte at a line end is '\n'" (step 4) raises FXG_V_BYTE_RANGE for it.
                hi_acc |= fma_add(x.x, one, 0x01010101u) | fma_add(x.y, one, 0x01010101u);
                hi_acc |= fma_add(x.z, one, 0x01010101u) | fma_add(x.w, one, 0x01010101u);
                const bool s_any = ((fma_add(x.x, one, kAdd41) & fma_add(x.y, one, kAdd41) & fma_add(x.z, one, kAdd41) &
                                     fma_add(x.w, one, kAdd41)) & kHi) != kHi;
                const bool q_any = ((fma_add(x.x, one, kAdd21) & fma_add(x.y, one, kAdd21) & fma_add(x.z, one, kAdd21) &
                                     fma_add(x.w, one, kAdd21)) & kHi) != kHi;
                const unsigned int sb = __ballot_sync(0xffffffffu, s_any);
                const unsigned int qb = __ballot_sync(0xffffffffu, q_any);
                S.sbits[c >> 5] = sb;                                 // warp-uniform address and value: every lane stores
                S.qbits[c >> 5] = qb;
                if (q_any) my_list[wcount + __popc(qb & lt_mask)] = (unsigned short)c;
                wcount += __popc(qb);
                if (LIT8 && !REDO) {
                    const unsigned int r0 = MODE == kModeLit9 ? x.x + x.y * lit_mul2 : x.x * lit_mul1 + x.y * lit_mul2;
                    const unsigned int r1 = MODE == kModeLit9 ? x.z + x.w * lit_mul2 : x.z * lit_mul1 + x.w * lit_mul2;
                    // the table words of this step are compared one step later: their latency hides behind the next
                    // step's class sums instead of stalling the warp in front of the vote
                    const unsigned int f0 = LS.fp[r0 >> lit_shift], f1 = LS.fp[r1 >> lit_shift];
                    if (j > 0) {
                        const bool c0 = pf0 == pr0, c1 = pf1 == pr1;
                        if (__any_sync(0xffffffffu, c0 || c1)) {
                            if (c0) push_candidate(LS, T, (c - 32) << 4, LS.masks[pr0 >> lit_shift]);
                            if (c1) push_candidate(LS, T, ((c - 32) << 4) + 8, LS.masks[pr1 >> lit_shift]);
                        }
                    }
                    pr0 = r0; pr1 = r1; pf0 = f0; pf1 = f1;
                    if (j == kChunksPerThread - 1) {
                        const bool c0 = pf0 == pr0, c1 = pf1 == pr1;
                        if (__any_sync(0xffffffffu, c0 || c1)) {
                            if (c0) push_candidate(LS, T, c << 4, LS.masks[pr0 >> lit_shift]);
                            if (c1) push_candidate(LS, T, (c << 4) + 8, LS.masks[pr1 >> lit_shift]);
                        }
                    }
                } else if (MODE == kModeLitSmall && !REDO) {
                    if ((c << 4) < tile_len) filter_small_units(LS, T, x.x, x.y, x.z, x.w, c << 4, lit_unit, lit_nkeys);
                } else if (KMER && !REDO) {
                    // windows of k bytes that start at byte 0, SB, 2 SB, ... of the lane's chunk; what lies past the
                    // chunk comes from the head of the next one (same shared-memory row; behind the tile's last chunk
                    // lie the 16 bytes staged after the tile)
                    unsigned int wd[7] = {x.x, x.y, x.z, x.w, 0u, 0u, 0u};
                    if (SB == 8) {
                        const uint2 n = *reinterpret_cast<const uint2 *>(&tile[c + 1]);
                        wd[4] = n.x; wd[5] = n.y;
                    } else if (SB == 4) {
                        const uint4 n = tile[c + 1];
                        wd[4] = n.x; wd[5] = n.y; wd[6] = n.z;
                    }
                    // small filter (two CTAs per SM): the filter words fetched in this step are tested one step later,
                    // their latency hides behind the next step's sums and hashes (c5c + 1 %; the two-group kernel of
                    // the large filter lost 1 % with it and tests at once)
                    if (KDEFER) {
                    unsigned int hs[NPROBE], fws[NPROBE];
#pragma unroll
                    for (int q = 0; q < NPROBE; ++q) {
                        const int o = q * (SB / 4);                   // first word of the window
                        hs[q] = window_hash(wd[o], wd[o + 1], wd[o + 2] & wmask2, wd[o + 3] & wmask3);
                        fws[q] = filt[hs[q] >> fshift];
                    }
                    if (j > 0) kmer_eval(khs, kfw, c - 32);
#pragma unroll
                    for (int q = 0; q < NPROBE; ++q) { khs[q] = hs[q]; kfw[q] = fws[q]; }
                    if (j == kChunksPerThread - 1) kmer_eval(khs, kfw, c);
                    } else {
                    unsigned int hs[NPROBE], hits = 0;
#pragma unroll
                    for (int q = 0; q < NPROBE; ++q) {
                        const int o = q * (SB / 4);                   // first word of the window
                        const unsigned int h = window_hash(wd[o], wd[o + 1], wd[o + 2] & wmask2, wd[o + 3] & wmask3);
                        hs[q] = h;
                        const unsigned int fw = filt[h >> fshift];
                        const unsigned int m = __funnelshift_l(0x00010001u, 0x00010001u, h >> kFiltShiftR) | __funnelshift_l(0x00000101u, 0x00000101u, h >> kFiltShiftS);
                        if ((m & ~fw) == 0u) hits |= 1u << q;
                    }
                    // filter hits go to the tile's list for kmer_resolve_kernel: one shared-memory atomic per warp
                    // and step (the slots of the step are dealt out by ballot rank); one vote when nothing hit
                    if (__any_sync(0xffffffffu, hits != 0u)) {
                        unsigned int bq[NPROBE], tot = 0;
#pragma unroll
                        for (int q = 0; q < NPROBE; ++q) { bq[q] = __ballot_sync(0xffffffffu, (hits >> q) & 1u); tot += __popc(bq[q]); }
                        unsigned int slot = 0;
                        if (lane == 0) slot = atomicAdd(&T.n_cand, tot);
                        slot = __shfl_sync(0xffffffffu, slot, 0);
                        uint2 *list = P.cand + t * (unsigned long long)P.cand_cap;
#pragma unroll
                        for (int q = 0; q < NPROBE; ++q) {
                            const unsigned int mine = slot + __popc(bq[q] & lt_mask);
                            if (((hits >> q) & 1u) && mine < (unsigned int)P.cand_cap)
                                list[mine] = make_uint2(hs[q], (unsigned int)((c << 4) + q * SB));
                            slot += __popc(bq[q]);
                        }
                    }
                    }
                }
            }
            if (__any_sync(0xffffffffu, (hi_acc & kHi) != 0) && lane == 0) note_violation(T, FXG_V_BYTE_RANGE);
        }

        // ---- 2. (warp local) masks of the bytes < 0x20 of the listed chunks, their count in the warp's segment
        //         (that every such byte is a '\n' is checked per line in step 4)
        {
            __syncwarp();
            unsigned int cnt = 0;
            for (unsigned int q = lane; q < wcount; q += 32) {
                const unsigned int m16 = lin_mask(tile[my_list[q]], kAdd20);
                my_mask[q] = (unsigned short)m16;
                cnt += __popc(m16);
            }
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (lane == 0) S.wtot[cur][warp] = cnt;
        }
        FXG_GROUP_SYNC();

        // ---- 3. (warp local) ordered '\n' position table
        unsigned int total;
        {
            const unsigned int wt = lane < kWarps ? S.wtot[cur][lane] : 0u;
            total = __reduce_add_sync(0xffffffffu, wt);
            unsigned int idx0 = __reduce_add_sync(0xffffffffu, lane < warp ? wt : 0u);
            for (unsigned int q0 = 0; q0 < wcount; q0 += 32) {
                const unsigned int q = q0 + lane;
                unsigned int m16 = q < wcount ? (unsigned int)my_mask[q] : 0u;
                const int cb = q < wcount ? ((int)my_list[q] << 4) - 1 : 0;
                const unsigned int n = __popc(m16);
                unsigned int idx, round_total;
                // exclusive prefix of the per-lane counts: two ballots while no chunk holds more than two '\n'
                const unsigned int b1 = __ballot_sync(0xffffffffu, n >= 1u), b2 = __ballot_sync(0xffffffffu, n >= 2u);
                if (!__any_sync(0xffffffffu, n >= 3u)) {
                    idx = idx0 + __popc(b1 & lt_mask) + __popc(b2 & lt_mask);
                    round_total = __popc(b1) + __popc(b2);
                } else {
                    unsigned int inc = n;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d);
                        if (lane >= d) inc += o;
                    }
                    idx = idx0 + inc - n;
                    round_total = __shfl_sync(0xffffffffu, inc, 31);
                }
                idx0 += round_total;
                // the table has 16 entries of slack: clamping the first index keeps every store inside it
                unsigned short *dst = &S.nlpos[idx < (unsigned int)kMaxTileLines ? idx : (unsigned int)kMaxTileLines];
                if (m16) {
                    dst[0] = (unsigned short)(cb + __ffs(m16));
                    m16 &= m16 - 1;
                    if (m16) {
                        dst[1] = (unsigned short)(cb + __ffs(m16));
                        m16 &= m16 - 1;
                        for (int k2 = 2; m16; ++k2) {
                            dst[k2] = (unsigned short)(cb + __ffs(m16));
                            m16 &= m16 - 1;
                        }
                    }
                }
            }
        }
        const int n_tab = total < (unsigned int)kMaxTileLines ? (int)total : kMaxTileLines;
        if (total > (unsigned int)kMaxTileLines && tid == 0) note_violation(T, FXG_V_DENSE_LINES);
        FXG_GROUP_SYNC();

        // ---- 3b. phase anchor: the first line among local lines 1..31 that is exactly "+" (every warp looks
        //          for itself -- cheaper than a block-wide minimum and one more barrier)
        unsigned int plus_idx = 0xffffffffu;
        if (!REDO) {
            bool is_plus = false;
            if (lane >= 1 && lane < n_tab) {
                const int e = S.nlpos[lane];
                is_plus = e - (int)S.nlpos[lane - 1] == 2 && tb[e - 1] == '+';
            }
            const unsigned int pb = __ballot_sync(0xffffffffu, is_plus);
            if (pb) plus_idx = (unsigned int)__ffs(pb) - 1u;
        }

        // ---- 4. record structure and character classes
        if (plus_idx != 0xffffffffu) {
            // fast path: the phase is taken from the "+" line (role 2): local line i has role (i + r) & 3.
            // tile_finalize_kernel checks the guess against the global line number and sends the tile to
            // the REDO launch if it was wrong (only a 1-base read whose quality line is "+" can cause that).
            // One thread per LINE, one line role per warp (role = warp & 3), so no lane of a warp diverges from
            // its neighbours and the longest dependent chain of this step is one class check.
            const int r = (int)((2u - plus_idx) & 3u);
            // 0 header, 1 sequence, 2 plus, 3 quality.  Warp w issues on scheduler w % 4, so with role = w % 4 each
            // role's work always lands on one scheduler; ROLEMIX rotates the roles from one group of four warps to
            // the next (measured: exact + 4.4 % -- the key hashing no longer sits on one scheduler --, small-filter
            // window matcher + 1.1 %, literal - 0.4 %, two-group window matcher - 5 %: those keep w % 4)
            const int role = ROLEMIX ? (warp + (warp >> 2)) & 3 : warp & 3;
            const int first = (role - r) & 3;                 // smallest local line index with that role
            unsigned int viol = 0;
            for (int i = first + 4 * ((warp >> 2) * 32 + lane); i <= n_tab; i += 4 * (kWarps / 4) * 32) {
                const int a = i >= 1 ? (int)S.nlpos[i - 1] + 1 : 0;
                const int e = i < n_tab ? (int)S.nlpos[i] : kTileBytes;
                const bool whole = i >= 1 && i < n_tab;       // line starts and ends inside the tile
                if (i < n_tab && tb[e] != '\n') viol |= FXG_V_BYTE_RANGE;   // a control byte posing as a line end
                if (role == 0) {
                    if (whole) {
                        if (e - a < 1 || tb[a] != '@') viol |= FXG_V_HEADER_AT;
                        else if (tb[e - 1] == ' ') viol |= FXG_V_HEADER_SPACE;
                    }
                } else if (role == 2) {
                    if (whole && (e - a != 1 || tb[a] != '+')) viol |= FXG_V_PLUS_LINE;
                } else {
                    const bool seq = role == 1;
                    if (class_bad(S, tile, a, e, seq)) viol |= seq ? FXG_V_SEQ_CLASS : FXG_V_QUAL_CLASS;
                    if (whole) {
                        if (seq) { if (e - a < 1) viol |= FXG_V_EMPTY_SEQ; }
                        else if (i >= 3 && e - a != (int)S.nlpos[i - 2] - (int)S.nlpos[i - 3] - 1) viol |= FXG_V_QUAL_LEN;
                    }
                }
            }
            viol = __reduce_or_sync(0xffffffffu, viol);
            if (lane == 0 && viol) atomicOr(&T.viol, viol);
            // exact mode: the (whole) lines of the target's role are hashed by the warps of the role two lines away
            // -- for a header target those are the "+"-line warps, which have next to nothing to check, so the
            // longest chain of this step stays one class check or one key hash, not both.  The line that straddles
            // the tile's left edge is left to tile_finalize_kernel.
            if (EXACT && role == ((trole + 2) & 3)) {
                const int hfirst = (trole - r) & 3;
                for (int i = hfirst + 4 * ((warp >> 2) * 32 + lane); i < n_tab; i += 4 * (kWarps / 4) * 32) {
                    if (i < 1) continue;
                    P.keyhash[t * (unsigned long long)(kMaxTileLines / 4) + (unsigned int)(i >> 2)] =
                        exact_line_hash_tile(P, tb, (int)S.nlpos[i - 1] + 1, (int)S.nlpos[i]);
                }
            }
        } else {
            // generic path (no "+" line in the tile, or REDO): one thread per line, flags OR-ed per
            // (local line index mod 4); tile_finalize_kernel maps them to roles once the phase is known.
            unsigned int f_own = 0, f_nxt = 0;
            for (int i = tid; i < n_tab; i += kThreads) {
                unsigned int own, nxt;
                line_flags(S, tile, i, n_tab, own, nxt);
                f_own |= own;
                f_nxt |= nxt;
            }
            unsigned int v = f_own | (f_nxt << 8);
            v = __reduce_or_sync(0x11111111u << (lane & 3), v);
            if (lane < 4 && v) atomicOr(&T.cls[lane], v);
            if (tid == kThreads - 1) {           // bytes after the last '\n': an unterminated line of index n_tab
                unsigned int ms, mq;
                range_any2(S, n_tab ? (S.nlpos[n_tab - 1] >> 4) + 1 : 0, kChunks, ms, mq);
                const unsigned int f = (ms ? kFS : 0u) | (mq ? kFQ : 0u);
                if (f) atomicOr(&T.cls[n_tab & 3], f);
            }
            if (EXACT && REDO) {
                // the true phase is known by now (prefix kernels ran before this launch): hash the whole lines of
                // the target's role; line 0 is left to tile_finalize_kernel like in the first launch
                const int r = (int)((P.lbase[t] + P.bsum[t / kPrefixSpan]) & 3ull);
                for (int i = 1 + tid; i < n_tab; i += kThreads) {
                    if (((i + r) & 3) != trole) continue;
                    P.keyhash[t * (unsigned long long)(kMaxTileLines / 4) + (unsigned int)(i >> 2)] =
                        exact_line_hash_tile(P, tb, (int)S.nlpos[i - 1] + 1, (int)S.nlpos[i]);
                }
            }
        }

        // ---- 5. verify the matcher's candidates (phase-free; line hits)
        if (LIT && !REDO) {
            const unsigned int nc_all = T.n_cand;
            if (nc_all <= (unsigned int)kCandCap) {
                for (unsigned int q = tid; q < nc_all; q += kThreads)
                    verify_candidate(S, T, tb, P, base, n_tab, LS.cand_off[q], LS.cand_mask[q]);
            } else {
                // queue overflow (pattern hits nearly everywhere): rescan the tile and verify in place
                for (int c = tid; c < kChunks && (c << 4) < tile_len; c += kThreads) {
                    const uint4 x = tile[c];
                    if (LIT8) {
                        const unsigned int r0 = MODE == kModeLit9 ? x.x + x.y * lit_mul2 : x.x * lit_mul1 + x.y * lit_mul2;
                        const unsigned int r1 = MODE == kModeLit9 ? x.z + x.w * lit_mul2 : x.z * lit_mul1 + x.w * lit_mul2;
                        if (LS.fp[r0 >> lit_shift] == r0) verify_candidate(S, T, tb, P, base, n_tab, c << 4, LS.masks[r0 >> lit_shift]);
                        if (LS.fp[r1 >> lit_shift] == r1) verify_candidate(S, T, tb, P, base, n_tab, (c << 4) + 8, LS.masks[r1 >> lit_shift]);
                    } else {
                        const unsigned int wv[4] = {x.x, x.y, x.z, x.w};
                        for (int u = 0; u < 16; u += lit_unit) {
                            unsigned int word = wv[u >> 2];
                            unsigned int val = lit_unit == 4 ? word : (lit_unit == 2 ? ((word >> ((u & 2) * 8)) & 0xffffu)
                                                                                     : ((word >> ((u & 3) * 8)) & 0xffu));
                            for (int q = 0; q < lit_nkeys; ++q)
                                if ((unsigned int)LS.keys[q] == val)
                                    verify_candidate(S, T, tb, P, base, n_tab, (c << 4) + u, LS.masks[q]);
                        }
                    }
                }
            }
        }
        FXG_GROUP_SYNC();

        // ---- 6. publish the tile: count, violations / line-class flags, '\n' table, line hits.  No barrier
        //         after it: the next writers of anything read here (the bulk copy into this buffer, the table,
        //         the hit words, this copy of the scalars) all come after the first barrier of the next tile.
        {
            uint4 *dst = reinterpret_cast<uint4 *>(P.tile_nl + t * (unsigned long long)kMaxTileLines);
            const uint4 *src = reinterpret_cast<const uint4 *>(S.nlpos);
            // the duties of this step start on different warps -- table copy: warp 4, scalars and hit words: warp 8;
            // thread 0 keeps the prefetch of the next tile -- so that no single warp enters the next tile's main pass
            // late by the sum of them: the whole group waits for its last warp at the next barrier (c2 + 1.9 %, c4
            // + 2.5 %).  All three are header-role warps, which have the least to do in step 4 in every mode.
            constexpr int kNlBase = ROLEMIX ? 224 : 128, kInfoTid = ROLEMIX ? 320 : 256;   // header-role warps: 0, 4, 8 / 0, 7, 10
            const int tid_nl = tid >= kNlBase ? tid - kNlBase : tid + kThreads - kNlBase;
            const int tid_hit = tid > kInfoTid ? tid - kInfoTid - 1 : tid + kThreads - kInfoTid - 1;
            const int tid_info = kInfoTid;
            for (int q = tid_nl; q * 8 < n_tab; q += kThreads) dst[q] = src[q];
            if (HITS) {
                unsigned int *hd = P.tile_hit + t * (unsigned long long)kHitWords;
                const int nw = (n_tab >> 5) + 1;               // local line indices 0..n_tab
                for (int w = tid_hit; w < nw; w += kThreads) {
                    const unsigned int v = S.linehit[w];
                    hd[w] = v;
                    if (v) S.linehit[w] = 0;
                }
            }
            if (tid == tid_info) {
                if (KMER && !REDO) P.tile_ncand[t] = T.n_cand <= (unsigned int)P.cand_cap ? T.n_cand : ((unsigned int)P.cand_cap | kCandOverflow);
                // the REDO launch re-checks structure only: the hit words of the first launch stand, and so does their flag
                const unsigned int kept = REDO ? (P.info[t].viol & kTileHasHit) : 0u;
                const TileInfo ti = {total, kept | T.viol | (plus_idx != 0xffffffffu ? (((2u - plus_idx) & 3u) + 1u) << 16 : 0u),
                                     {(unsigned short)T.cls[0], (unsigned short)T.cls[1], (unsigned short)T.cls[2], (unsigned short)T.cls[3]}};
                P.info[t] = ti;
                reset_state(T);                                // next used two tiles on, behind that tile's barriers
            }
        }
    }
#undef FXG_TILE_OF
#undef FXG_ISSUE_TILE
#undef FXG_GROUP_SYNC
}

// ---------------------------------------------------------------- launcher (one per instantiation)
// Persistent grid: one wave of CTAs (SMs x resident CTAs for this kernel and its shared memory).  The per-device
// attribute / occupancy cache is guarded: contexts on different host threads may launch concurrently.
cudaError_t scan_launch_config(const void *func, int threads, size_t smem, int dev, int *grid_cap);

template <class G, int MODE, bool REDO>
cudaError_t launch_scan_tiles(const ScanParams &P, size_t extra_smem, cudaStream_t stream, int dev, unsigned long long max_grid)
{
    const size_t smem = G::kGroups * sizeof(ScanSmem<G>) + extra_smem;
    int cap = 0;
    cudaError_t e = scan_launch_config((const void *)scan_tiles_kernel<G, MODE, REDO>, G::kCtaThreads, smem, dev, &cap);
    if (e != cudaSuccess) return e;
    const unsigned long long want = (P.ntiles + G::kGroups - 1) / G::kGroups;
    unsigned long long grid = want < (unsigned long long)cap ? want : (unsigned long long)cap;
    if (max_grid && grid > max_grid) grid = max_grid;
    scan_tiles_kernel<G, MODE, REDO><<<(unsigned int)grid, G::kCtaThreads, smem, stream>>>(P);
    return cudaGetLastError();
}

}  // namespace fxg
