// fxg_scan.cu -- raw FASTQ path (Path A): one fused pass over the batch.
//
// Replaces, for canonical 4-line FASTQ, the reference's per-record loop
//   kstream::read (kseq.hpp:120-164)  ->  record boundaries
//   prepare_data (main.cpp:291-304)   ->  target field
//   strstr x1..2 (main.cpp:568-588)   ->  literal hit
//   on_match (main.cpp:258-271)       ->  count / select
// with a position-centric kernel that has NO cross-tile dependency.  Every CTA takes one
// 24 KiB tile, pulls it into shared memory with one bulk async copy (TMA, mbarrier
// completion) and then
//   1. main pass, 16 bytes per lane per step, SWAR only: count of bytes < 0x20, byte
//      range check, per-chunk class summaries ("has a byte < 0x41", "has a byte < 0x21")
//      as ballot bit rows, and the literal filter over aligned 8-byte units;
//   2. block scan of the per-chunk counts -> ordered table of '\n' positions;
//   3. one thread per line: record structure and character classes, evaluated for ALL
//      FOUR possible line phases of the tile at once (4-bit "hypothesis" mask) -- FASTQ
//      line phase is not locally decidable ('@' may open a quality line), but the checks
//      are, once conditioned on the phase;
//   4. literal candidates are verified byte by byte and recorded as "local line i has a
//      match that is valid if line i has the target's role";
//   5. the tile publishes {count, hypothesis mask, '\n' table, line-hit bits}.
// tile_prefix_kernel then turns the counts into global line numbers (line phase =
// line# mod 4 decides header / sequence / plus / quality) and tile_finalize_kernel picks
// the hypothesis, checks the lines that straddle tile edges, converts line hits into
// record hits and, when records are emitted or a second-stage matcher follows, writes
// record starts / the target-field table.
// The canonical-FASTQ predicate is what makes kseq-parse-then-print
// the identity on the record bytes; a batch that violates it is reported (never silently
// mis-parsed) and the host re-routes it.
//
// HBM traffic: each input byte is read once; the scan writes 2 B per line ('\n' table) and
// a few words per tile; the finalize pass reads those back.
#include "fxg_common.cuh"

namespace fxg {

constexpr int kWarps = kThreads / 32;
constexpr int kWarpChunks = kChunks / kWarps;       // chunks one warp covers (contiguous)
static_assert(kWarps % 4 == 0, "step 4 gives every line role its own warps");

struct __align__(128) ScanSmem {
    uint4 tile[2][kChunks];                   // staged tile bytes, double buffered (bulk-copy destinations)
    unsigned short nlpos[kMaxTileLines + 16]; // tile-relative '\n' positions, in order (+ slack, see step 3)
    unsigned short wl[kWarps][kWarpChunks];   // per warp: its chunks that contain a byte < 0x21 ('\n' candidates)
    unsigned short wm[kWarps][kWarpChunks];   // per entry of wl: 16-bit mask of the bytes < 0x20 of that chunk
    unsigned int wtot[2][kWarps];             // '\n' count of each warp's 2 KiB segment (by tile parity)
    unsigned int below[17];                   // interleaved-layout mask of the bytes < p of a chunk
    unsigned int sbits[kChunks / 32];         // bit c: chunk c has a byte < 0x41 (not a sequence letter)
    unsigned int qbits[kChunks / 32];         // bit c: chunk c has a byte < 0x21 (not a quality char)
    unsigned int linehit[kHitWords];          // bit i: local line i has a (phase-conditional) match
    unsigned long long lit_keys[64];
    unsigned int lit_fp[64];
    unsigned short lit_masks[64];
    int cand_off[kCandCap];                   // tile-relative offset of a candidate unit
    unsigned short cand_mask[kCandCap];       // (pattern,shift) bits to verify
    unsigned long long mbar[2];               // bulk-copy completion barriers
    // per-tile scalars, double buffered by tile parity: the copy of the previous tile is reset by
    // thread 0 while the current one is in use (no extra barrier between "last read" and "first write")
    struct TileState {
        unsigned int cls[4];                  // generic path: line flags OR-ed per (local line index mod 4)
        unsigned int n_cand, viol;
    } st[2];
};
typedef ScanSmem::TileState TileState;
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
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned int bytes, unsigned long long *mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(mbar)) : "memory");
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
// il_mask of "bytes with index < p" (p in 0..16)
__host__ __device__ inline unsigned int il_below(int p)
{
    unsigned int m = 0;
    for (int b = 0; b < p; ++b) m |= 1u << (8 * (b & 3) + 7 - (b >> 2));
    return m;
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
__device__ __forceinline__ unsigned int text_byte(const unsigned char *tb, const ScanParams &P,
                                                  unsigned long long base, long long g)
{
    long long rel = g - (long long)base;
    if (rel >= 0 && rel < kTileBytes) return tb[rel];
    return P.buf[g];
}

// '\n' count in tile bytes [0, rel): lower bound in the finished position table
__device__ __forceinline__ unsigned int lines_before(const ScanSmem &S, int n_tab, int rel)
{
    int lo = 0, hi = n_tab;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if ((int)S.nlpos[mid] < rel) lo = mid + 1; else hi = mid;
    }
    return (unsigned int)lo;
}

// any bit set in rows[lo, hi) (chunk indices) of both class bit rows
__device__ __forceinline__ void range_any2(const ScanSmem &S, int lo, int hi, unsigned int &anyS, unsigned int &anyQ)
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
__device__ __forceinline__ void line_flags(const ScanSmem &S, const uint4 *tile, int i, int n_tab, unsigned int &own,
                                           unsigned int &nxt)
{
    const unsigned char *tb = reinterpret_cast<const unsigned char *>(tile);
    const int pos = S.nlpos[i];
    const bool ctl = tb[pos] != '\n';             // a control byte posing as a line end
    const int prev = i > 0 ? (int)S.nlpos[i - 1] : -1;
    const int next = i + 1 < n_tab ? (int)S.nlpos[i + 1] : kTileBytes;
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

// ---------------------------------------------------------------- verification of one literal candidate
// A verified occurrence is recorded on its local line; whether that line has the target's
// role (sequence / quality / header) is decided by tile_finalize_kernel.  Header targets add
// the geometric condition (inside the name / inside the comment), which is phase-free.
__device__ __noinline__ void verify_candidate(ScanSmem &S, TileState &T, const unsigned char *tb, const ScanParams &P,
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
            if (text_byte(tb, P, base, p + k) != pb[k]) { ok = false; break; }
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
                unsigned int ch = text_byte(tb, P, base, q);
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

__device__ __forceinline__ void push_candidate(ScanSmem &S, TileState &T, int unit_rel, unsigned int bits)
{
    unsigned int slot = atomicAdd(&T.n_cand, 1u);
    if (slot < kCandCap) { S.cand_off[slot] = unit_rel; S.cand_mask[slot] = (unsigned short)bits; }
}

// generic (short-pattern) literal filter: units of 4, 2 or 1 bytes compared against every key
__device__ __noinline__ void filter_small_units(ScanSmem &S, TileState &T, unsigned int x0, unsigned int x1, unsigned int x2,
                                                unsigned int x3, int off, int U, int nkeys)
{
    for (int u = 0; u < 16; u += U) {
        const int wi = u >> 2;
        unsigned int word = wi == 0 ? x0 : (wi == 1 ? x1 : (wi == 2 ? x2 : x3));
        unsigned int val = U == 4 ? word : (U == 2 ? ((word >> ((u & 2) * 8)) & 0xffffu) : ((word >> ((u & 3) * 8)) & 0xffu));
        for (int k = 0; k < nkeys; ++k)
            if ((unsigned int)S.lit_keys[k] == val) push_candidate(S, T, off + u, S.lit_masks[k]);
    }
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
__device__ __forceinline__ unsigned int class_bad(const ScanSmem &S, const uint4 *tile, int a, int b, bool seq)
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
// Persistent CTAs (3 per SM, 12 warps each): tile t = blockIdx.x, += gridDim.x.  The next tile
// is prefetched into the other shared-memory buffer while the current one is processed.
// LM = literal filter mode: 0 none (multi / exact strategies), 8 = aligned 8-byte units with the
// fingerprint table (9: same with first multiplier 1), 1 = short patterns (units of 4 / 2 / 1 bytes).
// REDO = second launch over the (normally empty) list of tiles whose locally guessed line phase
// turned out wrong: those are re-evaluated with the phase-free per-line path.
template <int LM, bool REDO>
__global__ void __launch_bounds__(kThreads, kMinCtas) scan_tiles_kernel(const ScanParams P)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ScanSmem &S = *reinterpret_cast<ScanSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const LiteralTable *lit = P.lit;
    const unsigned long long n_work = REDO ? (unsigned long long)*P.redo_count : P.ntiles;
    if ((unsigned long long)blockIdx.x >= n_work) return;
#define FXG_TILE_OF(k) (REDO ? (unsigned long long)P.redo_list[(k)] : (unsigned long long)(k))

    // ---- once per CTA: tables, barriers, first tile in flight
    if (tid == 0) {
        mbar_init(&S.mbar[0], 1);
        mbar_init(&S.mbar[1], 1);
        const unsigned long long t0 = FXG_TILE_OF(blockIdx.x);
        const unsigned long long rem0 = P.nbytes - t0 * (unsigned long long)kTileBytes;
        const unsigned int bulk0 = (unsigned int)(rem0 < (unsigned long long)kTileBytes ? rem0 : (unsigned long long)kTileBytes) & ~15u;
        if (bulk0) {
            mbar_expect_tx(&S.mbar[0], bulk0);
            bulk_g2s(S.tile[0], P.buf + t0 * (unsigned long long)kTileBytes, bulk0, &S.mbar[0]);
        }
        reset_state(S.st[0]);
        reset_state(S.st[1]);
    }
    if (tid >= 128 && tid < 145) S.below[tid - 128] = kBelowTab[tid - 128];
    if (LM != 0) for (int w = tid; w < kHitWords; w += kThreads) S.linehit[w] = 0;
    int lit_unit = 0, lit_nkeys = 0;
    unsigned int lit_mul1 = 0, lit_mul2 = 0, lit_shift = 26;
    if (LM != 0) {
        lit_unit = lit->unit; lit_nkeys = lit->n_keys; lit_mul1 = lit->mul1; lit_mul2 = lit->mul2; lit_shift = (unsigned int)lit->shift;
        if (tid >= 64 && tid < 128) {
            const int k = tid - 64;
            S.lit_keys[k] = lit->keys[k]; S.lit_fp[k] = lit->fp[k]; S.lit_masks[k] = lit->masks[k];
        }
    }
    const unsigned int one = P.one;
    const unsigned int lt_mask = (1u << lane) - 1u;
    unsigned short *my_list = S.wl[warp];
    unsigned short *my_mask = S.wm[warp];
    __syncthreads();

    unsigned int it = 0;
    for (unsigned long long k = blockIdx.x; k < n_work; k += gridDim.x, ++it) {
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
        //      previous iteration); reset the other copy of the per-tile scalars
        if (tid == 0) {
            if (it) reset_state(S.st[cur ^ 1]);
            const unsigned long long kn = k + gridDim.x;
            if (kn < n_work) {
                const unsigned long long tn = FXG_TILE_OF(kn);
                const unsigned long long remn = P.nbytes - tn * (unsigned long long)kTileBytes;
                const unsigned int bulkn = (unsigned int)(remn < (unsigned long long)kTileBytes ? remn : (unsigned long long)kTileBytes) & ~15u;
                if (bulkn) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    mbar_expect_tx(&S.mbar[cur ^ 1], bulkn);
                    bulk_g2s(S.tile[cur ^ 1], P.buf + tn * (unsigned long long)kTileBytes, bulkn, &S.mbar[cur ^ 1]);
                }
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
            __syncthreads();
        }
        if (bulk) mbar_wait(&S.mbar[cur], (it >> 1) & 1u);

        // ---- 1. main pass: 16 bytes per lane per step, SWAR only.  Per chunk: byte range, "has a
        //         byte < 0x41" / "< 0x21" ballot rows, list of the chunks with a byte < 0x21 (the only
        //         ones that can hold a '\n'), literal filter.
        unsigned int wcount = 0;
        {
            unsigned int hi_acc = 0;
#pragma unroll
            for (int j = 0; j < kChunksPerThread; ++j) {
                const int c = warp * kWarpChunks + j * 32 + lane;
                const uint4 x = tile[c];
                hi_acc |= x.x | fma_add(x.x, one, 0x01010101u);     // bytes >= 0x7f anywhere
                hi_acc |= x.y | fma_add(x.y, one, 0x01010101u);
                hi_acc |= x.z | fma_add(x.z, one, 0x01010101u);
                hi_acc |= x.w | fma_add(x.w, one, 0x01010101u);
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
                if (LM == 8 || LM == 9) {
                    const unsigned int r0 = LM == 9 ? x.x + x.y * lit_mul2 : x.x * lit_mul1 + x.y * lit_mul2;
                    const unsigned int r1 = LM == 9 ? x.z + x.w * lit_mul2 : x.z * lit_mul1 + x.w * lit_mul2;
                    const bool c0 = S.lit_fp[r0 >> lit_shift] == r0, c1 = S.lit_fp[r1 >> lit_shift] == r1;
                    if (__any_sync(0xffffffffu, c0 || c1)) {          // rare: keep the hot loop branch-free
                        if (c0) push_candidate(S, T, c << 4, S.lit_masks[r0 >> lit_shift]);
                        if (c1) push_candidate(S, T, (c << 4) + 8, S.lit_masks[r1 >> lit_shift]);
                    }
                } else if (LM == 1) {
                    if ((c << 4) < tile_len) filter_small_units(S, T, x.x, x.y, x.z, x.w, c << 4, lit_unit, lit_nkeys);
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
        __syncthreads();

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
                        for (int k = 2; m16; ++k) {
                            dst[k] = (unsigned short)(cb + __ffs(m16));
                            m16 &= m16 - 1;
                        }
                    }
                }
            }
        }
        const int n_tab = total < (unsigned int)kMaxTileLines ? (int)total : kMaxTileLines;
        if (total > (unsigned int)kMaxTileLines && tid == 0) note_violation(T, FXG_V_DENSE_LINES);
        __syncthreads();

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
            const int role = warp & 3;                        // 0 header, 1 sequence, 2 plus, 3 quality
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
        }

        // ---- 5. verify the literal candidates (phase-free; line hits)
        if (LM != 0 && !REDO) {
            const unsigned int nc_all = T.n_cand;
            if (nc_all <= (unsigned int)kCandCap) {
                for (unsigned int q = tid; q < nc_all; q += kThreads)
                    verify_candidate(S, T, tb, P, base, n_tab, S.cand_off[q], S.cand_mask[q]);
            } else {
                // queue overflow (pattern hits nearly everywhere): rescan the tile and verify in place
                for (int c = tid; c < kChunks && (c << 4) < tile_len; c += kThreads) {
                    const uint4 x = tile[c];
                    if (LM == 8 || LM == 9) {
                        const unsigned int r0 = LM == 9 ? x.x + x.y * lit_mul2 : x.x * lit_mul1 + x.y * lit_mul2;
                        const unsigned int r1 = LM == 9 ? x.z + x.w * lit_mul2 : x.z * lit_mul1 + x.w * lit_mul2;
                        if (S.lit_fp[r0 >> lit_shift] == r0) verify_candidate(S, T, tb, P, base, n_tab, c << 4, S.lit_masks[r0 >> lit_shift]);
                        if (S.lit_fp[r1 >> lit_shift] == r1) verify_candidate(S, T, tb, P, base, n_tab, (c << 4) + 8, S.lit_masks[r1 >> lit_shift]);
                    } else {
                        const unsigned int wv[4] = {x.x, x.y, x.z, x.w};
                        for (int u = 0; u < 16; u += lit_unit) {
                            unsigned int word = wv[u >> 2];
                            unsigned int val = lit_unit == 4 ? word : (lit_unit == 2 ? ((word >> ((u & 2) * 8)) & 0xffffu)
                                                                                     : ((word >> ((u & 3) * 8)) & 0xffu));
                            for (int q = 0; q < lit_nkeys; ++q)
                                if ((unsigned int)S.lit_keys[q] == val)
                                    verify_candidate(S, T, tb, P, base, n_tab, (c << 4) + u, S.lit_masks[q]);
                        }
                    }
                }
            }
        }
        __syncthreads();

        // ---- 6. publish the tile: count, violations / line-class flags, '\n' table, line hits.  No barrier
        //         after it: the next writers of anything read here (the bulk copy into this buffer, the table,
        //         the hit words, this copy of the scalars) all come after the first barrier of the next tile.
        {
            uint4 *dst = reinterpret_cast<uint4 *>(P.tile_nl + t * (unsigned long long)kNlStride);
            const uint4 *src = reinterpret_cast<const uint4 *>(S.nlpos);
            for (int q = tid; q * 8 < n_tab; q += kThreads) dst[q] = src[q];
            if (LM != 0 && !REDO) {
                unsigned int *hd = P.tile_hit + t * (unsigned long long)kHitWords;
                const int nw = (n_tab >> 5) + 1;               // local line indices 0..n_tab
                for (int w = tid; w < nw; w += kThreads) {
                    const unsigned int v = S.linehit[w];
                    hd[w] = v;
                    if (v) S.linehit[w] = 0;
                }
            }
            if (tid == 0) {
                // the REDO launch re-checks structure only: the hit words of the first launch stand, and so does their flag
                const unsigned int kept = REDO ? (P.info[t].viol & kTileHasHit) : 0u;
                const TileInfo ti = {total, kept | T.viol | (plus_idx != 0xffffffffu ? (((2u - plus_idx) & 3u) + 1u) << 16 : 0u),
                                     {(unsigned short)T.cls[0], (unsigned short)T.cls[1], (unsigned short)T.cls[2], (unsigned short)T.cls[3]}};
                P.info[t] = ti;
            }
        }
    }
#undef FXG_TILE_OF
}

// ---------------------------------------------------------------- kernel 2a: exclusive scan of the tile counts
// Two tiny launches: every CTA scans 4096 consecutive tiles (4 per thread) and leaves its
// block sum; one CTA then scans the block sums.  '\n' count before tile t =
// lbase[t] + bsum[t / kPrefixSpan].
constexpr int kPrefixThreads = 1024;
constexpr int kPrefixSpan = 4 * kPrefixThreads;
__global__ void __launch_bounds__(kPrefixThreads) tile_prefix_local_kernel(const ScanParams P)
{
    __shared__ unsigned long long wsum[kPrefixThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long t0 = (unsigned long long)blockIdx.x * kPrefixSpan + 4ull * tid;
    unsigned long long v[4], s = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        v[k] = t0 + k < P.ntiles ? (unsigned long long)P.info[t0 + k].total : 0ull;
        s += v[k];
    }
    unsigned long long inc = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    unsigned long long woff = 0, tot = 0;
    for (int w = 0; w < kPrefixThreads / 32; ++w) { unsigned long long x = wsum[w]; if (w < warp) woff += x; tot += x; }
    unsigned long long run = woff + inc - s;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (t0 + k < P.ntiles) P.lbase[t0 + k] = run;
        run += v[k];
    }
    if (tid == 0) P.bsum[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(kPrefixThreads) tile_prefix_blocks_kernel(const ScanParams P, unsigned int nblk)
{
    __shared__ unsigned long long wsum[kPrefixThreads / 32];
    __shared__ unsigned long long carry_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (unsigned int b0 = 0; b0 < nblk; b0 += kPrefixThreads) {
        const unsigned int i = b0 + tid;
        const unsigned long long v = i < nblk ? P.bsum[i] : 0ull;
        unsigned long long inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        unsigned long long woff = 0, tot = 0;
        for (int w = 0; w < kPrefixThreads / 32; ++w) { unsigned long long x = wsum[w]; if (w < warp) woff += x; tot += x; }
        if (i < nblk) P.bsum[i] = carry_s + woff + inc - v;
        __syncthreads();
        if (tid == 0) carry_s += tot;
        __syncthreads();
    }
}

// ---------------------------------------------------------------- kernels 2b / 2c: per-tile finalize
struct TileView {
    const ScanParams &P;
    __device__ TileView(const ScanParams &p) : P(p) {}
    __device__ unsigned long long lbase(unsigned long long t) const { return P.lbase[t] + P.bsum[t / kPrefixSpan]; }
    __device__ int lines(unsigned long long t) const
    {
        unsigned int n = P.info[t].total;
        return n < (unsigned int)kMaxTileLines ? (int)n : kMaxTileLines;
    }
    __device__ long long nl(unsigned long long t, int i) const
    {
        return (long long)(t * (unsigned long long)kTileBytes) + P.tile_nl[t * (unsigned long long)kNlStride + i];
    }
    // k-th '\n' before tile t (k = 1 is the nearest); a virtual '\n' at offset -1 opens line 0
    __device__ long long prev_nl(unsigned long long t, int k) const
    {
        for (long long tt = (long long)t - 1; tt >= 0; --tt) {
            const int n = lines((unsigned long long)tt);
            if (n >= k) return nl((unsigned long long)tt, n - k);
            k -= n;
        }
        return k == 1 ? -1 : kNoPos;
    }
    // global position of the '\n' with tile-local index i (i >= -3)
    __device__ long long at(unsigned long long t, int i) const { return i >= 0 ? nl(t, i) : prev_nl(t, -i); }
};

// 2a': tiles whose guessed phase disagrees with the global line number go to the REDO list
__global__ void __launch_bounds__(256) tile_phase_check_kernel(const ScanParams P)
{
    const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= P.ntiles) return;
    const unsigned int guess = (P.info[t].viol >> 16) & kGuessMask;
    if (!guess) return;
    const unsigned long long lb = P.lbase[t] + P.bsum[t / kPrefixSpan];
    if ((unsigned int)(lb & 3) != guess - 1) P.redo_list[atomicAdd(P.redo_count, 1u)] = (unsigned int)t;
}

// 2b: one THREAD per tile -- pick the phase, check the lines that straddle the tile's left edge,
// turn line hits into record hits.
__global__ void __launch_bounds__(256) tile_finalize_kernel(const ScanParams P)
{
    const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= P.ntiles) return;
    const TileView V(P);
    const TileInfo ti = P.info[t];
    const unsigned long long lb = V.lbase(t);
    const unsigned long long base = t * (unsigned long long)kTileBytes;
    const int r = (int)(lb & 3);
    const int n_tab = ti.total < (unsigned int)kMaxTileLines ? (int)ti.total : kMaxTileLines;
    unsigned int viol = ti.viol & 0xffffu;
    const unsigned int guess = (ti.viol >> 16) & kGuessMask;   // 0: flags are per (line index mod 4); else guessed phase + 1
    if (guess && (int)(guess - 1) != r) viol |= FXG_V_DENSE_LINES;   // cannot happen after the REDO launch; never pass silently
#pragma unroll
    for (int k = 0; k < 4; ++k)                  // lines with local index == k (mod 4) have role (k + r) & 3
        viol |= role_kinds(ti.cls[k] & 0x7fu, (k + r) & 3) | role_kinds(ti.cls[k] >> 8, (k + 1 + r) & 3) |
                ((ti.cls[k] & kFCTL) ? (unsigned int)FXG_V_BYTE_RANGE : 0u);
    // structure of the first lines of the tile, whose predecessors lie in earlier tiles
    if (n_tab > 0) {
        long long p[6];                          // '\n' positions with local index -3 .. 2
        p[2] = V.prev_nl(t, 1);
        const int ph0 = r;
        const long long pos0 = V.nl(t, 0);
        const long long llen0 = pos0 - p[2] - 1;
        if (ph0 == 0) {
            if (llen0 < 1 || P.buf[p[2] + 1] != '@') viol |= FXG_V_HEADER_AT;
            else if (P.buf[pos0 - 1] == ' ') viol |= FXG_V_HEADER_SPACE;
        } else if (ph0 == 1) { if (llen0 < 1) viol |= FXG_V_EMPTY_SEQ; }
        else if (ph0 == 2) { if (llen0 != 1 || P.buf[pos0 - 1] != '+') viol |= FXG_V_PLUS_LINE; }
        // the one quality line among local lines 0..2 (if it exists): |qual| == |seq|
        const int iq = (3 - r) & 3;              // local index of the first line with phase 3
        if (iq < 3 && iq < n_tab) {
            p[3] = pos0;
            if (iq >= 1) p[4] = V.nl(t, 1);
            if (iq >= 2) p[5] = V.nl(t, 2);
            if (iq < 2) p[1] = V.prev_nl(t, 2);
            if (iq < 1) p[0] = V.prev_nl(t, 3);
            const long long llen = p[3 + iq] - p[2 + iq] - 1;
            const long long slen = p[1 + iq] - p[iq] - 1;
            if (llen != slen) viol |= FXG_V_QUAL_LEN;
        }
    }
    if (viol) { atomicOr(&P.ctr->violation, viol); atomicMin(&P.ctr->first_bad, base); }

    // line hits -> record hits: only lines whose role is the target's count
    if (P.match_enabled && (ti.viol & kTileHasHit)) {
        const int want = P.target == FXG_TARGET_SEQ ? 1 : (P.target == FXG_TARGET_QUAL ? 3 : 0);
        const unsigned int role_mask = 0x11111111u << ((want - r) & 3);     // 32 % 4 == 0: same mask for every word
        const unsigned int *hw = P.tile_hit + t * (unsigned long long)kHitWords;
        unsigned int fresh = 0;
        for (int w = 0; w * 32 <= n_tab; ++w) {
            unsigned int bits = hw[w] & role_mask;
            while (bits) {
                const int b = __ffs(bits) - 1;
                bits &= bits - 1;
                const unsigned long long rec = (lb + (unsigned long long)(w * 32 + b)) >> 2;
                if (rec < P.rec_cap) {
                    const unsigned int bit = 1u << (rec & 31);
                    const unsigned int old = atomicOr(&P.hitbits[rec >> 5], bit);
                    if (!(old & bit)) ++fresh;
                }
            }
        }
        if (fresh) atomicAdd(&P.ctr->n_hits, (unsigned long long)fresh);
    }
    if (t + 1 == P.ntiles) {
        P.ctr->n_lines = lb + ti.total;
        P.ctr->last_byte = P.buf[P.nbytes - 1];
    }
}

// 2c: one WARP per tile -- record starts / target-field table (only when records are emitted or a
// second-stage matcher follows)
__global__ void __launch_bounds__(256) tile_emit_kernel(const ScanParams P)
{
    const int lane = threadIdx.x & 31;
    const unsigned long long t = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (t >= P.ntiles) return;
    const TileView V(P);
    const unsigned long long lb = V.lbase(t);
    const int n_tab = V.lines(t);
    for (int i = lane; i < n_tab; i += 32) {
        const unsigned long long j = lb + (unsigned long long)i;
        const int ph = (int)(j & 3);
        const unsigned long long rec = j >> 2;
        const long long pos = V.nl(t, i);
        if (P.fld_off && rec < P.rec_cap) {
            if ((ph == 3 && P.target == FXG_TARGET_QUAL) || (ph == 1 && P.target == FXG_TARGET_SEQ)) {
                const long long prev = V.at(t, i - 1);
                P.fld_off[rec] = (unsigned long long)(prev + 1); P.fld_len[rec] = (unsigned int)(pos - prev - 1);
            } else if (ph == 0 && (P.target == FXG_TARGET_NAME || P.target == FXG_TARGET_COMMENT)) {
                const long long prev = V.at(t, i - 1);
                if (P.header_whole) {                 // the exact look-up finds the end of the name while it hashes
                    P.fld_off[rec] = (unsigned long long)(prev + 2);
                    P.fld_len[rec] = (unsigned int)(pos > prev + 2 ? pos - (prev + 2) : 0);
                    continue;
                }
                long long q = prev + 2, sp = -1;
                for (; q < pos; ++q) if (P.buf[q] == ' ') { sp = q; break; }
                if (P.target == FXG_TARGET_NAME) {
                    const long long ne = sp >= 0 ? sp : pos;
                    P.fld_off[rec] = (unsigned long long)(prev + 2);
                    P.fld_len[rec] = (unsigned int)(ne > prev + 2 ? ne - (prev + 2) : 0);
                } else {
                    P.fld_off[rec] = (unsigned long long)(sp >= 0 ? sp + 1 : pos);
                    P.fld_len[rec] = (unsigned int)(sp >= 0 ? pos - (sp + 1) : 0);
                }
            }
        }
    }
}

size_t scan_smem_bytes() { return sizeof(ScanSmem); }

template <int LM, bool REDO> static cudaError_t launch_scan_tiles(const ScanParams &P, cudaStream_t stream, int dev)
{
    static bool attr_set[64] = {false};
    static int grid_cap[64] = {0};
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(scan_tiles_kernel<LM, REDO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScanSmem));
        if (e != cudaSuccess) return e;
        int sms = 0, per_sm = 0;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, scan_tiles_kernel<LM, REDO>, kThreads, sizeof(ScanSmem))) != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
        grid_cap[dev] = sms * per_sm;            // persistent CTAs: one wave, 148 SMs x 3 on B200
        attr_set[dev] = true;
    }
    unsigned long long grid = P.ntiles < (unsigned long long)grid_cap[dev] ? P.ntiles : (unsigned long long)grid_cap[dev];
    if (REDO && grid > 64) grid = 64;            // the list is normally empty: CTAs beyond its length exit at once
    scan_tiles_kernel<LM, REDO><<<(unsigned int)grid, kThreads, sizeof(ScanSmem), stream>>>(P);
    return cudaGetLastError();
}

cudaError_t launch_scan_raw(const ScanParams &P, cudaStream_t stream, int *launches, cudaEvent_t after_scan)
{
    int dev = 0;
    cudaGetDevice(&dev);
    if (((uintptr_t)P.buf & 15) != 0) return cudaErrorMisalignedAddress;     // bulk copies need 16-byte alignment
    if (P.ntiles == 0) return cudaSuccess;
    if (P.ntiles > 0xffffffffull) return cudaErrorInvalidValue;
    cudaError_t e;
    if (!P.match_enabled) e = launch_scan_tiles<0, false>(P, stream, dev);
    else if (P.lit_unit == 9) e = launch_scan_tiles<9, false>(P, stream, dev);
    else if (P.lit_unit == 8) e = launch_scan_tiles<8, false>(P, stream, dev);
    else e = launch_scan_tiles<1, false>(P, stream, dev);
    if (e != cudaSuccess) return e;
    if (after_scan) cudaEventRecord(after_scan, stream);
    const unsigned int nblk = (unsigned int)((P.ntiles + kPrefixSpan - 1) / kPrefixSpan);
    tile_prefix_local_kernel<<<nblk, kPrefixThreads, 0, stream>>>(P);
    tile_prefix_blocks_kernel<<<1, kPrefixThreads, 0, stream>>>(P, nblk);
    tile_phase_check_kernel<<<(unsigned int)((P.ntiles + 255) / 256), 256, 0, stream>>>(P);
    if ((e = launch_scan_tiles<0, true>(P, stream, dev)) != cudaSuccess) return e;
    tile_finalize_kernel<<<(unsigned int)((P.ntiles + 255) / 256), 256, 0, stream>>>(P);
    if (launches) *launches += 6;
    if (P.fld_off) {
        tile_emit_kernel<<<(unsigned int)((P.ntiles + 7) / 8), 256, 0, stream>>>(P);
        if (launches) *launches += 1;
    }
    return cudaGetLastError();
}

// ================================================================= K6: flags -> ordered runs
// sel(k) = hit(k) XOR invert for k < n_records; a run is a maximal block of
// consecutive selected records; runs[r] = {start(first), start(last+1) - start(first)}; record starts are
// looked up in the tile tables (offset_after_line), not stored per record.
// Replaces the per-record `pair.print(out)` of on_match (main.cpp:258-271, kseq.cpp:32-48)
// for canonical FASTQ, where print(parse(bytes)) == bytes.
// Byte offset just after the '\n' with global index L (L = -1: offset 0): binary search of the tile whose line
// range holds L (tiles without a '\n' share their base with the next tile: the LAST tile with base <= L holds
// it), then that tile's position table.
__device__ unsigned long long offset_after_line(const ScanParams &P, long long L)
{
    if (L < 0) return 0ull;
    unsigned long long lo = 0, hi = P.ntiles;
    while (hi - lo > 1) {
        const unsigned long long mid = (lo + hi) >> 1;
        if (P.lbase[mid] + P.bsum[mid / kPrefixSpan] <= (unsigned long long)L) lo = mid; else hi = mid;
    }
    const unsigned long long k = (unsigned long long)L - (P.lbase[lo] + P.bsum[lo / kPrefixSpan]);
    if (k >= (unsigned long long)kMaxTileLines) return 0ull;        // dense tile: the batch is rejected anyway
    return lo * (unsigned long long)kTileBytes + P.tile_nl[lo * (unsigned long long)kNlStride + k] + 1ull;
}

struct CompactParams {
    const unsigned int *hitbits;
    ScanParams P;                   // tile tables for the record-boundary look-ups
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
    // the grid covers the record capacity of the batch (bytes / smallest record); blocks past the records that
    // exist leave at once
    if ((unsigned long long)blockIdx.x * kCompactThreads * 32ull >= nrec) {
        if (threadIdx.x == 0) C.block_sums[blockIdx.x] = 0ull;
        return;
    }
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
    // only the blocks that hold records can have run starts (and only they read their prefix)
    const unsigned long long nrec = C.ctr->n_lines >> 2;
    const unsigned long long need = (nrec + 32ull * kCompactThreads - 1) / (32ull * kCompactThreads);
    if (need < nblocks) nblocks = (unsigned int)need;
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
    if ((unsigned long long)blockIdx.x * kCompactThreads * 32ull >= nrec) return;   // no record, no run (block-uniform)
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
            if (opened < C.run_cap) C.runs[opened].off = offset_after_line(C.P, 4ll * (long long)rec - 1);
            ++opened;
        }
        if (ends & (1u << b)) {
            unsigned long long r = opened - 1;
            if (r < C.run_cap) C.runs[r].len = offset_after_line(C.P, 4ll * (long long)rec + 3);   // end offset; fixed up below
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

cudaError_t launch_compact(const unsigned int *hitbits, const ScanParams &P, const ScanCounters *ctr,
                           unsigned long long rec_cap, unsigned long long *block_sums, unsigned long long *out_counts,
                           fxg_run *runs, unsigned long long run_cap, int invert, cudaStream_t stream, int *launches)
{
    CompactParams C;
    C.hitbits = hitbits; C.P = P; C.ctr = ctr; C.block_sums = block_sums;
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
