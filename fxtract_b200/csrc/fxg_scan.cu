// fxg_scan.cu -- raw FASTQ path (Path A): epilogue kernels of the fused scan (prefix, phase check, finalize,
// field table, order-preserving compaction) and the launcher that strings one batch together.  The per-tile scan
// kernel itself is the template of fxg_scan_kernel.cuh; its G24 instantiations (literal / exact / index-only) are
// compiled here, the G32 ones (fused k-mer matcher) in fxg_scan_kmer.cu.
//
// HBM traffic: each input byte is read once; the scan writes 2 B per line ('\n' table) and a few words per
// tile; the finalize pass reads those back.
#include "fxg_scan_kernel.cuh"
#include <map>
#include <mutex>
#include <tuple>

namespace fxg {

// ---------------------------------------------------------------- kernel 2a: exclusive scan of the tile counts
// Two tiny launches: every CTA scans 4096 consecutive tiles (4 per thread) and leaves its
// block sum; one CTA then scans the block sums.  '\n' count before tile t =
// lbase[t] + bsum[t / kPrefixSpan].
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
        return n < (unsigned int)P.nl_stride ? (int)n : P.nl_stride;
    }
    __device__ long long nl(unsigned long long t, int i) const
    {
        return (long long)(t * (unsigned long long)P.tile_bytes) + P.tile_nl[t * (unsigned long long)P.nl_stride + i];
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
    if (!guess) {
        // no anchor, no role-directed work: in exact mode the look-ups of such a tile are done by the REDO launch
        if (P.redo_unanchored && P.info[t].total > 1u) P.redo_list[atomicAdd(P.redo_count, 1u)] = (unsigned int)t;
        return;
    }
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
    const unsigned long long base = t * (unsigned long long)P.tile_bytes;
    const int r = (int)(lb & 3);
    const int n_tab = ti.total < (unsigned int)P.nl_stride ? (int)ti.total : P.nl_stride;
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
    if (t + 1 == P.ntiles && ((lb + ti.total) >> 2) > P.rec_cap) viol |= FXG_V_DENSE_LINES;   // more records than the per-record buffers hold
    if (viol) { atomicOr(&P.ctr->violation, viol); atomicMin(&P.ctr->first_bad, base); }

    // exact mode: exact_resolve_kernel looks up the lines that lie wholly inside a tile; the line that ends at this
    // tile's first '\n' (it starts in an earlier tile, or exactly at the tile's first byte) is done here
    if (P.mode == kModeExact && n_tab > 0 && r == target_role(P.target)) {
        const long long ls = V.prev_nl(t, 1) + 1;
        int pid = -1;
        if (exact_line_global(P, ls, V.nl(t, 0), pid)) {
            const unsigned long long rec = lb >> 2;
            if (rec < P.rec_cap) {
                if (P.rec_first) P.rec_first[rec] = (unsigned long long)(unsigned int)pid;
                const unsigned int bit = 1u << (rec & 31);
                atomicOr(&P.hitbits[rec >> 5], bit);
            }
        }
    }

    // line hits -> record hits: only lines whose role is the target's count
    if (P.match_enabled && (ti.viol & kTileHasHit)) {
        const int want = P.target == FXG_TARGET_SEQ ? 1 : (P.target == FXG_TARGET_QUAL ? 3 : 0);
        const unsigned int role_mask = 0x11111111u << ((want - r) & 3);     // 32 % 4 == 0: same mask for every word
        const unsigned int *hw = P.tile_hit + t * (unsigned long long)P.hit_words;
        for (int w = 0; w * 32 <= n_tab; ++w) {
            unsigned int bits = hw[w] & role_mask;
            while (bits) {
                const int b = __ffs(bits) - 1;
                bits &= bits - 1;
                const unsigned long long rec = (lb + (unsigned long long)(w * 32 + b)) >> 2;
                if (rec < P.rec_cap) atomicOr(&P.hitbits[rec >> 5], 1u << (rec & 31));
            }
        }
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

// ---------------------------------------------------------------- second half of the fused matchers
// The scan kernel does everything that reads the staged tile (hashing a key, packing and filtering windows) and
// leaves one small entry per look-up; the look-ups themselves -- dependent global loads into tables that do not fit
// shared memory -- run here, one thread per entry over the whole batch, where tens of thousands of them are in
// flight at once instead of stalling a tile of the lock-step scan.
//
// exact mode: one WARP per tile, one lane per line of the target's role; key hash from keyhash[], bucket probe,
// and for a fingerprint match the key compared word by word against the batch text (replaces manager.find(*data),
// main.cpp:315).  Line 0 of a tile (it may start in an earlier tile) is done by tile_finalize_kernel.
constexpr int kExactLanesPerTile = 8;      // four tiles per warp: four chains of dependent loads per wait (see kmer_resolve_kernel)
__global__ void __launch_bounds__(256) exact_resolve_kernel(const ScanParams P)
{
    constexpr int LPT = kExactLanesPerTile;
    const int lane = (threadIdx.x & 31) % LPT;
    const unsigned long long t = (((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 / LPT) + (threadIdx.x & 31) / LPT;
    if (t >= P.ntiles) return;
    const TileView V(P);
    const unsigned long long lb = V.lbase(t);
    const int n_tab = V.lines(t);
    const int first = (target_role(P.target) - (int)(lb & 3ull)) & 3;
    const unsigned long long *kh = P.keyhash + t * (unsigned long long)(P.nl_stride / 4);
    // four lines per lane and round: their key hashes, then their pre-filter words, are fetched together; ~99 % of
    // the lines end there
    const bool use_pre = P.exact.pre != nullptr && P.target != FXG_TARGET_COMMENT;
    for (int i0 = first + 4 * lane; i0 < n_tab; i0 += 16 * LPT) {
        unsigned long long hv[4];
        unsigned int pw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + 4 * LPT * u;
            hv[u] = (i >= 1 && i < n_tab) ? __ldg(kh + (i >> 2)) : 0ull;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) pw[u] = use_pre ? __ldg(P.exact.pre + exact_pre_word(hv[u], P.exact.pre_mask)) : 0xffffffffu;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + 4 * LPT * u;
            if (i < 1 || i >= n_tab) continue;
            const unsigned long long h = hv[u];
            if (use_pre) { const unsigned int m = exact_pre_mask(h); if ((pw[u] & m) != m) continue; }
            int pid = -1;
            if (P.target == FXG_TARGET_COMMENT) {          // the key starts behind the first ' ': found by the slow path
                if (!exact_fingerprint_present(P.exact, h)) continue;
                if (!exact_line_global(P, V.nl(t, i - 1) + 1, V.nl(t, i), pid)) continue;
            } else {
                const bool name = P.target == FXG_TARGET_NAME;
                if (!exact_verify_hashed(P.exact, h, P.buf, V.nl(t, i - 1) + 1 + (name ? 1 : 0), V.nl(t, i), name, pid)) continue;
            }
            const unsigned long long rec = (lb + (unsigned long long)i) >> 2;
            if (rec >= P.rec_cap) continue;
            if (P.rec_first) P.rec_first[rec] = (unsigned long long)(unsigned int)pid;
            atomicOr(&P.hitbits[rec >> 5], 1u << (rec & 31));
        }
    }
}

// window matcher: one WARP per tile, one lane per candidate {window hash, tile-relative offset of the window}:
// table bucket, every pattern record that holds a window with this hash at one of its first S offsets compared with
// the batch text, the occurrence's line found in the tile's '\n' table, kept if that line has the target's role.
// A tile whose list overflowed (kCandOverflow: pattern-dense text) has every probe position walked here instead.
// `best` (routing): every occurrence through this window is examined and the smallest (end offset, 0xffff - length,
// pattern id) triple is kept; without it the first verified occurrence ends the walk.
__device__ __forceinline__ bool window_lookup(const ScanParams &P, unsigned int h, long long win, long long *best_end, unsigned long long *best_low)
{
    bool any = false;
    const KmerTile &K = P.kmer;
    unsigned int bkt = kmer_tile_bucket(h, K.ht_mask);
    for (;;) {
        const uint4 kk = __ldg(K.ht + 2 * (size_t)bkt), vv = __ldg(K.ht + 2 * (size_t)bkt + 1);
        const unsigned int keys[4] = {kk.x, kk.y, kk.z, kk.w}, vals[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const unsigned int v = vals[e];
            if (!v || keys[e] != h) continue;
            const unsigned int o = (v - 1u) & 127u;
            const uint4 *rec = K.rec + ((v - 1u) >> 7);
            const unsigned int m = __ldg(reinterpret_cast<const unsigned int *>(rec));
            const long long p = win - (long long)o;
            if (p < 0 || (unsigned long long)p + m > P.nbytes) continue;
            // pattern bytes (8-byte aligned in the record) against the text, 8 bytes per step: three aligned text
            // words, two funnel shifts by the occurrence's byte misalignment (the batch has 16 readable bytes of slack)
            const unsigned long long *pw = reinterpret_cast<const unsigned long long *>(rec) + 1;
            const unsigned int *tw = reinterpret_cast<const unsigned int *>(P.buf + (p & ~3ll));
            const int sh = (int)(p & 3) * 8;
            bool eq = true;
            unsigned int g0 = __ldg(tw);
            for (unsigned int i = 0; i < m; i += 8, tw += 2) {
                const unsigned int g1 = __ldg(tw + 1), g2 = __ldg(tw + 2);
                unsigned long long w = (unsigned long long)__funnelshift_r(g0, g1, sh) | ((unsigned long long)__funnelshift_r(g1, g2, sh) << 32);
                unsigned long long q = __ldg(pw + (i >> 3));
                if (m - i < 8u) { const unsigned long long mk = (1ull << (8 * (m - i))) - 1ull; w &= mk; q &= mk; }
                if (w != q) { eq = false; break; }
                g0 = g2;
            }
            if (!eq) continue;
            if (!best_end) return true;
            any = true;
            const unsigned long long low = ((unsigned long long)(0xffffu - (m < 0xffffu ? m : 0xffffu)) << 24) |
                                           (unsigned long long)(__ldg(reinterpret_cast<const unsigned int *>(rec) + 1) & 0xffffffu);
            const long long end = p + (long long)m;
            if (end < *best_end || (end == *best_end && low < *best_low)) { *best_end = end; *best_low = low; }
        }
        if (vv.w == 0u) return any;                          // the bucket has a free slot: nothing spilled past it
        bkt = (bkt + 1) & K.ht_mask;
    }
}

// LPT = lanes per tile: 32 for pattern sets behind the 128 KiB filter (~90 false candidates per tile), 8 for the
// small sets (a handful per tile): what the kernel waits for is the chain of dependent loads per tile, so a warp
// that walks four tiles side by side gets four times the work out of each wait.
template <int LPT>
__global__ void __launch_bounds__(256) kmer_resolve_kernel(const ScanParams P)
{
    constexpr int kTilesPerWarp = 32 / LPT;
    const int lane = threadIdx.x & 31, gl = lane % LPT, grp = lane / LPT;
    const unsigned int gmask = LPT == 32 ? 0xffffffffu : (((1u << LPT) - 1u) << (grp * LPT));
    const unsigned long long t = (((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * kTilesPerWarp + (unsigned int)grp;
    if (t >= P.ntiles) return;
    const unsigned int nc_raw = P.tile_ncand[t];
    if (!nc_raw) return;
    const KmerTile &K = P.kmer;
    const TileView V(P);
    const int trole = target_role(P.target);
    const unsigned long long base = t * (unsigned long long)P.tile_bytes;
    const unsigned short *nl = P.tile_nl + t * (unsigned long long)P.nl_stride;
    const uint2 *list = P.cand + t * (unsigned long long)P.cand_cap;
    const bool overflow = (nc_raw & kCandOverflow) != 0u;
    const unsigned long long left = P.nbytes - base;
    const unsigned int tile_len = left < (unsigned long long)P.tile_bytes ? (unsigned int)left : (unsigned int)P.tile_bytes;
    const unsigned int nc = overflow ? (tile_len + (unsigned int)K.stride - 1u) / (unsigned int)K.stride : nc_raw;
    const unsigned long long lb = V.lbase(t);
    const int n_tab = V.lines(t);
    // two candidates per lane and round: their list entries and table buckets are fetched before either is examined
    for (unsigned int c0 = 0; c0 < nc; c0 += 2 * LPT) {         // uniform over the tile's lanes: the line look-up below is cooperative
        unsigned int hh[2];
        int wr[2];
        bool live[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const unsigned int c = c0 + (unsigned int)gl + (unsigned int)(LPT * u);
            live[u] = c < nc;
            hh[u] = 0; wr[u] = 0;
            if (!live[u]) continue;
            if (!overflow) {
                const uint2 cd = __ldg(list + c);
                hh[u] = cd.x; wr[u] = (int)cd.y;
            } else {
                wr[u] = (int)(c * (unsigned int)K.stride);
                const unsigned long long g = base + (unsigned long long)wr[u];
                if (g + (unsigned long long)K.k > P.nbytes) { live[u] = false; continue; }   // a window inside an occurrence lies inside the batch
                unsigned int w[4] = {0u, 0u, 0u, 0u};
                for (int i = 0; i < K.k; ++i) w[i >> 2] |= (unsigned int)P.buf[g + i] << (8 * (i & 3));
                hh[u] = window_hash(w[0], w[1], w[2], w[3]);
            }
        }
        uint4 kk[2];
        unsigned int full[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const uint4 *b = K.ht + 2 * (size_t)kmer_tile_bucket(hh[u], K.ht_mask);
            kk[u] = __ldg(b); full[u] = __ldg(reinterpret_cast<const unsigned int *>(b + 1) + 3);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            bool hit = false;
            long long best_end = 0x7fffffffffffffffll;       // batch offset of the end of the first reported occurrence
            unsigned long long best_low = ~0ull;
            // most candidates end here: no entry of the first bucket carries the hash and the bucket is not full
            if (live[u] && (kk[u].x == hh[u] || kk[u].y == hh[u] || kk[u].z == hh[u] || kk[u].w == hh[u] || full[u] != 0u))
                hit = window_lookup(P, hh[u], (long long)base + wr[u], P.rec_first ? &best_end : nullptr, &best_low);
            // local line of a verified occurrence = number of '\n' before its window (a pattern holds no '\n', so all
            // of it lies on the line of its window): counted by all lanes of the tile, eight table entries per lane
            // and step, instead of a binary search of dependent loads by the one lane that has the hit
            unsigned int hb = __ballot_sync(gmask, hit) & gmask;
            while (hb) {
                const int src = __ffs(hb) - 1;
                hb &= hb - 1;
                const int pos = __shfl_sync(gmask, wr[u], src);
                int lo = 0;
                for (int j = gl * 8; j < n_tab; j += LPT * 8) {
                    const uint4 v = __ldg(reinterpret_cast<const uint4 *>(nl + j));
                    const unsigned int w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        lo += (j + 2 * q < n_tab && (int)(w[q] & 0xffffu) < pos) ? 1 : 0;
                        lo += (j + 2 * q + 1 < n_tab && (int)(w[q] >> 16) < pos) ? 1 : 0;
                    }
                }
                lo = (int)__reduce_add_sync(gmask, (unsigned int)lo);
                if (lane != src) continue;
                const unsigned long long line = lb + (unsigned long long)lo;
                if ((int)(line & 3ull) != trole) continue;
                const unsigned long long r = line >> 2;
                if (r >= P.rec_cap) continue;
                atomicOr(&P.hitbits[r >> 5], 1u << (r & 31));
                if (P.rec_first) {
                    // end offset relative to the start of the line (which may lie in an earlier tile)
                    const long long ls = lo >= 1 ? (long long)base + __ldg(nl + lo - 1) + 1 : V.prev_nl(t, 1) + 1;
                    const unsigned long long end_rel = (unsigned long long)(best_end - ls);
                    atomicMin(&P.rec_first[r], ((end_rel < 0xffffffull ? end_rel : 0xffffffull) << 40) | (best_low & 0xffffffffffull));
                }
            }
        }
    }
}

// matched_records (main.cpp:80): the records with a hit are counted from the finished bit set -- one pass over
// 1 bit per record, one atomic per block -- instead of one same-address atomic per hit in the kernels that find them
__global__ void __launch_bounds__(256) count_hits_kernel(const ScanParams P)
{
    __shared__ unsigned int wsum[8];
    unsigned long long nrec = P.ctr->n_lines >> 2;
    if (nrec > P.rec_cap) nrec = P.rec_cap;
    const unsigned long long words = (nrec + 31) >> 5;
    unsigned int n = 0;
    for (unsigned long long w = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (unsigned long long)gridDim.x * blockDim.x) {
        unsigned int v = P.hitbits[w];
        const unsigned long long left = nrec - w * 32ull;
        if (left < 32ull) v &= (1u << left) - 1u;
        n += __popc(v);
    }
    n = __reduce_add_sync(0xffffffffu, n);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int tot = 0;
        for (int i = 0; i < 8; ++i) tot += wsum[i];
        if (tot) atomicAdd(&P.ctr->n_hits, (unsigned long long)tot);
    }
}

// ---------------------------------------------------------------- launch configuration cache
// cudaFuncSetAttribute + occupancy query once per (kernel, device, shared-memory size); guarded, because contexts
// driven from different host threads share it.
cudaError_t scan_launch_config(const void *func, int threads, size_t smem, int dev, int *grid_cap)
{
    static std::mutex mu;
    static std::map<std::tuple<const void *, int, size_t>, int> cache;
    std::lock_guard<std::mutex> lock(mu);
    const auto key = std::make_tuple(func, dev, smem);
    auto it = cache.find(key);
    if (it != cache.end()) { *grid_cap = it->second; return cudaSuccess; }
    cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int sms = 0, per_sm = 0;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, func, threads, smem)) != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    *grid_cap = sms * per_sm;                    // persistent CTAs: one wave (148 SMs x 3 / 2 / 1 on B200)
    cache[key] = *grid_cap;
    return cudaSuccess;
}

void scan_geometry(int mode, int filter_log2_words, int *tile_bytes, int *nl_stride, int *hit_words)
{
    if (mode_is_kmer(mode) && filter_log2_words > 13) { *tile_bytes = G16D::kTileBytes; *nl_stride = G16D::kMaxTileLines; *hit_words = G16D::kHitWords; }
    else if (mode_is_kmer(mode)) { *tile_bytes = G32::kTileBytes; *nl_stride = G32::kMaxTileLines; *hit_words = G32::kHitWords; }
    else { *tile_bytes = G24::kTileBytes; *nl_stride = G24::kMaxTileLines; *hit_words = G24::kHitWords; }
}

cudaError_t launch_scan_raw(const ScanParams &P, cudaStream_t stream, int *launches, cudaEvent_t after_scan)
{
    int dev = 0;
    cudaGetDevice(&dev);
    if (((uintptr_t)P.buf & 15) != 0) return cudaErrorMisalignedAddress;     // bulk copies need 16-byte alignment
    if (P.ntiles == 0) return cudaSuccess;
    if (P.ntiles > 0xffffffffull) return cudaErrorInvalidValue;
    cudaError_t e;
    const bool kmer = mode_is_kmer(P.mode);
    if (kmer) e = launch_scan_kmer(P, stream, dev);
    else if (P.mode == kModeExact) e = launch_scan_tiles<G24, kModeExact, false>(P, 0, stream, dev, 0);
    else if (P.mode == kModeLit9) e = launch_scan_tiles<G24, kModeLit9, false>(P, sizeof(LitSmem), stream, dev, 0);
    else if (P.mode == kModeLit8) e = launch_scan_tiles<G24, kModeLit8, false>(P, sizeof(LitSmem), stream, dev, 0);
    else if (P.mode == kModeLitSmall) e = launch_scan_tiles<G24, kModeLitSmall, false>(P, sizeof(LitSmem), stream, dev, 0);
    else e = launch_scan_tiles<G24, kModeIndex, false>(P, 0, stream, dev, 0);
    if (e != cudaSuccess) return e;
    if (after_scan) cudaEventRecord(after_scan, stream);
    const unsigned int nblk = (unsigned int)((P.ntiles + kPrefixSpan - 1) / kPrefixSpan);
    tile_prefix_local_kernel<<<nblk, kPrefixThreads, 0, stream>>>(P);
    tile_prefix_blocks_kernel<<<1, kPrefixThreads, 0, stream>>>(P, nblk);
    tile_phase_check_kernel<<<(unsigned int)((P.ntiles + 255) / 256), 256, 0, stream>>>(P);
    // REDO: the list is normally empty (CTAs beyond its length exit at once); in exact mode it also holds the
    // tiles without a phase anchor (long reads), so that launch gets the full persistent grid
    if (kmer) e = launch_scan_kmer_redo(P, stream, dev);
    else if (P.mode == kModeExact) e = launch_scan_tiles<G24, kModeExact, true>(P, 0, stream, dev, 0);
    else e = launch_scan_tiles<G24, kModeIndex, true>(P, 0, stream, dev, 64);
    if (e != cudaSuccess) return e;
    tile_finalize_kernel<<<(unsigned int)((P.ntiles + 255) / 256), 256, 0, stream>>>(P);
    if (launches) *launches += 6;
    if (kmer) {
        if (P.kmer.filter_log2_words > 13) kmer_resolve_kernel<32><<<(unsigned int)((P.ntiles + 7) / 8), 256, 0, stream>>>(P);
        else kmer_resolve_kernel<8><<<(unsigned int)((P.ntiles + 31) / 32), 256, 0, stream>>>(P);
        if (launches) *launches += 1;
    } else if (P.mode == kModeExact) {
        exact_resolve_kernel<<<(unsigned int)((P.ntiles + 8 * (32 / kExactLanesPerTile) - 1) / (8 * (32 / kExactLanesPerTile))), 256, 0, stream>>>(P);
        if (launches) *launches += 1;
    }
    if (P.mode != kModeIndex) {          // index-only: the field matcher that follows counts its own hits
        count_hits_kernel<<<148 * 4, 256, 0, stream>>>(P);
        if (launches) *launches += 1;
    }
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
    if (k >= (unsigned long long)P.nl_stride) return 0ull;          // dense tile: the batch is rejected anyway
    return lo * (unsigned long long)P.tile_bytes + P.tile_nl[lo * (unsigned long long)P.nl_stride + k] + 1ull;
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
    int *run_pid;                   // routing: pattern id per run (P.rec_first); implies per_record
    int per_record;                 // every selected record is its own run (routing; mates from two files)
    unsigned int nblocks;           // blocks of 1024 hit words the batch's record capacity covers (bounds every walk)
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
    // a fixed grid walks the blocks of 1024 hit words that hold records (the record CAPACITY of a batch -- bytes /
    // smallest record -- is 7 x what 150-base reads fill: one CTA per capacity block cost 40 us in empty CTAs)
    for (unsigned long long vb = blockIdx.x; vb < C.nblocks && vb * kCompactThreads * 32ull < nrec; vb += gridDim.x) {
    const unsigned long long w = vb * kCompactThreads + threadIdx.x;
    unsigned int s = sel_word(C, w, nrec);
    unsigned int prev_msb = (w > 0) ? (sel_word(C, w - 1, nrec) >> 31) : 0u;
    unsigned int starts = C.per_record ? s : s & ~((s << 1) | prev_msb);
    unsigned long long a = __popc(starts), b = __popc(s);
    for (int d = 16; d > 0; d >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, d); b += __shfl_xor_sync(0xffffffffu, b, d); }
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = a; red[1][threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long ta = 0, tb2 = 0;
        for (int i = 0; i < kCompactThreads / 32; ++i) { ta += red[0][i]; tb2 += red[1][i]; }
        C.block_sums[vb] = ta;
        if (tb2) atomicAdd(&C.out_counts[1], tb2);
    }
    __syncthreads();
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
    for (unsigned long long vb = blockIdx.x; vb < C.nblocks && vb * kCompactThreads * 32ull < nrec; vb += gridDim.x) {
    const unsigned long long w = vb * kCompactThreads + threadIdx.x;
    unsigned int s = sel_word(C, w, nrec);
    unsigned int prev_msb = (w > 0) ? (sel_word(C, w - 1, nrec) >> 31) : 0u;
    unsigned int next_lsb = sel_word(C, w + 1, nrec) & 1u;
    unsigned int starts = C.per_record ? s : s & ~((s << 1) | prev_msb);
    unsigned int ends = C.per_record ? s : s & ~((s >> 1) | (next_lsb << 31));
    unsigned long long n = __popc(starts), inc = n;
    for (int d = 1; d < 32; d <<= 1) { unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d); if ((threadIdx.x & 31) >= d) inc += o; }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
    __syncthreads();
    unsigned long long woff = 0;
    for (int k = 0; k < (int)(threadIdx.x >> 5); ++k) woff += wsum[k];
    unsigned long long excl = C.block_sums[vb] + woff + inc - n;   // run starts before this word
    // a run that is open at the start of this word has index excl-1
    unsigned int m = starts | ends;
    unsigned long long opened = excl;
    while (m) {
        int b = __ffs(m) - 1;
        m &= m - 1;
        unsigned long long rec = w * 32ull + b;
        if (starts & (1u << b)) {
            if (opened < C.run_cap) {
                C.runs[opened].off = offset_after_line(C.P, 4ll * (long long)rec - 1);
                if (C.run_pid) {
                    const unsigned long long f = C.P.rec_first[rec];
                    C.run_pid[opened] = f == ~0ull ? -1 : (int)(f & 0xffffffull);
                }
            }
            ++opened;
        }
        if (ends & (1u << b)) {
            unsigned long long r = opened - 1;
            if (r < C.run_cap) C.runs[r].len = offset_after_line(C.P, 4ll * (long long)rec + 3);   // end offset; fixed up below
        }
    }
    __syncthreads();                 // wsum is reused by the next block of words
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
                           fxg_run *runs, unsigned long long run_cap, int invert, int *run_pid, int per_record,
                           cudaStream_t stream, int *launches)
{
    CompactParams C;
    C.hitbits = hitbits; C.P = P; C.ctr = ctr; C.block_sums = block_sums;
    C.out_counts = out_counts; C.runs = runs; C.run_cap = run_cap; C.invert = invert; C.run_pid = run_pid;
    C.per_record = (per_record || run_pid) ? 1 : 0;
    unsigned long long words = (rec_cap + 31) / 32;
    unsigned int nblocks = (unsigned int)((words + kCompactThreads - 1) / kCompactThreads);
    if (nblocks == 0) nblocks = 1;
    C.nblocks = nblocks;
    const unsigned int grid = nblocks < 148u * 16u ? nblocks : 148u * 16u;   // enough CTAs to hide the look-ups of run-dense batches
    compact_count_kernel<<<grid, kCompactThreads, 0, stream>>>(C);
    compact_scan_kernel<<<1, 1024, 0, stream>>>(C, nblocks);
    compact_write_kernel<<<grid, kCompactThreads, 0, stream>>>(C);
    compact_fixup_kernel<<<148, 256, 0, stream>>>(C);
    if (launches) *launches += 4;
    return cudaGetLastError();
}

// ================================================================= mates: per-record hits -> per-pair selection
// Fxstream::read (fx.cpp:58-77) hands the strategies one pair at a time; what they do with the two mates' hits
// (m1, m2) reduces, for everything but -l, to one selection bit per pair:
//   simple_string_search (main.cpp:599-648)   select = -v ? !m1 : m1 | m2          (-v never looks at read 2, :618-619)
//   hash_search          (main.cpp:311-346)   select = -v ? !m1 : m1 | m2          (a read-2 hit prints even under -v, :331-337)
//   multipattern_search  (main.cpp:388-446)   select = -v ? !(m1 | m2) : m1 | m2   (:427-438)
// pair_select_kernel evaluates that rule 32 pairs per thread from the finished per-record hit bit set and counts the
// selected pairs; pair_spread_kernel writes the selection back as one bit per RECORD (both mates of a selected
// pair), so that the compaction emits their byte runs in input order.  Layouts: interleaved -- mates are records
// 2i, 2i+1; two files -- the batch is file 1's R records followed by file 2's R records, mates are i and R + i.
struct PairParams {
    unsigned int *hitbits;          // in: hit bit per record; out (pair_spread_kernel): selected bit per record
    unsigned int *pairsel;          // scratch: selected bit per pair
    ScanParams P;
    unsigned long long rec_cap;
    unsigned long long split;       // two files: byte offset of the second part; 0 = interleaved
    PairRule rule;
};

__device__ __forceinline__ unsigned int even_bits(unsigned int x)     // bits 0,2,4,... of x -> bits 0..15
{
    x &= 0x55555555u;
    x = (x | (x >> 1)) & 0x33333333u;
    x = (x | (x >> 2)) & 0x0f0f0f0fu;
    x = (x | (x >> 4)) & 0x00ff00ffu;
    return (x | (x >> 8)) & 0xffffu;
}
__device__ __forceinline__ unsigned int twice_bits(unsigned int x)    // bit i of x (i < 16) -> bits 2i and 2i+1
{
    x = (x | (x << 8)) & 0x00ff00ffu;
    x = (x | (x << 4)) & 0x0f0f0f0fu;
    x = (x | (x << 2)) & 0x33333333u;
    x = (x | (x << 1)) & 0x55555555u;
    return x | (x << 1);
}
// 32 bits of a bit set starting at bit `first` (the set is readable one word past its end)
__device__ __forceinline__ unsigned int bits_at(const unsigned int *set, unsigned long long first)
{
    const unsigned long long q = first >> 5;
    return __funnelshift_r(set[q], set[q + 1], (unsigned int)(first & 31ull));
}

__global__ void __launch_bounds__(256) pair_select_kernel(const PairParams C)
{
    __shared__ unsigned int wsum[8];
    unsigned long long nrec = C.P.ctr->n_lines >> 2;
    if (nrec > C.rec_cap) nrec = C.rec_cap;
    const unsigned long long npairs = nrec >> 1;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        // whole pairs only; two files: the second part starts exactly where record `npairs` starts
        bool bad = (nrec & 1ull) != 0ull;
        if (!bad && C.split && offset_after_line(C.P, 4ll * (long long)npairs - 1) != C.split) bad = true;
        if (bad) atomicOr(&C.P.ctr->violation, (unsigned int)FXG_V_PAIRING);
    }
    const unsigned long long words = (npairs + 31) >> 5;
    unsigned int n = 0;
    for (unsigned long long w = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (unsigned long long)gridDim.x * blockDim.x) {
        unsigned int a, b;
        if (C.rule.all_match) a = b = 0xffffffffu;
        else if (C.split) { a = C.hitbits[w]; b = bits_at(C.hitbits, npairs + w * 32ull); }
        else {
            const unsigned int h0 = C.hitbits[2 * w], h1 = C.hitbits[2 * w + 1];
            a = even_bits(h0) | (even_bits(h1) << 16);
            b = even_bits(h0 >> 1) | (even_bits(h1 >> 1) << 16);
        }
        unsigned int sel = C.rule.invert ? (C.rule.need_neither ? ~(a | b) : ~a) : (a | b);
        const unsigned long long left = npairs - w * 32ull;
        if (left < 32ull) sel &= (1u << left) - 1u;
        C.pairsel[w] = sel;
        n += __popc(sel);
    }
    // one zero word behind the set: pair_spread_kernel reads one word past the last
    if (blockIdx.x == 0 && threadIdx.x == 0) C.pairsel[words] = 0u;
    n = __reduce_add_sync(0xffffffffu, n);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int tot = 0;
        for (int i = 0; i < 8; ++i) tot += wsum[i];
        if (tot) atomicAdd(&C.P.ctr->n_pair_sel, (unsigned long long)tot);
    }
}

__global__ void __launch_bounds__(256) pair_spread_kernel(const PairParams C)
{
    unsigned long long nrec = C.P.ctr->n_lines >> 2;
    if (nrec > C.rec_cap) nrec = C.rec_cap;
    const unsigned long long npairs = nrec >> 1;
    const unsigned long long words = (nrec + 31) >> 5;       // an odd last record (rejected batch) stays unselected
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < words; k += (unsigned long long)gridDim.x * blockDim.x) {
        unsigned int out;
        if (!C.split) out = twice_bits((C.pairsel[k >> 1] >> ((unsigned int)(k & 1ull) * 16u)) & 0xffffu);
        else {
            const unsigned long long first = k * 32ull;       // records first .. first + 31
            out = 0u;
            if (first < npairs) {                              // file 1's part: pair i is record i
                out = C.pairsel[k];
                const unsigned long long left = npairs - first;
                if (left < 32ull) out &= (1u << left) - 1u;
            }
            if (first + 32ull > npairs) {                      // file 2's part: pair i is record npairs + i
                if (first >= npairs) out |= bits_at(C.pairsel, first - npairs);
                else out |= C.pairsel[0] << (unsigned int)(npairs - first);
            }
        }
        C.hitbits[k] = out;
    }
}

cudaError_t launch_pair_select(unsigned int *hitbits, unsigned int *pairsel, const ScanParams &P, unsigned long long rec_cap,
                               unsigned long long split, PairRule rule, cudaStream_t stream, int *launches)
{
    PairParams C;
    C.hitbits = hitbits; C.pairsel = pairsel; C.P = P; C.rec_cap = rec_cap; C.split = split; C.rule = rule;
    pair_select_kernel<<<148 * 2, 256, 0, stream>>>(C);
    pair_spread_kernel<<<148 * 2, 256, 0, stream>>>(C);
    if (launches) *launches += 2;
    return cudaGetLastError();
}

}  // namespace fxg
