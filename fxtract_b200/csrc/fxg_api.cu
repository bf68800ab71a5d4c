// fxg_api.cu -- context, pattern tables and the C ABI of libfxgpu.so (include/fxgpu.h).
#include "fxg_common.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <vector>

namespace fxg {

static std::string g_last_error;

struct Slot {
    void *pinned = nullptr;             // input staging (host path)
    uint8_t *d_in = nullptr; size_t d_in_cap = 0;   // device input buffer of this slot (host path: H2D target)
    cudaEvent_t h2d_done = nullptr;     // recorded on the copy stream after the slot's H2D
    ScanCounters *h_ctr = nullptr;      // pinned result staging
    unsigned long long *h_counts = nullptr;   // [2]: n_runs, n_selected (K6)
    fxg_run *h_runs = nullptr; size_t h_runs_cap = 0;
    fxg_run *d_runs = nullptr; size_t d_runs_cap = 0;
    int *d_run_pid = nullptr, *h_run_pid = nullptr; size_t h_run_pid_cap = 0;     // routing: pattern id per run
    uint8_t *h_flags = nullptr; int *h_pid = nullptr; uint32_t *h_off = nullptr; size_t fields_cap = 0;
    cudaEvent_t done = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};   // profiling: pipeline begin, scan begin, scan end, pipeline end
    bool timed = false;
    int state = 0;                      // 0 free, 1 lent to caller, 2 submitted
    int kind = 0;                       // 0 raw, 1 fields
    size_t nbytes = 0; uint32_t nrec = 0; int is_last = 0;
    size_t runs_copied = 0;
    int pair = 0;                       // raw batch of mates (fxg_submit_raw_interleaved_from / _pair_from)
    uint64_t split = 0;                 // ... two files: bytes of file 1's part; 0 = interleaved
};

}  // namespace fxg

struct fxg_ctx {
    fxg_config cfg;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    // patterns
    std::vector<uint8_t> blob; std::vector<uint32_t> off; uint32_t npat = 0;
    bool have_patterns = false;
    fxg::LiteralTable h_lit;
    fxg::LiteralTable *d_lit = nullptr;
    uint8_t *d_pat = nullptr; uint32_t *d_patoff = nullptr;
    uint8_t *d_cls = nullptr; int *d_next = nullptr; int *d_out = nullptr; int ac_ncls = 0, ac_nstates = 0, ac_wide = 0;
    unsigned int *d_kfilter = nullptr, *d_kkey = nullptr, *d_kval = nullptr;      // packed k-mer index of the field engine (ACGT pattern sets)
    unsigned int k_mask = 0, k_kmask = 0; int k_k = 0, k_stride = 0; bool kmer_ok = false;
    unsigned int *d_ktfilter = nullptr; uint4 *d_ktht = nullptr, *d_ktrec = nullptr;  // window index of the fused scan matcher (KmerTile)
    fxg::KmerTile kt = {}; bool kt_ok = false;
    int scan_mode = 0;                  // kMode* the raw path runs with (fixed by fxg_set_patterns)
    int tile_bytes = 0, nl_stride = 0, hit_words = 0;     // geometry of that mode
    bool two_stage = false;             // FXG_TWO_STAGE=1: multi / exact keep the index pass + field matcher (A/B measurements)
    cudaStream_t copy_stream = nullptr; // H2D of batch k+1 overlaps the kernels of batch k
    uint4 *d_bkt = nullptr; unsigned int exact_mask = 0;                          // bucketed exact table
    unsigned int *d_exact_pre = nullptr; unsigned int exact_pre_mask = 0;         // its Bloom pre-filter
    unsigned long long *d_keyhash = nullptr;                                      // exact mode: key hash per target line (scan -> resolve)
    unsigned long long *d_rec_first = nullptr;                                    // routing: first reported hit per record
    unsigned int *d_pairsel = nullptr; size_t pairsel_cap = 0;                    // mates: selected bit per pair (scratch)
    uint2 *d_cand = nullptr; unsigned int *d_tile_ncand = nullptr;                // k-mer mode: filter hits per tile (scan -> resolve)
    unsigned long long *d_key_words = nullptr;
    // scan resources (shared; everything is ordered on one stream)
    size_t cap_bytes = 0; size_t rec_cap = 0; size_t ntiles_cap = 0; size_t nblocks_cap = 0;
    fxg::TileInfo *d_info = nullptr; unsigned short *d_tile_nl = nullptr; unsigned int *d_tile_hit = nullptr;
    unsigned long long *d_lbase = nullptr, *d_bsum = nullptr;
    unsigned int *d_redo = nullptr;      // [ntiles] list + [1] count at index ntiles_cap
    fxg::ScanCounters *d_ctr = nullptr; unsigned int *d_hitbits = nullptr;
    unsigned long long *d_fld_off = nullptr, *d_block_sums = nullptr, *d_counts = nullptr;
    unsigned int *d_fld_len = nullptr;
    // host path
    uint8_t *d_flags = nullptr; int *d_pid = nullptr; uint32_t *d_off32 = nullptr; size_t d_fields_cap = 0;
    fxg::Slot slots[3];                 // 0,1: host path; 2: device-resident path
    std::deque<int> inflight;
    uint64_t total_selected = 0;
    int last_launches = 0;
    uint64_t last_d2h_bytes = 0;        // bytes the last collected raw batch brought back (counters + runs)
    bool profiling = false;
    float last_scan_ms = 0.f, last_pipeline_ms = 0.f;
};

namespace fxg {

int set_error(fxg_ctx *ctx, int code, const std::string &msg)
{
    g_last_error = msg;
    if (ctx) ctx->err = msg;
    return code;
}
int set_cuda_error(fxg_ctx *ctx, cudaError_t e, const char *what)
{
    return set_error(ctx, FXG_ERR_CUDA, std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what);
}

template <typename T> static cudaError_t dev_realloc(T *&p, size_t n)
{
    if (p) { cudaFree(p); p = nullptr; }
    if (n == 0) return cudaSuccess;
    return cudaMalloc((void **)&p, n * sizeof(T));
}
template <typename T> static cudaError_t host_realloc(T *&p, size_t n)
{
    if (p) { cudaFreeHost(p); p = nullptr; }
    if (n == 0) return cudaSuccess;
    return cudaHostAlloc((void **)&p, n * sizeof(T), cudaHostAllocDefault);
}

// ---------------------------------------------------------------- literal table (host)
static void build_literal_table(fxg_ctx *ctx)
{
    LiteralTable &T = ctx->h_lit;
    memset(&T, 0, sizeof T);
    T.n_patterns = (int)ctx->npat;
    int minlen = 1 << 30;
    bool any_valid = false;
    for (uint32_t q = 0; q < ctx->npat && q < 2; ++q) {
        int len = (int)(ctx->off[q + 1] - ctx->off[q]);
        T.pat_len[q] = len; T.pat_off[q] = (int)ctx->off[q];
        if (len == 0) T.all_match = 1;
        bool has_nl = memchr(ctx->blob.data() + ctx->off[q], '\n', (size_t)len) != nullptr;
        if (has_nl) { T.pat_len[q] = -1; continue; }      // can never occur inside a field
        any_valid = true;
        if (len < minlen) minlen = len;
    }
    if (!any_valid) { T.never = 1; T.unit = 8; return; }
    if (T.all_match) { T.unit = 8; return; }
    T.unit = minlen >= 15 ? 8 : (minlen >= 7 ? 4 : (minlen >= 3 ? 2 : 1));
    const int U = T.unit;
    std::vector<unsigned long long> keys; std::vector<unsigned short> masks;
    for (uint32_t q = 0; q < ctx->npat && q < 2; ++q) {
        if (T.pat_len[q] < 0) continue;
        const uint8_t *p = ctx->blob.data() + ctx->off[q];
        for (int s = 0; s < U; ++s) {
            unsigned long long k = 0;
            for (int b = 0; b < U; ++b) k |= (unsigned long long)p[s + b] << (8 * b);
            size_t i = 0;
            for (; i < keys.size(); ++i) if (keys[i] == k) break;
            if (i == keys.size()) { keys.push_back(k); masks.push_back(0); }
            masks[i] |= (unsigned short)(1u << (q * 8 + s));
        }
    }
    T.n_keys = (int)keys.size();
    if (U == 8) {
        unsigned long long rng = 0x243F6A8885A308D3ull;
        // Perfect hash of the keys into the fingerprint table, cheapest form first:
        //   attempts    0.. 8191: mul1 == 1 (fingerprint = lo + hi * mul2: one IMAD per unit, kernel variant LM 9),
        //                         32 slots (one per shared-memory bank: the look-up never conflicts);
        //   attempts 8192..12287: mul1 == 1, 64 slots;   later: two multipliers, 64 slots.
        T.shift = 26;
        for (int attempt = 0; attempt < 1000000; ++attempt) {
            rng = rng * 6364136223846793005ull + 1442695040888963407ull;
            unsigned int m1 = attempt < 12288 ? 1u : ((unsigned int)(rng >> 32) | 1u);
            rng = rng * 6364136223846793005ull + 1442695040888963407ull;
            unsigned int m2 = (unsigned int)(rng >> 32) | 1u;
            const int shift = attempt < 8192 && keys.size() <= 32 ? 27 : 26;
            bool used[64] = {false}; bool ok = true;
            for (size_t i = 0; i < keys.size() && ok; ++i) {
                unsigned int lo = (unsigned int)keys[i], hi = (unsigned int)(keys[i] >> 32);
                unsigned int h = (lo * m1 + hi * m2) >> shift;
                if (used[h]) ok = false; else used[h] = true;
            }
            if (ok) { T.mul1 = m1; T.mul2 = m2; T.shift = shift; break; }
        }
        for (unsigned int h = 0; h < 64; ++h) T.fp[h] = (h ^ 1u) << T.shift;      // top bits != h: can never match
        for (size_t i = 0; i < keys.size(); ++i) {
            unsigned int lo = (unsigned int)keys[i], hi = (unsigned int)(keys[i] >> 32);
            unsigned int r = lo * T.mul1 + hi * T.mul2;
            T.keys[r >> T.shift] = keys[i]; T.fp[r >> T.shift] = r; T.masks[r >> T.shift] = masks[i];
        }
    } else {
        for (size_t i = 0; i < keys.size() && i < 64; ++i) { T.keys[i] = keys[i]; T.masks[i] = masks[i]; }
    }
    // patterns that cannot match keep pat_len -1: verification skips them (no key bits set)
}

// ---------------------------------------------------------------- Aho-Corasick DFA (host build)
static int build_ac(fxg_ctx *ctx)
{
    const uint32_t n = ctx->npat;
    std::vector<uint8_t> cls(256, 0);
    int ncls = 1;
    {
        bool used[256] = {false};
        for (size_t i = 0; i < ctx->blob.size(); ++i) used[ctx->blob[i]] = true;
        for (int b = 0; b < 256; ++b) if (used[b]) cls[b] = (uint8_t)(ncls++);
    }
    size_t max_states = ctx->blob.size() + 1;
    if ((double)max_states * ncls * 4.0 > 24e9) return set_error(ctx, FXG_ERR_UNSUPPORTED, "pattern set too large for the DFA engine");
    std::vector<int> next(max_states * (size_t)ncls, -1), term(max_states, -1);
    size_t ns = 1;
    for (uint32_t i = 0; i < n; ++i) {
        size_t s = 0;
        for (uint32_t j = ctx->off[i]; j < ctx->off[i + 1]; ++j) {
            int c = cls[ctx->blob[j]];
            int t = next[s * ncls + c];
            if (t < 0) { t = (int)ns++; next[s * ncls + c] = t; }
            s = (size_t)t;
        }
        if (term[s] < 0) term[s] = (int)i;
    }
    next.resize(ns * (size_t)ncls); term.resize(ns);
    std::vector<int> fail(ns, 0), out(ns, -1), queue; queue.reserve(ns);
    out[0] = term[0];
    for (int c = 0; c < ncls; ++c) {
        int t = next[c];
        if (t < 0) next[c] = 0; else { fail[t] = 0; queue.push_back(t); }
    }
    for (size_t qh = 0; qh < queue.size(); ++qh) {
        int s = queue[qh], f = fail[s];
        out[s] = term[s] >= 0 ? term[s] : out[f];     // longest pattern that is a suffix of this state
        for (int c = 0; c < ncls; ++c) {
            int t = next[(size_t)s * ncls + c];
            if (t < 0) next[(size_t)s * ncls + c] = next[(size_t)f * ncls + c];
            else { fail[t] = next[(size_t)f * ncls + c]; queue.push_back(t); }
        }
    }
    // transition entries: next state | high bit "the next state reports a pattern"; 16 bits while the states fit
    const bool wide = ns >= 32768;
    std::vector<unsigned short> n16; std::vector<unsigned int> n32;
    if (wide) {
        n32.resize(next.size() + 4, 0u);
        for (size_t i = 0; i < next.size(); ++i) n32[i] = (unsigned int)next[i] | (out[next[i]] >= 0 ? 0x80000000u : 0u);
    } else {
        n16.resize(next.size() + 8, 0);
        for (size_t i = 0; i < next.size(); ++i) n16[i] = (unsigned short)((unsigned int)next[i] | (out[next[i]] >= 0 ? 0x8000u : 0u));
    }
    cudaError_t e;
    if ((e = dev_realloc(ctx->d_cls, 256)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc cls");
    if ((e = dev_realloc(ctx->d_next, wide ? n32.size() : (n16.size() + 1) / 2)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc dfa");
    if ((e = dev_realloc(ctx->d_out, out.size())) != cudaSuccess) return set_cuda_error(ctx, e, "alloc dfa out");
    if ((e = cudaMemcpy(ctx->d_cls, cls.data(), 256, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload dfa classes");
    if ((e = cudaMemcpy(ctx->d_next, wide ? (const void *)n32.data() : (const void *)n16.data(), wide ? n32.size() * 4 : n16.size() * 2,
                        cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload dfa");
    if ((e = cudaMemcpy(ctx->d_out, out.data(), out.size() * sizeof(int), cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload dfa");
    ctx->ac_wide = wide ? 1 : 0;
    ctx->ac_ncls = ncls; ctx->ac_nstates = (int)ns;
    return FXG_OK;
}

// ---------------------------------------------------------------- packed k-mer index (host build)
// Usable when every pattern is a string over {A,C,G,T} of length >= 12 (fxg_common.cuh: KmerDevice).
static int build_kmer(fxg_ctx *ctx)
{
    ctx->kmer_ok = false;
    const uint32_t n = ctx->npat;
    uint32_t minlen = 0xffffffffu;
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t len = ctx->off[i + 1] - ctx->off[i];
        if (len < minlen) minlen = len;
        for (uint32_t j = ctx->off[i]; j < ctx->off[i + 1]; ++j) if (!fxg_is_acgt(ctx->blob[j])) return FXG_OK;
    }
    if (n == 0 || minlen < 12) return FXG_OK;
    const int k = minlen < 16 ? (int)minlen : 16;
    int stride = (int)minlen - k + 1;                 // any stride <= min length - k + 1 is complete; the kernel has 1, 2 and 4
    stride = stride >= 4 ? 4 : (stride >= 2 ? 2 : 1);
    const unsigned int kmask = k == 16 ? 0xffffffffu : ((1u << (2 * k)) - 1u);
    size_t entries = (size_t)n * (size_t)stride, nslots = 16;
    while (nslots < entries * 2) nslots <<= 1;
    if (nslots > ((size_t)1 << 30)) return FXG_OK;
    std::vector<unsigned int> filter(kKmerFilterWords, 0u), hk(nslots, 0u), hv(nslots, 0u);
    const unsigned int mask = (unsigned int)(nslots - 1);
    for (uint32_t i = 0; i < n; ++i) {
        const uint8_t *p = ctx->blob.data() + ctx->off[i];
        for (int o = 0; o < stride; ++o) {
            unsigned int key = 0;
            for (int b = 0; b < k; ++b) key = ((key << 2) | fxg_base_code(p[o + b])) & kmask;
            const unsigned int h = fxg_kmer_hash(key);
            const unsigned int fb = h >> (32 - kKmerFilterBits);
            filter[fb >> 5] |= 1u << (fb & 31);
            const unsigned int gb = fxg_kmer_hash2(key) >> (32 - kKmerFilterBits);
            filter[gb >> 5] |= 1u << (gb & 31);
            unsigned int slot = (h >> 7) & mask;
            while (hv[slot]) slot = (slot + 1) & mask;
            hk[slot] = key; hv[slot] = ((i << 5) | (unsigned int)o) + 1u;
        }
    }
    if (n >= (1u << 26)) return FXG_OK;
    cudaError_t e;
    if ((e = dev_realloc(ctx->d_kfilter, filter.size())) != cudaSuccess) return set_cuda_error(ctx, e, "alloc k-mer filter");
    if ((e = dev_realloc(ctx->d_kkey, nslots)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc k-mer table");
    if ((e = dev_realloc(ctx->d_kval, nslots)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc k-mer table");
    if ((e = cudaMemcpy(ctx->d_kfilter, filter.data(), filter.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload k-mer filter");
    if ((e = cudaMemcpy(ctx->d_kkey, hk.data(), nslots * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload k-mer table");
    if ((e = cudaMemcpy(ctx->d_kval, hv.data(), nslots * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload k-mer table");
    ctx->k_mask = mask; ctx->k_kmask = kmask; ctx->k_k = k; ctx->k_stride = stride; ctx->kmer_ok = true;
    return FXG_OK;
}

// ---------------------------------------------------------------- window index of the fused scan matcher (host build)
// Usable when no pattern holds a '\n' and the shortest one is long enough for a probe stride S of 16, 8 or 4 bytes
// with a window of k <= 16 bytes, k + S - 1 <= shortest pattern (fxg_common.cuh: KmerTile): every pattern contributes
// the S windows at its first S offsets, so every occurrence holds a window at a probed text position.  The largest
// S is taken whose k keeps the share of text windows that equal an index entry by chance (entries / 4^k for a
// four-letter text) under 1/64 -- those cost a table look-up and a byte comparison each.  Any byte values: the
// windows are hashed as raw bytes.
static int build_kmer_tile(fxg_ctx *ctx)
{
    ctx->kt_ok = false;
    const uint32_t n = ctx->npat;
    if (n == 0 || n >= (1u << 24)) return FXG_OK;
    uint32_t minlen = 0xffffffffu;
    for (uint32_t i = 0; i < n; ++i) minlen = std::min(minlen, ctx->off[i + 1] - ctx->off[i]);
    if (!ctx->blob.empty() && memchr(ctx->blob.data(), '\n', ctx->blob.size())) return FXG_OK;
    int S = 0, k = 0;
    for (int s = 16; s >= 4; s >>= 1) {
        if ((int)minlen - s + 1 < 11) continue;
        const int kk = std::min(16, (int)minlen - s + 1);
        const double space = kk >= 16 ? 4294967296.0 : (double)(1ull << (2 * kk));
        if ((double)n * s / space > 1.0 / 64) continue;
        S = s; k = kk;
        break;
    }
    if (!S) return FXG_OK;
    const size_t entries = (size_t)n * (size_t)S;
    if (entries > 400000) return FXG_OK;              // the 128 KiB filter would pass more than ~ 1 window in 6
    size_t nbuckets = 16;                             // 4 entries per bucket, <= 1 on average
    while (nbuckets < entries) nbuckets <<= 1;
    int flw = entries <= 32768 ? 13 : 15;             // 32 KiB filter (2 CTAs / SM) or 128 KiB (1 CTA / SM); 4 bits per entry
    if (const char *f = getenv("FXG_FORCE_FLW")) { if (atoi(f) == 15 || (atoi(f) == 13 && flw == 13)) flw = atoi(f); }   // experiments
    std::vector<unsigned int> filter((size_t)1 << flw, 0u), ht(nbuckets * 8, 0u);
    std::vector<unsigned char> fill(nbuckets, 0);
    const unsigned int mask = (unsigned int)(nbuckets - 1);
    // pattern records {length, id, bytes}, 16-byte aligned
    std::vector<unsigned int> rec;
    std::vector<unsigned int> rec_at(n, 0);
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t len = ctx->off[i + 1] - ctx->off[i];
        rec_at[i] = (unsigned int)(rec.size() / 4);
        rec.push_back(len); rec.push_back(i);
        const size_t words = (len + 3) / 4;
        for (size_t w = 0; w < words; ++w) {
            unsigned int v = 0;
            for (uint32_t b = 0; b < 4 && w * 4 + b < len; ++b) v |= (unsigned int)ctx->blob[ctx->off[i] + w * 4 + b] << (8 * b);
            rec.push_back(v);
        }
        while (rec.size() % 4) rec.push_back(0u);
    }
    for (int pad = 0; pad < 12; ++pad) rec.push_back(0u);
    if (rec.size() / 4 >= ((size_t)1 << 25)) return FXG_OK;          // record offsets are stored in 25 bits
    for (uint32_t i = 0; i < n; ++i) {
        const uint8_t *p = ctx->blob.data() + ctx->off[i];
        for (int o = 0; o < S; ++o) {                                 // offset of the window inside the pattern
            unsigned int w[4] = {0u, 0u, 0u, 0u};
            for (int b = 0; b < k; ++b) w[b >> 2] |= (unsigned int)p[o + b] << (8 * (b & 3));
            const unsigned int h = window_hash(w[0], w[1], w[2], w[3]);
            filter[kmer_tile_filter_word(h, flw)] |= kmer_tile_filter_mask(h);
            unsigned int bkt = kmer_tile_bucket(h, mask);
            while (fill[bkt] == 4) bkt = (bkt + 1) & mask;
            const int e = fill[bkt]++;
            ht[(size_t)bkt * 8 + e] = h; ht[(size_t)bkt * 8 + 4 + e] = ((rec_at[i] << 7) | (unsigned int)o) + 1u;
        }
    }
    cudaError_t e;
    if ((e = dev_realloc(ctx->d_ktfilter, filter.size())) != cudaSuccess) return set_cuda_error(ctx, e, "alloc window filter");
    if ((e = dev_realloc(ctx->d_ktht, nbuckets * 2)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc window table");
    if ((e = cudaMemcpy(ctx->d_ktfilter, filter.data(), filter.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload window filter");
    if ((e = cudaMemcpy(ctx->d_ktht, ht.data(), ht.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload window table");
    if ((e = dev_realloc(ctx->d_ktrec, rec.size() / 4)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pattern records");
    if ((e = cudaMemcpy(ctx->d_ktrec, rec.data(), rec.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload pattern records");
    KmerTile &K = ctx->kt;
    K.filter = ctx->d_ktfilter; K.ht = ctx->d_ktht; K.rec = ctx->d_ktrec;
    K.ht_mask = mask; K.k = k; K.stride = S; K.filter_log2_words = flw;
    K.mask2 = k >= 12 ? 0xffffffffu : ((1u << (8 * (k - 8))) - 1u);
    K.mask3 = k >= 16 ? 0xffffffffu : (k > 12 ? ((1u << (8 * (k - 12))) - 1u) : 0u);
    ctx->kt_ok = true;
    return FXG_OK;
}

// ---------------------------------------------------------------- exact table (host build)
static int build_exact(fxg_ctx *ctx)
{
    const uint32_t n = ctx->npat;
    // key storage: meta word (pattern id << 32 | length) + the key as zero-padded little-endian 8-byte words
    std::vector<unsigned long long> kw; std::vector<unsigned int> kwoff(n ? n : 1, 0u);
    for (uint32_t i = 0; i < n; ++i) {
        const uint8_t *kp = ctx->blob.data() + ctx->off[i];
        const uint32_t len = ctx->off[i + 1] - ctx->off[i];
        kwoff[i] = (unsigned int)kw.size();
        kw.push_back(((unsigned long long)i << 32) | len);
        for (uint32_t o = 0; o < len; o += 8) {
            unsigned long long w = 0;
            for (uint32_t b = 0; b < 8 && o + b < len; ++b) w |= (unsigned long long)kp[o + b] << (8 * b);
            kw.push_back(w);
        }
        if (kw.size() > 0xfffffff0ull) return set_error(ctx, FXG_ERR_UNSUPPORTED, "key set too large for the exact table");
    }
    for (int pad = 0; pad < 8; ++pad) kw.push_back(0);
    size_t nbuckets = 16;                             // 8 entries per bucket, <= 2 keys on average
    while (nbuckets * 2 < (size_t)n) nbuckets <<= 1;
    if (nbuckets > ((size_t)1 << 27)) return set_error(ctx, FXG_ERR_UNSUPPORTED, "key set too large for the exact table");
    std::vector<unsigned int> bk(nbuckets * 16, 0u);  // per bucket: {fp, off} x 8
    std::vector<unsigned char> fill(nbuckets, 0);
    const unsigned int mask = (unsigned int)(nbuckets - 1);
    size_t pre_words = 64;                            // >= 16 bits per key
    while (pre_words * 2 < (size_t)n) pre_words <<= 1;
    std::vector<unsigned int> pre(pre_words, 0u);
    const unsigned int pre_mask = (unsigned int)(pre_words - 1);
    for (uint32_t i = 0; i < n; ++i) {
        const unsigned long long h = fxg_hash_bytes(ctx->blob.data() + ctx->off[i], ctx->off[i + 1] - ctx->off[i]);
        pre[exact_pre_word(h, pre_mask)] |= exact_pre_mask(h);
        unsigned int bkt = (unsigned int)h & mask;
        while (fill[bkt] == 8) bkt = (bkt + 1) & mask;
        const int e = fill[bkt]++;
        bk[(size_t)bkt * 16 + 2 * e] = (unsigned int)(h >> 32) | 1u; bk[(size_t)bkt * 16 + 2 * e + 1] = kwoff[i];
    }
    cudaError_t e;
    if ((e = dev_realloc(ctx->d_bkt, nbuckets * 4)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc table");
    if ((e = cudaMemcpy(ctx->d_bkt, bk.data(), bk.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload table");
    ctx->exact_mask = mask;
    if ((e = dev_realloc(ctx->d_exact_pre, pre.size())) != cudaSuccess) return set_cuda_error(ctx, e, "alloc key pre-filter");
    if ((e = cudaMemcpy(ctx->d_exact_pre, pre.data(), pre.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload key pre-filter");
    ctx->exact_pre_mask = pre_mask;
    if ((e = dev_realloc(ctx->d_key_words, kw.size())) != cudaSuccess) return set_cuda_error(ctx, e, "alloc key words");
    if ((e = cudaMemcpy(ctx->d_key_words, kw.data(), kw.size() * 8, cudaMemcpyHostToDevice)) != cudaSuccess) return set_cuda_error(ctx, e, "upload key words");
    return FXG_OK;
}

// ---------------------------------------------------------------- resources
static int ensure_scan_capacity(fxg_ctx *ctx, size_t nbytes)
{
    if (nbytes <= ctx->cap_bytes && ctx->d_info) return FXG_OK;
    const size_t cap = std::max(nbytes, (size_t)1 << 20);
    const size_t ntiles = (cap + ctx->tile_bytes - 1) / ctx->tile_bytes;
    // a tile holds at most nl_stride lines (denser ones raise FXG_V_DENSE_LINES and the batch is re-routed), so an
    // accepted batch has at most ntiles * nl_stride / 4 records: every per-record buffer is sized with that
    const size_t rec_cap = ntiles * (size_t)(ctx->nl_stride / 4) + 64;
    const size_t words = (rec_cap + 31) / 32;
    const size_t nblocks = (words + 1023) / 1024 + 1;
    const bool field_stage = ctx->scan_mode == kModeIndex && ctx->cfg.mode != FXG_MODE_LITERAL;
    cudaError_t e;
    if ((e = dev_realloc(ctx->d_info, ntiles)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc tile info");
    if ((e = dev_realloc(ctx->d_tile_nl, ntiles * (size_t)ctx->nl_stride)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc tile line tables");
    if ((e = dev_realloc(ctx->d_lbase, ntiles + 1)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc tile prefix");
    if ((e = dev_realloc(ctx->d_bsum, ntiles / 4096 + 2)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc span prefix");
    if ((e = dev_realloc(ctx->d_redo, ntiles + 1)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc redo list");
    if (mode_is_literal(ctx->scan_mode) &&
        (e = dev_realloc(ctx->d_tile_hit, ntiles * (size_t)ctx->hit_words)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc tile line hits");
    if ((e = dev_realloc(ctx->d_hitbits, words + 4)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc hit bitset");
    if (ctx->scan_mode == kModeExact &&
        (e = dev_realloc(ctx->d_keyhash, ntiles * (size_t)(ctx->nl_stride / 4))) != cudaSuccess) return set_cuda_error(ctx, e, "alloc key hashes");
    if (mode_is_kmer(ctx->scan_mode)) {
        if ((e = dev_realloc(ctx->d_cand, ntiles * (size_t)(ctx->tile_bytes / kKmerCandDiv))) != cudaSuccess) return set_cuda_error(ctx, e, "alloc window candidates");
        if ((e = dev_realloc(ctx->d_tile_ncand, ntiles)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc window candidate counts");
    }
    if (ctx->cfg.route_runs && (e = dev_realloc(ctx->d_rec_first, rec_cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc first-hit table");
    if (!ctx->d_ctr && (e = dev_realloc(ctx->d_ctr, 1)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc counters");
    if (!ctx->d_counts && (e = dev_realloc(ctx->d_counts, 2)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc counts");
    if (!ctx->cfg.count_only) {
        if ((e = dev_realloc(ctx->d_block_sums, nblocks)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc block sums");
    }
    if (field_stage) {
        if ((e = dev_realloc(ctx->d_fld_off, rec_cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc field table");
        if ((e = dev_realloc(ctx->d_fld_len, rec_cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc field table");
    }
    ctx->cap_bytes = cap; ctx->ntiles_cap = ntiles; ctx->rec_cap = rec_cap; ctx->nblocks_cap = nblocks;
    return FXG_OK;
}

static int ensure_slot_result(fxg_ctx *ctx, Slot &s)
{
    cudaError_t e;
    if (!s.h_ctr && (e = host_realloc(s.h_ctr, 1)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned counters");
    if (!s.h_counts && (e = host_realloc(s.h_counts, 2)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned counts");
    if (!s.done && (e = cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming)) != cudaSuccess) return set_cuda_error(ctx, e, "event");
    return FXG_OK;
}

// Selected runs come back in two parts: the first kRunPrefix of them ride with the counters (one small copy,
// enqueued before the count is known); a batch with more runs fetches exactly n_runs once its count has arrived.
constexpr size_t kRunPrefix = 1024;
static bool routed(const fxg_ctx *ctx)
{
    return ctx->cfg.route_runs && !ctx->cfg.invert && !ctx->cfg.count_only &&
           (ctx->scan_mode == kModeExact || mode_is_kmer(ctx->scan_mode));
}

static int ensure_slot_runs(fxg_ctx *ctx, Slot &s, bool per_record)
{
    // plain: a run needs a selected record and an unselected one behind it; routed / mates from two files: every
    // selected record is a run
    size_t want = std::min<size_t>(per_record ? ctx->rec_cap + 2 : ctx->rec_cap / 2 + 2, (size_t)1 << 26);
    cudaError_t e;
    if (s.d_runs_cap < want) {
        if ((e = dev_realloc(s.d_runs, want)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc runs");
        if (routed(ctx) && (e = dev_realloc(s.d_run_pid, want)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc run pattern ids");
        s.d_runs_cap = want;
    }
    if (routed(ctx) && s.h_run_pid_cap < kRunPrefix) {
        if ((e = host_realloc(s.h_run_pid, kRunPrefix)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned run pattern ids");
        s.h_run_pid_cap = kRunPrefix;
    }
    if (s.h_runs_cap < kRunPrefix) {
        if ((e = host_realloc(s.h_runs, kRunPrefix)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned runs");
        s.h_runs_cap = kRunPrefix;
    }
    return FXG_OK;
}

// Enqueue the whole raw pipeline for one batch that is already in device memory.
static int enqueue_raw(fxg_ctx *ctx, const uint8_t *d_buf, size_t nbytes, Slot &s, int pair = 0, uint64_t split = 0)
{
    int rc;
    if (!ctx->have_patterns) return set_error(ctx, FXG_ERR_STATE, "fxg_set_patterns has not been called");
    if (ctx->cfg.route_runs && !ctx->cfg.invert && !ctx->cfg.count_only && !fxg_can_route(ctx))
        return set_error(ctx, FXG_ERR_UNSUPPORTED, "route_runs: this pattern set is not served by a fused matcher (see fxg_can_route)");
    if ((rc = ensure_scan_capacity(ctx, nbytes)) != FXG_OK) return rc;
    if ((rc = ensure_slot_result(ctx, s)) != FXG_OK) return rc;
    const bool emit = !ctx->cfg.count_only;
    if (pair && split >= nbytes && nbytes) return set_error(ctx, FXG_ERR_ARG, "pair batch: the second file's part is empty");
    if (emit && (rc = ensure_slot_runs(ctx, s, routed(ctx) || (pair && split))) != FXG_OK) return rc;
    cudaStream_t st = ctx->stream;
    ctx->last_launches = 0;
    s.nbytes = nbytes; s.kind = 0; s.runs_copied = 0; s.pair = pair; s.split = pair ? split : 0;
    memset(s.h_ctr, 0, sizeof(ScanCounters)); s.h_counts[0] = s.h_counts[1] = 0;
    s.timed = false;
    if (nbytes == 0) { FXG_CUDA_OK(cudaEventRecord(s.done, st)); return FXG_OK; }
    if (ctx->profiling) {
        for (cudaEvent_t &e : s.ev) if (!e) FXG_CUDA_OK(cudaEventCreate(&e));
        FXG_CUDA_OK(cudaEventRecord(s.ev[0], st));
    }

    const size_t ntiles = (nbytes + ctx->tile_bytes - 1) / ctx->tile_bytes;
    const size_t rec_cap = std::min(ctx->rec_cap, ntiles * (size_t)(ctx->nl_stride / 4) + 64);
    const size_t words = (rec_cap + 31) / 32;
    FXG_CUDA_OK(cudaMemsetAsync(ctx->d_hitbits, 0, (words + 4) * sizeof(unsigned int), st));
    FXG_CUDA_OK(cudaMemsetAsync(ctx->d_ctr, 0, sizeof(ScanCounters), st));
    FXG_CUDA_OK(cudaMemsetAsync(ctx->d_counts, 0, 2 * sizeof(unsigned long long), st));
    FXG_CUDA_OK(cudaMemsetAsync(&ctx->d_ctr->first_bad, 0xff, 8, st));      // min-reduced by the kernel
    FXG_CUDA_OK(cudaMemsetAsync(ctx->d_redo + ctx->ntiles_cap, 0, sizeof(unsigned int), st));
    ScanParams P;
    memset(&P, 0, sizeof P);
    P.buf = d_buf; P.nbytes = nbytes; P.ntiles = ntiles;
    P.tile_bytes = ctx->tile_bytes; P.nl_stride = ctx->nl_stride; P.hit_words = ctx->hit_words;
    P.info = ctx->d_info; P.tile_nl = ctx->d_tile_nl;
    P.tile_hit = ctx->d_tile_hit; P.lbase = ctx->d_lbase; P.bsum = ctx->d_bsum; P.ctr = ctx->d_ctr; P.hitbits = ctx->d_hitbits;
    P.rec_cap = rec_cap; P.one = 1;
    { const char *h = getenv("FXG_L2_HINT"); P.l2_evict_first = !(h && !strcmp(h, "0")); }
    P.redo_list = ctx->d_redo; P.redo_count = ctx->d_redo + ctx->ntiles_cap;
    P.lit = ctx->d_lit; P.pat_bytes = ctx->d_pat; P.target = ctx->cfg.target;
    const bool literal = ctx->cfg.mode == FXG_MODE_LITERAL;
    P.mode = ctx->scan_mode;
    P.match_enabled = mode_is_literal(P.mode);
    P.lit_unit = ctx->h_lit.unit;
    const bool field_stage = P.mode == kModeIndex && !literal;
    if (field_stage) { P.fld_off = ctx->d_fld_off; P.fld_len = ctx->d_fld_len; }
    P.header_whole = ctx->cfg.mode == FXG_MODE_EXACT && ctx->cfg.target == FXG_TARGET_NAME;
    if (mode_is_kmer(P.mode)) { P.kmer = ctx->kt; P.cand = ctx->d_cand; P.tile_ncand = ctx->d_tile_ncand; P.cand_cap = ctx->tile_bytes / kKmerCandDiv; }
    const bool route = routed(ctx);
    if (route) {
        FXG_CUDA_OK(cudaMemsetAsync(ctx->d_rec_first, 0xff, rec_cap * sizeof(unsigned long long), st));
        P.rec_first = ctx->d_rec_first;
    }
    if (P.mode == kModeExact) {
        P.exact = ExactDevice{ctx->d_bkt, ctx->d_key_words, ctx->exact_mask, ctx->d_exact_pre, ctx->exact_pre_mask};
        P.keyhash = ctx->d_keyhash;
        P.redo_unanchored = 1;
    }
    if (ctx->profiling) FXG_CUDA_OK(cudaEventRecord(s.ev[1], st));
    FXG_CUDA_OK(launch_scan_raw(P, st, &ctx->last_launches, ctx->profiling ? s.ev[2] : nullptr));
    if (field_stage) {
        FieldSrc src; memset(&src, 0, sizeof src);
        src.base = d_buf; src.off64 = ctx->d_fld_off; src.len32 = ctx->d_fld_len; src.n_lines_dev = &ctx->d_ctr->n_lines;
        src.stop_at_space = P.header_whole;
        FieldOut O; memset(&O, 0, sizeof O);
        O.hitbits = ctx->d_hitbits; O.n_hits = &ctx->d_ctr->n_hits;
        if (ctx->cfg.mode == FXG_MODE_MULTI && ctx->kmer_ok) {
            KmerDevice K{ctx->d_kfilter, ctx->d_kkey, ctx->d_kval, ctx->d_pat, ctx->d_patoff, ctx->k_mask, ctx->k_kmask, ctx->k_k, ctx->k_stride};
            FXG_CUDA_OK(launch_fields_kmer(src, rec_cap, K, O, st));
        } else if (ctx->cfg.mode == FXG_MODE_MULTI) {
            AcDevice A{ctx->d_cls, ctx->d_next, ctx->d_out, ctx->ac_ncls, ctx->ac_nstates, ctx->ac_wide};
            FXG_CUDA_OK(launch_fields_ac(src, rec_cap, A, O, st));
        } else {
            ExactDevice E{ctx->d_bkt, ctx->d_key_words, ctx->exact_mask, ctx->d_exact_pre, ctx->exact_pre_mask};
            FXG_CUDA_OK(launch_fields_exact(src, rec_cap, E, O, st));
        }
        ctx->last_launches += 1;
    }
    if (pair) {
        // mates: the strategy's pair rule (invert included) turns the hit bits into selected bits of both mates
        if (ctx->pairsel_cap < words / 2 + 8) {
            cudaError_t e = dev_realloc(ctx->d_pairsel, words / 2 + 8);
            if (e != cudaSuccess) return set_cuda_error(ctx, e, "alloc pair selection");
            ctx->pairsel_cap = words / 2 + 8;
        }
        PairRule rule;
        rule.invert = ctx->cfg.invert ? 1 : 0;
        rule.need_neither = ctx->cfg.mode == FXG_MODE_MULTI ? 1 : 0;
        rule.all_match = (literal && ctx->h_lit.all_match) ? 1 : 0;
        FXG_CUDA_OK(launch_pair_select(ctx->d_hitbits, ctx->d_pairsel, P, rec_cap, split, rule, st, &ctx->last_launches));
    }
    if (emit) {
        int inv = ctx->cfg.invert ? 1 : 0;
        if (literal && ctx->h_lit.all_match) inv ^= 1;     // every record matches: hit bits stay 0
        if (pair) inv = 0;                                 // the pair rule has applied it
        FXG_CUDA_OK(launch_compact(ctx->d_hitbits, P, ctx->d_ctr, rec_cap, ctx->d_block_sums, ctx->d_counts,
                                   s.d_runs, s.d_runs_cap, inv, route ? s.d_run_pid : nullptr, (pair && split) ? 1 : 0,
                                   st, &ctx->last_launches));
        if (ctx->profiling) { FXG_CUDA_OK(cudaEventRecord(s.ev[3], st)); s.timed = true; }
        FXG_CUDA_OK(cudaMemcpyAsync(s.h_counts, ctx->d_counts, 16, cudaMemcpyDeviceToHost, st));
        s.runs_copied = std::min<size_t>(kRunPrefix, s.d_runs_cap);
        FXG_CUDA_OK(cudaMemcpyAsync(s.h_runs, s.d_runs, s.runs_copied * sizeof(fxg_run), cudaMemcpyDeviceToHost, st));
        if (route) FXG_CUDA_OK(cudaMemcpyAsync(s.h_run_pid, s.d_run_pid, s.runs_copied * sizeof(int), cudaMemcpyDeviceToHost, st));
    }
    if (ctx->profiling && !s.timed) { FXG_CUDA_OK(cudaEventRecord(s.ev[3], st)); s.timed = true; }
    FXG_CUDA_OK(cudaMemcpyAsync(s.h_ctr, ctx->d_ctr, sizeof(ScanCounters), cudaMemcpyDeviceToHost, st));
    FXG_CUDA_OK(cudaEventRecord(s.done, st));
    return FXG_OK;
}

static int finish_raw(fxg_ctx *ctx, Slot &s, fxg_result *res)
{
    FXG_CUDA_OK(cudaEventSynchronize(s.done));
    if (s.timed) {
        FXG_CUDA_OK(cudaEventElapsedTime(&ctx->last_scan_ms, s.ev[1], s.ev[2]));
        FXG_CUDA_OK(cudaEventElapsedTime(&ctx->last_pipeline_ms, s.ev[0], s.ev[3]));
        s.timed = false;
    }
    memset(res, 0, sizeof *res);
    const ScanCounters &c = *s.h_ctr;
    res->n_lines = c.n_lines;
    res->n_records = c.n_lines >> 2;
    unsigned int viol = c.violation;
    if (s.nbytes && (c.last_byte != '\n' || (c.n_lines & 3))) viol |= FXG_V_TRUNCATED;
    res->violation = viol;
    res->canonical_ok = viol == 0;
    res->first_bad_offset = c.first_bad;
    res->consumed_bytes = viol == 0 ? s.nbytes : 0;
    const bool literal = ctx->cfg.mode == FXG_MODE_LITERAL;
    res->n_matched = (literal && ctx->h_lit.all_match) ? res->n_records : c.n_hits;
    res->n_selected = ctx->cfg.invert ? res->n_records - res->n_matched : res->n_matched;
    if (s.pair) {                                          // mates: the counts are in pairs
        res->n_records >>= 1;
        res->n_matched = res->n_selected = c.n_pair_sel;
    }
    uint64_t d2h = sizeof(ScanCounters);
    if (!ctx->cfg.count_only && s.nbytes) {
        d2h += 16 + s.runs_copied * sizeof(fxg_run);
        unsigned long long nr = s.h_counts[0];
        if (nr > s.d_runs_cap) return set_error(ctx, FXG_ERR_NOMEM, "selected-run list overflow; use smaller batches");
        if (nr > s.runs_copied) {
            // more runs than the prefix that rode with the counters: fetch the rest (d_runs of this slot is not
            // touched again until the slot is resubmitted)
            size_t from = s.runs_copied;
            if (s.h_runs_cap < nr) {
                cudaError_t e;
                size_t want = std::max<size_t>((size_t)nr, s.h_runs_cap * 2);
                if ((e = host_realloc(s.h_runs, want)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned runs");
                s.h_runs_cap = want;
                from = 0;
            }
            FXG_CUDA_OK(cudaMemcpy(s.h_runs + from, s.d_runs + from, ((size_t)nr - from) * sizeof(fxg_run), cudaMemcpyDeviceToHost));
            d2h += ((size_t)nr - from) * sizeof(fxg_run);
            if (routed(ctx)) {
                size_t pfrom = s.runs_copied;
                if (s.h_run_pid_cap < nr) {
                    cudaError_t e;
                    size_t want = std::max<size_t>((size_t)nr, s.h_run_pid_cap * 2);
                    if ((e = host_realloc(s.h_run_pid, want)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned run pattern ids");
                    s.h_run_pid_cap = want;
                    pfrom = 0;
                }
                FXG_CUDA_OK(cudaMemcpy(s.h_run_pid + pfrom, s.d_run_pid + pfrom, ((size_t)nr - pfrom) * sizeof(int), cudaMemcpyDeviceToHost));
                d2h += ((size_t)nr - pfrom) * sizeof(int);
            }
            s.runs_copied = (size_t)nr;
        }
        res->n_runs = nr; res->runs = s.h_runs;
        if (routed(ctx)) { res->run_pattern_id = s.h_run_pid; d2h += std::min<size_t>(kRunPrefix, (size_t)nr) * sizeof(int); }
    }
    ctx->last_d2h_bytes = d2h;
    return FXG_OK;
}

}  // namespace fxg

using namespace fxg;

// ================================================================= C ABI
extern "C" int fxg_device_count(int *n)
{
    if (!n) return FXG_ERR_ARG;
    cudaError_t e = cudaGetDeviceCount(n);
    if (e != cudaSuccess) { *n = 0; return set_cuda_error(nullptr, e, "cudaGetDeviceCount"); }
    return FXG_OK;
}

extern "C" const char *fxg_last_error(const fxg_ctx *ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

extern "C" int fxg_create(const fxg_config *cfg, fxg_ctx **out)
{
    if (!cfg || !out) return set_error(nullptr, FXG_ERR_ARG, "null argument");
    if (cfg->mode < 0 || cfg->mode > 2 || cfg->target < 0 || cfg->target > 3) return set_error(nullptr, FXG_ERR_ARG, "bad mode/target");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0)
        return set_error(nullptr, FXG_ERR_CUDA, std::string("no usable CUDA device (there is no CPU fallback): ") +
                                                    (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    if (cfg->device < 0 || cfg->device >= ndev) return set_error(nullptr, FXG_ERR_ARG, "bad device ordinal");
    if ((e = cudaSetDevice(cfg->device)) != cudaSuccess) return set_cuda_error(nullptr, e, "cudaSetDevice");
    fxg_ctx *ctx = new (std::nothrow) fxg_ctx();
    if (!ctx) return set_error(nullptr, FXG_ERR_NOMEM, "out of memory");
    ctx->cfg = *cfg;
    if (ctx->cfg.batch_bytes == 0) ctx->cfg.batch_bytes = 256ull << 20;
    if (cfg->stream) ctx->stream = (cudaStream_t)cfg->stream;
    else {
        if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) { delete ctx; return set_cuda_error(nullptr, e, "stream"); }
        ctx->own_stream = true;
    }
    *out = ctx;
    return FXG_OK;
}

extern "C" void fxg_destroy(fxg_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->cfg.device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    for (Slot &s : ctx->slots) {
        if (s.pinned) cudaFreeHost(s.pinned);
        if (s.h_ctr) cudaFreeHost(s.h_ctr);
        if (s.h_counts) cudaFreeHost(s.h_counts);
        if (s.h_runs) cudaFreeHost(s.h_runs);
        if (s.h_run_pid) cudaFreeHost(s.h_run_pid);
        if (s.d_run_pid) cudaFree(s.d_run_pid);
        if (s.h_flags) cudaFreeHost(s.h_flags);
        if (s.h_pid) cudaFreeHost(s.h_pid);
        if (s.h_off) cudaFreeHost(s.h_off);
        if (s.d_runs) cudaFree(s.d_runs);
        if (s.d_in) cudaFree(s.d_in);
        if (s.done) cudaEventDestroy(s.done);
        if (s.h2d_done) cudaEventDestroy(s.h2d_done);
        for (cudaEvent_t e : s.ev) if (e) cudaEventDestroy(e);
    }
    void *dev[] = {ctx->d_pairsel, ctx->d_rec_first, ctx->d_exact_pre, ctx->d_keyhash, ctx->d_cand, ctx->d_tile_ncand, ctx->d_ktfilter, ctx->d_ktht, ctx->d_ktrec, ctx->d_kfilter, ctx->d_kkey, ctx->d_kval, ctx->d_lit, ctx->d_pat, ctx->d_patoff, ctx->d_cls, ctx->d_next, ctx->d_out, ctx->d_bkt, ctx->d_key_words,
                   ctx->d_info, ctx->d_tile_nl, ctx->d_tile_hit, ctx->d_lbase, ctx->d_bsum, ctx->d_redo, ctx->d_ctr, ctx->d_hitbits, ctx->d_fld_off, ctx->d_block_sums, ctx->d_counts,
                   ctx->d_fld_len, ctx->d_flags, ctx->d_pid, ctx->d_off32};
    for (void *p : dev) if (p) cudaFree(p);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int fxg_set_patterns(fxg_ctx *ctx, const uint8_t *blob, const uint32_t *off, uint32_t n)
{
    if (!ctx || !off || (n && !blob && off[n])) return set_error(ctx, FXG_ERR_ARG, "null argument");
    if (n == 0) return set_error(ctx, FXG_ERR_ARG, "empty pattern set");
    if (ctx->cfg.mode == FXG_MODE_LITERAL && n > 2) return set_error(ctx, FXG_ERR_ARG, "literal mode takes 1 or 2 patterns");
    cudaSetDevice(ctx->cfg.device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    ctx->npat = n;
    ctx->off.assign(off, off + n + 1);
    ctx->blob.assign(blob, blob + off[n]);
    cudaError_t e;
    if ((e = dev_realloc(ctx->d_pat, ctx->blob.size() + 16)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc patterns");
    if ((e = dev_realloc(ctx->d_patoff, (size_t)n + 1)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc patterns");
    if (!ctx->blob.empty() && (e = cudaMemcpy(ctx->d_pat, ctx->blob.data(), ctx->blob.size(), cudaMemcpyHostToDevice)) != cudaSuccess)
        return set_cuda_error(ctx, e, "upload patterns");
    if ((e = cudaMemcpy(ctx->d_patoff, ctx->off.data(), ((size_t)n + 1) * 4, cudaMemcpyHostToDevice)) != cudaSuccess)
        return set_cuda_error(ctx, e, "upload patterns");
    int rc = FXG_OK;
    if (ctx->cfg.mode == FXG_MODE_LITERAL) {
        build_literal_table(ctx);
        if (!ctx->d_lit && (e = dev_realloc(ctx->d_lit, 1)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc literal table");
        if ((e = cudaMemcpy(ctx->d_lit, &ctx->h_lit, sizeof(LiteralTable), cudaMemcpyHostToDevice)) != cudaSuccess)
            return set_cuda_error(ctx, e, "upload literal table");
    } else if (ctx->cfg.mode == FXG_MODE_MULTI) {
        rc = build_ac(ctx);
        if (rc == FXG_OK) rc = build_kmer(ctx);
        if (rc == FXG_OK) rc = build_kmer_tile(ctx);
    } else rc = build_exact(ctx);
    if (rc != FXG_OK) return rc;
    // the scan kernel variant of the raw path
    const char *ts = getenv("FXG_TWO_STAGE");
    ctx->two_stage = ts && *ts && strcmp(ts, "0") != 0;
    int mode = kModeIndex;
    if (ctx->cfg.mode == FXG_MODE_LITERAL) {
        if (!ctx->h_lit.all_match && !ctx->h_lit.never)
            mode = ctx->h_lit.unit == 8 ? (ctx->h_lit.mul1 == 1u ? kModeLit9 : kModeLit8) : kModeLitSmall;
    } else if (ctx->cfg.mode == FXG_MODE_MULTI) {
        // fused window matcher: any-hit per line, so only whole-line targets (sequence / quality) use it;
        // everything else keeps the index pass + field matcher (DFA)
        if (ctx->kt_ok && !ctx->two_stage && (ctx->cfg.target == FXG_TARGET_SEQ || ctx->cfg.target == FXG_TARGET_QUAL)) {
            int lg = 0;
            while ((1 << lg) < ctx->kt.stride) ++lg;
            mode = kModeKmer + lg;
        }
    } else if (!ctx->two_stage) mode = kModeExact;
    {
        int tb = 0, ns = 0, hw = 0;
        scan_geometry(mode, ctx->kt_ok ? ctx->kt.filter_log2_words : 0, &tb, &ns, &hw);
        if (mode != ctx->scan_mode || tb != ctx->tile_bytes) {   // geometry may change: tile tables are re-made on the next batch
            ctx->scan_mode = mode;
            ctx->tile_bytes = tb; ctx->nl_stride = ns; ctx->hit_words = hw;
            ctx->cap_bytes = 0;
        }
    }
    ctx->have_patterns = true;
    return FXG_OK;
}

// ---------------------------------------------------------------- host path
static int free_slot(fxg_ctx *ctx)
{
    for (int i = 0; i < 2; ++i) if (ctx->slots[i].state == 0) return i;
    return -1;
}

extern "C" int fxg_acquire_buffer(fxg_ctx *ctx, void **pinned, size_t *cap)
{
    if (!ctx || !pinned || !cap) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    for (int i = 0; i < 2; ++i) if (ctx->slots[i].state == 1) return set_error(ctx, FXG_ERR_STATE, "a buffer is already lent out");
    int i = free_slot(ctx);
    if (i < 0) return set_error(ctx, FXG_ERR_STATE, "both buffers are in flight: call fxg_collect first");
    Slot &s = ctx->slots[i];
    if (!s.pinned) {
        cudaError_t e = cudaHostAlloc(&s.pinned, ctx->cfg.batch_bytes + 64, cudaHostAllocDefault);
        if (e != cudaSuccess) return set_cuda_error(ctx, e, "cudaHostAlloc(staging)");
    }
    s.state = 1;
    *pinned = s.pinned; *cap = ctx->cfg.batch_bytes;
    return FXG_OK;
}

static int ensure_d_in(fxg_ctx *ctx, Slot &s, size_t n)
{
    cudaError_t e;
    if (!s.h2d_done && (e = cudaEventCreateWithFlags(&s.h2d_done, cudaEventDisableTiming)) != cudaSuccess) return set_cuda_error(ctx, e, "event");
    if (!ctx->copy_stream && (e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return set_cuda_error(ctx, e, "copy stream");
    if (s.d_in_cap >= n + 64) return FXG_OK;
    size_t cap = std::max<size_t>(n, ctx->cfg.batch_bytes) + 64;
    if ((e = dev_realloc(s.d_in, cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc device input buffer");
    s.d_in_cap = cap;
    return FXG_OK;
}

// Host path, one batch: H2D on the copy stream into the slot's own device buffer, kernels on the compute stream
// behind an event.  With two slots the copy of batch k+1 runs while the kernels of batch k do (the slot's buffer
// is free: its previous batch has been collected).
static int submit_raw_common(fxg_ctx *ctx, int si, const void *src, size_t n1, int is_last, int pair = 0,
                             const void *src2 = nullptr, size_t n2 = 0)
{
    Slot &s = ctx->slots[si];
    int rc;
    const size_t nbytes = n1 + n2;
    if (nbytes > ctx->cfg.batch_bytes) return set_error(ctx, FXG_ERR_ARG, "batch larger than batch_bytes");
    if (!ctx->have_patterns) return set_error(ctx, FXG_ERR_STATE, "fxg_set_patterns has not been called");
    if ((rc = ensure_d_in(ctx, s, nbytes)) != FXG_OK) return rc;
    if (nbytes) {
        if (n1) FXG_CUDA_OK(cudaMemcpyAsync(s.d_in, src, n1, cudaMemcpyHostToDevice, ctx->copy_stream));
        if (n2) FXG_CUDA_OK(cudaMemcpyAsync(s.d_in + n1, src2, n2, cudaMemcpyHostToDevice, ctx->copy_stream));
        FXG_CUDA_OK(cudaEventRecord(s.h2d_done, ctx->copy_stream));
        FXG_CUDA_OK(cudaStreamWaitEvent(ctx->stream, s.h2d_done, 0));
    }
    s.is_last = is_last;
    if ((rc = enqueue_raw(ctx, s.d_in, nbytes, s, pair, src2 ? (uint64_t)n1 : 0)) != FXG_OK) return rc;
    s.state = 2;
    ctx->inflight.push_back(si);
    return FXG_OK;
}

extern "C" int fxg_submit_raw(fxg_ctx *ctx, size_t nbytes, int is_last)
{
    if (!ctx) return set_error(ctx, FXG_ERR_ARG, "null ctx");
    cudaSetDevice(ctx->cfg.device);
    int si = -1;
    for (int i = 0; i < 2; ++i) if (ctx->slots[i].state == 1) si = i;
    if (si < 0) return set_error(ctx, FXG_ERR_STATE, "no acquired buffer");
    return submit_raw_common(ctx, si, ctx->slots[si].pinned, nbytes, is_last);
}

extern "C" int fxg_submit_raw_from(fxg_ctx *ctx, const void *host_buf, size_t nbytes, int is_last)
{
    if (!ctx || (!host_buf && nbytes)) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    for (int i = 0; i < 2; ++i) if (ctx->slots[i].state == 1) return set_error(ctx, FXG_ERR_STATE, "a buffer is lent out");
    int si = free_slot(ctx);
    if (si < 0) return set_error(ctx, FXG_ERR_STATE, "both slots are in flight: call fxg_collect first");
    return submit_raw_common(ctx, si, host_buf, nbytes, is_last);
}

extern "C" int fxg_submit_raw_interleaved_from(fxg_ctx *ctx, const void *host_buf, size_t nbytes, int is_last)
{
    if (!ctx || (!host_buf && nbytes)) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    for (int i = 0; i < 2; ++i) if (ctx->slots[i].state == 1) return set_error(ctx, FXG_ERR_STATE, "a buffer is lent out");
    int si = free_slot(ctx);
    if (si < 0) return set_error(ctx, FXG_ERR_STATE, "both slots are in flight: call fxg_collect first");
    return submit_raw_common(ctx, si, host_buf, nbytes, is_last, 1);
}

extern "C" int fxg_submit_raw_pair_from(fxg_ctx *ctx, const void *host1, size_t n1, const void *host2, size_t n2, int is_last)
{
    if (!ctx || !host1 || !host2 || !n1 || !n2) return set_error(ctx, FXG_ERR_ARG, "null argument or empty part");
    cudaSetDevice(ctx->cfg.device);
    for (int i = 0; i < 2; ++i) if (ctx->slots[i].state == 1) return set_error(ctx, FXG_ERR_STATE, "a buffer is lent out");
    int si = free_slot(ctx);
    if (si < 0) return set_error(ctx, FXG_ERR_STATE, "both slots are in flight: call fxg_collect first");
    return submit_raw_common(ctx, si, host1, n1, is_last, 1, host2, n2);
}

extern "C" int fxg_submit_fields(fxg_ctx *ctx, size_t nbytes, const uint32_t *off, uint32_t nrec)
{
    if (!ctx || (!off && nrec)) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    if (!ctx->have_patterns) return set_error(ctx, FXG_ERR_STATE, "fxg_set_patterns has not been called");
    int si = -1;
    for (int i = 0; i < 2; ++i) if (ctx->slots[i].state == 1) si = i;
    if (si < 0) return set_error(ctx, FXG_ERR_STATE, "no acquired buffer");
    if (nbytes > ctx->cfg.batch_bytes) return set_error(ctx, FXG_ERR_ARG, "batch larger than batch_bytes");
    Slot &s = ctx->slots[si];
    int rc;
    cudaError_t e;
    if ((rc = ensure_slot_result(ctx, s)) != FXG_OK) return rc;
    if ((rc = ensure_d_in(ctx, s, nbytes)) != FXG_OK) return rc;
    if (s.fields_cap < (size_t)nrec + 1) {
        size_t cap = std::max<size_t>((size_t)nrec + 1, 1 << 16);
        if ((e = host_realloc(s.h_flags, cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned flags");
        if ((e = host_realloc(s.h_pid, cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned ids");
        if ((e = host_realloc(s.h_off, cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc pinned offsets");
        s.fields_cap = cap;
    }
    if (ctx->d_fields_cap < (size_t)nrec + 1) {
        size_t cap = std::max<size_t>((size_t)nrec + 1, 1 << 16);
        if ((e = dev_realloc(ctx->d_flags, cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc flags");
        if ((e = dev_realloc(ctx->d_pid, cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc ids");
        if ((e = dev_realloc(ctx->d_off32, cap)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc offsets");
        ctx->d_fields_cap = cap;
    }
    if (!ctx->d_ctr && (e = dev_realloc(ctx->d_ctr, 1)) != cudaSuccess) return set_cuda_error(ctx, e, "alloc counters");
    cudaStream_t st = ctx->stream;
    ctx->last_launches = 0;
    s.kind = 1; s.nrec = nrec; s.nbytes = nbytes;
    memset(s.h_ctr, 0, sizeof(ScanCounters));
    if (nrec) {
        memcpy(s.h_off, off, ((size_t)nrec + 1) * 4);
        if (nbytes) FXG_CUDA_OK(cudaMemcpyAsync(s.d_in, s.pinned, nbytes, cudaMemcpyHostToDevice, st));
        FXG_CUDA_OK(cudaMemcpyAsync(ctx->d_off32, s.h_off, ((size_t)nrec + 1) * 4, cudaMemcpyHostToDevice, st));
        FXG_CUDA_OK(cudaMemsetAsync(ctx->d_ctr, 0, sizeof(ScanCounters), st));
        FieldSrc src; memset(&src, 0, sizeof src);
        src.base = s.d_in; src.off32 = ctx->d_off32;
        FieldOut O; memset(&O, 0, sizeof O);
        O.flags = ctx->d_flags; O.pattern_id = ctx->cfg.want_pattern_id ? ctx->d_pid : nullptr; O.n_hits = &ctx->d_ctr->n_hits;
        if (ctx->cfg.mode == FXG_MODE_LITERAL) {
            LiteralFieldParams L; memset(&L, 0, sizeof L);
            L.n_patterns = (int)std::min<uint32_t>(ctx->npat, 2);
            for (int q = 0; q < L.n_patterns; ++q) { L.pat[q] = ctx->d_pat + ctx->off[q]; L.len[q] = (int)(ctx->off[q + 1] - ctx->off[q]); }
            FXG_CUDA_OK(launch_fields_literal(src, nrec, L, O, st));
        } else if (ctx->cfg.mode == FXG_MODE_MULTI && ctx->kmer_ok && !ctx->cfg.want_pattern_id) {
            KmerDevice K{ctx->d_kfilter, ctx->d_kkey, ctx->d_kval, ctx->d_pat, ctx->d_patoff, ctx->k_mask, ctx->k_kmask, ctx->k_k, ctx->k_stride};
            FXG_CUDA_OK(launch_fields_kmer(src, nrec, K, O, st));
        } else if (ctx->cfg.mode == FXG_MODE_MULTI) {
            AcDevice A{ctx->d_cls, ctx->d_next, ctx->d_out, ctx->ac_ncls, ctx->ac_nstates, ctx->ac_wide};
            FXG_CUDA_OK(launch_fields_ac(src, nrec, A, O, st));
        } else {
            ExactDevice E{ctx->d_bkt, ctx->d_key_words, ctx->exact_mask, ctx->d_exact_pre, ctx->exact_pre_mask};
            FXG_CUDA_OK(launch_fields_exact(src, nrec, E, O, st));
        }
        ctx->last_launches = 1;
        FXG_CUDA_OK(cudaMemcpyAsync(s.h_flags, ctx->d_flags, nrec, cudaMemcpyDeviceToHost, st));
        if (ctx->cfg.want_pattern_id) FXG_CUDA_OK(cudaMemcpyAsync(s.h_pid, ctx->d_pid, (size_t)nrec * 4, cudaMemcpyDeviceToHost, st));
        FXG_CUDA_OK(cudaMemcpyAsync(s.h_ctr, ctx->d_ctr, sizeof(ScanCounters), cudaMemcpyDeviceToHost, st));
    }
    FXG_CUDA_OK(cudaEventRecord(s.done, st));
    s.state = 2;
    ctx->inflight.push_back(si);
    return FXG_OK;
}

extern "C" int fxg_collect(fxg_ctx *ctx, fxg_result *res)
{
    if (!ctx || !res) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    if (ctx->inflight.empty()) return set_error(ctx, FXG_ERR_STATE, "nothing submitted");
    int si = ctx->inflight.front();
    ctx->inflight.pop_front();
    Slot &s = ctx->slots[si];
    int rc = FXG_OK;
    if (s.kind == 0) rc = finish_raw(ctx, s, res);
    else {
        FXG_CUDA_OK(cudaEventSynchronize(s.done));
        memset(res, 0, sizeof *res);
        res->n_records = s.nrec;
        res->n_matched = res->n_selected = s.h_ctr->n_hits;
        res->canonical_ok = 1;
        res->flags = s.h_flags;
        res->pattern_id = ctx->cfg.want_pattern_id ? s.h_pid : nullptr;
    }
    s.state = 0;
    if (rc == FXG_OK) ctx->total_selected += res->n_selected;
    return rc;
}

extern "C" int fxg_count(fxg_ctx *ctx, uint64_t *total)
{
    if (!ctx || !total) return FXG_ERR_ARG;
    *total = ctx->total_selected;
    return FXG_OK;
}

extern "C" size_t fxg_find_record_cut(const uint8_t *buf, size_t n)
{
    // A record starts at c iff buf[c]=='@' at a line start and the line after next starts
    // with '+'.  For canonical FASTQ this is exact (a quality line may start with '@', but
    // then the line two below it is a sequence line, which cannot start with '+'); for
    // anything else the GPU predicate rejects the batch and the caller re-routes it.
    if (!buf || n == 0) return 0;
    size_t pos = n;
    for (int tries = 0; tries < 64 && pos > 0; ++tries) {
        // line start at or before pos-1
        size_t c = pos - 1;
        while (c > 0 && buf[c - 1] != '\n') --c;
        if (buf[c] == '@') {
            const uint8_t *e1 = (const uint8_t *)memchr(buf + c, '\n', n - c);
            if (e1) {
                const uint8_t *e2 = (const uint8_t *)memchr(e1 + 1, '\n', n - (size_t)(e1 + 1 - buf));
                if (e2 && (size_t)(e2 + 1 - buf) < n && e2[1] == '+') return c;
            }
        }
        pos = c;        // continue with the previous line
    }
    return 0;
}

// ---------------------------------------------------------------- device-resident path
extern "C" int fxg_scan_device_async(fxg_ctx *ctx, const void *d_buf, uint64_t nbytes)
{
    if (!ctx || (!d_buf && nbytes)) return set_error(ctx, FXG_ERR_ARG, "null argument");
    if (((uintptr_t)d_buf & 15) != 0) return set_error(ctx, FXG_ERR_ARG, "d_buf must be 16-byte aligned (the tiles are fetched with bulk copies)");
    cudaSetDevice(ctx->cfg.device);
    Slot &s = ctx->slots[2];
    int rc = enqueue_raw(ctx, (const uint8_t *)d_buf, (size_t)nbytes, s);
    if (rc == FXG_OK) s.state = 2;
    return rc;
}

extern "C" int fxg_scan_device_pairs_async(fxg_ctx *ctx, const void *d_buf, uint64_t nbytes, uint64_t split)
{
    if (!ctx || (!d_buf && nbytes)) return set_error(ctx, FXG_ERR_ARG, "null argument");
    if (((uintptr_t)d_buf & 15) != 0) return set_error(ctx, FXG_ERR_ARG, "d_buf must be 16-byte aligned (the tiles are fetched with bulk copies)");
    cudaSetDevice(ctx->cfg.device);
    Slot &s = ctx->slots[2];
    int rc = enqueue_raw(ctx, (const uint8_t *)d_buf, (size_t)nbytes, s, 1, split);
    if (rc == FXG_OK) s.state = 2;
    return rc;
}

extern "C" int fxg_scan_device_finish(fxg_ctx *ctx, fxg_result *res)
{
    if (!ctx || !res) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    Slot &s = ctx->slots[2];
    if (s.state != 2) return set_error(ctx, FXG_ERR_STATE, "no device scan in flight");
    int rc = finish_raw(ctx, s, res);
    s.state = 0;
    return rc;
}

extern "C" int fxg_can_route(const fxg_ctx *ctx)
{
    return ctx && ctx->have_patterns && (ctx->scan_mode == kModeExact || mode_is_kmer(ctx->scan_mode)) &&
           ctx->npat < (1u << 24) ? 1 : 0;
}
extern "C" int fxg_last_launch_count(const fxg_ctx *ctx) { return ctx ? ctx->last_launches : 0; }
extern "C" uint64_t fxg_last_d2h_bytes(const fxg_ctx *ctx) { return ctx ? ctx->last_d2h_bytes : 0; }

extern "C" int fxg_set_profiling(fxg_ctx *ctx, int enabled)
{
    if (!ctx) return FXG_ERR_ARG;
    ctx->profiling = enabled != 0;
    return FXG_OK;
}

extern "C" int fxg_last_kernel_ms(fxg_ctx *ctx, float *scan_ms, float *pipeline_ms)
{
    if (!ctx) return FXG_ERR_ARG;
    if (scan_ms) *scan_ms = ctx->last_scan_ms;
    if (pipeline_ms) *pipeline_ms = ctx->last_pipeline_ms;
    return FXG_OK;
}

extern "C" int fxg_synth_device(fxg_ctx *ctx, const fxg_synth_spec *s, void *d_out)
{
    if (!ctx || !s || !d_out) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    FXG_CUDA_OK(launch_synth(*s, (uint8_t *)d_out, ctx->stream));
    FXG_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    return FXG_OK;
}

extern "C" int fxg_host_alloc(void **p, size_t n)
{
    if (!p) return FXG_ERR_ARG;
    cudaError_t e = cudaHostAlloc(p, n, cudaHostAllocPortable);      // pinned for every device of the process
    if (e != cudaSuccess) { *p = nullptr; return set_cuda_error(nullptr, e, "cudaHostAlloc"); }
    return FXG_OK;
}
extern "C" int fxg_host_free(void *p)
{
    if (p) cudaFreeHost(p);
    return FXG_OK;
}

extern "C" int fxg_copy_to_device(fxg_ctx *ctx, void *d_dst, const void *host_src, size_t n)
{
    if (!ctx || !d_dst || !host_src) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    FXG_CUDA_OK(cudaMemcpyAsync(d_dst, host_src, n, cudaMemcpyHostToDevice, ctx->stream));
    FXG_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    return FXG_OK;
}

extern "C" int fxg_copy_to_host(fxg_ctx *ctx, void *host_dst, const void *d_src, size_t n)
{
    if (!ctx || !host_dst || !d_src) return set_error(ctx, FXG_ERR_ARG, "null argument");
    cudaSetDevice(ctx->cfg.device);
    FXG_CUDA_OK(cudaMemcpyAsync(host_dst, d_src, n, cudaMemcpyDeviceToHost, ctx->stream));
    FXG_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    return FXG_OK;
}

// ================================================================= allocation-free literal count (fxg_count_device_noalloc)
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops).  Every scratch table the raw scan needs lives in ONE
// file-scope __device__ array, carved below for the largest batch the entry point accepts.  In host code the name
// `g_noalloc_scratch` is only a host-side stand-in for the symbol, never a device address: its address is taken
// once per device with cudaGetSymbolAddress (that device current) and the kernels get pointers derived from it.
namespace fxg {

constexpr size_t na_align(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t kNaMaxBytes = FXG_NOALLOC_MAX_BYTES;
constexpr size_t kNaTiles = (kNaMaxBytes + G24::kTileBytes - 1) / G24::kTileBytes;
constexpr size_t kNaRecCap = kNaTiles * (size_t)(G24::kMaxTileLines / 4) + 64;
constexpr size_t kNaWords = (kNaRecCap + 31) / 32 + 4;
// offsets of the tables inside the array (each 256-byte aligned)
constexpr size_t kNaOffCtr = 0;
constexpr size_t kNaOffInfo = kNaOffCtr + na_align(sizeof(ScanCounters));
constexpr size_t kNaOffLbase = kNaOffInfo + na_align(kNaTiles * sizeof(TileInfo));
constexpr size_t kNaOffBsum = kNaOffLbase + na_align((kNaTiles + 1) * sizeof(unsigned long long));
constexpr size_t kNaOffRedo = kNaOffBsum + na_align((kNaTiles / 4096 + 2) * sizeof(unsigned long long));
constexpr size_t kNaOffHitbits = kNaOffRedo + na_align((kNaTiles + 1) * sizeof(unsigned int));
constexpr size_t kNaOffTileHit = kNaOffHitbits + na_align(kNaWords * sizeof(unsigned int));
constexpr size_t kNaOffTileNl = kNaOffTileHit + na_align(kNaTiles * (size_t)G24::kHitWords * sizeof(unsigned int));
constexpr size_t kNaScratchBytes = kNaOffTileNl + na_align(kNaTiles * (size_t)G24::kMaxTileLines * sizeof(unsigned short));
static_assert(kNaScratchBytes < (2u << 20), "the no-alloc scratch is meant to stay modest (< 2 MiB per GPU)");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaScratchBytes];

// counters -> the caller's result record (what finish_raw does on the host for the ordinary path)
__global__ void noalloc_publish_kernel(const ScanCounters *ctr, unsigned long long nbytes, int all_match, fxg_noalloc_counts *out)
{
    if (blockIdx.x || threadIdx.x) return;
    fxg_noalloc_counts r;
    r.n_lines = ctr->n_lines;
    r.n_records = ctr->n_lines >> 2;
    unsigned int viol = ctr->violation;
    if (nbytes && (ctr->last_byte != '\n' || (ctr->n_lines & 3ull))) viol |= FXG_V_TRUNCATED;
    r.n_matched = all_match ? r.n_records : ctr->n_hits;
    r.first_bad_offset = ctr->first_bad;
    r.violation = viol;
    r.canonical_ok = viol == 0u ? 1u : 0u;
    *out = r;
}

// device address of the scratch array on `dev` (looked up once per device, with that device current)
static int noalloc_scratch_base(fxg_ctx *ctx, int dev, unsigned char **base)
{
    static std::mutex mu;
    static unsigned char *ptr[64];
    std::lock_guard<std::mutex> lock(mu);
    if (dev < 0 || dev >= 64) return set_error(ctx, FXG_ERR_ARG, "fxg_count_device_noalloc: device ordinal out of range");
    if (!ptr[dev]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return set_cuda_error(ctx, e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        if (!p || ((uintptr_t)p & 255) != 0) return set_error(ctx, FXG_ERR_CUDA, "fxg_count_device_noalloc: scratch symbol has no usable device address");
        ptr[dev] = (unsigned char *)p;
    }
    *base = ptr[dev];
    return FXG_OK;
}

}  // namespace fxg

extern "C" int fxg_count_device_noalloc(fxg_ctx *ctx, const void *d_buf, uint64_t nbytes, fxg_noalloc_counts *d_out)
{
    if (!ctx || !d_out || (!d_buf && nbytes)) return set_error(ctx, FXG_ERR_ARG, "fxg_count_device_noalloc: null argument");
    if (nbytes > kNaMaxBytes)
        return set_error(ctx, FXG_ERR_ARG, "fxg_count_device_noalloc: batch of " + std::to_string(nbytes) + " bytes exceeds FXG_NOALLOC_MAX_BYTES (" +
                                               std::to_string(kNaMaxBytes) + "); use fxg_scan_device_async or smaller batches");
    if (((uintptr_t)d_buf & 15) != 0) return set_error(ctx, FXG_ERR_ARG, "d_buf must be 16-byte aligned (the tiles are fetched with bulk copies)");
    if (!ctx->have_patterns) return set_error(ctx, FXG_ERR_STATE, "fxg_set_patterns has not been called");
    if (ctx->cfg.mode != FXG_MODE_LITERAL)
        return set_error(ctx, FXG_ERR_UNSUPPORTED, "fxg_count_device_noalloc serves FXG_MODE_LITERAL contexts only");
    if (ctx->tile_bytes != G24::kTileBytes || ctx->nl_stride != G24::kMaxTileLines || ctx->hit_words != G24::kHitWords)
        return set_error(ctx, FXG_ERR_STATE, "fxg_count_device_noalloc: unexpected scan geometry");
    FXG_CUDA_OK(cudaSetDevice(ctx->cfg.device));
    unsigned char *base = nullptr;
    int rc = noalloc_scratch_base(ctx, ctx->cfg.device, &base);
    if (rc != FXG_OK) return rc;
    cudaStream_t st = ctx->stream;
    ScanCounters *ctr = reinterpret_cast<ScanCounters *>(base + kNaOffCtr);
    unsigned int *hitbits = reinterpret_cast<unsigned int *>(base + kNaOffHitbits);
    unsigned int *redo = reinterpret_cast<unsigned int *>(base + kNaOffRedo);
    const size_t ntiles = ((size_t)nbytes + G24::kTileBytes - 1) / G24::kTileBytes;
    const size_t rec_cap = ntiles * (size_t)(G24::kMaxTileLines / 4) + 64;
    const size_t words = (rec_cap + 31) / 32;
    // the array keeps the previous call's contents: everything the kernels accumulate into starts at zero again
    FXG_CUDA_OK(cudaMemsetAsync(ctr, 0, sizeof(ScanCounters), st));
    FXG_CUDA_OK(cudaMemsetAsync(&ctr->first_bad, 0xff, 8, st));
    FXG_CUDA_OK(cudaMemsetAsync(hitbits, 0, (words + 4) * sizeof(unsigned int), st));
    FXG_CUDA_OK(cudaMemsetAsync(redo + kNaTiles, 0, sizeof(unsigned int), st));
    ctx->last_launches = 0;
    if (nbytes) {
        ScanParams P;
        memset(&P, 0, sizeof P);
        P.buf = (const uint8_t *)d_buf; P.nbytes = nbytes; P.ntiles = ntiles;
        P.tile_bytes = G24::kTileBytes; P.nl_stride = G24::kMaxTileLines; P.hit_words = G24::kHitWords;
        P.info = reinterpret_cast<TileInfo *>(base + kNaOffInfo);
        P.tile_nl = reinterpret_cast<unsigned short *>(base + kNaOffTileNl);
        P.tile_hit = reinterpret_cast<unsigned int *>(base + kNaOffTileHit);
        P.lbase = reinterpret_cast<unsigned long long *>(base + kNaOffLbase);
        P.bsum = reinterpret_cast<unsigned long long *>(base + kNaOffBsum);
        P.ctr = ctr; P.hitbits = hitbits; P.rec_cap = rec_cap; P.one = 1; P.l2_evict_first = 1;
        P.redo_list = redo; P.redo_count = redo + kNaTiles;
        P.lit = ctx->d_lit; P.pat_bytes = ctx->d_pat; P.target = ctx->cfg.target;
        P.mode = ctx->scan_mode; P.match_enabled = mode_is_literal(P.mode); P.lit_unit = ctx->h_lit.unit;
        FXG_CUDA_OK(launch_scan_raw(P, st, &ctx->last_launches, nullptr));
    }
    noalloc_publish_kernel<<<1, 32, 0, st>>>(ctr, nbytes, ctx->h_lit.all_match ? 1 : 0, d_out);
    ctx->last_launches += 1;
    FXG_CUDA_OK(cudaGetLastError());
    return FXG_OK;
}
