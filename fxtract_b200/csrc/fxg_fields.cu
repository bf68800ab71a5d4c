// fxg_fields.cu -- field engine: the three matchers over (offset,length) fields.
//
// Serves (a) the general path, where the host's kseq reader has already parsed
// records (FASTA, wrapped lines, gz, pairs...) and packed the target fields of a
// batch back to back, and (b) the raw path's second stage for the multi-pattern
// and exact strategies, where the fused scan kernel has produced the per-record
// field table over the raw bytes in HBM.
//
//   literal : strstr(field, pattern) [|| strstr(field, revcomp)]   main.cpp:568-588
//   multi   : ssearch_scan(ssp, {ptr,len}, cb)                     main.cpp:406,426
//             byte-class compressed Aho-Corasick DFA; first report = smallest end
//             offset, longest pattern (see oracle/fxo_ac.h for why this rule)
//   exact   : manager.find(field) on std::map<string,int>         fileManager.cpp:171-177
//             open-addressing table, 64-bit hash + byte verification
#include "fxg_common.cuh"
#include "fxg_exact.cuh"

namespace fxg {

__device__ __forceinline__ void emit_warp_per_field(const FieldOut &O, unsigned long long k, int lane, bool hit, int pid)
{
    if (lane == 0) {
        if (O.flags) O.flags[k] = hit ? 1 : 0;
        if (O.pattern_id) O.pattern_id[k] = hit ? pid : -1;
        if (hit) {
            if (O.hitbits) atomicOr(&O.hitbits[k >> 5], 1u << (k & 31));
            if (O.n_hits) atomicAdd(O.n_hits, 1ull);
        }
    }
}

// thread-per-field result: one ballot per warp keeps the bitset writes atomic-free
__device__ __forceinline__ void emit_thread_per_field(const FieldOut &O, unsigned long long k, unsigned long long n,
                                                      bool hit, int pid)
{
    bool valid = k < n;
    if (valid) {
        if (O.flags) O.flags[k] = hit ? 1 : 0;
        if (O.pattern_id) O.pattern_id[k] = hit ? pid : -1;
    }
    unsigned int b = __ballot_sync(0xffffffffu, valid && hit);
    if ((threadIdx.x & 31) == 0 && b) {
        if (O.hitbits) O.hitbits[k >> 5] = b;        // k is a multiple of 32 for lane 0
        if (O.n_hits) atomicAdd(O.n_hits, (unsigned long long)__popc(b));
    }
}

// ---------------------------------------------------------------- literal: warp per field
__global__ void __launch_bounds__(256) fields_literal_kernel(FieldSrc src, unsigned long long n, LiteralFieldParams L, FieldOut O)
{
    const int lane = threadIdx.x & 31;
    const unsigned long long warp0 = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    n = src.count(n);
    for (unsigned long long k = warp0; k < n; k += nwarps) {
        const uint8_t *p; unsigned int len;
        src.get(k, p, len);
        bool hit = false;
        int pid = -1;
        for (int q = 0; q < L.n_patterns && !hit; ++q) {
            int m = L.len[q];
            if (m == 0) { hit = true; pid = q; break; }        // strstr(hay, "") != NULL
            if ((unsigned int)m > len) continue;
            const uint8_t *pp = L.pat[q];
            unsigned int first = pp[0];
            unsigned int last = len - (unsigned int)m;          // last start position
            for (unsigned int s0 = 0; s0 <= last; s0 += 32) {
                unsigned int s = s0 + lane;
                bool ok = false;
                if (s <= last && p[s] == first) {
                    ok = true;
                    for (int i = 1; i < m; ++i) if (p[s + i] != pp[i]) { ok = false; break; }
                }
                if (__any_sync(0xffffffffu, ok)) { hit = true; pid = q; break; }
            }
        }
        emit_warp_per_field(O, k, lane, hit, pid);
    }
}

// ---------------------------------------------------------------- multi: thread per field, DFA walk
// ENTRY = unsigned short / unsigned int (transition entry width), SMEM = the table is staged in shared memory.
// The field is read 8 bytes per load; per byte: class look-up (shared memory, independent of the state) and one
// dependent transition look-up; the walk stops at the first entry whose high bit says "reports a pattern" = smallest
// end offset, longest pattern (oracle/fxo_ac.h).
template <typename ENTRY, bool SMEM>
__global__ void __launch_bounds__(512) fields_ac_kernel(FieldSrc src, unsigned long long n, AcDevice A, FieldOut O)
{
    extern __shared__ __align__(16) unsigned char ac_smem[];
    uint8_t *cls = ac_smem;                                                    // [256]
    const ENTRY *next = reinterpret_cast<const ENTRY *>(A.next);
    if (threadIdx.x < 256) cls[threadIdx.x] = A.cls[threadIdx.x];
    if (SMEM) {
        ENTRY *tab = reinterpret_cast<ENTRY *>(ac_smem + 256);
        const size_t words = ((size_t)A.nstates * A.ncls * sizeof(ENTRY) + 3) / 4;
        const unsigned int *g = reinterpret_cast<const unsigned int *>(A.next);      // allocation is padded to whole words
        for (size_t w = threadIdx.x; w < words; w += blockDim.x) reinterpret_cast<unsigned int *>(tab)[w] = __ldg(g + w);
        next = tab;
    }
    __syncthreads();
    constexpr ENTRY kFlag = (ENTRY)1 << (sizeof(ENTRY) * 8 - 1);
    const unsigned int ncls = (unsigned int)A.ncls;
    n = src.count(n);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long nround = (n + 31) & ~31ull;      // keep whole warps in the loop for the ballot
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < nround; k += stride) {
        bool hit = false;
        int pid = -1;
        if (k < n) {
            const uint8_t *p; unsigned int len;
            src.get(k, p, len);
            pid = __ldg(&A.out[0]);                  // the empty pattern ends at offset 0
            if (pid < 0 && len) {
                const uint8_t *end = p + len;
                const uint8_t *ga = (const uint8_t *)((uintptr_t)p & ~(uintptr_t)7);
                unsigned int skip = (unsigned int)((uintptr_t)p & 7);        // bytes of the first word in front of the field
                unsigned int s = 0;
                bool found = false;
                while (ga < end && !found) {
                    unsigned long long w = __ldg(reinterpret_cast<const unsigned long long *>(ga)) >> (8 * skip);
                    const unsigned int avail = 8 - skip;
                    const unsigned int left = (unsigned int)(end - (ga + skip));
                    const unsigned int cnt = left < avail ? left : avail;
                    for (unsigned int b = 0; b < cnt; ++b) {
                        const unsigned int c = cls[(unsigned int)w & 0xffu];
                        w >>= 8;
                        const ENTRY t = SMEM ? next[s * ncls + c] : __ldg(&next[(size_t)s * ncls + c]);
                        s = (unsigned int)(t & (ENTRY)~kFlag);
                        if (t & kFlag) { found = true; break; }
                    }
                    ga += 8;
                    skip = 0;
                }
                if (found) pid = __ldg(&A.out[s]);
            }
            hit = pid >= 0;
        }
        emit_thread_per_field(O, k, n, hit, pid);
    }
}

// ---------------------------------------------------------------- multi, ACGT pattern sets: packed 2-bit k-mer index
// Any-hit only (no first-hit pattern id): used when the caller does not route records per pattern.
// One thread per field rolls a 2-bit packed k-mer over the field from 16-byte aligned loads; every
// stride-th position probes a 2^20-bit filter in shared memory; filter hits go to the open-addressing
// table (L2) and candidates are verified byte by byte, so there are no false positives.
__device__ __noinline__ bool kmer_verify(const KmerDevice &K, unsigned int key, const uint8_t *kstart, const uint8_t *p,
                                            const uint8_t *end)
{
    const unsigned int h = fxg_kmer_hash(key);
    unsigned int slot = (h >> 7) & K.ht_mask;
    for (;;) {
        const unsigned int v = __ldg(&K.ht_val[slot]);
        if (!v) return false;
        if (__ldg(&K.ht_key[slot]) == key) {
            const unsigned int q = (v - 1u) >> 5, o = (v - 1u) & 31u;
            const unsigned int a = K.off[q], m = K.off[q + 1] - a;
            const uint8_t *ts = kstart - o;
            if (ts >= p && ts + m <= end) {
                bool eq = true;
                for (unsigned int i = 0; i < m; ++i) if (ts[i] != K.blob[a + i]) { eq = false; break; }
                if (eq) return true;
            }
        }
        slot = (slot + 1) & K.ht_mask;
    }
}

// One thread per field.  The field is read with 16-byte aligned loads (one chunk prefetched ahead) and
// walked in FIELD-RELATIVE 4-byte words -- funnel shift of two neighbouring aligned words by the byte
// misalignment of the field start -- so that all lanes run the same instruction stream whatever the
// alignment of their field: per word a SWAR "four bases?" test, one multiply to pack the four 2-bit
// codes, and S-strided probes of a two-hash filter in shared memory.  S = probe stride (1, 2 or 4).
__device__ __forceinline__ unsigned int pick_word(const uint4 &x, int i)
{
    return i == 0 ? x.x : (i == 1 ? x.y : (i == 2 ? x.z : x.w));
}
__device__ __forceinline__ bool kmer_probe(const unsigned int *filt, const KmerDevice &K, unsigned int key, const uint8_t *kend,
                                           const uint8_t *p, const uint8_t *end)
{
    const unsigned int h = fxg_kmer_hash(key) >> (32 - kKmerFilterBits);
    const unsigned int g = fxg_kmer_hash2(key) >> (32 - kKmerFilterBits);
    if (!((filt[h >> 5] >> (h & 31)) & (filt[g >> 5] >> (g & 31)) & 1u)) return false;
    return kmer_verify(K, key, kend - (K.k - 1), p, end);
}

#ifndef FXG_KM_THREADS
#define FXG_KM_THREADS 1024
#endif
#ifndef FXG_KM_CTAS
#define FXG_KM_CTAS 1
#endif
// One field-relative word: `w` = bytes 4r..4r+3 of the field, r in [0, nwords] (r == nwords: the last len & 3 bytes).
template <int S>
__device__ __forceinline__ void kmer_word(const unsigned int *filt, const KmerDevice &K, unsigned int w, int r, int nwords,
                                          unsigned int len, const uint8_t *p, const uint8_t *end, int k, unsigned int kmask,
                                          int j0, unsigned int &key, int &run, bool &hit)
{
    const uint8_t *wa = p + 4 * r;
    const unsigned int codes = (w >> 1) & 0x03030303u;
    // four "bases": a SUPERSET test -- A, C, G, T differ from 0x40 in bits 0, 1, 2, 4 only, so the test admits
    // 0x40..0x47 and 0x50..0x57 (not N) -- is enough for the fast path: a letter that is no base only makes a k-mer
    // that no pattern has (every candidate is compared byte by byte); it can never hide an occurrence, whose bytes
    // are all real bases and whose k-mer is probed whatever preceded it
    if (r < nwords && ((w ^ 0x40404040u) & 0xe8e8e8e8u) == 0u) {
        const unsigned int c4 = (codes * 0x40100401u) >> 24;
        if (S == 4) {
            if (run + j0 + 1 >= k &&
                kmer_probe(filt, K, ((key << (2 * j0 + 2)) | (c4 >> (6 - 2 * j0))) & kmask, wa + j0, p, end)) hit = true;
        } else if (S == 2) {
            if (run + j0 + 1 >= k &&
                kmer_probe(filt, K, ((key << (2 * j0 + 2)) | (c4 >> (6 - 2 * j0))) & kmask, wa + j0, p, end)) hit = true;
            if (run + j0 + 3 >= k &&
                kmer_probe(filt, K, ((key << (2 * j0 + 6)) | (c4 >> (2 - 2 * j0))) & kmask, wa + j0 + 2, p, end)) hit = true;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (run + j + 1 >= k &&
                    kmer_probe(filt, K, ((key << (2 * j + 2)) | (c4 >> (6 - 2 * j))) & kmask, wa + j, p, end)) hit = true;
        }
        key = ((key << 8) | c4) & kmask;
        run += 4;
    } else {                                                // a non-base byte, or the field's last 0..3 bytes
        const int lim = r < nwords ? 4 : (int)(len & 3);
        for (int j = 0; j < lim; ++j) {
            const unsigned int b = (w >> (8 * j)) & 0xffu;
            if (!fxg_is_acgt(b)) { run = 0; continue; }
            key = ((key << 2) | fxg_base_code(b)) & kmask;
            ++run;
            if (((j - j0) & (S - 1)) == 0 && run >= k && kmer_probe(filt, K, key, wa + j, p, end)) hit = true;
        }
    }
}

// The walk steps through the ALIGNED 16-byte chunks that hold the field, all lanes in lock step (one 16-byte load
// per chunk, no per-lane chunk bookkeeping): with G[g] the aligned words from the first chunk on and c0 the index
// of the word that holds the field's first byte, field word r = funnelshift(G[c0 + r], G[c0 + r + 1], 8 * (p & 3)),
// so aligned word g completes field word r = g - 1 - c0.  Only the range test of r differs between lanes.
template <int S>
__global__ void __launch_bounds__(FXG_KM_THREADS, FXG_KM_CTAS) fields_kmer_kernel(FieldSrc src, unsigned long long n, KmerDevice K, FieldOut O)
{
    extern __shared__ __align__(16) unsigned int filt[];
    for (int i = threadIdx.x; i < kKmerFilterWords; i += blockDim.x) filt[i] = K.filter[i];
    __syncthreads();
    n = src.count(n);
    const unsigned long long stride_n = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long nround = (n + 31) & ~31ull;
    const int k = K.k;
    const unsigned int kmask = K.kmask;
    const int j0 = (k - 1) & (S - 1);                  // probe positions inside a field-relative word: j0, j0 + S, ...
    for (unsigned long long f = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; f < nround; f += stride_n) {
        bool hit = false;
        const uint8_t *p = src.base; unsigned int len = 0;
        if (f < n) src.get(f, p, len);
        if (len >= (unsigned int)(k + S - 1)) {
            const uint8_t *end = p + len;
            const uint8_t *ca = (const uint8_t *)((uintptr_t)p & ~(uintptr_t)15);      // first aligned chunk
            const int sh = (int)((uintptr_t)p & 3) * 8;
            const int nwords = (int)(len >> 2);
            const int c0 = (int)(((uintptr_t)p & 15) >> 2);
            const int glast = nwords + 1 + c0;                                          // aligned word that completes word `nwords`
            unsigned int prev = 0, key = 0;
            int run = 0;
            uint4 x = __ldg(reinterpret_cast<const uint4 *>(ca));
            for (int g = 0; g <= glast && !hit; g += 4) {
                ca += 16;
                uint4 xn = make_uint4(0, 0, 0, 0);                                     // chunks past the field are not read
                if (ca < end) xn = __ldg(reinterpret_cast<const uint4 *>(ca));
                const int r = g - 1 - c0;                                               // field word completed by x.x
                if (r >= 0 && r <= nwords)
                    kmer_word<S>(filt, K, __funnelshift_r(prev, x.x, sh), r, nwords, len, p, end, k, kmask, j0, key, run, hit);
                if (r + 1 >= 0 && r + 1 <= nwords && !hit)
                    kmer_word<S>(filt, K, __funnelshift_r(x.x, x.y, sh), r + 1, nwords, len, p, end, k, kmask, j0, key, run, hit);
                if (r + 2 >= 0 && r + 2 <= nwords && !hit)
                    kmer_word<S>(filt, K, __funnelshift_r(x.y, x.z, sh), r + 2, nwords, len, p, end, k, kmask, j0, key, run, hit);
                if (r + 3 >= 0 && r + 3 <= nwords && !hit)
                    kmer_word<S>(filt, K, __funnelshift_r(x.z, x.w, sh), r + 3, nwords, len, p, end, k, kmask, j0, key, run, hit);
                prev = x.w;
                x = xn;
            }
        }
        emit_thread_per_field(O, f, n, hit, -1);
    }
}

// ---------------------------------------------------------------- exact: thread per field, hash probe + verify
// (exact_lookup of fxg_exact.cuh: fxg_hash_bytes of the key 8 bytes per step from aligned loads, nothing past the
// field's last word is read; with stop_at_space the key ends at the first ' ' of the field)
__global__ void __launch_bounds__(256) fields_exact_kernel(FieldSrc src, unsigned long long n, ExactDevice E, FieldOut O)
{
    n = src.count(n);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long nround = (n + 31) & ~31ull;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < nround; k += stride) {
        bool hit = false;
        int pid = -1;
        if (k < n) {
            const uint8_t *p; unsigned int len;
            src.get(k, p, len);
            hit = exact_lookup_global(E, p, len, src.stop_at_space != 0, pid);
            if (!hit) pid = -1;
        }
        emit_thread_per_field(O, k, n, hit, pid);
    }
}

static unsigned int grid_for(unsigned long long work_items, int per_block)
{
    unsigned long long b = (work_items + per_block - 1) / per_block;
    unsigned long long cap = 148ull * 32;          // persistent-ish: 32 CTAs per SM worth of blocks
    if (b > cap) b = cap;
    if (b == 0) b = 1;
    return (unsigned int)b;
}

cudaError_t launch_fields_literal(const FieldSrc &src, unsigned long long n, const LiteralFieldParams &L,
                                  const FieldOut &O, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    fields_literal_kernel<<<grid_for(n, 8), 256, 0, st>>>(src, n, L, O);
    return cudaGetLastError();
}
template <typename ENTRY, bool SMEM>
static cudaError_t launch_fields_ac_t(const FieldSrc &src, unsigned long long n, const AcDevice &A, const FieldOut &O, cudaStream_t st, size_t smem)
{
    int dev = 0;
    cudaGetDevice(&dev);
    int cap = 0;
    cudaError_t e = scan_launch_config((const void *)fields_ac_kernel<ENTRY, SMEM>, 512, smem, dev, &cap);
    if (e != cudaSuccess) return e;
    unsigned long long b = (n + 511) / 512;
    if (SMEM && b > (unsigned long long)cap) b = (unsigned long long)cap;       // persistent CTAs: the table is staged once per CTA
    if (!SMEM && b > 148ull * 16) b = 148ull * 16;
    if (b == 0) b = 1;
    fields_ac_kernel<ENTRY, SMEM><<<(unsigned int)b, 512, smem, st>>>(src, n, A, O);
    return cudaGetLastError();
}
cudaError_t launch_fields_ac(const FieldSrc &src, unsigned long long n, const AcDevice &A, const FieldOut &O, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    const size_t table = (size_t)A.nstates * A.ncls * (A.wide ? 4 : 2);
    const bool smem = table + 256 + 16 <= 200u * 1024u;
    const size_t bytes = 256 + (smem ? ((table + 15) & ~(size_t)15) : 0);
    if (A.wide) return smem ? launch_fields_ac_t<unsigned int, true>(src, n, A, O, st, bytes) : launch_fields_ac_t<unsigned int, false>(src, n, A, O, st, bytes);
    return smem ? launch_fields_ac_t<unsigned short, true>(src, n, A, O, st, bytes) : launch_fields_ac_t<unsigned short, false>(src, n, A, O, st, bytes);
}
template <int S> static cudaError_t launch_fields_kmer_s(const FieldSrc &src, unsigned long long n, const KmerDevice &K,
                                                         const FieldOut &O, cudaStream_t st)
{
    const int smem = kKmerFilterWords * (int)sizeof(unsigned int);
    cudaError_t e = cudaFuncSetAttribute(fields_kmer_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   // per device
    if (e != cudaSuccess) return e;
    unsigned long long b = (n + FXG_KM_THREADS - 1) / FXG_KM_THREADS;
    if (b > 148ull * FXG_KM_CTAS) b = 148ull * FXG_KM_CTAS;            // persistent CTAs: the filter is loaded once per CTA
    fields_kmer_kernel<S><<<(unsigned int)b, FXG_KM_THREADS, smem, st>>>(src, n, K, O);
    return cudaGetLastError();
}
cudaError_t launch_fields_kmer(const FieldSrc &src, unsigned long long n, const KmerDevice &K, const FieldOut &O, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    if (K.stride >= 4) return launch_fields_kmer_s<4>(src, n, K, O, st);
    if (K.stride >= 2) return launch_fields_kmer_s<2>(src, n, K, O, st);
    return launch_fields_kmer_s<1>(src, n, K, O, st);
}
cudaError_t launch_fields_exact(const FieldSrc &src, unsigned long long n, const ExactDevice &E, const FieldOut &O, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    fields_exact_kernel<<<grid_for(n, 256), 256, 0, st>>>(src, n, E, O);
    return cudaGetLastError();
}

}  // namespace fxg
