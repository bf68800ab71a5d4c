// fxg_exact.cuh -- whole-field key look-up, shared by the field engine (fxg_fields.cu: packed fields / field table in
// global memory) and the fused scan kernel (fxg_scan_kernel.cuh: keys read from the staged tile in shared memory).
//
// Replaces manager.find(*data) of hash_search (main.cpp:315, fileManager.cpp:171-177: std::map<string,int>::find).
// Bucketed table (ExactDevice, fxg_common.cuh) over the 64-bit key hash of fxg_hash_bytes: one 32-byte bucket per
// look-up; a fingerprint match is verified word by word against a zero-padded copy of the key, so no parity claim
// rests on the hash.
#pragma once
#include "fxg_common.cuh"

namespace fxg {

// Word loaders: aligned 32-bit word i of the byte space the key lives in.
struct GlobalWords {
    const unsigned int *w;          // 4-byte aligned base
    long long limit;                // words with index >= limit are not read (they lie past the last byte of interest)
    __device__ __forceinline__ unsigned int operator()(long long i) const { return i < limit ? __ldg(w + i) : 0u; }
};
struct TileWords {
    const unsigned int *w;          // staged tile (shared memory); reads past the tile stay inside the CTA's allocation
    __device__ __forceinline__ unsigned int operator()(long long i) const { return w[i]; }
};

// Key = bytes [start, start + len) of the loader's byte space; with stop_at_space it ends at the first ' ' before that
// (the read name of a header line).  16 bytes per round: five aligned words are fetched at once (the loads of a round do
// not depend on its data, so their latency is paid once per 16 bytes), four funnel shifts by the key's byte
// misalignment give two 8-byte key words.  Returns fxg_hash_bytes of the key; `keylen` = its length.
__device__ __forceinline__ unsigned long long exact_space_mask(unsigned long long w)
{
    const unsigned long long x = w ^ 0x2020202020202020ull;
    return (x - 0x0101010101010101ull) & ~x & 0x8080808080808080ull;
}
template <class LD>
__device__ __forceinline__ unsigned long long exact_hash(const LD &ld, long long start, unsigned int len, bool stop_at_space,
                                                         unsigned int &keylen)
{
    const int sh = (int)(start & 3) * 8;
    unsigned long long h = kHashSeed;
    unsigned int used = 0;
    if (len) {
        long long wi = start >> 2;
        unsigned int g0 = ld(wi);
        for (;;) {
            const unsigned int g1 = ld(wi + 1), g2 = ld(wi + 2), g3 = ld(wi + 3), g4 = ld(wi + 4);
            const unsigned long long wa = (unsigned long long)__funnelshift_r(g0, g1, sh) | ((unsigned long long)__funnelshift_r(g1, g2, sh) << 32);
            const unsigned long long wb = (unsigned long long)__funnelshift_r(g2, g3, sh) | ((unsigned long long)__funnelshift_r(g3, g4, sh) << 32);
            const unsigned int left = len - used;              // key bytes still to come, this round included
            unsigned int n = left < 16u ? left : 16u;          // ... of which this round holds n
            if (stop_at_space) {
                const unsigned long long za = exact_space_mask(wa), zb = exact_space_mask(wb);
                unsigned int sp = 16u;
                if (za) sp = (unsigned int)(__ffsll((long long)za) - 1) >> 3;
                else if (zb) sp = 8u + ((unsigned int)(__ffsll((long long)zb) - 1) >> 3);
                if (sp < n) n = sp;
            }
            if (n >= 8u) {
                h = fxg_hash_word(h, wa);
                if (n >= 16u) h = fxg_hash_word(h, wb);
                else if (n > 8u) h = fxg_hash_word(h, wb & ((1ull << (8 * (n - 8u))) - 1ull));
            } else if (n) h = fxg_hash_word(h, wa & ((1ull << (8 * n)) - 1ull));
            used += n;
            if (n < 16u || used == len) break;
            g0 = g4;
            wi += 4;
        }
    }
    keylen = used;
    return fxg_hash_finish(h, used);
}

// true iff the bucket chain of hash h holds its fingerprint (one memory round trip unless a bucket is full)
__device__ __forceinline__ bool exact_fingerprint_present(const ExactDevice &E, unsigned long long h)
{
    if (E.pre) {
        const unsigned int m = exact_pre_mask(h);
        if ((__ldg(E.pre + exact_pre_word(h, E.pre_mask)) & m) != m) return false;
    }
    unsigned int bkt = (unsigned int)h & E.mask;
    const unsigned int fp = (unsigned int)(h >> 32) | 1u;
    for (;;) {
        const uint4 *B = E.bkt + 4 * (size_t)bkt;
        const uint4 q0 = __ldg(B), q1 = __ldg(B + 1), q2 = __ldg(B + 2), q3 = __ldg(B + 3);
        if (q0.x == fp || q0.z == fp || q1.x == fp || q1.z == fp || q2.x == fp || q2.z == fp || q3.x == fp || q3.z == fp) return true;
        if (q3.z == 0u) return false;
        bkt = (bkt + 1) & E.mask;
    }
}

// table probe + verification of the key [start, start + len) whose hash is h
template <class LD>
__device__ __forceinline__ bool exact_probe(const ExactDevice &E, unsigned long long h, const LD &ld, long long start, unsigned int len,
                                            int &pid)
{
    if (E.pre) {
        const unsigned int m = exact_pre_mask(h);
        if ((__ldg(E.pre + exact_pre_word(h, E.pre_mask)) & m) != m) return false;
    }
    const long long w0 = start >> 2;
    const int sh = (int)(start & 3) * 8;
    unsigned int bkt = (unsigned int)h & E.mask;
    const unsigned int fp = (unsigned int)(h >> 32) | 1u;
    for (;;) {
        const uint4 *B = E.bkt + 4 * (size_t)bkt;
        const uint4 q0 = __ldg(B), q1 = __ldg(B + 1), q2 = __ldg(B + 2), q3 = __ldg(B + 3);
        const unsigned int fps[8] = {q0.x, q0.z, q1.x, q1.z, q2.x, q2.z, q3.x, q3.z};
        const unsigned int offs[8] = {q0.y, q0.w, q1.y, q1.w, q2.y, q2.w, q3.y, q3.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            if (fps[e] != fp) continue;
            // meta word and the first four key words in one round trip (the array has slack behind the last key)
            const unsigned long long *kw = E.kw + offs[e];
            const unsigned long long meta = __ldg(kw);
            const unsigned long long pre[4] = {__ldg(kw + 1), __ldg(kw + 2), __ldg(kw + 3), __ldg(kw + 4)};
            if ((unsigned int)meta != len) continue;
            bool eq = true;
            long long wi = w0;
            unsigned int g0 = ld(wi);
            for (unsigned int o = 0; o < len; o += 8) {
                const unsigned int g1 = ld(wi + 1), g2 = ld(wi + 2);
                unsigned long long w = (unsigned long long)__funnelshift_r(g0, g1, sh) |
                                       ((unsigned long long)__funnelshift_r(g1, g2, sh) << 32);
                if (len - o < 8) w &= (1ull << (8 * (len - o))) - 1ull;
                const unsigned int i = o >> 3;
                const unsigned long long kv = i == 0 ? pre[0] : (i == 1 ? pre[1] : (i == 2 ? pre[2] : (i == 3 ? pre[3] : __ldg(kw + 1 + i))));
                if (w != kv) { eq = false; break; }
                g0 = g2;
                wi += 2;
            }
            if (eq) { pid = (int)(meta >> 32); return true; }
        }
        if (q3.z == 0u) return false;                      // the bucket has a free slot: nothing spilled past it
        bkt = (bkt + 1) & E.mask;
    }
}

// Look-up of a key whose hash h is already known (exact_resolve_kernel: the scan kernel hashed it from the staged
// tile).  The key is text[ks, ke) -- or, with `name`, text[ks, first ' ' in [ks, ke)) -- of the batch in global
// memory.  Pre-filter, bucket, and for a fingerprint match the candidate key (its length comes from the table) is
// compared with the text in place, 8 bytes per step; nothing is hashed again.
__device__ __forceinline__ bool exact_verify_hashed(const ExactDevice &E, unsigned long long h, const uint8_t *buf, long long ks,
                                                    long long ke, bool name, int &pid)
{
    if (E.pre) {
        const unsigned int m = exact_pre_mask(h);
        if ((__ldg(E.pre + exact_pre_word(h, E.pre_mask)) & m) != m) return false;
    }
    unsigned int bkt = (unsigned int)h & E.mask;
    const unsigned int fp = (unsigned int)(h >> 32) | 1u;
    for (;;) {
        const uint4 *B = E.bkt + 4 * (size_t)bkt;
        const uint4 q0 = __ldg(B), q1 = __ldg(B + 1), q2 = __ldg(B + 2), q3 = __ldg(B + 3);
        const unsigned int fps[8] = {q0.x, q0.z, q1.x, q1.z, q2.x, q2.z, q3.x, q3.z};
        const unsigned int offs[8] = {q0.y, q0.w, q1.y, q1.w, q2.y, q2.w, q3.y, q3.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            if (fps[e] != fp) continue;
            const unsigned long long *kw = E.kw + offs[e];
            const unsigned long long meta = __ldg(kw);
            const unsigned int len = (unsigned int)meta;
            if (name) {
                if (ks + (long long)len > ke) continue;
                if (ks + (long long)len < ke && buf[ks + len] != ' ') continue;      // the name goes on
            } else if ((long long)len != ke - ks) continue;
            const uintptr_t a = (uintptr_t)(buf + ks);
            const unsigned int *w32 = reinterpret_cast<const unsigned int *>(a & ~(uintptr_t)3);
            const int sh = (int)(a & 3) * 8;
            const unsigned int limit = (unsigned int)(((a & 3) + len + 3) >> 2);      // words that hold key bytes
            bool eq = true;
            unsigned int wi = 0;
            unsigned int g0 = limit ? __ldg(w32) : 0u;
            for (unsigned int o = 0; o < len; o += 8, wi += 2) {
                const unsigned int g1 = wi + 1 < limit ? __ldg(w32 + wi + 1) : 0u, g2 = wi + 2 < limit ? __ldg(w32 + wi + 2) : 0u;
                unsigned long long w = (unsigned long long)__funnelshift_r(g0, g1, sh) | ((unsigned long long)__funnelshift_r(g1, g2, sh) << 32);
                if (len - o < 8) w &= (1ull << (8 * (len - o))) - 1ull;
                if (w != __ldg(kw + 1 + (o >> 3))) { eq = false; break; }
                if (name) {                                                           // a key with a ' ' is no read name
                    const unsigned long long x = w ^ 0x2020202020202020ull;
                    if ((x - 0x0101010101010101ull) & ~x & 0x8080808080808080ull) { eq = false; break; }
                }
                g0 = g2;
            }
            if (eq) { pid = (int)(meta >> 32); return true; }
        }
        if (q3.z == 0u) return false;                      // the bucket has a free slot: nothing spilled past it
        bkt = (bkt + 1) & E.mask;
    }
}

template <class LD>
__device__ __forceinline__ bool exact_lookup(const ExactDevice &E, const LD &ld, long long start, unsigned int len,
                                             bool stop_at_space, int &pid)
{
    unsigned int keylen;
    const unsigned long long h = exact_hash(ld, start, len, stop_at_space, keylen);
    return exact_probe(E, h, ld, start, keylen, pid);
}

// key in global memory: bytes [p, p + len) of a buffer whose last readable byte is at or after p + len - 1
// rounded up to a 4-byte word (all library buffers carry 64 bytes of slack)
__device__ __forceinline__ bool exact_lookup_global(const ExactDevice &E, const uint8_t *p, unsigned int len,
                                                    bool stop_at_space, int &pid)
{
    const uintptr_t a = (uintptr_t)p;
    GlobalWords ld;
    ld.w = reinterpret_cast<const unsigned int *>(a & ~(uintptr_t)3);
    ld.limit = (long long)(((a & 3) + len + 3) >> 2);
    return exact_lookup(E, ld, (long long)(a & 3), len, stop_at_space, pid);
}

}  // namespace fxg
