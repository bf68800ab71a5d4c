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
__global__ void __launch_bounds__(256) fields_ac_kernel(FieldSrc src, unsigned long long n, AcDevice A, FieldOut O)
{
    __shared__ uint8_t cls[256];
    cls[threadIdx.x & 255] = A.cls[threadIdx.x & 255];
    __syncthreads();
    n = src.count(n);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long nround = (n + 31) & ~31ull;      // keep whole warps in the loop for the ballot
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < nround; k += stride) {
        bool hit = false;
        int pid = -1;
        if (k < n) {
            const uint8_t *p; unsigned int len;
            src.get(k, p, len);
            int s = 0;
            pid = A.out[0];                   // empty pattern ends at offset 0
            unsigned int i = 0;
            while (pid < 0 && i < len) {
                s = __ldg(&A.next[(size_t)s * A.ncls + cls[p[i]]]);
                pid = __ldg(&A.out[s]);
                ++i;
            }
            hit = pid >= 0;
        }
        emit_thread_per_field(O, k, n, hit, pid);
    }
}

// ---------------------------------------------------------------- exact: thread per field, hash probe + verify
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
            unsigned long long h = fxg_hash_bytes(p, len);
            unsigned int slot = (unsigned int)(h >> 20) & E.mask;
            for (;;) {
                unsigned long long sh = __ldg(&E.slot_hash[slot]);
                if (sh == 0) break;
                if (sh == h) {
                    int cand = __ldg(&E.slot_pat[slot]);
                    unsigned int a = E.off[cand], b = E.off[cand + 1];
                    if (b - a == len) {
                        bool eq = true;
                        for (unsigned int i = 0; i < len; ++i) if (E.blob[a + i] != p[i]) { eq = false; break; }
                        if (eq) { hit = true; pid = cand; break; }
                    }
                }
                slot = (slot + 1) & E.mask;
            }
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
cudaError_t launch_fields_ac(const FieldSrc &src, unsigned long long n, const AcDevice &A, const FieldOut &O, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    fields_ac_kernel<<<grid_for(n, 256), 256, 0, st>>>(src, n, A, O);
    return cudaGetLastError();
}
cudaError_t launch_fields_exact(const FieldSrc &src, unsigned long long n, const ExactDevice &E, const FieldOut &O, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    fields_exact_kernel<<<grid_for(n, 256), 256, 0, st>>>(src, n, E, O);
    return cudaGetLastError();
}

}  // namespace fxg
