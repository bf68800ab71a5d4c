// fxg_scan_kmer.cu -- G32 instantiations of the fused scan kernel (fxg_scan_kernel.cuh): the packed 2-bit k-mer
// matcher for pattern sets over {A,C,G,T} (replaces ssearch_scan, main.cpp:406,426, on the raw path), one kernel
// per probe stride.  The filter is staged behind the kernel's fixed shared-memory block: 16 KiB (2 CTAs of 16
// warps per SM) or 128 KiB (1 CTA per SM) -- same code, the host picks the size from the number of index entries.
#include "fxg_scan_kernel.cuh"

namespace fxg {

cudaError_t launch_scan_kmer(const ScanParams &P, cudaStream_t stream, int dev)
{
    const size_t filter_bytes = (size_t)4 << P.kmer.filter_log2_words;
    switch (P.mode) {
    case kModeKmer + 1: return launch_scan_tiles<G32, kModeKmer + 1, false>(P, filter_bytes, stream, dev, 0);
    case kModeKmer + 2: return launch_scan_tiles<G32, kModeKmer + 2, false>(P, filter_bytes, stream, dev, 0);
    case kModeKmer + 3: return launch_scan_tiles<G32, kModeKmer + 3, false>(P, filter_bytes, stream, dev, 0);
    case kModeKmer + 4: return launch_scan_tiles<G32, kModeKmer + 4, false>(P, filter_bytes, stream, dev, 0);
    default: return cudaErrorInvalidValue;
    }
}

// structure-only re-check of the tiles whose guessed phase was wrong (same geometry as the first launch)
cudaError_t launch_scan_kmer_redo(const ScanParams &P, cudaStream_t stream, int dev)
{
    return launch_scan_tiles<G32, kModeIndex, true>(P, 0, stream, dev, 64);
}

}  // namespace fxg
