// fxg_scan_kmer.cu -- window-matcher instantiations of the fused scan kernel (fxg_scan_kernel.cuh): the multi-pattern
// matcher of the raw path (replaces ssearch_scan, main.cpp:406,426), one kernel per probe stride and filter size.
// The Bloom filter is staged behind the kernel's fixed shared-memory blocks: 32 KiB (G32: 2 CTAs of 16 warps per SM,
// 32 KiB tiles) or 128 KiB (G16D: 1 CTA per SM whose 32 warps form two independent groups with 16 KiB tiles) -- same
// code; the host picks the size from the number of index entries.
#include "fxg_scan_kernel.cuh"

namespace fxg {

template <class G> static cudaError_t launch_mode(const ScanParams &P, size_t filter_bytes, cudaStream_t stream, int dev)
{
    switch (P.mode) {
    case kModeKmer + 2: return launch_scan_tiles<G, kModeKmer + 2, false>(P, filter_bytes, stream, dev, 0);
    case kModeKmer + 3: return launch_scan_tiles<G, kModeKmer + 3, false>(P, filter_bytes, stream, dev, 0);
    case kModeKmer + 4: return launch_scan_tiles<G, kModeKmer + 4, false>(P, filter_bytes, stream, dev, 0);
    default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_scan_kmer(const ScanParams &P, cudaStream_t stream, int dev)
{
    const size_t filter_bytes = (size_t)4 << P.kmer.filter_log2_words;
    if (P.kmer.filter_log2_words > 13) return launch_mode<G16D>(P, filter_bytes, stream, dev);
    return launch_mode<G32>(P, filter_bytes, stream, dev);
}

// structure-only re-check of the tiles whose guessed phase was wrong (same tile geometry as the first launch)
cudaError_t launch_scan_kmer_redo(const ScanParams &P, cudaStream_t stream, int dev)
{
    if (P.kmer.filter_log2_words > 13) return launch_scan_tiles<G16, kModeIndex, true>(P, 0, stream, dev, 64);
    return launch_scan_tiles<G32, kModeIndex, true>(P, 0, stream, dev, 64);
}

}  // namespace fxg
