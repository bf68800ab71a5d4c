/*
 * include/fxgpu.h -- C ABI of libfxgpu.so, the B200 (sm_100a) implementation of
 * fxtract's per-record search path.
 *
 * The reference (ctSkennerton/fxtract @ 958619a) has no plugin / FFI interface;
 * the seam this library sits behind is the three strategy functions that
 * main() dispatches to once per input (main.cpp:749-757):
 *     int simple_string_search(Fxstream&, const char*)   main.cpp:596
 *     int multipattern_search(Fxstream&)                 main.cpp:351
 *     int hash_search(Fxstream&)                         main.cpp:306
 * plus the pieces they call per record: prepare_data (main.cpp:291-304),
 * simple_string_search_core / strstr (main.cpp:568-588), ssearch_scan
 * (main.cpp:406,426; missing util/ submodule), FileManager::find's lookup role
 * (fileManager.cpp:171-177), on_match / on_match2 counting (main.cpp:258-289)
 * and the tokenizer kstream::read (kseq.hpp:120-164) whose record boundaries the
 * raw path recomputes on the GPU.
 *
 * Plain C: pointers and sizes only, no C++/torch types, no exceptions, never
 * writes to stdout.  All functions return 0 (FXG_OK) or a negative error code;
 * fxg_last_error() gives text.  A context is single-owner (not thread-safe).
 * There is NO CPU fallback: every entry point that computes fails with
 * FXG_ERR_CUDA when no CUDA device is usable.
 */
#ifndef FXGPU_H
#define FXGPU_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define FXG_OK 0
#define FXG_ERR_ARG (-1)
#define FXG_ERR_CUDA (-2)
#define FXG_ERR_NOMEM (-3)
#define FXG_ERR_STATE (-4)
#define FXG_ERR_UNSUPPORTED (-5)

/* strategy: which of the three reference functions is being replaced */
enum { FXG_MODE_LITERAL = 0,   /* simple_string_search  main.cpp:596 */
       FXG_MODE_MULTI = 1,     /* multipattern_search   main.cpp:351 */
       FXG_MODE_EXACT = 2 };   /* hash_search           main.cpp:306 */

/* target field, prepare_data main.cpp:291-304 (precedence H > Q > C > seq) */
enum { FXG_TARGET_SEQ = 0, FXG_TARGET_NAME = 1, FXG_TARGET_COMMENT = 2, FXG_TARGET_QUAL = 3 };

/* canonical-FASTQ predicate violations reported by the raw path (bitmask) */
enum { FXG_V_BYTE_RANGE = 1,      /* byte outside {'\n'} U [0x20,0x7e] */
       FXG_V_SEQ_CLASS = 2,       /* sequence byte kseq would drop or stop at (kseq.hpp:138-144) */
       FXG_V_QUAL_CLASS = 4,      /* quality byte outside 33..126 (kseq.hpp:156-159) */
       FXG_V_HEADER_AT = 8,       /* record does not start with '@' */
       FXG_V_EMPTY_SEQ = 16,      /* empty sequence line (record would print as FASTA, kseq.cpp:22) */
       FXG_V_PLUS_LINE = 32,      /* third line is not exactly "+" (kseq.cpp:45 prints bare '+') */
       FXG_V_QUAL_LEN = 64,       /* |qual| != |seq| (kseq.hpp:161) */
       FXG_V_HEADER_SPACE = 128,  /* header separator would not round-trip (kseq.cpp:35-37) */
       FXG_V_DENSE_LINES = 256,   /* pathological newline density */
       FXG_V_TRUNCATED = 512,     /* line count not a multiple of 4 / no final newline */
       FXG_V_PAIRING = 1024 };    /* pair submissions: the mates do not pair up inside the batch (odd record
                                   * count, or the two files' parts do not hold the same number of records) */

typedef struct fxg_ctx fxg_ctx;

typedef struct fxg_config {
    int32_t device;           /* CUDA ordinal */
    int32_t mode;             /* FXG_MODE_* */
    int32_t target;           /* FXG_TARGET_* */
    int32_t invert;           /* -v : raw path selects non-matching records */
    int32_t count_only;       /* -c : raw path returns counts only (no runs) */
    int32_t want_pattern_id;  /* fields path: also return first-hit pattern id per field */
    uint64_t batch_bytes;     /* capacity of each pinned staging buffer; 0 = 256 MiB */
    void *stream;             /* cudaStream_t to launch on; NULL = context-owned stream */
    int32_t route_runs;       /* raw path, MULTI / EXACT, !invert, !count_only: per-pattern output routing
                               * (on_match2 main.cpp:285-287, hash_search main.cpp:315-324): every selected record
                               * is its own run and run_pattern_id[k] names the pattern that selected run k */
    int32_t reserved;
} fxg_config;

typedef struct fxg_run { uint64_t off, len; } fxg_run;   /* selected byte range of the submitted batch */

typedef struct fxg_result {
    uint64_t n_records;       /* complete records (raw) / fields (fields path) in the batch */
    uint64_t n_matched;       /* records/fields with >= 1 pattern hit (before invert) */
    uint64_t n_selected;      /* raw path: after invert; fields path: == n_matched */
    uint64_t n_lines;         /* raw path: '\n' count */
    uint64_t consumed_bytes;  /* raw path: bytes covered by complete records */
    int32_t canonical_ok;     /* raw path: 1 iff the batch satisfies the canonical-FASTQ predicate */
    uint32_t violation;       /* FXG_V_* bitmask when canonical_ok == 0 */
    uint64_t first_bad_offset;/* smallest offset at which a violation was seen (diagnostic) */
    uint64_t n_runs;          /* raw path, !count_only */
    const fxg_run *runs;      /* selected byte ranges in input order; owned by ctx until next collect */
    const uint8_t *flags;     /* fields path: 1 byte per field, 1 = hit (NOT inverted) */
    const int32_t *pattern_id;/* fields path, want_pattern_id: first-hit pattern (sorted order) or -1 */
    const int32_t *run_pattern_id; /* raw path, route_runs: pattern id per run (EXACT: the key; MULTI: the first
                               * reported hit = smallest end offset, then longest pattern -- oracle/fxo_ac.h) */
} fxg_result;

/* ---- lifecycle ---------------------------------------------------------- */
int fxg_device_count(int *n);
int fxg_create(const fxg_config *cfg, fxg_ctx **out);
void fxg_destroy(fxg_ctx *ctx);
const char *fxg_last_error(const fxg_ctx *ctx);   /* ctx may be NULL: last global error */

/* Patterns: caller has already applied -r expansion, de-duplication and the
 * std::map sort (main.cpp:356-360) -- pattern ids are indices into this list,
 * because on_match2 maps strnum -> output file (main.cpp:285-287).
 * blob = concatenated pattern bytes, off[n+1] = offsets. LITERAL mode takes
 * 1 or 2 patterns (pattern and optional reverse complement, main.cpp:557-565). */
int fxg_set_patterns(fxg_ctx *ctx, const uint8_t *blob, const uint32_t *off, uint32_t n);

/* ---- host (streaming) path: pinned double buffers + cudaMemcpyAsync ----- */
/* Lends the next free pinned staging buffer (capacity cfg.batch_bytes). */
int fxg_acquire_buffer(fxg_ctx *ctx, void **pinned, size_t *cap);
/* Raw path: the acquired buffer holds nbytes of uncompressed FASTQ made of whole
 * records (cut with fxg_find_record_cut); asynchronous. */
int fxg_submit_raw(fxg_ctx *ctx, size_t nbytes, int is_last);
/* Same, but the batch lives in caller-owned host memory (ideally pinned with
 * fxg_host_alloc): copied straight from there, no staging copy.  nbytes must be
 * <= cfg.batch_bytes; the memory must stay valid until the batch is collected. */
int fxg_submit_raw_from(fxg_ctx *ctx, const void *host_buf, size_t nbytes, int is_last);
/* Mates on the raw path (Fxstream::read, fx.cpp:58-77; the pair rules of simple_string_search main.cpp:599-648,
 * hash_search main.cpp:311-346 and multipattern_search main.cpp:388-446).  Both calls take caller-owned (ideally
 * pinned) memory like fxg_submit_raw_from and are collected with fxg_collect.  The per-strategy rule is applied on
 * the device: a pair is selected when either mate has a hit; under cfg.invert when read 1 has none (LITERAL, EXACT:
 * main.cpp:618-619, :331-343) or when neither mate has one (MULTI: main.cpp:427-438).  In the result n_records,
 * n_matched and n_selected count PAIRS (n_matched == n_selected), and a batch whose mates do not pair up raises
 * FXG_V_PAIRING (canonical_ok = 0: the caller re-routes it like any other rejected batch).  route_runs: every
 * record of a selected pair carries the first-hit pattern id of that record (-1 when the record itself has none);
 * the pair's file is read 1's pattern if it has one, else read 2's (main.cpp:408-435, :315-337).
 *   interleaved (-I): mates are records 2i and 2i+1 of the batch, which must hold whole pairs.  Runs as in
 *     fxg_submit_raw_from (a selected pair is 8 consecutive lines).
 *   two files: the batch is n1 bytes of file 1 and n2 bytes of file 2 (n1 + n2 <= cfg.batch_bytes), each made of
 *     the same number of whole records; mate i of file 1 pairs with mate i of file 2.  Every selected record is
 *     its own run: the first n_runs / 2 runs are the selected records of file 1 (offsets into host1), the second
 *     half their mates in the same order, with offsets relative to the concatenation (off - n1 indexes host2). */
int fxg_submit_raw_interleaved_from(fxg_ctx *ctx, const void *host_buf, size_t nbytes, int is_last);
int fxg_submit_raw_pair_from(fxg_ctx *ctx, const void *host1, size_t n1, const void *host2, size_t n2, int is_last);
/* Fields path: the acquired buffer holds nrec target fields back to back
 * (nbytes total); off[nrec+1] are their offsets; asynchronous. */
int fxg_submit_fields(fxg_ctx *ctx, size_t nbytes, const uint32_t *off, uint32_t nrec);
/* Waits for the oldest submitted batch and returns its result (input order). */
int fxg_collect(fxg_ctx *ctx, fxg_result *res);
/* Sum of n_selected over everything collected so far (-c, main.cpp:759-761). */
int fxg_count(fxg_ctx *ctx, uint64_t *total);

/* Largest prefix length <= n of buf that ends on a FASTQ record boundary
 * (host helper for cutting batches / GPU shards; verified by the GPU predicate). */
size_t fxg_find_record_cut(const uint8_t *buf, size_t n);

/* ---- device-resident path (inputs already in HBM) ------------------------ */
/* Enqueues the raw scan over d_buf[0..nbytes) (whole records, ends in '\n') on the
 * context's stream; fxg_scan_device_finish() synchronises and fills res.
 * d_buf must be 16-byte aligned (the tiles are fetched with bulk copies; FXG_ERR_ARG otherwise) and readable for
 * 16 bytes past nbytes (edge look-ups and candidate verification read whole aligned words). */
int fxg_scan_device_async(fxg_ctx *ctx, const void *d_buf, uint64_t nbytes);
/* The same for mates already in HBM: split == 0: interleaved; else d_buf[0..split) holds file 1's records and
 * d_buf[split..nbytes) their mates (see fxg_submit_raw_pair_from). */
int fxg_scan_device_pairs_async(fxg_ctx *ctx, const void *d_buf, uint64_t nbytes, uint64_t split);
int fxg_scan_device_finish(fxg_ctx *ctx, fxg_result *res);
/* ---- allocation-free count over a device-resident batch ------------------ */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The literal search of fxg_scan_device_async (FXG_MODE_LITERAL contexts only; invert / count_only / route_runs
 * are ignored: counts only) over d_buf[0..nbytes), with every scratch table of the scan -- per-tile line tables,
 * prefix sums, line-hit words, hit bit set, counters: about 1.6 MB -- in a file-scope __device__ array of the
 * library instead of cudaMalloc'ed memory.  Every GPU that loads the library holds that array until the process
 * exits.  The call allocates nothing on the device (no cudaMalloc / cudaFree, no allocating library routine); it
 * only enqueues memsets and kernels on the context's stream and returns without synchronising.  The result is
 * written to d_out, a caller-owned DEVICE buffer of one fxg_noalloc_counts, in stream order.
 *   limit : nbytes <= FXG_NOALLOC_MAX_BYTES (8 MiB); larger batches are refused with FXG_ERR_ARG and a message.
 *   d_buf : 16-byte aligned, readable for 16 bytes past nbytes (like fxg_scan_device_async).
 *   sharing: all calls on one GPU -- from any context -- share the one scratch array.  What must start at zero is
 *            cleared by the call itself (stream-ordered), but two calls that overlap on the GPU (different streams,
 *            or different contexts with their own streams) RACE on it: serialise them. */
#define FXG_NOALLOC_MAX_BYTES (8u << 20)
typedef struct fxg_noalloc_counts {
    uint64_t n_lines;         /* '\n' count */
    uint64_t n_records;       /* n_lines / 4 */
    uint64_t n_matched;       /* records with >= 1 hit (before invert) */
    uint64_t first_bad_offset;/* as in fxg_result (all ones: none) */
    uint32_t violation;       /* FXG_V_* bitmask; 0 = the batch satisfies the canonical-FASTQ predicate */
    uint32_t canonical_ok;
} fxg_noalloc_counts;
int fxg_count_device_noalloc(fxg_ctx *ctx, const void *d_buf, uint64_t nbytes, fxg_noalloc_counts *d_out);

/* 1 iff the context can serve cfg.route_runs on the raw path with the current pattern set (EXACT always; MULTI
 * when the set is served by the fused window matcher); otherwise the caller keeps such inputs on the fields path. */
int fxg_can_route(const fxg_ctx *ctx);
/* Number of kernel launches the last *_async / submit_* call enqueued. */
int fxg_last_launch_count(const fxg_ctx *ctx);
/* Bytes the last collected / finished raw batch brought back to the host (counters + selected runs). */
uint64_t fxg_last_d2h_bytes(const fxg_ctx *ctx);
/* Measurement aid: when enabled, every raw-path enqueue brackets its kernels with
 * CUDA events on the context's stream.  After the batch has been collected /
 * finished, fxg_last_kernel_ms() returns the device time of the dominant kernel
 * (scan_ms: the fused scan kernel alone) and of the whole kernel pipeline of that
 * batch (pipeline_ms: memsets + scan + second-stage matcher + compaction). */
int fxg_set_profiling(fxg_ctx *ctx, int enabled);
int fxg_last_kernel_ms(fxg_ctx *ctx, float *scan_ms, float *pipeline_ms);

/* ---- synthetic workload (deterministic, counter-based) */
typedef struct fxg_synth_spec {
    uint64_t seed;
    uint64_t first_record;    /* global index of record 0 of this buffer */
    uint64_t n_records;
    uint32_t read_len;        /* bases per read (150) */
    uint32_t plant_mode;      /* 0 none, 1 = plant patterns of the synthetic pattern set */
    uint64_t pattern_seed;    /* seed of the synthetic pattern set */
    uint32_t n_patterns;      /* size of that set */
    uint32_t pattern_len;     /* 20 */
    uint32_t plant_per_million; /* planted reads per million (10000 = 1 %) */
    uint32_t plant_revcomp;   /* 1 = half of the planted copies are reverse-complemented */
} fxg_synth_spec;
uint64_t fxg_synth_bytes(const fxg_synth_spec *s);                 /* exact byte size */
uint64_t fxg_synth_record_offset(const fxg_synth_spec *s, uint64_t i); /* offset of record i (local index) */
int fxg_synth_host(const fxg_synth_spec *s, uint8_t *out);         /* CPU generator */
int fxg_synth_device(fxg_ctx *ctx, const fxg_synth_spec *s, void *d_out); /* same bytes, generated on the GPU */
int fxg_synth_pattern(const fxg_synth_spec *s, uint32_t p, uint8_t *out /* pattern_len bytes */);
int fxg_synth_name(const fxg_synth_spec *s, uint64_t i, uint8_t *out, uint32_t *len); /* read name (no '@') */

/* pinned host allocation helpers (bench / tests) */
int fxg_host_alloc(void **p, size_t n);
int fxg_host_free(void *p);
/* synchronous device -> host copy on the context's stream (bench / tests: materialise a
 * device-generated synthetic shard in host memory) */
int fxg_copy_to_host(fxg_ctx *ctx, void *host_dst, const void *d_src, size_t n);
/* synchronous host -> device copy on the context's stream (bench: the pinned-H2D probe that gives the end-to-end
 * figure its PCIe / host-ingest bound) */
int fxg_copy_to_device(fxg_ctx *ctx, void *d_dst, const void *host_src, size_t n);

#ifdef __cplusplus
}
#endif
#endif
