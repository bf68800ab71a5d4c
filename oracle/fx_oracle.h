/*
 * oracle/fx_oracle.h -- TEST INFRASTRUCTURE ONLY (never linked into, imported by
 * or executed from the product path; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may use it).
 *
 * Plain-C restatement of fxtract's per-record search path (reference @ 958619a):
 * tokenizer, pair stream, target selection, reverse complement, the three
 * literal strategies (single / multi / exact), invert / count / -l and the
 * FileManager routing.  Every function cites the reference file:line it follows.
 *
 * Parity status: PINNED.  Checked (tests/test_oracle.py) against every
 * non-PCRE golden vector of the reference's own fxtract_test suite and, in this
 * container, against the reference binary itself (oracle/_ref/fxtract, built by
 * oracle/Makefile from the unmodified sources) on randomised inputs.  The one
 * unpinned corner is the multi-pattern engine's report order when several
 * patterns hit one record with different output files (see fxo_ac.h).
 */
#ifndef FX_ORACLE_H
#define FX_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* main.cpp:36-76 (struct Options), flags that reach the literal strategies */
typedef struct {
    int H, Q, C, X, r, I, v, c, l, S;
    int have_f;   /* -f given (pattern file) */
} fxo_opts;

typedef struct fxo_job fxo_job;

fxo_job *fxo_job_new(const fxo_opts *o);
void fxo_job_free(fxo_job *j);
/* main.cpp:667-674: single command-line pattern (+ revcomp if -r) */
void fxo_job_set_pattern(fxo_job *j, const char *pat);
/* main.cpp:222-256 tokenizePatternFile over an in-memory pattern file */
void fxo_job_load_patterns(fxo_job *j, const char *text, size_t n);
/* one iteration of main.cpp:691-768 over in-memory (already inflated) files;
 * f2 == NULL for -S / -I / lone trailing file */
int fxo_job_run(fxo_job *j, const char *name1, const uint8_t *f1, size_t n1,
                const char *name2, const uint8_t *f2, size_t n2);
/* outputs */
const uint8_t *fxo_job_stdout(const fxo_job *j, size_t *len);
int fxo_job_nfiles(const fxo_job *j);
const char *fxo_job_file(const fxo_job *j, int i, const uint8_t **data, size_t *len, int *created);
long fxo_job_matched(const fxo_job *j);

/* util.cpp:3-17 + util.h:11-20 */
void fxo_revcomp(char *s, size_t n);

/* kseq.hpp:120-164 over a whole in-memory file: parses every record kstream::read
 * accepts (stops at the first negative return) and packs the four fields.
 * Offsets arrays must hold max_rec+1 entries.  Returns number of records, or -1
 * if max_rec / blob capacity is too small. */
long fxo_parse_buffer(const uint8_t *buf, size_t n, long max_rec,
                      uint8_t *name_blob, uint64_t *name_off, size_t name_cap,
                      uint8_t *comm_blob, uint64_t *comm_off, size_t comm_cap,
                      uint8_t *seq_blob, uint64_t *seq_off, size_t seq_cap,
                      uint8_t *qual_blob, uint64_t *qual_off, size_t qual_cap);

/* main.cpp:568-588 on one field (NUL-terminated), rc may be NULL */
int fxo_literal(const char *pat, const char *rc, const char *hay);

#ifdef __cplusplus
}
#endif
#endif
