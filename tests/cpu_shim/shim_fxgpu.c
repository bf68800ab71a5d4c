/*
 * tests/cpu_shim/shim_fxgpu.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A CPU stand-in for the *fields path* of libfxgpu's C ABI (include/fxgpu.h) so that the
 * host CLI's own logic -- reader, pair rules, output routing, option surface, raw->fields
 * re-routing -- can be exercised by the `-m "not gpu"` tests in a container without a GPU.
 * It is built by tests/test_cli_host.py into tests/cpu_shim/_build/ and reaches the CLI only
 * through an explicit LD_LIBRARY_PATH set by that test.  It is never installed, never on the
 * product's rpath, and proves nothing about the CUDA kernels: GPU parity is tests -m gpu.
 * The raw path answers "not canonical", which forces (and thereby tests) the re-route -- unless FXSHIM_RAW=1 is
 * set: then literal / sequence-target batches are walked on the CPU so that the host's threaded ingest (windows,
 * carried partial records, ordering over several contexts) can be tested without a GPU.
 */
#include "../../include/fxgpu.h"
#include "../../oracle/fxo_ac.h"
#include <stdlib.h>
#include <stdio.h>
#include <string.h>

struct fxg_ctx {
    fxg_config cfg;
    uint8_t *blob; uint32_t *off; uint32_t n;
    fxo_ac *ac;
    uint8_t *buf[2]; int state[2];          /* 0 free 1 lent 2 submitted */
    uint8_t *flags[2]; int32_t *pid[2]; uint64_t nrec[2], nhit[2]; int raw[2];
    int fifo[2]; int nfifo;
    const uint8_t *rawp[2]; size_t rawn[2]; fxg_run *runs[2];   /* FXSHIM_RAW=1: a plain CPU walk over canonical FASTQ */
    int pair[2]; const uint8_t *rawp2[2]; size_t rawn2[2];      /* mates: 1 interleaved, 2 two files */
};

static const char *g_err = "";

int fxg_device_count(int *n) { if (n) *n = 1; return FXG_OK; }
const char *fxg_last_error(const fxg_ctx *c) { (void)c; return g_err; }

int fxg_create(const fxg_config *cfg, fxg_ctx **out)
{
    fxg_ctx *c = (fxg_ctx *)calloc(1, sizeof *c);
    c->cfg = *cfg;
    if (!c->cfg.batch_bytes) c->cfg.batch_bytes = 1u << 20;
    *out = c;
    return FXG_OK;
}

void fxg_destroy(fxg_ctx *c)
{
    if (!c) return;
    for (int i = 0; i < 2; ++i) { free(c->buf[i]); free(c->flags[i]); free(c->pid[i]); }
    if (c->ac) fxo_ac_free(c->ac);
    free(c->blob); free(c->off); free(c);
}

int fxg_set_patterns(fxg_ctx *c, const uint8_t *blob, const uint32_t *off, uint32_t n)
{
    if (n == 0) { g_err = "empty pattern set"; return FXG_ERR_ARG; }
    c->n = n;
    c->off = (uint32_t *)malloc(sizeof(uint32_t) * (n + 1));
    memcpy(c->off, off, sizeof(uint32_t) * (n + 1));
    c->blob = (uint8_t *)malloc(off[n] + 1);
    memcpy(c->blob, blob, off[n]);
    if (c->cfg.mode == FXG_MODE_MULTI) {
        const char **p = (const char **)malloc(sizeof(char *) * n);
        size_t *l = (size_t *)malloc(sizeof(size_t) * n);
        for (uint32_t i = 0; i < n; ++i) { p[i] = (const char *)c->blob + off[i]; l[i] = off[i + 1] - off[i]; }
        c->ac = fxo_ac_build(p, l, (int)n);
        free(p); free(l);
    }
    return FXG_OK;
}

int fxg_acquire_buffer(fxg_ctx *c, void **pinned, size_t *cap)
{
    for (int i = 0; i < 2; ++i) if (c->state[i] == 1) { g_err = "a buffer is already lent out"; return FXG_ERR_STATE; }
    for (int i = 0; i < 2; ++i) if (c->state[i] == 0) {
        if (!c->buf[i]) c->buf[i] = (uint8_t *)malloc(c->cfg.batch_bytes + 64);
        c->state[i] = 1; *pinned = c->buf[i]; *cap = c->cfg.batch_bytes;
        return FXG_OK;
    }
    g_err = "both buffers are in flight";
    return FXG_ERR_STATE;
}

static const uint8_t *find_bytes(const uint8_t *h, size_t hn, const uint8_t *p, size_t pn);
static int lent(fxg_ctx *c) { for (int i = 0; i < 2; ++i) if (c->state[i] == 1) return i; return -1; }

int fxg_can_route(const fxg_ctx *c) { (void)c; return 0; }

int fxg_submit_raw(fxg_ctx *c, size_t nbytes, int is_last)
{
    (void)is_last;
    int s = lent(c);
    if (s < 0) { g_err = "no acquired buffer"; return FXG_ERR_STATE; }
    c->raw[s] = nbytes ? 1 : 2; c->state[s] = 2; c->fifo[c->nfifo++] = s;
    return FXG_OK;
}

int fxg_submit_raw_from(fxg_ctx *c, const void *h, size_t n, int l)
{
    (void)l;
    for (int i = 0; i < 2; ++i) if (c->state[i] == 0) {
        c->raw[i] = n ? 1 : 2; c->state[i] = 2; c->fifo[c->nfifo++] = i;
        c->rawp[i] = (const uint8_t *)h; c->rawn[i] = n; c->pair[i] = 0;
        return FXG_OK;
    }
    g_err = "both slots are in flight";
    return FXG_ERR_STATE;
}
int fxg_submit_raw_interleaved_from(fxg_ctx *c, const void *h, size_t n, int l)
{
    int rc = fxg_submit_raw_from(c, h, n, l);
    if (rc == FXG_OK) c->pair[c->fifo[c->nfifo - 1]] = 1;
    return rc;
}
int fxg_submit_raw_pair_from(fxg_ctx *c, const void *h1, size_t n1, const void *h2, size_t n2, int l)
{
    int rc = fxg_submit_raw_from(c, h1, n1, l);
    if (rc == FXG_OK) { int s = c->fifo[c->nfifo - 1]; c->pair[s] = 2; c->rawp2[s] = (const uint8_t *)h2; c->rawn2[s] = n2; }
    return rc;
}

/* FXSHIM_RAW: records of a part of whole 4-line records; returns the record count or -1 */
static long shim_walk(fxg_ctx *c, const uint8_t *b, size_t n, size_t base, uint8_t **hits, fxg_run **recs, long have)
{
    size_t pos = 0;
    while (pos < n) {
        size_t e[4], q = pos;
        for (int k = 0; k < 4; ++k) {
            const uint8_t *nl = (const uint8_t *)memchr(b + q, '\n', n - q);
            if (!nl) return -1;
            e[k] = (size_t)(nl - b); q = e[k] + 1;
        }
        if (b[pos] != '@') return -1;
        int hit = 0;
        for (uint32_t p = 0; p < c->n && p < 2 && !hit; ++p)
            hit = find_bytes(b + e[0] + 1, e[1] - e[0] - 1, c->blob + c->off[p], c->off[p + 1] - c->off[p]) != NULL;
        (*hits)[have] = (uint8_t)hit; (*recs)[have].off = base + pos; (*recs)[have].len = q - pos;
        ++have; pos = q;
    }
    return have;
}
int fxg_host_alloc(void **p, size_t n) { *p = malloc(n); return *p ? FXG_OK : FXG_ERR_NOMEM; }
int fxg_host_free(void *p) { free(p); return FXG_OK; }

static const uint8_t *find_bytes(const uint8_t *h, size_t hn, const uint8_t *p, size_t pn)
{
    if (pn == 0) return h;
    if (pn > hn) return NULL;
    for (size_t i = 0; i + pn <= hn; ++i) if (h[i] == p[0] && !memcmp(h + i, p, pn)) return h + i;
    return NULL;
}

int fxg_submit_fields(fxg_ctx *c, size_t nbytes, const uint32_t *off, uint32_t nrec)
{
    (void)nbytes;
    int s = lent(c);
    if (s < 0) { g_err = "no acquired buffer"; return FXG_ERR_STATE; }
    free(c->flags[s]); free(c->pid[s]);
    c->flags[s] = (uint8_t *)calloc(nrec + 1, 1);
    c->pid[s] = (int32_t *)malloc(sizeof(int32_t) * (nrec + 1));
    c->nhit[s] = 0; c->nrec[s] = nrec; c->raw[s] = 0;
    for (uint32_t k = 0; k < nrec; ++k) {
        const uint8_t *f = c->buf[s] + off[k];
        size_t fl = off[k + 1] - off[k];
        int id = -1;
        if (c->cfg.mode == FXG_MODE_LITERAL) {
            for (uint32_t q = 0; q < c->n && q < 2 && id < 0; ++q)
                if (find_bytes(f, fl, c->blob + c->off[q], c->off[q + 1] - c->off[q])) id = (int)q;
        } else if (c->cfg.mode == FXG_MODE_MULTI) {
            size_t end;
            id = fxo_ac_first(c->ac, (const char *)f, fl, &end);
        } else {
            for (uint32_t q = 0; q < c->n && id < 0; ++q)
                if (c->off[q + 1] - c->off[q] == fl && !memcmp(c->blob + c->off[q], f, fl)) id = (int)q;
        }
        c->flags[s][k] = id >= 0; c->pid[s][k] = id; c->nhit[s] += id >= 0;
    }
    c->state[s] = 2; c->fifo[c->nfifo++] = s;
    return FXG_OK;
}

int fxg_collect(fxg_ctx *c, fxg_result *r)
{
    if (!c->nfifo) { g_err = "nothing submitted"; return FXG_ERR_STATE; }
    int s = c->fifo[0];
    c->fifo[0] = c->fifo[1]; c->nfifo--;
    memset(r, 0, sizeof *r);
    if (c->raw[s] == 1 && c->pair[s] && getenv("FXSHIM_RAW") && c->cfg.mode == FXG_MODE_LITERAL && c->cfg.target == FXG_TARGET_SEQ) {
        /* mates (host-ingest tests): literal rule select = -v ? !m1 : m1 | m2; counts in pairs */
        const size_t cap = (c->rawn[s] + (c->pair[s] == 2 ? c->rawn2[s] : 0)) / 8 + 2;
        uint8_t *hits = (uint8_t *)malloc(cap); fxg_run *recs = (fxg_run *)malloc(sizeof(fxg_run) * cap);
        long n1 = shim_walk(c, c->rawp[s], c->rawn[s], 0, &hits, &recs, 0), nrec = n1;
        if (n1 >= 0 && c->pair[s] == 2) nrec = shim_walk(c, c->rawp2[s], c->rawn2[s], c->rawn[s], &hits, &recs, n1);
        r->canonical_ok = nrec >= 0 && (nrec & 1) == 0 && (c->pair[s] == 1 || nrec == 2 * n1);
        if (!r->canonical_ok) { r->violation = FXG_V_PAIRING; free(hits); free(recs); c->state[s] = 0; return FXG_OK; }
        const long np = nrec / 2;
        free(c->runs[s]); c->runs[s] = (fxg_run *)malloc(sizeof(fxg_run) * (size_t)(nrec + 2));
        uint8_t *sel = (uint8_t *)calloc((size_t)np + 1, 1);
        for (long i = 0; i < np; ++i) {
            const int m1 = c->pair[s] == 1 ? hits[2 * i] : hits[i], m2 = c->pair[s] == 1 ? hits[2 * i + 1] : hits[np + i];
            sel[i] = (uint8_t)(c->cfg.invert ? !m1 : (m1 | m2));
            r->n_selected += sel[i];
        }
        r->n_records = (uint64_t)np; r->n_matched = r->n_selected;
        if (!c->cfg.count_only) {
            if (c->pair[s] == 1) {                       /* merged runs of consecutive selected pairs */
                for (long i = 0; i < np; ++i) {
                    if (!sel[i]) continue;
                    const uint64_t off = recs[2 * i].off, len = recs[2 * i].len + recs[2 * i + 1].len;
                    if (r->n_runs && c->runs[s][r->n_runs - 1].off + c->runs[s][r->n_runs - 1].len == off && i && sel[i - 1])
                        c->runs[s][r->n_runs - 1].len += len;
                    else { c->runs[s][r->n_runs].off = off; c->runs[s][r->n_runs++].len = len; }
                }
            } else {                                     /* one run per record: file 1's, then their mates */
                for (int which = 0; which < 2; ++which)
                    for (long i = 0; i < np; ++i) if (sel[i]) c->runs[s][r->n_runs++] = recs[which * np + i];
            }
        }
        r->runs = c->runs[s];
        r->consumed_bytes = c->rawn[s] + (c->pair[s] == 2 ? c->rawn2[s] : 0);
        free(sel); free(hits); free(recs);
    } else if (c->raw[s] == 1 && !c->pair[s] && getenv("FXSHIM_RAW") && c->cfg.mode == FXG_MODE_LITERAL && c->cfg.target == FXG_TARGET_SEQ) {
        /* host-ingest tests: the batch must be whole 4-line records; literal search on the sequence line, runs of
         * consecutive selected records (NOT the product's algorithm, and no validation beyond the line count) */
        const uint8_t *b = c->rawp[s]; size_t n = c->rawn[s], pos = 0, run_start = 0; int in_run = 0;
        free(c->runs[s]); c->runs[s] = (fxg_run *)malloc(sizeof(fxg_run) * (n / 8 + 2));
        r->canonical_ok = 1;
        while (pos < n) {
            size_t e[4], q = pos; int ok = 1;
            for (int k = 0; k < 4; ++k) {
                const uint8_t *nl = (const uint8_t *)memchr(b + q, '\n', n - q);
                if (!nl) { ok = 0; break; }
                e[k] = (size_t)(nl - b); q = e[k] + 1;
            }
            if (!ok || b[pos] != '@') { r->canonical_ok = 0; r->violation = FXG_V_TRUNCATED; break; }
            int hit = 0;
            for (uint32_t p = 0; p < c->n && p < 2 && !hit; ++p)
                hit = find_bytes(b + e[0] + 1, e[1] - e[0] - 1, c->blob + c->off[p], c->off[p + 1] - c->off[p]) != NULL;
            r->n_records++; r->n_matched += hit;
            int sel = hit ^ (c->cfg.invert ? 1 : 0);
            r->n_selected += sel;
            if (sel && !in_run) { in_run = 1; run_start = pos; }
            if (!sel && in_run) { in_run = 0; c->runs[s][r->n_runs].off = run_start; c->runs[s][r->n_runs++].len = pos - run_start; }
            pos = q;
        }
        if (in_run) { c->runs[s][r->n_runs].off = run_start; c->runs[s][r->n_runs++].len = pos - run_start; }
        if (c->cfg.count_only) r->n_runs = 0;
        r->runs = c->runs[s];
        r->consumed_bytes = r->canonical_ok ? n : 0;
    } else if (c->raw[s]) {                  /* raw path: "not canonical" (empty batches are fine) */
        r->canonical_ok = c->raw[s] == 2;
        r->violation = c->raw[s] == 2 ? 0 : FXG_V_DENSE_LINES;
    } else {
        r->canonical_ok = 1; r->n_records = c->nrec[s]; r->n_matched = r->n_selected = c->nhit[s];
        r->flags = c->flags[s]; r->pattern_id = c->cfg.want_pattern_id ? c->pid[s] : NULL;
    }
    c->state[s] = 0;
    return FXG_OK;
}

int fxg_count(fxg_ctx *c, uint64_t *t) { (void)c; *t = 0; return FXG_OK; }

size_t fxg_find_record_cut(const uint8_t *buf, size_t n)
{
    /* last line start that opens a record: '@' there and '+' at the start of the line after next */
    size_t pos = n;
    for (int tries = 0; tries < 64 && pos > 0; ++tries) {
        size_t c = pos - 1;
        while (c > 0 && buf[c - 1] != '\n') --c;
        if (buf[c] == '@') {
            const uint8_t *e1 = (const uint8_t *)memchr(buf + c, '\n', n - c);
            if (e1) {
                const uint8_t *e2 = (const uint8_t *)memchr(e1 + 1, '\n', n - (size_t)(e1 + 1 - buf));
                if (e2 && (size_t)(e2 + 1 - buf) < n && e2[1] == '+') return c;
            }
        }
        pos = c;
    }
    return 0;
}
