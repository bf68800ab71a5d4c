/*
 * oracle/fx_oracle.c -- TEST INFRASTRUCTURE ONLY.  See fx_oracle.h.
 *
 * CPU restatement (plain C) of the reference's per-record search path.  Written
 * from the reference's observable behaviour; file:line citations are relative to
 * /root/reference.  Not a copy: the reference is C++ (std::string / std::map /
 * templates over gzFile); this works on in-memory files with its own containers.
 */
#include "fx_oracle.h"
#include "fxo_ac.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- strings */
typedef struct { char *p; size_t n, cap; } str_t;

static void s_reserve(str_t *s, size_t need)
{
    if (need + 1 <= s->cap) return;
    size_t c = s->cap ? s->cap * 2 : 64;
    while (c < need + 1) c *= 2;
    s->p = (char *)realloc(s->p, c);
    s->cap = c;
}
static void s_clear(str_t *s) { s_reserve(s, 0); s->n = 0; s->p[0] = 0; }
static void s_push(str_t *s, char c) { s_reserve(s, s->n + 1); s->p[s->n++] = c; s->p[s->n] = 0; }
static void s_append(str_t *s, const char *d, size_t n)
{
    s_reserve(s, s->n + n);
    if (n) memcpy(s->p + s->n, d, n);
    s->n += n; s->p[s->n] = 0;
}
static void s_free(str_t *s) { free(s->p); s->p = NULL; s->n = s->cap = 0; }

/* libc ctype in the "C" locale, on the int the tokenizer sees (sign-extended char,
 * kseq.hpp:180,203,211) */
static int c_isspace(int c) { return c == ' ' || (c >= 9 && c <= 13); }
static int c_isgraph(int c) { return c >= 33 && c <= 126; }

/* ---------------------------------------------------------------- record */
/* kseq.hpp:79-97 (class kseq) */
typedef struct { str_t name, comment, seq, qual; int last_char; } rec_t;

static void rec_init(rec_t *r) { memset(r, 0, sizeof *r); s_clear(&r->name); s_clear(&r->comment); s_clear(&r->seq); s_clear(&r->qual); }
static void rec_free(rec_t *r) { s_free(&r->name); s_free(&r->comment); s_free(&r->seq); s_free(&r->qual); }
/* kseq.cpp:13-15 */
static int rec_empty(const rec_t *r) { return !r->name.n && !r->comment.n && !r->seq.n && !r->qual.n; }
/* kseq.cpp:22-24 */
static int rec_is_fasta(const rec_t *r) { return r->qual.n == 0; }
/* kseq.cpp:25-30: clear() keeps last_char */
static void rec_clear(rec_t *r) { s_clear(&r->name); s_clear(&r->comment); s_clear(&r->seq); s_clear(&r->qual); }

/* ---------------------------------------------------------------- stream */
/* kseq.hpp:101-118,167-181: 4096-byte refill buffer; EOF == short read */
typedef struct {
    const uint8_t *file; size_t fsize, fpos;
    const uint8_t *buf; int begin, end, is_eof;
} ks_t;

static void ks_open(ks_t *k, const uint8_t *f, size_t n)
{
    k->file = f; k->fsize = n; k->fpos = 0; k->buf = f; k->begin = 0; k->end = 0; k->is_eof = 0;
}
static void ks_refill(ks_t *k)
{
    size_t left = k->fsize - k->fpos;
    size_t chunk = left < 4096 ? left : 4096;
    k->buf = k->file + k->fpos; k->fpos += chunk;
    k->begin = 0; k->end = (int)chunk;
    if (k->end < 4096) k->is_eof = 1;
}
/* kseq.hpp:167-181 */
static int ks_getc(ks_t *k)
{
    if (k->is_eof && k->begin >= k->end) return -1;
    if (k->begin >= k->end) {
        ks_refill(k);
        if (k->end == 0) return -1;
    }
    return (int)(signed char)k->buf[k->begin++];
}
/* kseq.hpp:183-245 (delimiter 0 = any isspace, >1 = that byte) */
static int ks_getuntil(ks_t *k, int delimiter, str_t *str, int *dret)
{
    if (dret) *dret = 0;
    s_clear(str);
    if (k->begin >= k->end && k->is_eof) return -1;
    for (;;) {
        int i;
        if (k->begin >= k->end) {
            if (!k->is_eof) {
                ks_refill(k);
                if (k->end == 0) break;
            } else break;
        }
        if (delimiter > 1) {
            for (i = k->begin; i < k->end; ++i) if ((int)(signed char)k->buf[i] == delimiter) break;
        } else if (delimiter == 0) {
            for (i = k->begin; i < k->end; ++i) if (c_isspace((int)(signed char)k->buf[i])) break;
        } else {
            for (i = k->begin; i < k->end; ++i) {
                int c = (int)(signed char)k->buf[i];
                if (c_isspace(c) && c != ' ') break;
            }
        }
        s_append(str, (const char *)k->buf + k->begin, (size_t)(i - k->begin));
        k->begin = i + 1;
        if (i < k->end) {
            if (dret) *dret = (int)(signed char)k->buf[i];
            break;
        }
    }
    return (int)str->n;
}
/* kseq.hpp:120-164 */
static int ks_read(ks_t *k, rec_t *r)
{
    int c;
    if (r->last_char == 0) {
        while ((c = ks_getc(k)) != -1 && c != '>' && c != '@') { }
        if (c == -1) return -1;
        r->last_char = c;
    }
    s_clear(&r->comment); s_clear(&r->seq); s_clear(&r->qual);
    if (ks_getuntil(k, 0, &r->name, &c) < 0) return -1;
    if (c != '\n') ks_getuntil(k, '\n', &r->comment, NULL);
    while ((c = ks_getc(k)) != -1 && c != '>' && c != '+' && c != '@') {
        if (c_isgraph(c)) s_push(&r->seq, (char)c);
    }
    if (c == '>' || c == '@') r->last_char = c;
    if (c != '+') return (int)r->seq.n;
    while ((c = ks_getc(k)) != -1 && c != '\n') { }
    if (c == -1) return -2;
    while ((c = ks_getc(k)) != -1 && r->qual.n < r->seq.n) {
        if (c >= 33 && c <= 127) s_push(&r->qual, (char)c);
    }
    r->last_char = 0;
    if (r->seq.n != r->qual.n) return -2;
    return (int)r->seq.n;
}

/* ---------------------------------------------------------------- output sink */
typedef struct { char *name; str_t data; int opened; int is_stdout; size_t records; } sink_t;

/* kseq.cpp:32-48: fprintf("%s") => every field is cut at its first NUL */
static void rec_print(const rec_t *r, str_t *out)
{
    if (rec_empty(r)) return;
    if (rec_is_fasta(r)) {
        s_push(out, '>'); s_append(out, r->name.p, strlen(r->name.p));
        if (r->comment.n) { s_push(out, ' '); s_append(out, r->comment.p, strlen(r->comment.p)); }
        s_push(out, '\n'); s_append(out, r->seq.p, strlen(r->seq.p)); s_push(out, '\n');
    } else {
        s_push(out, '@'); s_append(out, r->name.p, strlen(r->name.p));
        if (r->comment.n) { s_push(out, ' '); s_append(out, r->comment.p, strlen(r->comment.p)); }
        s_push(out, '\n'); s_append(out, r->seq.p, strlen(r->seq.p));
        s_append(out, "\n+\n", 3); s_append(out, r->qual.p, strlen(r->qual.p)); s_push(out, '\n');
    }
}

/* ---------------------------------------------------------------- job */
typedef struct { char *pat; size_t len; int file; } pent_t;   /* fileManager.h:55 patternMapping */

struct fxo_job {
    fxo_opts o;
    pent_t *pats; int npats, cap;   /* kept sorted like std::map<std::string,int> */
    sink_t *files; int nfiles;      /* fileManager.h:57 files; index 0.. in creation order */
    char *pattern;                  /* main.cpp:657 */
    long matched;                   /* main.cpp:80 matched_records (cumulative) */
};

static int key_cmp(const char *a, size_t na, const char *b, size_t nb)
{
    size_t m = na < nb ? na : nb;
    int c = m ? memcmp(a, b, m) : 0;
    if (c) return c;
    return (na > nb) - (na < nb);
}
static int pat_find(const fxo_job *j, const char *k, size_t n)
{
    int lo = 0, hi = j->npats - 1;
    while (lo <= hi) {
        int mid = (lo + hi) / 2;
        int c = key_cmp(j->pats[mid].pat, j->pats[mid].len, k, n);
        if (c == 0) return mid;
        if (c < 0) lo = mid + 1; else hi = mid - 1;
    }
    return -1 - lo;
}
static int file_index(fxo_job *j, const char *name)   /* filenameMapping, fileManager.h:56 */
{
    for (int i = 0; i < j->nfiles; ++i) if (!strcmp(j->files[i].name, name)) return i;
    j->files = (sink_t *)realloc(j->files, sizeof(sink_t) * (size_t)(j->nfiles + 1));
    sink_t *s = &j->files[j->nfiles];
    memset(s, 0, sizeof *s);
    s->name = strdup(name);
    s_clear(&s->data);
    s->is_stdout = (name[0] == 0);
    s->opened = s->is_stdout;       /* fileManager.cpp:141-147 */
    return j->nfiles++;
}
/* fileManager.cpp:76-120 and :123-169: first mapping of a pattern wins */
static void mgr_add(fxo_job *j, const char *pat, size_t len, const char *file)
{
    int pos = pat_find(j, pat, len);
    if (pos >= 0) return;
    pos = -1 - pos;
    if (j->npats == j->cap) {
        j->cap = j->cap ? j->cap * 2 : 16;
        j->pats = (pent_t *)realloc(j->pats, sizeof(pent_t) * (size_t)j->cap);
    }
    memmove(j->pats + pos + 1, j->pats + pos, sizeof(pent_t) * (size_t)(j->npats - pos));
    j->pats[pos].pat = (char *)malloc(len + 1);
    memcpy(j->pats[pos].pat, pat, len); j->pats[pos].pat[len] = 0;
    j->pats[pos].len = len;
    j->pats[pos].file = file_index(j, file ? file : "");
    j->npats++;
}
/* fileManager.cpp:171-202: lookup with the lazy-open side effect */
static sink_t *mgr_find(fxo_job *j, const char *key, size_t n)
{
    int pos = pat_find(j, key, n);
    if (pos < 0) return NULL;
    sink_t *s = &j->files[j->pats[pos].file];
    s->opened = 1;
    s->records++;
    return s;
}

fxo_job *fxo_job_new(const fxo_opts *o)
{
    fxo_job *j = (fxo_job *)calloc(1, sizeof *j);
    j->o = *o;
    return j;
}
void fxo_job_free(fxo_job *j)
{
    if (!j) return;
    for (int i = 0; i < j->npats; ++i) free(j->pats[i].pat);
    for (int i = 0; i < j->nfiles; ++i) { free(j->files[i].name); s_free(&j->files[i].data); }
    free(j->pats); free(j->files); free(j->pattern); free(j);
}

/* util.h:11-20 comp_tab, restated as rules: letters map to their IUPAC
 * complement (U->A, case kept), '`' (96) maps to '@' (64) exactly like the
 * reference table's entry, everything else is unchanged. */
static unsigned char comp_byte(unsigned char c)
{
    static const char up[] = "TVGHEFCDIJMLKNOPQYSAABWXRZ";   /* complement of 'A'..'Z' */
    if (c >= 'A' && c <= 'Z') return (unsigned char)up[c - 'A'];
    if (c >= 'a' && c <= 'z') return (unsigned char)(up[c - 'a'] + 32);
    if (c == 96) return 64;
    return c;
}
/* util.cpp:3-17 */
void fxo_revcomp(char *s, size_t n)
{
    size_t i;
    for (i = 0; i < n / 2; ++i) {
        unsigned char a = comp_byte((unsigned char)s[i]);
        unsigned char b = comp_byte((unsigned char)s[n - 1 - i]);
        s[i] = (char)b; s[n - 1 - i] = (char)a;
    }
    if (n & 1) s[n / 2] = (char)comp_byte((unsigned char)s[n / 2]);
}

/* main.cpp:667-674 */
void fxo_job_set_pattern(fxo_job *j, const char *pat)
{
    free(j->pattern);
    j->pattern = strdup(pat);
    mgr_add(j, pat, strlen(pat), NULL);
    if (j->o.r) {
        char *rc = strdup(pat);
        fxo_revcomp(rc, strlen(rc));
        mgr_add(j, rc, strlen(rc), NULL);
        free(rc);
    }
}

/* main.cpp:222-256 (+ split, :200-220): getline, skip empty, TAB-split */
void fxo_job_load_patterns(fxo_job *j, const char *text, size_t n)
{
    size_t pos = 0;
    while (pos < n) {
        size_t e = pos;
        while (e < n && text[e] != '\n') ++e;
        size_t len = e - pos;
        if (len) {
            const char *line = text + pos;
            size_t t1 = 0;
            while (t1 < len && line[t1] != '\t') ++t1;
            char *p = (char *)malloc(t1 + 1);
            memcpy(p, line, t1); p[t1] = 0;
            char *file = NULL;
            if (t1 < len) {
                size_t s2 = t1 + 1, t2 = s2;
                while (t2 < len && line[t2] != '\t') ++t2;
                file = (char *)malloc(t2 - s2 + 1);
                memcpy(file, line + s2, t2 - s2); file[t2 - s2] = 0;
            }
            mgr_add(j, p, t1, file);
            if (j->o.r) {
                fxo_revcomp(p, t1);
                mgr_add(j, p, t1, file);
            }
            free(p); free(file);
        }
        pos = e + 1;
    }
}

/* ---------------------------------------------------------------- pair stream */
/* fx.h:8-21 ReadPair, fx.cpp:30-81 Fxstream */
typedef struct { rec_t first, second; } pair_t;
typedef struct { ks_t k1, k2; int have2, interleaved; const char *n1, *n2; } fxs_t;

static int fxs_read(fxs_t *s, pair_t *p)
{
    int l = ks_read(&s->k1, &p->first);
    if (l < 0) return 1;
    if (s->interleaved) {
        if (ks_read(&s->k1, &p->second) < 0) return 1;
    } else if (s->have2) {
        if (ks_read(&s->k2, &p->second) < 0) return 1;
    }
    return 0;
}
static void pair_print(const pair_t *p, str_t *out) { rec_print(&p->first, out); rec_print(&p->second, out); }
static void pair_clear(pair_t *p) { rec_clear(&p->first); rec_clear(&p->second); }

/* main.cpp:291-304 */
static const str_t *prepare_data(const fxo_opts *o, const rec_t *r)
{
    if (o->H) return &r->name;
    else if (o->Q) { if (!rec_is_fasta(r)) return &r->qual; }
    else if (o->C) return &r->comment;
    return &r->seq;
}

/* main.cpp:258-271 */
static int on_match(fxo_job *j, str_t *out, const pair_t *p)
{
    if (j->o.l) return 1;
    if (j->o.c) { ++j->matched; return 0; }
    pair_print(p, out);
    return 0;
}

static void print_name(fxo_job *j, const char *name)
{
    str_t *so = &j->files[file_index(j, "")].data;
    s_append(so, name, strlen(name)); s_push(so, '\n');
}

/* main.cpp:568-588 */
int fxo_literal(const char *pat, const char *rc, const char *hay)
{
    if (strstr(hay, pat)) return 1;
    if (rc && strstr(hay, rc)) return 1;
    return 0;
}

/* main.cpp:596-652 (+ prepare_revcomp :557-565) */
static int simple_string_search(fxo_job *j, fxs_t *s)
{
    const fxo_opts *o = &j->o;
    str_t *so = &j->files[file_index(j, "")].data;
    char *rc = NULL;
    if ((o->r ? 1 : 0) ^ ((o->H | o->Q | o->C) ? 1 : 0)) {
        rc = strdup(j->pattern);
        fxo_revcomp(rc, strlen(rc));
    }
    const char *rcp = (rc && rc[0]) ? rc : NULL;    /* rcpattern.empty() test, :602 */
    pair_t p; rec_init(&p.first); rec_init(&p.second);
    int ret = 0;
    while (fxs_read(s, &p) == 0) {
        int m = fxo_literal(j->pattern, rcp, prepare_data(o, &p.first)->p);
        if (m) {
            if (!o->v) {
                if (on_match(j, so, &p) == 1) { print_name(j, s->n1); goto done; }
            }
        } else {
            if (o->v) on_match(j, so, &p);
            else if (!rec_empty(&p.second)) {
                m = fxo_literal(j->pattern, rcp, prepare_data(o, &p.second)->p);
                if (m) {
                    if (on_match(j, so, &p) == 1) { print_name(j, s->n1); goto done; }   /* :636 file 1's name */
                }
            }
        }
        pair_clear(&p);
    }
done:
    free(rc); rec_free(&p.first); rec_free(&p.second);
    return ret;
}

/* main.cpp:306-349 */
static int hash_search(fxo_job *j, fxs_t *s)
{
    const fxo_opts *o = &j->o;
    str_t *so = &j->files[file_index(j, "")].data;
    pair_t p; rec_init(&p.first); rec_init(&p.second);
    while (fxs_read(s, &p) == 0) {
        const str_t *d = prepare_data(o, &p.first);
        sink_t *out = mgr_find(j, d->p, d->n);
        if (out) {
            if (!o->v) {
                if (on_match(j, &out->data, &p) == 1) { print_name(j, s->n1); goto done; }
            }
        } else {
            if (!rec_empty(&p.second)) {
                d = prepare_data(o, &p.second);
                out = mgr_find(j, d->p, d->n);
                if (out) {
                    if (on_match(j, &out->data, &p) == 1) { print_name(j, s->n2 ? s->n2 : ""); goto done; }
                } else if (o->v) on_match(j, so, &p);
            } else if (o->v) on_match(j, so, &p);
        }
        pair_clear(&p);
    }
done:
    rec_free(&p.first); rec_free(&p.second);
    return 0;
}

/* main.cpp:273-289 */
struct m2ctx { fxo_job *j; const pair_t *p; };
static int on_match2(int strnum, const char *textp, void *ctx)
{
    struct m2ctx *c = (struct m2ctx *)ctx;
    fxo_job *j = c->j;
    (void)textp;
    if (j->o.v) return 1;
    if (j->o.l) return 1;
    if (j->o.c) { ++j->matched; return 1; }
    /* manager.find(pattv[strnum].ptr): std::string built from a C string */
    sink_t *out = mgr_find(j, j->pats[strnum].pat, strlen(j->pats[strnum].pat));
    if (out) pair_print(c->p, &out->data);
    return 1;
}

/* main.cpp:389-405 / :410-425: (ptr,len) of the target; NB read 2 under -Q scans seq (:416) */
static void multi_target(const fxo_opts *o, const rec_t *r, int second, const char **ptr, size_t *len, int *valid)
{
    *valid = 1;
    if (o->H) { *ptr = r->name.p; *len = r->name.n; }
    else if (o->Q) {
        if (!rec_is_fasta(r)) { *ptr = second ? r->seq.p : r->qual.p; *len = r->seq.n; }
        else *valid = 0;   /* reference leaves `data` unchanged/uninitialised (UB for read 1) */
    }
    else if (o->C) { *ptr = r->comment.p; *len = r->comment.n; }
    else { *ptr = r->seq.p; *len = r->seq.n; }
}

/* main.cpp:351-452 */
static int multipattern_search(fxo_job *j, fxs_t *s)
{
    const fxo_opts *o = &j->o;
    str_t *so = &j->files[file_index(j, "")].data;
    if (j->npats <= 0) return 1;
    const char **pp = (const char **)malloc(sizeof(char *) * (size_t)j->npats);
    size_t *pl = (size_t *)malloc(sizeof(size_t) * (size_t)j->npats);
    for (int i = 0; i < j->npats; ++i) { pp[i] = j->pats[i].pat; pl[i] = j->pats[i].len; }
    fxo_ac *ac = fxo_ac_build(pp, pl, j->npats);
    free(pp); free(pl);
    if (!ac) { fprintf(stderr, "problem initalizing multi-search\n"); return 1; }
    pair_t p; rec_init(&p.first); rec_init(&p.second);
    struct m2ctx ctx = { j, &p };
    const char *ptr = ""; size_t len = 0; int valid;
    while (fxs_read(s, &p) == 0) {
        multi_target(o, &p.first, 0, &ptr, &len, &valid);
        if (!valid) { ptr = ""; len = 0; }
        int ret = fxo_ac_scan(ac, ptr, len, on_match2, &ctx);
        if (!ret) {
            if (!rec_empty(&p.second)) {
                multi_target(o, &p.second, 1, &ptr, &len, &valid);   /* invalid => previous data kept */
                ret = fxo_ac_scan(ac, ptr, len, on_match2, &ctx);
                if (!ret && o->v) on_match(j, so, &p);
                else if (o->l) { print_name(j, s->n2 ? s->n2 : ""); goto done; }
            } else if (o->v) on_match(j, so, &p);
        } else if (o->l) { print_name(j, s->n1); goto done; }
        pair_clear(&p);
    }
done:
    rec_free(&p.first); rec_free(&p.second);
    fxo_ac_free(ac);
    return 0;
}

/* one iteration of main.cpp:691-768 (literal strategies only; dispatch order :749-757) */
int fxo_job_run(fxo_job *j, const char *name1, const uint8_t *f1, size_t n1,
                const char *name2, const uint8_t *f2, size_t n2)
{
    fxs_t s; memset(&s, 0, sizeof s);
    ks_open(&s.k1, f1, n1); s.n1 = name1;
    if (f2) { ks_open(&s.k2, f2, n2); s.have2 = 1; s.n2 = name2; }
    s.interleaved = j->o.I;
    file_index(j, "");   /* stdout sink always exists */
    if (j->o.X) hash_search(j, &s);
    else if (j->o.have_f) multipattern_search(j, &s);
    else simple_string_search(j, &s);
    if (j->o.c) {
        char tmp[32];
        int k = snprintf(tmp, sizeof tmp, "%d\n", (int)j->matched);   /* main.cpp:759-761 */
        s_append(&j->files[file_index(j, "")].data, tmp, (size_t)k);
    }
    return 0;
}

const uint8_t *fxo_job_stdout(const fxo_job *j, size_t *len)
{
    for (int i = 0; i < j->nfiles; ++i)
        if (j->files[i].is_stdout) { *len = j->files[i].data.n; return (const uint8_t *)j->files[i].data.p; }
    *len = 0; return (const uint8_t *)"";
}
int fxo_job_nfiles(const fxo_job *j) { return j->nfiles; }
const char *fxo_job_file(const fxo_job *j, int i, const uint8_t **data, size_t *len, int *created)
{
    const sink_t *s = &j->files[i];
    *data = (const uint8_t *)s->data.p; *len = s->data.n; *created = s->opened && !s->is_stdout;
    return s->name;
}
long fxo_job_matched(const fxo_job *j) { return j->matched; }

/* kseq.hpp:120-164 over a whole buffer, packed output (test helper) */
long fxo_parse_buffer(const uint8_t *buf, size_t n, long max_rec,
                      uint8_t *name_blob, uint64_t *name_off, size_t name_cap,
                      uint8_t *comm_blob, uint64_t *comm_off, size_t comm_cap,
                      uint8_t *seq_blob, uint64_t *seq_off, size_t seq_cap,
                      uint8_t *qual_blob, uint64_t *qual_off, size_t qual_cap)
{
    ks_t k; ks_open(&k, buf, n);
    rec_t r; rec_init(&r);
    long cnt = 0;
    uint64_t no = 0, co = 0, so = 0, qo = 0;
    name_off[0] = comm_off[0] = seq_off[0] = qual_off[0] = 0;
    while (ks_read(&k, &r) >= 0) {
        if (cnt >= max_rec || no + r.name.n > name_cap || co + r.comment.n > comm_cap ||
            so + r.seq.n > seq_cap || qo + r.qual.n > qual_cap) { rec_free(&r); return -1; }
        memcpy(name_blob + no, r.name.p, r.name.n); no += r.name.n;
        memcpy(comm_blob + co, r.comment.p, r.comment.n); co += r.comment.n;
        memcpy(seq_blob + so, r.seq.p, r.seq.n); so += r.seq.n;
        memcpy(qual_blob + qo, r.qual.p, r.qual.n); qo += r.qual.n;
        ++cnt;
        name_off[cnt] = no; comm_off[cnt] = co; seq_off[cnt] = so; qual_off[cnt] = qo;
        rec_clear(&r);
    }
    rec_free(&r);
    return cnt;
}

#ifdef FXO_MAIN
/* Command-line front end with the reference's argv surface for the literal
 * strategies (main.cpp:110-198 parseOptions, :654-770 main), so the restatement
 * can be diffed against oracle/_ref/fxtract on files.  Regex flags are rejected. */
#include <unistd.h>
#include <zlib.h>

static uint8_t *slurp(const char *path, size_t *n)
{
    /* gzopen failures are unchecked in the reference (fx.cpp:32-33): missing file => no records */
    gzFile g = gzopen(path, "r");
    uint8_t *b = (uint8_t *)malloc(1 << 16); size_t cap = 1 << 16, len = 0;
    if (g) {
        for (;;) {
            if (len + 65536 > cap) { cap *= 2; b = (uint8_t *)realloc(b, cap); }
            int k = gzread(g, b + len, 65536);
            if (k <= 0) break;
            len += (size_t)k;
        }
        gzclose(g);
    }
    *n = len; return b;
}

int main(int argc, char **argv)
{
    fxo_opts o; memset(&o, 0, sizeof o);
    const char *ffile = NULL;
    int c;
    while ((c = getopt(argc, argv, "HhCcIf:zjqVvrQGEPX12SlL")) != -1) {
        switch (c) {
        case 'c': o.c = 1; break;
        case 'f': ffile = optarg; o.have_f = 1; break;
        case 'I': o.I = 1; break;
        case 'S': o.S = 1; break;
        case 'Q': o.Q = 1; break;
        case 'V': puts("2.3"); exit(1);
        case 'X': o.X = 1; break;
        case 'H': o.H = 1; break;
        case 'l': o.l = 1; break;
        case 'L': break;
        case 'C': o.C = 1; break;
        case 'r': o.r = 1; break;
        case 'v': o.v = 1; break;
        case '1': case '2': case 'z': break;
        default: fprintf(stderr, "fxoracle: unsupported flag\n"); exit(2);
        }
    }
    fxo_job *j = fxo_job_new(&o);
    int idx = optind;
    if (!ffile) {
        if (idx >= argc - 1) { fprintf(stderr, "fxoracle: need pattern and input\n"); exit(2); }
        fxo_job_set_pattern(j, argv[idx++]);
    } else {
        if (idx > argc - 1) { fprintf(stderr, "fxoracle: need input\n"); exit(2); }
        size_t n; FILE *f = fopen(ffile, "rb");
        if (!f) { fprintf(stderr, "problem opening pattern file\n"); exit(1); }
        char *t = NULL; size_t cap = 0; n = 0;
        for (;;) { if (n + 65536 > cap) { cap = cap ? cap * 2 : 1 << 17; t = (char *)realloc(t, cap); }
                   size_t k = fread(t + n, 1, 65536, f); if (!k) break; n += k; }
        fclose(f);
        fxo_job_load_patterns(j, t, n); free(t);
    }
    while (idx < argc) {
        size_t n1 = 0, n2 = 0; uint8_t *b1, *b2 = NULL;
        const char *f2 = NULL;
        if (!o.S && !o.I) f2 = argv[idx + 1];
        b1 = slurp(argv[idx], &n1);
        if (f2) b2 = slurp(f2, &n2);
        fxo_job_run(j, argv[idx], b1, n1, f2, b2, n2);
        free(b1); free(b2);
        idx += (!o.S && !o.I) ? 2 : 1;
    }
    size_t n; const uint8_t *so = fxo_job_stdout(j, &n);
    fwrite(so, 1, n, stdout);
    for (int i = 0; i < fxo_job_nfiles(j); ++i) {
        const uint8_t *d; size_t len; int created;
        const char *name = fxo_job_file(j, i, &d, &len, &created);
        if (created) { FILE *f = fopen(name, "w"); if (f) { fwrite(d, 1, len, f); fclose(f); } }
    }
    fxo_job_free(j);
    return 0;
}
#endif
