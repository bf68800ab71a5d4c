/*
 * oracle/fxo_ac.c -- TEST INFRASTRUCTURE ONLY.  See fxo_ac.h for the contract
 * and the reference call sites (main.cpp:371-373,406,426,273-289) it restates.
 *
 * Textbook Aho-Corasick: byte-class compressed trie -> BFS failure links ->
 * dense DFA (next[state*ncls+cls]) + per-state "longest pattern ending here"
 * (own terminal, else inherited through the failure chain) + dictionary link
 * to the next shorter pattern ending at the same position.
 */
#include "fxo_ac.h"
#include <stdlib.h>
#include <string.h>

struct fxo_ac {
    int ncls;
    unsigned char cls[256];
    size_t nstates;
    int *next;      /* nstates * ncls */
    int *out;       /* longest pattern id ending at this state, or -1 */
    int *dict;      /* state of the next shorter pattern ending here, or -1 */
    int *term;      /* pattern id terminal exactly at this state (smallest id wins), or -1 */
};

fxo_ac *fxo_ac_build(const char *const *pats, const size_t *lens, int n)
{
    if (n <= 0) return NULL;
    fxo_ac *ac = (fxo_ac *)calloc(1, sizeof *ac);
    if (!ac) return NULL;
    /* byte classes: 0 = "byte not in any pattern" */
    int used[256]; memset(used, 0, sizeof used);
    size_t total = 1;
    for (int i = 0; i < n; ++i) {
        total += lens[i];
        for (size_t j = 0; j < lens[i]; ++j) used[(unsigned char)pats[i][j]] = 1;
    }
    ac->ncls = 1;
    for (int b = 0; b < 256; ++b) ac->cls[b] = used[b] ? (unsigned char)(ac->ncls++) : 0;
    if (ac->ncls > 255) { /* cannot happen with 256 bytes + class 0, keep uchar safe */ }
    int ncls = ac->ncls;
    /* trie (upper bound on states = total) */
    int *next = (int *)malloc(sizeof(int) * total * (size_t)ncls);
    int *term = (int *)malloc(sizeof(int) * total);
    if (!next || !term) { free(next); free(term); free(ac); return NULL; }
    for (int c = 0; c < ncls; ++c) next[c] = -1;
    term[0] = -1;
    size_t ns = 1;
    for (int i = 0; i < n; ++i) {
        size_t s = 0;
        for (size_t j = 0; j < lens[i]; ++j) {
            int c = ac->cls[(unsigned char)pats[i][j]];
            int t = next[s * ncls + c];
            if (t < 0) {
                t = (int)ns++;
                for (int k = 0; k < ncls; ++k) next[(size_t)t * ncls + k] = -1;
                term[t] = -1;
                next[s * ncls + c] = t;
            }
            s = (size_t)t;
        }
        if (term[s] < 0) term[s] = i;   /* duplicates: first id wins */
    }
    int *fail = (int *)malloc(sizeof(int) * ns);
    int *out = (int *)malloc(sizeof(int) * ns);
    int *dict = (int *)malloc(sizeof(int) * ns);
    int *queue = (int *)malloc(sizeof(int) * ns);
    if (!fail || !out || !dict || !queue) {
        free(fail); free(out); free(dict); free(queue); free(next); free(term); free(ac);
        return NULL;
    }
    size_t qh = 0, qt = 0;
    fail[0] = 0; out[0] = term[0]; dict[0] = -1;
    for (int c = 0; c < ncls; ++c) {
        int t = next[c];
        if (t < 0) next[c] = 0;
        else { fail[t] = 0; queue[qt++] = t; }
    }
    while (qh < qt) {
        int s = queue[qh++];
        int f = fail[s];
        /* dictionary link: nearest proper suffix state that is terminal */
        int d = (term[f] >= 0) ? f : dict[f];
        dict[s] = d;
        out[s] = (term[s] >= 0) ? term[s] : (d >= 0 ? term[d] : -1);
        for (int c = 0; c < ncls; ++c) {
            int t = next[(size_t)s * ncls + c];
            if (t < 0) next[(size_t)s * ncls + c] = next[(size_t)f * ncls + c];
            else { fail[t] = next[(size_t)f * ncls + c]; queue[qt++] = t; }
        }
    }
    free(fail); free(queue);
    ac->nstates = ns; ac->next = next; ac->out = out; ac->dict = dict; ac->term = term;
    return ac;
}

int fxo_ac_scan(const fxo_ac *ac, const char *text, size_t len, fxo_ac_cb cb, void *ctx)
{
    size_t s = 0;
    int ncls = ac->ncls;
    /* the empty pattern (root terminal) ends at offset 0 */
    for (size_t i = 0;; ++i) {
        if (ac->out[s] >= 0) {
            int t = (ac->term[s] >= 0) ? (int)s : ac->dict[s];
            while (t >= 0) {
                int r = cb(ac->term[t], text + i, ctx);
                if (r) return r;
                t = ac->dict[t];
            }
        }
        if (i >= len) break;
        s = (size_t)ac->next[s * ncls + ac->cls[(unsigned char)text[i]]];
    }
    return 0;
}

struct first_ctx { int id; const char *p; };
static int first_cb(int strnum, const char *textp, void *c)
{
    struct first_ctx *f = (struct first_ctx *)c;
    f->id = strnum; f->p = textp;
    return 1;
}

int fxo_ac_first(const fxo_ac *ac, const char *text, size_t len, size_t *end_out)
{
    struct first_ctx f = { -1, NULL };
    if (fxo_ac_scan(ac, text, len, first_cb, &f)) {
        if (end_out) *end_out = (size_t)(f.p - text);
        return f.id;
    }
    return -1;
}

size_t fxo_ac_nstates(const fxo_ac *ac) { return ac->nstates; }

void fxo_ac_free(fxo_ac *ac)
{
    if (!ac) return;
    free(ac->next); free(ac->out); free(ac->dict); free(ac->term); free(ac);
}
