/*
 * oracle/standin/standin.c -- TEST INFRASTRUCTURE ONLY.
 * The five symbols the reference needs from its missing util/ submodule
 * (SURVEY.md App. B), so that the reference's own .cpp files link unmodified.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "util/msutil.h"
#include "util/ssearch.h"
#include "../fxo_ac.h"

MEMREF *refsplit(char *text, char sep, int *pcount)
{
    /* the caller passes a NUL-terminated buffer (main.cpp:362-371) */
    size_t n = strlen(text);
    int count = 1;
    for (size_t i = 0; i < n; ++i) if (text[i] == sep) ++count;
    MEMREF *v = (MEMREF *)malloc(sizeof(MEMREF) * (size_t)count);
    int k = 0;
    char *start = text;
    for (size_t i = 0; i <= n; ++i) {
        if (i == n || text[i] == sep) {
            v[k].ptr = start; v[k].len = (size_t)(text + i - start); ++k;
            if (i < n) text[i] = '\0';
            start = text + i + 1;
        }
    }
    *pcount = count;
    return v;
}

void usage(char const *msg)
{
    fprintf(stderr, "Usage: fxtract %s\n", msg);
    exit(2);
}

SSEARCH *ssearch_create(MEMREF const *strv, int nstrs)
{
    if (nstrs <= 0) return NULL;
    const char **p = (const char **)malloc(sizeof(char *) * (size_t)nstrs);
    size_t *l = (size_t *)malloc(sizeof(size_t) * (size_t)nstrs);
    for (int i = 0; i < nstrs; ++i) { p[i] = strv[i].ptr; l[i] = strv[i].len; }
    fxo_ac *ac = fxo_ac_build(p, l, nstrs);
    free(p); free(l);
    return ac;
}

int ssearch_scan(SSEARCH const *s, MEMREF text, SSEARCH_CB cb, void *context)
{
    return fxo_ac_scan(s, text.ptr, text.len, (fxo_ac_cb)cb, context);
}

void ssearch_destroy(SSEARCH *s) { fxo_ac_free(s); }
