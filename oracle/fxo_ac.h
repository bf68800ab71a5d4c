/*
 * oracle/fxo_ac.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C Aho-Corasick multi-string matcher.  It restates the *contract* of the
 * reference's missing third-party engine `ssearch` (module ctSkennerton/util,
 * pinned aa50dc2cd37e9fd07602b06dda22c5cbb177fd08, not vendored in
 * /root/reference) as fixed by the reference's own call sites:
 *   main.cpp:371-373  refsplit + ssearch_create(pattv, npatts)
 *   main.cpp:406,426  ssearch_scan(ssp, data, on_match2, &pair)
 *   main.cpp:273-289  on_match2 always returns non-zero => scan stops at the
 *                     first reported occurrence.
 * Semantics: exact, case-sensitive byte-substring search of every pattern in a
 * (ptr,len) text.  Report order (UNPINNED by the reference, our rule): smallest
 * end offset first; among patterns ending at the same offset, longest first.
 */
#ifndef FXO_AC_H
#define FXO_AC_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct fxo_ac fxo_ac;

/* callback: return non-zero to stop the scan; that value is returned. */
typedef int (*fxo_ac_cb)(int strnum, const char *textp, void *context);

fxo_ac *fxo_ac_build(const char *const *pats, const size_t *lens, int n);
/* calls cb(pattern id, pointer one past the occurrence's last byte, ctx) per hit */
int fxo_ac_scan(const fxo_ac *ac, const char *text, size_t len, fxo_ac_cb cb, void *ctx);
/* first hit under the report order above: returns pattern id or -1; *end_out = end offset */
int fxo_ac_first(const fxo_ac *ac, const char *text, size_t len, size_t *end_out);
size_t fxo_ac_nstates(const fxo_ac *ac);
void fxo_ac_free(fxo_ac *ac);

#ifdef __cplusplus
}
#endif
#endif
