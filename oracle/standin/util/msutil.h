/*
 * oracle/standin/util/msutil.h -- TEST INFRASTRUCTURE ONLY.
 * Stand-in for the reference's missing `util/` submodule (ctSkennerton/util @
 * aa50dc2cd37e9fd07602b06dda22c5cbb177fd08; /root/reference/util is empty).
 * Declares only what the reference uses: MEMREF (main.cpp:81,389), refsplit
 * (main.cpp:371), usage (main.cpp:180,195,662,665,680).
 */
#ifndef FXO_STANDIN_MSUTIL_H
#define FXO_STANDIN_MSUTIL_H
#include <stddef.h>
typedef struct { char const *ptr; size_t len; } MEMREF;
/* split `text` in place on `sep` (each sep byte becomes NUL); malloc'ed array */
MEMREF *refsplit(char *text, char sep, int *pcount);
/* print "Usage: fxtract <msg>" to stderr and exit(2) (stream/code unpinned) */
void usage(char const *msg);
#endif
