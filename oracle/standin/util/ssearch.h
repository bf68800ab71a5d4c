/*
 * oracle/standin/util/ssearch.h -- TEST INFRASTRUCTURE ONLY.
 * Stand-in for the reference's missing multi-string engine (call sites
 * main.cpp:373,406,426,450).  Backed by oracle/fxo_ac.c (Aho-Corasick).
 */
#ifndef FXO_STANDIN_SSEARCH_H
#define FXO_STANDIN_SSEARCH_H
#include "msutil.h"
typedef struct fxo_ac SSEARCH;
typedef int (*SSEARCH_CB)(int strnum, char const *textp, void *context);
SSEARCH *ssearch_create(MEMREF const *strv, int nstrs);
int ssearch_scan(SSEARCH const *, MEMREF text, SSEARCH_CB cb, void *context);
void ssearch_destroy(SSEARCH *);
#endif
