/*
 * fixseed.h -- force-included (-include) when building the reference so that
 * its clock-derived RNG seed (lda-estimate.c:414-416: time(&t1); seedMT(t1);)
 * becomes reproducible WITHOUT editing the reference sources.
 * TEST INFRASTRUCTURE ONLY.
 *
 * time(p) is redirected to oni_fixed_time(p), which returns the value of the
 * environment variable ONI_LDA_SEED (default 4357) and stores it through p.
 */
#ifndef ONI_REF_FIXSEED_H
#define ONI_REF_FIXSEED_H
#define time(p) oni_fixed_time(p)
#endif
