/* fixseed.c -- see fixseed.h.  TEST INFRASTRUCTURE ONLY. */
#include <stdlib.h>

long oni_fixed_time(long *out)
{
    const char *e = getenv("ONI_LDA_SEED");
    long v = e ? atol(e) : 4357L;
    if (out) *out = v;
    return v;
}
