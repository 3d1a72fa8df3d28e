/*
 * lda_main.c -- the `lda` executable: a thin main over liboni_lda.so with the
 * argv of oni-lda-c (lda-estimate.c:403-468).  ml_ops.sh:80 runs it as
 *     ./lda est 2.5 20 settings.txt <ngpus> ../<day>/model.dat random ../<day>
 * without mpiexec.  Exit status is 0 on success (the reference always exits 0;
 * nothing upstream checks it) and the oni_status on failure, with the reason
 * on stderr.
 */
#include <stdio.h>

#include "../../include/oni_lda.h"

int main(int argc, char **argv)
{
    int rc = oni_lda_main(argc, argv);
    if (rc != ONI_OK) fprintf(stderr, "lda: error %d: %s\n", rc, oni_last_error());
    return rc;
}
