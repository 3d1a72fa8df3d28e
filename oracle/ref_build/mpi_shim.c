/*
 * mpi_shim.c -- forked-process backing for oracle/ref_build/mpi.h.
 *
 * TEST INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * MPI_Init() forks ONI_MPI_NP-1 children; every process then continues through
 * the reference's run_em() as one rank.  The reference parses model.dat and
 * seeds its RNG *before* MPI_Init (lda-estimate.c:414-441), so the children
 * inherit an identical corpus and RNG state, which is what a real mpiexec
 * launch gives when all ranks start within the same second.
 *
 *   Barrier    sense-reversing counter in shared memory
 *   Allreduce  every rank copies its operand into its own slot, barrier, every
 *              rank adds the slots in rank order 0..np-1, barrier
 *              (deterministic; only MPI_SUM over MPI_DOUBLE / MPI_INT is used)
 *   Send/Recv  one bounded FIFO per destination rank holding (source, tag,
 *              8-byte payload); the reference only ever sends one double per
 *              message (lda-estimate.c:287) and receives with ANY_SOURCE /
 *              ANY_TAG (lda-estimate.c:297)
 */
#define _GNU_SOURCE
#include "mpi.h"

#include <pthread.h>
#include <sched.h>
#include <stdatomic.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#define SLOT_BYTES ((size_t)64 << 20) /* per-rank allreduce operand slot */
#define QCAP 65536                    /* messages per mailbox             */

typedef struct {
    int source;
    int tag;
    unsigned char payload[8];
} shim_msg;

typedef struct {
    pthread_mutex_t mu;
    pthread_cond_t nonempty;
    pthread_cond_t nonfull;
    size_t head, tail; /* head = next to pop, tail = next to push */
    shim_msg ring[QCAP];
} shim_mailbox;

typedef struct {
    int np;
    atomic_int arrived;
    atomic_int sense;
} shim_header;

static shim_header *g_hdr;
static shim_mailbox *g_box;      /* np mailboxes                 */
static unsigned char *g_slots;   /* np slots of SLOT_BYTES       */
static int g_rank = 0, g_np = 1;
static int g_local_sense = 0;
static pid_t *g_children;

static size_t type_size(MPI_Datatype t)
{
    if (t == MPI_DOUBLE) return sizeof(double);
    if (t == MPI_INT) return sizeof(int);
    fprintf(stderr, "mpi_shim: unsupported datatype %d\n", t);
    abort();
}

int MPI_Init(int *argc, char ***argv)
{
    (void)argc; (void)argv;
    const char *e = getenv("ONI_MPI_NP");
    g_np = e ? atoi(e) : 1;
    if (g_np < 1) g_np = 1;

    size_t bytes = sizeof(shim_header) + (size_t)g_np * sizeof(shim_mailbox) +
                   (size_t)g_np * SLOT_BYTES + 4096;
    unsigned char *base = mmap(NULL, bytes, PROT_READ | PROT_WRITE,
                               MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (base == MAP_FAILED) { perror("mpi_shim mmap"); abort(); }
    g_hdr = (shim_header *)base;
    g_box = (shim_mailbox *)(base + ((sizeof(shim_header) + 63) & ~(size_t)63));
    g_slots = (unsigned char *)(g_box + g_np);
    g_hdr->np = g_np;
    atomic_init(&g_hdr->arrived, 0);
    atomic_init(&g_hdr->sense, 0);

    pthread_mutexattr_t ma;
    pthread_condattr_t ca;
    pthread_mutexattr_init(&ma);
    pthread_mutexattr_setpshared(&ma, PTHREAD_PROCESS_SHARED);
    pthread_condattr_init(&ca);
    pthread_condattr_setpshared(&ca, PTHREAD_PROCESS_SHARED);
    for (int r = 0; r < g_np; r++) {
        pthread_mutex_init(&g_box[r].mu, &ma);
        pthread_cond_init(&g_box[r].nonempty, &ca);
        pthread_cond_init(&g_box[r].nonfull, &ca);
        g_box[r].head = g_box[r].tail = 0;
    }

    fflush(NULL);
    g_children = calloc((size_t)g_np, sizeof(pid_t));
    for (int r = 1; r < g_np; r++) {
        pid_t p = fork();
        if (p < 0) { perror("mpi_shim fork"); abort(); }
        if (p == 0) { g_rank = r; break; }
        g_children[r] = p;
    }
    return MPI_SUCCESS;
}

int MPI_Finalize(void)
{
    fflush(NULL);
    if (g_rank != 0) _exit(0);
    for (int r = 1; r < g_np; r++) {
        int st;
        waitpid(g_children[r], &st, 0);
    }
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm comm, int *rank) { (void)comm; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm comm, int *size) { (void)comm; *size = g_np; return MPI_SUCCESS; }

int MPI_Barrier(MPI_Comm comm)
{
    (void)comm;
    if (g_np == 1) return MPI_SUCCESS;
    g_local_sense = !g_local_sense;
    if (atomic_fetch_add(&g_hdr->arrived, 1) == g_np - 1) {
        atomic_store(&g_hdr->arrived, 0);
        atomic_store(&g_hdr->sense, g_local_sense);
    } else {
        unsigned spins = 0;
        while (atomic_load(&g_hdr->sense) != g_local_sense) {
            if (++spins > 2000) sched_yield();
        }
    }
    return MPI_SUCCESS;
}

int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count,
                  MPI_Datatype type, MPI_Op op, MPI_Comm comm)
{
    if (op != MPI_SUM) { fprintf(stderr, "mpi_shim: only MPI_SUM\n"); abort(); }
    size_t bytes = (size_t)count * type_size(type);
    if (bytes > SLOT_BYTES) { fprintf(stderr, "mpi_shim: operand too large\n"); abort(); }
    if (g_np == 1) {
        if (recvbuf != sendbuf) memmove(recvbuf, sendbuf, bytes);
        return MPI_SUCCESS;
    }
    memcpy(g_slots + (size_t)g_rank * SLOT_BYTES, sendbuf, bytes);
    MPI_Barrier(comm);
    if (type == MPI_DOUBLE) {
        double *out = recvbuf;
        const double *s0 = (const double *)g_slots;
        for (int i = 0; i < count; i++) out[i] = s0[i];
        for (int r = 1; r < g_np; r++) {
            const double *s = (const double *)(g_slots + (size_t)r * SLOT_BYTES);
            for (int i = 0; i < count; i++) out[i] += s[i];
        }
    } else {
        int *out = recvbuf;
        const int *s0 = (const int *)g_slots;
        for (int i = 0; i < count; i++) out[i] = s0[i];
        for (int r = 1; r < g_np; r++) {
            const int *s = (const int *)(g_slots + (size_t)r * SLOT_BYTES);
            for (int i = 0; i < count; i++) out[i] += s[i];
        }
    }
    MPI_Barrier(comm);
    return MPI_SUCCESS;
}

int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag,
             MPI_Comm comm)
{
    (void)comm;
    size_t bytes = (size_t)count * type_size(type);
    if (bytes > 8) { fprintf(stderr, "mpi_shim: message too large\n"); abort(); }
    shim_mailbox *b = &g_box[dest];
    pthread_mutex_lock(&b->mu);
    while (b->tail - b->head == QCAP) pthread_cond_wait(&b->nonfull, &b->mu);
    shim_msg *m = &b->ring[b->tail % QCAP];
    m->source = g_rank;
    m->tag = tag;
    memcpy(m->payload, buf, bytes);
    b->tail++;
    pthread_cond_signal(&b->nonempty);
    pthread_mutex_unlock(&b->mu);
    return MPI_SUCCESS;
}

int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int tag,
             MPI_Comm comm, MPI_Status *status)
{
    (void)comm;
    if (source != MPI_ANY_SOURCE || tag != MPI_ANY_TAG) {
        fprintf(stderr, "mpi_shim: only ANY_SOURCE/ANY_TAG receives\n");
        abort();
    }
    size_t bytes = (size_t)count * type_size(type);
    shim_mailbox *b = &g_box[g_rank];
    pthread_mutex_lock(&b->mu);
    while (b->tail == b->head) pthread_cond_wait(&b->nonempty, &b->mu);
    shim_msg *m = &b->ring[b->head % QCAP];
    memcpy(buf, m->payload, bytes);
    if (status) {
        status->MPI_SOURCE = m->source;
        status->MPI_TAG = m->tag;
        status->MPI_ERROR = MPI_SUCCESS;
    }
    b->head++;
    pthread_cond_signal(&b->nonfull);
    pthread_mutex_unlock(&b->mu);
    return MPI_SUCCESS;
}
