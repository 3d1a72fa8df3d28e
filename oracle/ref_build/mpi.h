/*
 * mpi.h -- minimal stand-in for the MPI subset that oni-lda-c calls.
 *
 * TEST INFRASTRUCTURE ONLY.  No MPI implementation exists in this image
 * (mpicc/mpirun/libmpi absent), so the UNMODIFIED reference sources under
 * /root/reference/oni-lda-c are compiled against this header and linked with
 * mpi_shim.c, which runs the ranks as forked processes that exchange data
 * through a shared anonymous mapping.  Only the calls the reference makes are
 * declared (call sites: lda-estimate.c:183-184,211,214,219,235,245,251,269,
 * 287,297,309,318,441,443).
 *
 * Rank count comes from the environment variable ONI_MPI_NP (default 1);
 * the reference takes its stride from argv[5], so pass the same number there.
 */
#ifndef ONI_REF_MPI_STUB_H
#define ONI_REF_MPI_STUB_H

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;

typedef struct {
    int MPI_SOURCE;
    int MPI_TAG;
    int MPI_ERROR;
} MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_DOUBLE 1
#define MPI_INT 2
#define MPI_SUM 1
#define MPI_ANY_SOURCE (-1)
#define MPI_ANY_TAG (-1)

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Barrier(MPI_Comm comm);
int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count,
                  MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag,
             MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int tag,
             MPI_Comm comm, MPI_Status *status);

#endif
