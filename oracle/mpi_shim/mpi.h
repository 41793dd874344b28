/* oracle/mpi_shim/mpi.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A minimal single-node stand-in for the 11 MPI-1 entry points that the
 * reference's collective CG program uses (parallel_cg.c:76-82,109-117,278-291,
 * 294,313,328-329) and the two more its point-to-point twin needs (MPI_Send,
 * MPI_Recv), so that the UNMODIFIED reference source can be compiled and
 * run in an image without any MPI implementation (no mpicc / mpi.h / libmpi
 * here or on the GPU box).  Ranks are forked processes, messages go through one
 * anonymous shared mapping, MPI_Allreduce sums in rank order 0..P-1.
 * The number of ranks comes from the environment variable CGSHIM_NP (default 1).
 */
#ifndef CG_MPI_SHIM_H
#define CG_MPI_SHIM_H

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS    0
#define MPI_FLOAT      4
#define MPI_DOUBLE     8
#define MPI_SUM        1

int    MPI_Init(int *argc, char ***argv);
int    MPI_Finalize(void);
int    MPI_Abort(MPI_Comm comm, int code);
int    MPI_Comm_size(MPI_Comm comm, int *size);
int    MPI_Comm_rank(MPI_Comm comm, int *rank);
int    MPI_Barrier(MPI_Comm comm);
double MPI_Wtime(void);
int    MPI_Bcast(void *buf, int count, MPI_Datatype dt, int root, MPI_Comm comm);
int    MPI_Scatter(const void *sendbuf, int sendcount, MPI_Datatype sdt,
                   void *recvbuf, int recvcount, MPI_Datatype rdt, int root, MPI_Comm comm);
int    MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype dt,
                     MPI_Op op, MPI_Comm comm);
int    MPI_Allgather(const void *sendbuf, int sendcount, MPI_Datatype sdt,
                     void *recvbuf, int recvcount, MPI_Datatype rdt, MPI_Comm comm);

/* point-to-point_cg.c replaces the four collectives above by blocking sends and receives through rank 0
 * (:183-256, :339-394) */
int    MPI_Send(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int    MPI_Recv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status *status);

/* Test driver hook (not MPI): run `fn(local_A, local_b, local_x, m, rank, np)`
 * -- the reference's SPMD conjugrad signature, parallel_cg.c:46-48 -- on np
 * forked ranks over in-memory inputs; rank 0's x (n elements of elem_size
 * bytes) is copied back into x.  Returns 0 on success. */
int cgshim_run_conjugrad(void (*fn)(void *, void *, void *, int, int, int),
                         const void *A, const void *b, void *x,
                         int n, int np, int elem_size);
#endif
