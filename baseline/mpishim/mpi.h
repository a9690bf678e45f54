/* mpi.h -- minimal MPI subset for running the reference's MPI programs on a box without MPI.
 *
 * The image has no MPI implementation (no mpicxx/mpirun/libmpi).  The reference's stage 2/3/4 sources use
 * 20 MPI symbols; this shim provides exactly those over POSIX shared memory, one
 * process per rank on one node.  Rank and size come from the launcher's environment (torchrun's
 * RANK / WORLD_SIZE, or PB_MPI_RANK / PB_MPI_SIZE), so the same binary runs under
 * `python -m torch.distributed.run` or under baseline/run_ref.py.
 *
 * Semantics kept honest: MPI_Sendrecv / MPI_Isend+MPI_Irecv+MPI_Waitall really move the bytes between
 * processes and block until they arrived; MPI_Allreduce / MPI_Reduce / MPI_Barrier really synchronise all
 * ranks.  MPI_Init additionally creates the CUDA context (cudaFree(0)) when the program links the CUDA
 * runtime, so the reference's own "Solver time" does not include context creation -- a favour to the
 * reference, since a production MPI launcher would also have the device initialised by then.
 */
#ifndef PB_MPISHIM_MPI_H
#define PB_MPISHIM_MPI_H

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct MPI_Status {
  int MPI_SOURCE, MPI_TAG, MPI_ERROR;
} MPI_Status;
typedef struct pb_mpi_request* MPI_Request;

#define MPI_COMM_WORLD 0
#define MPI_DOUBLE 1
#define MPI_INT 2
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3
#define MPI_PROC_NULL (-2)
#define MPI_SUCCESS 0
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_REQUEST_NULL ((MPI_Request)0)

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
double MPI_Wtime(void);
int MPI_Barrier(MPI_Comm comm);
int MPI_Sendrecv(const void* sendbuf, int sendcount, MPI_Datatype sendtype, int dest, int sendtag, void* recvbuf,
                 int recvcount, MPI_Datatype recvtype, int source, int recvtag, MPI_Comm comm, MPI_Status* status);
int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status* status);
int MPI_Isend(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status* statuses);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm);
int MPI_Abort(MPI_Comm comm, int code);

#ifdef __cplusplus
}
#endif
#endif
