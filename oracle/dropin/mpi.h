/* Stub mpi.h for LINKING AND RUNNING the reference's MPI drivers (impl/pingpong.c, impl/himenoBMTxps.c) in one process.
 * TEST INFRASTRUCTURE ONLY (tests/test_link_dropin.py).  The functions live in mpistub.c: the process is rank 0 of
 * MPISTUB_SIZE ranks (default 2) and its partner is emulated as an ECHO: a receive returns what was last sent with the same
 * tag; the ping-pong counter (one MPI_INT on tag 0) comes back incremented, as the partner of impl/pingpong.c:226-232 would
 * send it; a receive nobody sent to returns 'y' bytes (the ack of :267-273). */
#ifndef DCB_DROPIN_STUB_MPI_H
#define DCB_DROPIN_STUB_MPI_H
typedef int MPI_Datatype;
typedef int MPI_Comm;
typedef int MPI_Op;
typedef int MPI_Request;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
#define MPI_CHAR 1
#define MPI_UNSIGNED_CHAR 2
#define MPI_INT 3
#define MPI_UNSIGNED 4
#define MPI_DOUBLE 5
#define MPI_FLOAT 6
#define MPI_COMM_WORLD 0
#define MPI_PROC_NULL (-2)
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_SUCCESS 0
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int code);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
double MPI_Wtime(void);
int MPI_Barrier(MPI_Comm comm);
int MPI_Send(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Status* st);
int MPI_Isend(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Waitall(int n, MPI_Request* reqs, MPI_Status* sts);
int MPI_Bcast(void* buf, int count, MPI_Datatype dt, int root, MPI_Comm comm);
int MPI_Gather(const void* sb, int sc, MPI_Datatype sd, void* rb, int rc, MPI_Datatype rd, int root, MPI_Comm comm);
int MPI_Allreduce(const void* sb, void* rb, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm);
int MPI_Cart_create(MPI_Comm comm, int ndims, const int* dims, const int* periods, int reorder, MPI_Comm* cart);
int MPI_Cart_get(MPI_Comm comm, int maxdims, int* dims, int* periods, int* coords);
int MPI_Cart_shift(MPI_Comm comm, int dir, int disp, int* src, int* dst);
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype* newtype);
int MPI_Type_commit(MPI_Datatype* dt);
#endif
