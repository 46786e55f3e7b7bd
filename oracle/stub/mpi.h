/* Stub mpi.h: TEST INFRASTRUCTURE ONLY.
 * Lets /root/reference/impl/dataCompression.c compile without an MPI install.
 * The float32 hot path never calls MPI; only the L2 wrappers
 * (dataCompression.c:24-353, 800-1197) do, and they abort() here. */
#ifndef DC_ORACLE_STUB_MPI_H
#define DC_ORACLE_STUB_MPI_H
#include <stdlib.h>
typedef int MPI_Datatype;
typedef int MPI_Comm;
typedef struct { int dummy; } MPI_Status;
#define MPI_CHAR 1
#define MPI_UNSIGNED_CHAR 2
#define MPI_INT 3
#define MPI_UNSIGNED 4
#define MPI_DOUBLE 5
#define MPI_COMM_WORLD 0
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
static inline int MPI_Send(const void* b, int c, MPI_Datatype d, int dst, int tag, MPI_Comm cm) { (void)b;(void)c;(void)d;(void)dst;(void)tag;(void)cm; abort(); return 0; }
static inline int MPI_Recv(void* b, int c, MPI_Datatype d, int src, int tag, MPI_Comm cm, MPI_Status* s) { (void)b;(void)c;(void)d;(void)src;(void)tag;(void)cm;(void)s; abort(); return 0; }
static inline int MPI_Bcast(void* b, int c, MPI_Datatype d, int root, MPI_Comm cm) { (void)b;(void)c;(void)d;(void)root;(void)cm; abort(); return 0; }
static inline int MPI_Gather(const void* sb, int sc, MPI_Datatype sd, void* rb, int rc, MPI_Datatype rd, int root, MPI_Comm cm) { (void)sb;(void)sc;(void)sd;(void)rb;(void)rc;(void)rd;(void)root;(void)cm; abort(); return 0; }
static inline int MPI_Comm_rank(MPI_Comm cm, int* r) { (void)cm; *r = 0; return 0; }
static inline double MPI_Wtime(void) { return 0.0; }
#endif
