/* Single-process MPI stand-in, see mpi.h next to this file.  TEST INFRASTRUCTURE ONLY. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "mpi.h"

#define NTAGS 64
static struct { void* data; size_t bytes; } g_box[NTAGS];
static size_t dt_size(MPI_Datatype dt) {
  switch (dt) {
    case MPI_CHAR: case MPI_UNSIGNED_CHAR: return 1;
    case MPI_INT: case MPI_UNSIGNED: case MPI_FLOAT: return 4;
    case MPI_DOUBLE: return 8;
    default: return 0;                       /* derived (vector) types: the halo echo is not emulated */
  }
}
static int stub_size(void) { const char* e = getenv("MPISTUB_SIZE"); return e ? atoi(e) : 2; }

int MPI_Init(int* argc, char*** argv) { (void)argc; (void)argv; return MPI_SUCCESS; }
int MPI_Finalize(void) { return MPI_SUCCESS; }
int MPI_Abort(MPI_Comm comm, int code) { (void)comm; exit(code); }
int MPI_Comm_rank(MPI_Comm comm, int* rank) { (void)comm; *rank = 0; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm comm, int* size) { (void)comm; *size = stub_size(); return MPI_SUCCESS; }
double MPI_Wtime(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
int MPI_Barrier(MPI_Comm comm) { (void)comm; return MPI_SUCCESS; }

int MPI_Send(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm) {
  (void)dest; (void)comm;
  const size_t bytes = (size_t)count * dt_size(dt);
  if (tag < 0 || tag >= NTAGS || dest == MPI_PROC_NULL) return MPI_SUCCESS;
  free(g_box[tag].data);
  g_box[tag].data = malloc(bytes ? bytes : 1);
  memcpy(g_box[tag].data, buf, bytes);
  g_box[tag].bytes = bytes;
  return MPI_SUCCESS;
}
int MPI_Recv(void* buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Status* st) {
  (void)comm; (void)st;
  const size_t bytes = (size_t)count * dt_size(dt);
  if (src == MPI_PROC_NULL || tag < 0 || tag >= NTAGS) return MPI_SUCCESS;
  if (!g_box[tag].data) { memset(buf, 'y', bytes); return MPI_SUCCESS; }      /* nobody sent: the partner's ack */
  memcpy(buf, g_box[tag].data, bytes < g_box[tag].bytes ? bytes : g_box[tag].bytes);
  if (tag == 0 && dt == MPI_INT && count == 1) *(int*)buf += 1;               /* the partner increments the hop counter */
  return MPI_SUCCESS;
}
int MPI_Isend(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request* req) {
  if (req) *req = 0;
  return dt_size(dt) ? MPI_Send(buf, count, dt, dest, tag, comm) : MPI_SUCCESS;
}
int MPI_Irecv(void* buf, int count, MPI_Datatype dt, int src, int tag, MPI_Comm comm, MPI_Request* req) {
  (void)buf; (void)count; (void)dt; (void)src; (void)tag; (void)comm;
  if (req) *req = 0;
  return MPI_SUCCESS;                        /* the halo partner is absent: buffers stay as they are */
}
int MPI_Waitall(int n, MPI_Request* reqs, MPI_Status* sts) { (void)n; (void)reqs; (void)sts; return MPI_SUCCESS; }
int MPI_Bcast(void* buf, int count, MPI_Datatype dt, int root, MPI_Comm comm) { (void)buf; (void)count; (void)dt; (void)root; (void)comm; return MPI_SUCCESS; }
int MPI_Gather(const void* sb, int sc, MPI_Datatype sd, void* rb, int rc, MPI_Datatype rd, int root, MPI_Comm comm) {
  (void)rc; (void)rd; (void)root; (void)comm;
  memcpy(rb, sb, (size_t)sc * dt_size(sd));
  return MPI_SUCCESS;
}
int MPI_Allreduce(const void* sb, void* rb, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm) {
  (void)op; (void)comm;
  memcpy(rb, sb, (size_t)count * dt_size(dt));
  return MPI_SUCCESS;
}
int MPI_Cart_create(MPI_Comm comm, int ndims, const int* dims, const int* periods, int reorder, MPI_Comm* cart) {
  (void)comm; (void)ndims; (void)dims; (void)periods; (void)reorder; *cart = 1; return MPI_SUCCESS;
}
int MPI_Cart_get(MPI_Comm comm, int maxdims, int* dims, int* periods, int* coords) {
  (void)comm; (void)dims; (void)periods;
  for (int i = 0; i < maxdims; i++) coords[i] = 0;
  return MPI_SUCCESS;
}
int MPI_Cart_shift(MPI_Comm comm, int dir, int disp, int* src, int* dst) {
  (void)comm; (void)dir; (void)disp; *src = MPI_PROC_NULL; *dst = MPI_PROC_NULL; return MPI_SUCCESS;
}
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype* newtype) {
  (void)count; (void)blocklen; (void)stride; (void)old; *newtype = 100; return MPI_SUCCESS;
}
int MPI_Type_commit(MPI_Datatype* dt) { (void)dt; return MPI_SUCCESS; }
