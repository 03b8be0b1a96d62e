/* ORACLE (test infrastructure) — single-rank MPI shim, just enough to compile and run the reference's own
 * src/psolve.c + lib/parms serially.  One process, communicator size 1: collectives copy
 * sendbuf to recvbuf, point-to-point is never reached (no neighbours).  Not a general MPI. */
#ifndef FB200_MPI_STUB_H
#define FB200_MPI_STUB_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef int MPI_Comm;
typedef int MPI_Fint;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
typedef void(MPI_User_function)(void*, void*, int*, MPI_Datatype*);

#define MPI_SUCCESS 0
#define MPI_COMM_WORLD 1
#define MPI_COMM_NULL 0
#define MPI_INT 4
#define MPI_DOUBLE 8
#define MPI_CHAR 1
#define MPI_BYTE 1
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3
#define MPI_IN_PLACE ((void*)1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_REQUEST_NULL 0

int MPI_Init(int*, char***);
int MPI_Initialized(int* flag);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm, int);
double MPI_Wtime(void);
MPI_Comm MPI_Comm_f2c(MPI_Fint);
int MPI_Comm_rank(MPI_Comm, int*);
int MPI_Comm_size(MPI_Comm, int*);
int MPI_Comm_dup(MPI_Comm, MPI_Comm*);
int MPI_Comm_free(MPI_Comm*);
int MPI_Allreduce(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm);
int MPI_Iallreduce(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm, MPI_Request*);
int MPI_Allgather(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, MPI_Comm);
int MPI_Allgatherv(const void*, int, MPI_Datatype, void*, const int*, const int*, MPI_Datatype, MPI_Comm);
int MPI_Gather(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, int, MPI_Comm);
int MPI_Gatherv(const void*, int, MPI_Datatype, void*, const int*, const int*, MPI_Datatype, int, MPI_Comm);
int MPI_Bcast(void*, int, MPI_Datatype, int, MPI_Comm);
int MPI_Isend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Irecv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Wait(MPI_Request*, MPI_Status*);
int MPI_Waitall(int, MPI_Request*, MPI_Status*);
int MPI_Type_contiguous(int, MPI_Datatype, MPI_Datatype*);
int MPI_Type_indexed(int, const int*, const int*, MPI_Datatype, MPI_Datatype*);
int MPI_Type_commit(MPI_Datatype*);
int MPI_Type_free(MPI_Datatype*);
int MPI_Op_create(MPI_User_function*, int, MPI_Op*);
#ifdef __cplusplus
}
#endif
#endif
