/* ORACLE (test infrastructure) — one-process MPI shim, see mpi.h.  Datatype handles are their byte sizes. */
#include "mpi.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static void cp(const void* s, void* r, size_t bytes) { if (s != MPI_IN_PLACE && s != r && bytes) memmove(r, s, bytes); }
static void p2p(const char* what) { fprintf(stderr, "mpi_stub: %s reached on a single-rank communicator\n", what); abort(); }

int MPI_Init(int* a, char*** b) { (void)a; (void)b; return 0; }
int MPI_Initialized(int* flag) { *flag = 1; return 0; }
int MPI_Finalize(void) { return 0; }
int MPI_Abort(MPI_Comm c, int code) { (void)c; fprintf(stderr, "mpi_stub: MPI_Abort(%d)\n", code); abort(); return 0; }
double MPI_Wtime(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec; }
MPI_Comm MPI_Comm_f2c(MPI_Fint f) { (void)f; return MPI_COMM_WORLD; }
int MPI_Comm_rank(MPI_Comm c, int* r) { (void)c; *r = 0; return 0; }
/* pARMS switches to a serial matrix type without RAS support when the communicator has one rank
 * (parms_map.c:675, parms_mat_vcsr.c:96-107: no `extend` op), and FESOM itself never runs on fewer than two
 * partitions (fvom_init.F90:1522-1525).  The shim therefore reports TWO ranks: rank 0 is this process and owns
 * every row, rank 1 is a phantom that owns nothing, contributes zeros to collectives and never sends.
 * That keeps the reference on its distributed code path (dvcsr + RAS + ILU(k)) with a single subdomain. */
#define NPES 2
int MPI_Comm_size(MPI_Comm c, int* s) { (void)c; *s = NPES; return 0; }
int MPI_Comm_dup(MPI_Comm c, MPI_Comm* n) { *n = c; return 0; }
int MPI_Comm_free(MPI_Comm* c) { *c = MPI_COMM_NULL; return 0; }
int MPI_Allreduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op o, MPI_Comm c) { (void)o; (void)c; cp(s, r, (size_t)n * (size_t)t); return 0; }
int MPI_Iallreduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op o, MPI_Comm c, MPI_Request* q) { (void)o; (void)c; cp(s, r, (size_t)n * (size_t)t); *q = MPI_REQUEST_NULL; return 0; }
int MPI_Allgather(const void* s, int n, MPI_Datatype t, void* r, int rn, MPI_Datatype rt, MPI_Comm c) {
    (void)c; cp(s, r, (size_t)n * (size_t)t);
    memset((char*)r + (size_t)rn * (size_t)rt, 0, (size_t)(NPES - 1) * (size_t)rn * (size_t)rt);      /* phantom ranks */
    return 0;
}
int MPI_Allgatherv(const void* s, int n, MPI_Datatype t, void* r, const int* rc, const int* d, MPI_Datatype rt, MPI_Comm c) { (void)rc; (void)c; cp(s, (char*)r + (size_t)d[0] * (size_t)rt, (size_t)n * (size_t)t); return 0; }
int MPI_Gather(const void* s, int n, MPI_Datatype t, void* r, int rn, MPI_Datatype rt, int root, MPI_Comm c) {
    (void)c; if (root != 0) return 0;
    cp(s, r, (size_t)n * (size_t)t);
    memset((char*)r + (size_t)rn * (size_t)rt, 0, (size_t)(NPES - 1) * (size_t)rn * (size_t)rt);
    return 0;
}
int MPI_Gatherv(const void* s, int n, MPI_Datatype t, void* r, const int* rc, const int* d, MPI_Datatype rt, int root, MPI_Comm c) { (void)rc; (void)c; if (root != 0) return 0; cp(s, (char*)r + (size_t)d[0] * (size_t)rt, (size_t)n * (size_t)t); return 0; }
int MPI_Bcast(void* b, int n, MPI_Datatype t, int root, MPI_Comm c) { (void)b; (void)n; (void)t; (void)root; (void)c; return 0; }
int MPI_Isend(const void* b, int n, MPI_Datatype t, int d, int tag, MPI_Comm c, MPI_Request* q) { (void)b; (void)n; (void)t; (void)d; (void)tag; (void)c; (void)q; p2p("MPI_Isend"); return 0; }
int MPI_Irecv(void* b, int n, MPI_Datatype t, int d, int tag, MPI_Comm c, MPI_Request* q) { (void)b; (void)n; (void)t; (void)d; (void)tag; (void)c; (void)q; p2p("MPI_Irecv"); return 0; }
int MPI_Wait(MPI_Request* q, MPI_Status* s) { (void)q; (void)s; return 0; }
int MPI_Waitall(int n, MPI_Request* q, MPI_Status* s) { (void)n; (void)q; (void)s; return 0; }
int MPI_Type_contiguous(int n, MPI_Datatype t, MPI_Datatype* nt) { *nt = n * t; return 0; }
int MPI_Type_indexed(int n, const int* bl, const int* d, MPI_Datatype t, MPI_Datatype* nt) { (void)n; (void)bl; (void)d; *nt = t; return 0; }
int MPI_Type_commit(MPI_Datatype* t) { (void)t; return 0; }
int MPI_Type_free(MPI_Datatype* t) { (void)t; return 0; }
int MPI_Op_create(MPI_User_function* f, int commute, MPI_Op* op) { (void)f; (void)commute; *op = MPI_SUM; return 0; }

/* symbols only the ARMS/Schur preconditioners reference (BLAS/LAPACK, timer); never reached on the
 * BiCGStab + RAS + ILU(k) path FESOM selects (src/oce_ale.F90:2330) */
#define DEAD(name) void name(void) { fprintf(stderr, "mpi_stub: " #name " is not available in the serial reference build\n"); abort(); }
DEAD(dgemm_) DEAD(dgesvd_) DEAD(dgetrf_) DEAD(dgetri_) DEAD(dscal_)
double sys_timer(void) { return MPI_Wtime(); }
