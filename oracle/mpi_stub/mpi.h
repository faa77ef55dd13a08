/* oracle/mpi_stub/mpi.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Single-process stand-in for <mpi.h> (there is no MPI in this image) so that the reference's vendored HPDDM headers
 * (/root/reference/lib/hpddm/include) compile and its Krylov solvers run with ONE rank: every collective is the identity
 * on one rank, point-to-point calls are never reached with one subdomain.  Written from the MPI standard's signatures;
 * a datatype handle is simply its size in bytes.
 */
#ifndef ORACLE_MPI_STUB_H
#define ORACLE_MPI_STUB_H
#include <string.h>
#include <time.h>

#define MPI_VERSION 3
#define MPI_SUBVERSION 1
#define MPI_SUCCESS 0

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Group;
typedef int MPI_Request;
typedef int MPI_Fint;
typedef struct MPI_Status { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
typedef void(MPI_User_function)(void *, void *, int *, MPI_Datatype *);

#define MPI_COMM_NULL 0
#define MPI_COMM_SELF 1
#define MPI_COMM_WORLD 2
#define MPI_GROUP_NULL 0
#define MPI_REQUEST_NULL 0
#define MPI_DATATYPE_NULL 0
#define MPI_IN_PLACE ((void *)1)
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)
#define MPI_ANY_SOURCE (-1)
#define MPI_ANY_TAG (-1)
#define MPI_UNDEFINED (-32766)
#define MPI_IDENT 0
#define MPI_CONGRUENT 1
#define MPI_SIMILAR 2
#define MPI_UNEQUAL 3

#define MPI_CHAR 1
#define MPI_BYTE 1
#define MPI_UNSIGNED_CHAR 1
#define MPI_UNSIGNED_SHORT 2
#define MPI_INT 4
#define MPI_UNSIGNED 4
#define MPI_FLOAT 4
#define MPI_DOUBLE 8
#define MPI_LONG_LONG 8
#define MPI_UNSIGNED_LONG_LONG 8
#define MPI_C_COMPLEX 8
#define MPI_C_DOUBLE_COMPLEX 16

#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3
#define MPI_BOR 4

#ifdef __cplusplus
#define MPISTUB_INLINE inline
#else
#define MPISTUB_INLINE static inline
#endif

MPISTUB_INLINE int mpistub_copy(const void *s, void *r, int count, MPI_Datatype t) {
    if (s != MPI_IN_PLACE && s != r && count > 0) memcpy(r, s, (size_t)count * (size_t)t);
    return MPI_SUCCESS;
}
MPISTUB_INLINE int MPI_Comm_size(MPI_Comm c, int *n) { (void)c; *n = 1; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Comm_rank(MPI_Comm c, int *r) { (void)c; *r = 0; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Comm_dup(MPI_Comm c, MPI_Comm *o) { *o = c; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Comm_free(MPI_Comm *c) { *c = MPI_COMM_NULL; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Comm_group(MPI_Comm c, MPI_Group *g) { (void)c; *g = 1; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Comm_create(MPI_Comm c, MPI_Group g, MPI_Comm *o) { (void)g; *o = c; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Comm_compare(MPI_Comm a, MPI_Comm b, int *r) { *r = a == b ? MPI_IDENT : MPI_CONGRUENT; return MPI_SUCCESS; }
MPISTUB_INLINE MPI_Fint MPI_Comm_c2f(MPI_Comm c) { return c; }
MPISTUB_INLINE int MPI_Group_incl(MPI_Group g, int n, const int *r, MPI_Group *o) { (void)n; (void)r; *o = g; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Group_excl(MPI_Group g, int n, const int *r, MPI_Group *o) { (void)n; (void)r; *o = g; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Group_range_incl(MPI_Group g, int n, int r[][3], MPI_Group *o) { (void)n; (void)r; *o = g; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Group_free(MPI_Group *g) { *g = MPI_GROUP_NULL; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Op_create(MPI_User_function *f, int commute, MPI_Op *op) { (void)f; (void)commute; *op = 99; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Op_free(MPI_Op *op) { *op = 0; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Finalized(int *f) { *f = 0; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Finalize(void) { return MPI_SUCCESS; }
MPISTUB_INLINE double MPI_Wtime(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }
MPISTUB_INLINE int MPI_Get_count(const MPI_Status *s, MPI_Datatype t, int *c) { (void)s; (void)t; *c = 0; return MPI_SUCCESS; }

/* collectives on one rank */
MPISTUB_INLINE int MPI_Allreduce(const void *s, void *r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c) { (void)op; (void)c; return mpistub_copy(s, r, n, t); }
MPISTUB_INLINE int MPI_Reduce(const void *s, void *r, int n, MPI_Datatype t, MPI_Op op, int root, MPI_Comm c) { (void)op; (void)root; (void)c; return mpistub_copy(s, r, n, t); }
MPISTUB_INLINE int MPI_Exscan(const void *s, void *r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c) { (void)s; (void)r; (void)n; (void)t; (void)op; (void)c; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Bcast(void *b, int n, MPI_Datatype t, int root, MPI_Comm c) { (void)b; (void)n; (void)t; (void)root; (void)c; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Gather(const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, int root, MPI_Comm c) { (void)rn; (void)rt; (void)root; (void)c; return mpistub_copy(s, r, sn, st); }
MPISTUB_INLINE int MPI_Gatherv(const void *s, int sn, MPI_Datatype st, void *r, const int *rn, const int *d, MPI_Datatype rt, int root, MPI_Comm c) { (void)rn; (void)root; (void)c; return mpistub_copy(s, (char *)r + (d ? (size_t)d[0] * rt : 0), sn, st); }
MPISTUB_INLINE int MPI_Allgather(const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, MPI_Comm c) { (void)rn; (void)rt; (void)c; return mpistub_copy(s, r, sn, st); }
MPISTUB_INLINE int MPI_Allgatherv(const void *s, int sn, MPI_Datatype st, void *r, const int *rn, const int *d, MPI_Datatype rt, MPI_Comm c) { (void)rn; (void)c; return mpistub_copy(s, (char *)r + (d ? (size_t)d[0] * rt : 0), sn, st); }
MPISTUB_INLINE int MPI_Alltoallv(const void *s, const int *sn, const int *sd, MPI_Datatype st, void *r, const int *rn, const int *rd, MPI_Datatype rt, MPI_Comm c) { (void)rn; (void)c; return mpistub_copy((const char *)s + (sd ? (size_t)sd[0] * st : 0), (char *)r + (rd ? (size_t)rd[0] * rt : 0), sn[0], st); }
MPISTUB_INLINE int MPI_Barrier(MPI_Comm c) { (void)c; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Scatter(const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, int root, MPI_Comm c) { (void)rn; (void)rt; (void)root; (void)c; return (r == MPI_IN_PLACE) ? MPI_SUCCESS : mpistub_copy(s, r, sn, st); }
MPISTUB_INLINE int MPI_Scatterv(const void *s, const int *sn, const int *d, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, int root, MPI_Comm c) { (void)sn; (void)root; (void)c; return (r == MPI_IN_PLACE) ? MPI_SUCCESS : mpistub_copy((const char *)s + (d ? (size_t)d[0] * st : 0), r, rn, rt); }
MPISTUB_INLINE int MPI_Igather(const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, int root, MPI_Comm c, MPI_Request *q) { *q = MPI_REQUEST_NULL; return MPI_Gather(s, sn, st, r, rn, rt, root, c); }
MPISTUB_INLINE int MPI_Igatherv(const void *s, int sn, MPI_Datatype st, void *r, const int *rn, const int *d, MPI_Datatype rt, int root, MPI_Comm c, MPI_Request *q) { *q = MPI_REQUEST_NULL; return MPI_Gatherv(s, sn, st, r, rn, d, rt, root, c); }
MPISTUB_INLINE int MPI_Iscatter(const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, int root, MPI_Comm c, MPI_Request *q) { *q = MPI_REQUEST_NULL; return MPI_Scatter(s, sn, st, r, rn, rt, root, c); }
MPISTUB_INLINE int MPI_Iscatterv(const void *s, const int *sn, const int *d, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, int root, MPI_Comm c, MPI_Request *q) { *q = MPI_REQUEST_NULL; return MPI_Scatterv(s, sn, d, st, r, rn, rt, root, c); }

/* point-to-point: unreachable with one rank and no neighbours */
MPISTUB_INLINE int MPI_Send(const void *b, int n, MPI_Datatype t, int dst, int tag, MPI_Comm c) { (void)b; (void)n; (void)t; (void)dst; (void)tag; (void)c; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Isend(const void *b, int n, MPI_Datatype t, int dst, int tag, MPI_Comm c, MPI_Request *q) { (void)b; (void)n; (void)t; (void)dst; (void)tag; (void)c; *q = MPI_REQUEST_NULL; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Irecv(void *b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request *q) { (void)b; (void)n; (void)t; (void)src; (void)tag; (void)c; *q = MPI_REQUEST_NULL; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Wait(MPI_Request *q, MPI_Status *s) { (void)s; *q = MPI_REQUEST_NULL; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Waitall(int n, MPI_Request *q, MPI_Status *s) { (void)n; (void)q; (void)s; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Waitany(int n, MPI_Request *q, int *idx, MPI_Status *s) { (void)n; (void)q; (void)s; *idx = MPI_UNDEFINED; return MPI_SUCCESS; }
MPISTUB_INLINE int MPI_Testall(int n, MPI_Request *q, int *flag, MPI_Status *s) { (void)n; (void)q; (void)s; *flag = 1; return MPI_SUCCESS; }
#endif
