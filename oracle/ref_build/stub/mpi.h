/* Single-rank MPI stub (ORACLE BUILD ONLY — test infrastructure, not product code).
 *
 * Lets the reference's own CPU sources (/root/reference, unmodified, compiled where they lie)
 * build and run as ONE rank on a box without any MPI implementation: every communicator has
 * size 1, every collective is a memcpy, point-to-point messages to self go through a mailbox.
 * Only what the reference's src/core + include/El/core reference is provided.
 */
#ifndef ELB200_ORACLE_MPI_STUB_H
#define ELB200_ORACLE_MPI_STUB_H

#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#ifdef __cplusplus
#include <deque>
#include <vector>
extern "C++" {
#endif

typedef int MPI_Comm;
typedef int MPI_Group;
typedef int MPI_Op;
typedef int MPI_Errhandler;
typedef long MPI_Datatype; /* == extent in bytes */
typedef ptrdiff_t MPI_Aint;
typedef int MPI_Request;
typedef struct MPI_Status { int MPI_SOURCE, MPI_TAG, MPI_ERROR, count_bytes; } MPI_Status;
typedef void(MPI_User_function)(void*, void*, int*, MPI_Datatype*);

#define MPI_SUCCESS 0
#define MPI_COMM_NULL 0
#define MPI_COMM_WORLD 1
#define MPI_COMM_SELF 2
#define MPI_GROUP_NULL 0
#define MPI_GROUP_EMPTY 1
#define MPI_DATATYPE_NULL 0L
#define MPI_IN_PLACE ((void*)1)
#define MPI_UNDEFINED (-32766)
#define MPI_ANY_TAG (-1)
#define MPI_ANY_SOURCE (-2)
#define MPI_IDENT 0
#define MPI_CONGRUENT 1
#define MPI_SIMILAR 2
#define MPI_UNEQUAL 3
#define MPI_MAX_ERROR_STRING 256
#define MPI_ERRORS_RETURN 1
#define MPI_ERRORS_ARE_FATAL 2
#define MPI_THREAD_SINGLE 0
#define MPI_THREAD_FUNNELED 1
#define MPI_THREAD_SERIALIZED 2
#define MPI_THREAD_MULTIPLE 3
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_REQUEST_NULL 0

#define MPI_UNSIGNED_CHAR 1L
#define MPI_CHAR 1L
#define MPI_BYTE 1L
#define MPI_SHORT 2L
#define MPI_INT 4L
#define MPI_UNSIGNED 4L
#define MPI_FLOAT 4L
#define MPI_LONG 8L
#define MPI_UNSIGNED_LONG 8L
#define MPI_LONG_LONG_INT 8L
#define MPI_LONG_LONG 8L
#define MPI_UNSIGNED_LONG_LONG 8L
#define MPI_DOUBLE 8L
#define MPI_LONG_DOUBLE 16L
#define MPI_C_FLOAT_COMPLEX 8L
#define MPI_C_DOUBLE_COMPLEX 16L
#define MPI_FLOAT_INT 8L
#define MPI_DOUBLE_INT 16L
#define MPI_LONG_INT 16L

#define MPI_SUM 1
#define MPI_PROD 2
#define MPI_MIN 3
#define MPI_MAX 4
#define MPI_MINLOC 5
#define MPI_MAXLOC 6
#define MPI_LAND 7
#define MPI_LOR 8
#define MPI_LXOR 9
#define MPI_BAND 10
#define MPI_BOR 11
#define MPI_BXOR 12

#ifdef __cplusplus
namespace elb_mpi_stub {
struct Msg { int tag; std::vector<char> data; };
inline std::deque<Msg>& mailbox() { static std::deque<Msg> q; return q; }
inline int& counter() { static int c = 16; return c; }
inline int& inited() { static int i = 0; return i; }
inline int& finalized() { static int f = 0; return f; }
inline void cp(const void* s, void* d, long count, MPI_Datatype t) {
    if (s != MPI_IN_PLACE && s != d && count > 0 && t > 0) memcpy(d, s, (size_t)count * (size_t)t);
}
}  // namespace elb_mpi_stub
#define ELB_STUB_INLINE inline

ELB_STUB_INLINE int MPI_Init(int*, char***) { elb_mpi_stub::inited() = 1; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Init_thread(int*, char***, int required, int* provided) { elb_mpi_stub::inited() = 1; if (provided) *provided = required; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Query_thread(int* provided) { *provided = MPI_THREAD_MULTIPLE; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Initialized(int* flag) { *flag = elb_mpi_stub::inited(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Finalized(int* flag) { *flag = elb_mpi_stub::finalized(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Finalize(void) { elb_mpi_stub::finalized() = 1; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Abort(MPI_Comm, int code) { exit(code ? code : 1); return MPI_SUCCESS; }
ELB_STUB_INLINE double MPI_Wtime(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }
ELB_STUB_INLINE int MPI_Error_string(int, char* s, int* len) { strcpy(s, "mpi stub error"); *len = (int)strlen(s); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Comm_set_errhandler(MPI_Comm, MPI_Errhandler) { return MPI_SUCCESS; }

/* communicators and groups: every one has exactly one member, rank 0 */
ELB_STUB_INLINE int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Comm_dup(MPI_Comm, MPI_Comm* n) { *n = ++elb_mpi_stub::counter(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Comm_split(MPI_Comm, int color, int, MPI_Comm* n) { *n = (color == MPI_UNDEFINED) ? MPI_COMM_NULL : ++elb_mpi_stub::counter(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Comm_create(MPI_Comm, MPI_Group grp, MPI_Comm* n) { *n = (grp == MPI_GROUP_EMPTY) ? MPI_COMM_NULL : ++elb_mpi_stub::counter(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Comm_free(MPI_Comm* c) { *c = MPI_COMM_NULL; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Comm_group(MPI_Comm, MPI_Group* g) { *g = ++elb_mpi_stub::counter(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Comm_compare(MPI_Comm a, MPI_Comm b, int* r) { *r = (a == b) ? MPI_IDENT : MPI_CONGRUENT; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Cart_create(MPI_Comm, int, const int*, const int*, int, MPI_Comm* n) { *n = ++elb_mpi_stub::counter(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Cart_sub(MPI_Comm, const int*, MPI_Comm* n) { *n = ++elb_mpi_stub::counter(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_size(MPI_Group g, int* s) { *s = (g == MPI_GROUP_EMPTY) ? 0 : 1; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_rank(MPI_Group g, int* r) { *r = (g == MPI_GROUP_EMPTY) ? MPI_UNDEFINED : 0; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_free(MPI_Group* g) { *g = MPI_GROUP_NULL; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_dup(MPI_Group g, MPI_Group* n) { *n = g; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_incl(MPI_Group, int n, const int*, MPI_Group* o) { *o = n > 0 ? ++elb_mpi_stub::counter() : MPI_GROUP_EMPTY; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_excl(MPI_Group g, int n, const int*, MPI_Group* o) { *o = n > 0 ? MPI_GROUP_EMPTY : g; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_union(MPI_Group a, MPI_Group b, MPI_Group* o) { *o = (a == MPI_GROUP_EMPTY) ? b : a; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_difference(MPI_Group a, MPI_Group b, MPI_Group* o) { *o = (b == MPI_GROUP_EMPTY) ? a : MPI_GROUP_EMPTY; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_compare(MPI_Group a, MPI_Group b, int* r) { *r = ((a == MPI_GROUP_EMPTY) == (b == MPI_GROUP_EMPTY)) ? MPI_IDENT : MPI_UNEQUAL; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Group_translate_ranks(MPI_Group, int n, const int* r1, MPI_Group, int* r2) { for (int i = 0; i < n; ++i) r2[i] = r1[i]; return MPI_SUCCESS; }

/* datatypes: a handle IS its extent in bytes */
ELB_STUB_INLINE int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype* n) { *n = count * old; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Type_create_struct(int count, const int* bl, const MPI_Aint* disp, const MPI_Datatype* types, MPI_Datatype* n) {
    long ext = 0;
    for (int i = 0; i < count; ++i) { long e = (long)disp[i] + bl[i] * types[i]; if (e > ext) ext = e; }
    *n = ext; return MPI_SUCCESS;
}
ELB_STUB_INLINE int MPI_Type_create_resized(MPI_Datatype, MPI_Aint, MPI_Aint extent, MPI_Datatype* n) { *n = (long)extent; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Type_commit(MPI_Datatype*) { return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Type_free(MPI_Datatype* t) { *t = MPI_DATATYPE_NULL; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Get_address(const void* p, MPI_Aint* a) { *a = (MPI_Aint)p; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Op_create(MPI_User_function*, int, MPI_Op* op) { *op = ++elb_mpi_stub::counter(); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Op_free(MPI_Op* op) { *op = 0; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Get_count(const MPI_Status* st, MPI_Datatype t, int* c) { *c = (st && t > 0) ? (int)(st->count_bytes / t) : 0; return MPI_SUCCESS; }

/* point to point: only self-messages exist */
ELB_STUB_INLINE int MPI_Send(const void* buf, int count, MPI_Datatype t, int, int tag, MPI_Comm) {
    elb_mpi_stub::Msg m; m.tag = tag; m.data.assign((const char*)buf, (const char*)buf + (size_t)count * (size_t)t);
    elb_mpi_stub::mailbox().push_back(m); return MPI_SUCCESS;
}
ELB_STUB_INLINE int MPI_Recv(void* buf, int count, MPI_Datatype t, int, int tag, MPI_Comm, MPI_Status* st) {
    std::deque<elb_mpi_stub::Msg>& q = elb_mpi_stub::mailbox();
    for (size_t i = 0; i < q.size(); ++i) {
        if (tag == MPI_ANY_TAG || q[i].tag == tag) {
            size_t n = q[i].data.size(), cap = (size_t)count * (size_t)t;
            memcpy(buf, q[i].data.data(), n < cap ? n : cap);
            if (st) { st->MPI_SOURCE = 0; st->MPI_TAG = q[i].tag; st->MPI_ERROR = 0; st->count_bytes = (int)n; }
            q.erase(q.begin() + i);
            return MPI_SUCCESS;
        }
    }
    abort(); /* a receive with no matching self-send would deadlock under real MPI too */
}
ELB_STUB_INLINE int MPI_Isend(const void* b, int c, MPI_Datatype t, int d, int tag, MPI_Comm cm, MPI_Request* r) { *r = 1; return MPI_Send(b, c, t, d, tag, cm); }
ELB_STUB_INLINE int MPI_Issend(const void* b, int c, MPI_Datatype t, int d, int tag, MPI_Comm cm, MPI_Request* r) { *r = 1; return MPI_Send(b, c, t, d, tag, cm); }
ELB_STUB_INLINE int MPI_Irsend(const void* b, int c, MPI_Datatype t, int d, int tag, MPI_Comm cm, MPI_Request* r) { *r = 1; return MPI_Send(b, c, t, d, tag, cm); }
namespace elb_mpi_stub {
struct Pending { void* buf; int count; MPI_Datatype t; int tag; bool live; };
inline std::vector<Pending>& pending() { static std::vector<Pending> p; return p; }
}
ELB_STUB_INLINE int MPI_Irecv(void* b, int c, MPI_Datatype t, int, int tag, MPI_Comm, MPI_Request* r) {
    elb_mpi_stub::Pending p = {b, c, t, tag, true};
    elb_mpi_stub::pending().push_back(p);
    *r = 1000 + (int)elb_mpi_stub::pending().size() - 1; return MPI_SUCCESS;
}
ELB_STUB_INLINE int MPI_Wait(MPI_Request* r, MPI_Status* st) {
    if (*r >= 1000) {
        elb_mpi_stub::Pending& p = elb_mpi_stub::pending()[*r - 1000];
        if (p.live) { MPI_Recv(p.buf, p.count, p.t, 0, p.tag, MPI_COMM_SELF, st); p.live = false; }
    }
    *r = MPI_REQUEST_NULL; return MPI_SUCCESS;
}
ELB_STUB_INLINE int MPI_Waitall(int n, MPI_Request* r, MPI_Status* st) { for (int i = 0; i < n; ++i) MPI_Wait(&r[i], st ? &st[i] : 0); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Test(MPI_Request* r, int* flag, MPI_Status* st) { MPI_Wait(r, st); *flag = 1; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Iprobe(int, int tag, MPI_Comm, int* flag, MPI_Status* st) {
    *flag = 0;
    std::deque<elb_mpi_stub::Msg>& q = elb_mpi_stub::mailbox();
    for (size_t i = 0; i < q.size(); ++i)
        if (tag == MPI_ANY_TAG || q[i].tag == tag) { *flag = 1; if (st) { st->MPI_SOURCE = 0; st->MPI_TAG = q[i].tag; st->count_bytes = (int)q[i].data.size(); } break; }
    return MPI_SUCCESS;
}
ELB_STUB_INLINE int MPI_Sendrecv(const void* sb, int sc, MPI_Datatype st_, int, int, void* rb, int rc, MPI_Datatype rt, int, int, MPI_Comm, MPI_Status*) {
    size_t n = (size_t)sc * (size_t)st_, cap = (size_t)rc * (size_t)rt;
    if (sb != rb) memmove(rb, sb, n < cap ? n : cap);
    return MPI_SUCCESS;
}
ELB_STUB_INLINE int MPI_Sendrecv_replace(void*, int, MPI_Datatype, int, int, int, int, MPI_Comm, MPI_Status*) { return MPI_SUCCESS; }

/* collectives over one rank */
ELB_STUB_INLINE int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Bcast(void*, int, MPI_Datatype, int, MPI_Comm) { return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Ibcast(void*, int, MPI_Datatype, int, MPI_Comm, MPI_Request* r) { *r = 1; return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Gather(const void* s, int sc, MPI_Datatype st, void* r, int, MPI_Datatype, int, MPI_Comm) { elb_mpi_stub::cp(s, r, sc, st); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Igather(const void* s, int sc, MPI_Datatype st, void* r, int, MPI_Datatype, int, MPI_Comm, MPI_Request* q) { *q = 1; elb_mpi_stub::cp(s, r, sc, st); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Gatherv(const void* s, int sc, MPI_Datatype st, void* r, const int*, const int* displs, MPI_Datatype rt, int, MPI_Comm) { elb_mpi_stub::cp(s, (char*)r + (displs ? (long)displs[0] * rt : 0), sc, st); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Allgather(const void* s, int sc, MPI_Datatype st, void* r, int, MPI_Datatype, MPI_Comm) { elb_mpi_stub::cp(s, r, sc, st); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Allgatherv(const void* s, int sc, MPI_Datatype st, void* r, const int*, const int* displs, MPI_Datatype rt, MPI_Comm) { elb_mpi_stub::cp(s, (char*)r + (displs ? (long)displs[0] * rt : 0), sc, st); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Scatter(const void* s, int sc, MPI_Datatype st, void* r, int, MPI_Datatype, int, MPI_Comm) { if (r != MPI_IN_PLACE) elb_mpi_stub::cp(s, r, sc, st); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Scatterv(const void* s, const int* sc, const int* displs, MPI_Datatype st, void* r, int, MPI_Datatype, int, MPI_Comm) { if (r != MPI_IN_PLACE) elb_mpi_stub::cp((const char*)s + (displs ? (long)displs[0] * st : 0), r, sc[0], st); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Alltoall(const void* s, int sc, MPI_Datatype st, void* r, int, MPI_Datatype, MPI_Comm) { elb_mpi_stub::cp(s, r, sc, st); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Alltoallv(const void* s, const int* sc, const int* sd, MPI_Datatype st, void* r, const int*, const int* rd, MPI_Datatype rt, MPI_Comm) {
    elb_mpi_stub::cp((const char*)s + (sd ? (long)sd[0] * st : 0), (char*)r + (rd ? (long)rd[0] * rt : 0), sc[0], st); return MPI_SUCCESS;
}
ELB_STUB_INLINE int MPI_Reduce(const void* s, void* r, int c, MPI_Datatype t, MPI_Op, int, MPI_Comm) { elb_mpi_stub::cp(s, r, c, t); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Allreduce(const void* s, void* r, int c, MPI_Datatype t, MPI_Op, MPI_Comm) { elb_mpi_stub::cp(s, r, c, t); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Scan(const void* s, void* r, int c, MPI_Datatype t, MPI_Op, MPI_Comm) { elb_mpi_stub::cp(s, r, c, t); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Reduce_scatter(const void* s, void* r, const int* rc, MPI_Datatype t, MPI_Op, MPI_Comm) { elb_mpi_stub::cp(s, r, rc[0], t); return MPI_SUCCESS; }
ELB_STUB_INLINE int MPI_Reduce_scatter_block(const void* s, void* r, int rc, MPI_Datatype t, MPI_Op, MPI_Comm) { elb_mpi_stub::cp(s, r, rc, t); return MPI_SUCCESS; }

}  /* extern "C++" */
#endif /* __cplusplus */
#endif
