/* Minimal MPI for the ORACLE BUILD ONLY (test infrastructure, not product code).
 *
 * Lets the reference's own CPU sources (/root/reference, unmodified, compiled where they lie)
 * build and run on a box without any MPI implementation.  Implemented in mpi_stub.cpp:
 *   - one rank by default (collectives are memcpy);
 *   - or ELB_STUB_NP separate processes (ELB_STUB_RANK = 0..NP-1) that share one
 *     /dev/shm file (ELB_STUB_SHM) holding a byte ring per ordered rank pair — enough of MPI
 *     (communicators, groups, cartesian split, point-to-point, the collectives) for the
 *     reference's Grid / DistMatrix / Gemm to run multi-rank, so that the oracle's per-rank
 *     data placement can be pinned against the real thing.
 * Only what the reference's src/core + include/El/core use is provided.
 */
#ifndef ELB200_ORACLE_MPI_STUB_H
#define ELB200_ORACLE_MPI_STUB_H

#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

typedef int MPI_Comm;
typedef int MPI_Group;
typedef int MPI_Op;
typedef int MPI_Errhandler;
/* a datatype handle = (kind << 32) | extent in bytes; kind 0 = opaque bytes (derived types) */
typedef long MPI_Datatype;
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

#define ELB_DT(kind, bytes) (((long)(kind) << 32) | (long)(bytes))
#define MPI_CHAR ELB_DT(1, 1)
#define MPI_UNSIGNED_CHAR ELB_DT(2, 1)
#define MPI_BYTE ELB_DT(2, 1)
#define MPI_SHORT ELB_DT(3, 2)
#define MPI_INT ELB_DT(4, 4)
#define MPI_UNSIGNED ELB_DT(5, 4)
#define MPI_LONG ELB_DT(6, 8)
#define MPI_LONG_LONG_INT ELB_DT(6, 8)
#define MPI_LONG_LONG ELB_DT(6, 8)
#define MPI_UNSIGNED_LONG ELB_DT(7, 8)
#define MPI_UNSIGNED_LONG_LONG ELB_DT(7, 8)
#define MPI_FLOAT ELB_DT(8, 4)
#define MPI_DOUBLE ELB_DT(9, 8)
#define MPI_LONG_DOUBLE ELB_DT(10, 16)
#define MPI_C_FLOAT_COMPLEX ELB_DT(11, 8)
#define MPI_C_DOUBLE_COMPLEX ELB_DT(12, 16)
#define MPI_FLOAT_INT ELB_DT(13, 8)
#define MPI_DOUBLE_INT ELB_DT(14, 16)
#define MPI_LONG_INT ELB_DT(15, 16)

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
extern "C" {
#endif
int MPI_Init(int*, char***);
int MPI_Init_thread(int*, char***, int required, int* provided);
int MPI_Query_thread(int* provided);
int MPI_Initialized(int* flag);
int MPI_Finalized(int* flag);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm, int code);
double MPI_Wtime(void);
int MPI_Error_string(int, char* s, int* len);
int MPI_Comm_set_errhandler(MPI_Comm, MPI_Errhandler);

int MPI_Comm_rank(MPI_Comm, int*);
int MPI_Comm_size(MPI_Comm, int*);
int MPI_Comm_dup(MPI_Comm, MPI_Comm*);
int MPI_Comm_split(MPI_Comm, int color, int key, MPI_Comm*);
int MPI_Comm_create(MPI_Comm, MPI_Group, MPI_Comm*);
int MPI_Comm_free(MPI_Comm*);
int MPI_Comm_group(MPI_Comm, MPI_Group*);
int MPI_Comm_compare(MPI_Comm, MPI_Comm, int*);
int MPI_Cart_create(MPI_Comm, int ndims, const int* dims, const int* periods, int reorder, MPI_Comm*);
int MPI_Cart_sub(MPI_Comm, const int* remain, MPI_Comm*);
int MPI_Group_size(MPI_Group, int*);
int MPI_Group_rank(MPI_Group, int*);
int MPI_Group_free(MPI_Group*);
int MPI_Group_dup(MPI_Group, MPI_Group*);
int MPI_Group_incl(MPI_Group, int n, const int* ranks, MPI_Group*);
int MPI_Group_excl(MPI_Group, int n, const int* ranks, MPI_Group*);
int MPI_Group_union(MPI_Group, MPI_Group, MPI_Group*);
int MPI_Group_difference(MPI_Group, MPI_Group, MPI_Group*);
int MPI_Group_compare(MPI_Group, MPI_Group, int*);
int MPI_Group_translate_ranks(MPI_Group, int n, const int* r1, MPI_Group, int* r2);

int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype* n);
int MPI_Type_create_struct(int count, const int* bl, const MPI_Aint* disp, const MPI_Datatype* types, MPI_Datatype* n);
int MPI_Type_create_resized(MPI_Datatype, MPI_Aint lb, MPI_Aint extent, MPI_Datatype* n);
int MPI_Type_commit(MPI_Datatype*);
int MPI_Type_free(MPI_Datatype*);
int MPI_Get_address(const void* p, MPI_Aint* a);
int MPI_Op_create(MPI_User_function*, int commute, MPI_Op*);
int MPI_Op_free(MPI_Op*);
int MPI_Get_count(const MPI_Status*, MPI_Datatype, int*);

int MPI_Send(const void*, int, MPI_Datatype, int dest, int tag, MPI_Comm);
int MPI_Recv(void*, int, MPI_Datatype, int src, int tag, MPI_Comm, MPI_Status*);
int MPI_Isend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Issend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Irsend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Irecv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Wait(MPI_Request*, MPI_Status*);
int MPI_Waitall(int, MPI_Request*, MPI_Status*);
int MPI_Test(MPI_Request*, int* flag, MPI_Status*);
int MPI_Iprobe(int src, int tag, MPI_Comm, int* flag, MPI_Status*);
int MPI_Sendrecv(const void*, int, MPI_Datatype, int dest, int stag, void*, int, MPI_Datatype, int src, int rtag,
                 MPI_Comm, MPI_Status*);
int MPI_Sendrecv_replace(void*, int, MPI_Datatype, int dest, int stag, int src, int rtag, MPI_Comm, MPI_Status*);

int MPI_Barrier(MPI_Comm);
int MPI_Bcast(void*, int, MPI_Datatype, int root, MPI_Comm);
int MPI_Ibcast(void*, int, MPI_Datatype, int root, MPI_Comm, MPI_Request*);
int MPI_Gather(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, int root, MPI_Comm);
int MPI_Igather(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, int root, MPI_Comm, MPI_Request*);
int MPI_Gatherv(const void*, int, MPI_Datatype, void*, const int* rc, const int* displs, MPI_Datatype, int root, MPI_Comm);
int MPI_Allgather(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, MPI_Comm);
int MPI_Allgatherv(const void*, int, MPI_Datatype, void*, const int* rc, const int* displs, MPI_Datatype, MPI_Comm);
int MPI_Scatter(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, int root, MPI_Comm);
int MPI_Scatterv(const void*, const int* sc, const int* displs, MPI_Datatype, void*, int, MPI_Datatype, int root, MPI_Comm);
int MPI_Alltoall(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, MPI_Comm);
int MPI_Alltoallv(const void*, const int* sc, const int* sd, MPI_Datatype, void*, const int* rc, const int* rd,
                  MPI_Datatype, MPI_Comm);
int MPI_Reduce(const void*, void*, int, MPI_Datatype, MPI_Op, int root, MPI_Comm);
int MPI_Allreduce(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm);
int MPI_Scan(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm);
int MPI_Reduce_scatter(const void*, void*, const int* rc, MPI_Datatype, MPI_Op, MPI_Comm);
int MPI_Reduce_scatter_block(const void*, void*, int rc, MPI_Datatype, MPI_Op, MPI_Comm);
#ifdef __cplusplus
}
#endif
#endif
