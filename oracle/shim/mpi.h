/* oracle/shim/mpi.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * An in-process stand-in for the ~30 MPI entry points the reference's hot path
 * uses (no MPI exists in this image).  Ranks are std::threads of one process:
 * fake_mpi_run(n, fn) starts n threads, each with a thread-local rank;
 * collectives rendezvous on a shared barrier, point-to-point messages go
 * through eager mailboxes keyed by (src, dst, tag) with FIFO matching.
 * Implemented in fake_mpi.cpp.  Written from the MPI standard's semantics;
 * nothing is copied from an MPI implementation.
 */
#ifndef AMRVR_ORACLE_FAKE_MPI_H
#define AMRVR_ORACLE_FAKE_MPI_H

#ifdef __cplusplus
#include <functional>
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct fake_mpi_group* MPI_Group;
typedef struct fake_mpi_request* MPI_Request;
typedef struct {
  int MPI_SOURCE, MPI_TAG, MPI_ERROR;
} MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_UNDEFINED (-32766)
#define MPI_GROUP_NULL ((MPI_Group)0)
#define MPI_REQUEST_NULL ((MPI_Request)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)

/* datatypes: value = tag, size looked up in fake_mpi.cpp */
#define MPI_BYTE 1
#define MPI_INT 2
#define MPI_FLOAT 3
#define MPI_DOUBLE 4
#define MPI_LONG_LONG 5
#define MPI_UINT64_T 6
#define MPI_UNSIGNED_LONG_LONG 6

#define MPI_MIN 1
#define MPI_MAX 2
#define MPI_SUM 3

int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_group(MPI_Comm comm, MPI_Group* group);
int MPI_Group_size(MPI_Group group, int* size);
int MPI_Group_rank(MPI_Group group, int* rank);
int MPI_Group_free(MPI_Group* group);
int MPI_Group_incl(MPI_Group group, int n, const int ranks[], MPI_Group* out);
int MPI_Group_range_incl(MPI_Group group, int n, int ranges[][3],
                         MPI_Group* out);
int MPI_Group_translate_ranks(MPI_Group g1, int n, const int ranks1[],
                              MPI_Group g2, int ranks2[]);

int MPI_Isend(const void* buf, int count, MPI_Datatype type, int dest, int tag,
              MPI_Comm comm, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int tag,
              MPI_Comm comm, MPI_Request* req);
int MPI_Waitall(int count, MPI_Request reqs[], MPI_Status statuses[]);
int MPI_Waitany(int count, MPI_Request reqs[], int* index, MPI_Status* status);

int MPI_Barrier(MPI_Comm comm);
int MPI_Allgather(const void* sbuf, int scount, MPI_Datatype stype, void* rbuf,
                  int rcount, MPI_Datatype rtype, MPI_Comm comm);
int MPI_Allgatherv(const void* sbuf, int scount, MPI_Datatype stype, void* rbuf,
                   const int rcounts[], const int displs[], MPI_Datatype rtype,
                   MPI_Comm comm);
int MPI_Gather(const void* sbuf, int scount, MPI_Datatype stype, void* rbuf,
               int rcount, MPI_Datatype rtype, int root, MPI_Comm comm);
int MPI_Gatherv(const void* sbuf, int scount, MPI_Datatype stype, void* rbuf,
                const int rcounts[], const int displs[], MPI_Datatype rtype,
                int root, MPI_Comm comm);
int MPI_Allreduce(const void* sbuf, void* rbuf, int count, MPI_Datatype type,
                  MPI_Op op, MPI_Comm comm);
int MPI_Reduce(const void* sbuf, void* rbuf, int count, MPI_Datatype type,
               MPI_Op op, int root, MPI_Comm comm);
int MPI_Abort(MPI_Comm comm, int code);

#ifdef __cplusplus
}  /* extern "C" */

/* Run fn(rank) on nranks threads that share one fake MPI_COMM_WORLD.  Any
 * exception thrown by a rank is rethrown on the caller after all threads end.
 * pinCores: pin rank r to the r-th core of the caller's affinity mask. */
void fake_mpi_run(int nranks, const std::function<void(int)>& fn,
                  bool pinCores = false);
#endif

#endif /* AMRVR_ORACLE_FAKE_MPI_H */
