/*
 * mpi.h -- minimal stand-in for the 19 MPI entry points the reference's hot
 * path uses, backed by fork() + one shared anonymous mapping
 * (mpi_shim.cpp).  TEST INFRASTRUCTURE: exists only so that the unmodified
 * reference sources under /root/reference compile and run here, where no MPI
 * is installed.  Not an MPI implementation; semantics cover exactly what the
 * reference's call sites need.
 */
#ifndef ORACLE_STUB_MPI_H_
#define ORACLE_STUB_MPI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Info;
typedef int MPI_Win;
typedef long MPI_Aint;
typedef struct { int source, tag, error; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_BYTE 1
#define MPI_UINT64_T 8
#define MPI_SUM 1
#define MPI_INFO_NULL 0
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_SUCCESS 0

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Barrier(MPI_Comm comm);
int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Exscan(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Alloc_mem(MPI_Aint size, MPI_Info info, void *baseptr);
int MPI_Free_mem(void *base);
int MPI_Win_create(void *base, MPI_Aint size, int disp_unit, MPI_Info info, MPI_Comm comm, MPI_Win *win);
int MPI_Win_free(MPI_Win *win);
int MPI_Win_lock_all(int assert_, MPI_Win win);
int MPI_Win_unlock_all(MPI_Win win);
int MPI_Put(const void *origin, int origin_count, MPI_Datatype origin_type, int target_rank,
            MPI_Aint target_disp, int target_count, MPI_Datatype target_type, MPI_Win win);
int MPI_Win_flush_local(int rank, MPI_Win win);
int MPI_Win_flush_local_all(MPI_Win win);

/* shim control (called by the harness, not by the reference) */
int shim_setup(int nranks, size_t shared_bytes); /* before fork */
void shim_set_rank(int rank);                     /* in each child */
void *shim_result_area(void);                     /* shared, 4096 bytes per rank */

#ifdef __cplusplus
}
#endif
#endif
