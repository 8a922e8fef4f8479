/* Single-node stand-in for <mpi.h>, used only to build the UNMODIFIED reference (nkkko/TCLB) for the reference arm
 * of bench.py (baseline/build_reference.py).  The image has no MPI implementation; this header + mpi_stub.cpp provide
 * the entry points the reference calls (src/Lattice.cu.Rt:327-389,1279-1292, src/Solver.cpp.Rt:362-365,447-701,
 * src/MPMD.hpp:142-246, src/main.cpp:151,179, handlers) for ranks that are processes on ONE machine:
 * rank/size come from the environment (TCLB_MPI_RANK/TCLB_MPI_SIZE, else RANK/WORLD_SIZE as torchrun sets them),
 * messages and collectives go through a memory-mapped file.  Not a general MPI: one communicator (every split of the
 * world by program/host name gives the world back), contiguous basic datatypes, no one-sided, no I/O. */
#ifndef TCLB_MPI_STUB_H
#define TCLB_MPI_STUB_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef int MPI_Group;
typedef int MPI_Info;
typedef int MPI_Errhandler;
typedef long MPI_Aint;
typedef long long MPI_Offset;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR, count_bytes; } MPI_Status;

#define MPI_SUCCESS 0
#define MPI_COMM_NULL 0
#define MPI_COMM_WORLD 1
#define MPI_COMM_SELF 2
#define MPI_GROUP_NULL 0
#define MPI_GROUP_EMPTY 3
#define MPI_INFO_NULL 0
#define MPI_REQUEST_NULL (-1)
#define MPI_UNDEFINED (-32766)
#define MPI_ANY_SOURCE (-2)
#define MPI_ANY_TAG (-1)
#define MPI_PROC_NULL (-3)
#define MPI_ROOT (-4)
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_MAX_ERROR_STRING 256
#define MPI_UNIVERSE_SIZE 9
#define MPI_TAG_UB 10
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)
#define MPI_ARGV_NULL ((char **)0)
#define MPI_ERRCODES_IGNORE ((int *)0)
#define MPI_IN_PLACE ((void *)1)
#define MPI_ERRORS_RETURN 1
#define MPI_ERRORS_ARE_FATAL 0
#define MPI_THREAD_SINGLE 0
#define MPI_THREAD_MULTIPLE 3

/* basic datatypes: handle = kind; sizes in mpi_stub.cpp */
#define MPI_DATATYPE_NULL 0
#define MPI_BYTE 1
#define MPI_CHAR 2
#define MPI_UNSIGNED_CHAR 3
#define MPI_SHORT 4
#define MPI_UNSIGNED_SHORT 5
#define MPI_INT 6
#define MPI_UNSIGNED 7
#define MPI_LONG 8
#define MPI_UNSIGNED_LONG 9
#define MPI_LONG_LONG 10
#define MPI_UNSIGNED_LONG_LONG 11
#define MPI_FLOAT 12
#define MPI_DOUBLE 13
#define MPI_LONG_LONG_INT MPI_LONG_LONG
#define MPI_SIGNED_CHAR MPI_CHAR
#define MPI_C_BOOL MPI_UNSIGNED_CHAR

#define MPI_OP_NULL 0
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3
#define MPI_LOR 4
#define MPI_LAND 5
#define MPI_PROD 6
#define MPI_BOR 7
#define MPI_BAND 8

int MPI_Init(int *argc, char ***argv);
int MPI_Init_thread(int *argc, char ***argv, int required, int *provided);
int MPI_Initialized(int *flag);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int code);
double MPI_Wtime(void);
int MPI_Get_processor_name(char *name, int *len);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Comm_remote_size(MPI_Comm comm, int *size);
int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm *newcomm);
int MPI_Comm_dup(MPI_Comm comm, MPI_Comm *newcomm);
int MPI_Comm_free(MPI_Comm *comm);
int MPI_Comm_group(MPI_Comm comm, MPI_Group *group);
int MPI_Comm_create(MPI_Comm comm, MPI_Group group, MPI_Comm *newcomm);
int MPI_Comm_get_attr(MPI_Comm comm, int keyval, void *attr, int *flag);
int MPI_Comm_get_parent(MPI_Comm *parent);
int MPI_Comm_spawn(const char *cmd, char *argv[], int maxprocs, MPI_Info info, int root, MPI_Comm comm, MPI_Comm *inter, int errcodes[]);
int MPI_Comm_set_errhandler(MPI_Comm comm, MPI_Errhandler e);
int MPI_Intercomm_create(MPI_Comm local, int local_leader, MPI_Comm peer, int remote_leader, int tag, MPI_Comm *inter);
int MPI_Group_excl(MPI_Group group, int n, const int ranks[], MPI_Group *newgroup);
int MPI_Group_rank(MPI_Group group, int *rank);
int MPI_Group_size(MPI_Group group, int *size);
int MPI_Group_free(MPI_Group *group);
int MPI_Barrier(MPI_Comm comm);
int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Reduce(const void *send, void *recv, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm);
int MPI_Allreduce(const void *send, void *recv, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Gather(const void *send, int scount, MPI_Datatype stype, void *recv, int rcount, MPI_Datatype rtype, int root, MPI_Comm comm);
int MPI_Allgather(const void *send, int scount, MPI_Datatype stype, void *recv, int rcount, MPI_Datatype rtype, MPI_Comm comm);
int MPI_Gatherv(const void *send, int scount, MPI_Datatype stype, void *recv, const int *rcounts, const int *displs, MPI_Datatype rtype, int root, MPI_Comm comm);
int MPI_Scatterv(const void *send, const int *scounts, const int *displs, MPI_Datatype stype, void *recv, int rcount, MPI_Datatype rtype, int root, MPI_Comm comm);
int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Isend(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Irecv(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Wait(MPI_Request *req, MPI_Status *status);
int MPI_Test(MPI_Request *req, int *flag, MPI_Status *status);
int MPI_Waitall(int n, MPI_Request reqs[], MPI_Status statuses[]);
int MPI_Waitany(int n, MPI_Request reqs[], int *index, MPI_Status *status);
int MPI_Get_count(const MPI_Status *status, MPI_Datatype type, int *count);
int MPI_Type_size(MPI_Datatype type, int *size);
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *newtype);
int MPI_Type_indexed(int count, const int *blocklens, const int *displs, MPI_Datatype old, MPI_Datatype *newtype);
int MPI_Type_create_resized(MPI_Datatype old, MPI_Aint lb, MPI_Aint extent, MPI_Datatype *newtype);
int MPI_Type_get_extent(MPI_Datatype type, MPI_Aint *lb, MPI_Aint *extent);
int MPI_Type_commit(MPI_Datatype *type);
int MPI_Type_free(MPI_Datatype *type);
int MPI_Error_string(int code, char *str, int *len);

#ifdef __cplusplus
}
#endif
#endif
