/* hypre_b200/host/mpishim/mpi.h -- the MPI stand-in of the native multi-rank mode (our own code).
 *
 * A small message-passing library with the MPI-1/2 subset the reference calls through its
 * hypre_MPI_* wrappers (src/utilities/mpistubs.c:919-1530), so that the reference's host code can
 * run on N ranks (N processes on one machine) on boxes that have no MPI installation.  Two builds
 * use this source: the product's (hypre_b200/host/Makefile -> hypre_b200/lib/libmpishim.so: the ranks
 * of hypre's own distributed setup, one per GPU, ij_b200_mp / bench.py --native; with a real MPI
 * installation those binaries link the real library instead), and the checker's own (oracle/Makefile
 * -> oracle/_ref/libmpishim.so: reproduces the reference's multi-rank golden results, src/test/TEST_ij
 * .saved files, all recorded with mpirun -np 2..4).
 * Ranks are processes started by mpirun_shim (or by torchrun, see mpishim.c), connected pairwise by Unix stream
 * sockets; sends are buffered (eager), receives match on (communicator, source, tag) in arrival
 * order, collectives are linear algorithms over point-to-point messages with a fixed
 * combination order (rank 0, 1, 2, ...), hence deterministic.
 *
 * Not a general MPI: one-sided operations, I/O, topologies, inter-communicators, error handlers
 * and threads are not provided.
 */
#ifndef ORACLE_MPISHIM_MPI_H
#define ORACLE_MPISHIM_MPI_H

#ifdef __cplusplus
extern "C" {
#endif

#define MPI_VERSION    3
#define MPI_SUBVERSION 1

typedef int  MPI_Comm;
typedef int  MPI_Group;
typedef int  MPI_Request;
typedef int  MPI_Datatype;
typedef int  MPI_Op;
typedef int  MPI_Info;
typedef int  MPI_Fint;
typedef long MPI_Aint;

typedef struct
{
   int MPI_SOURCE;
   int MPI_TAG;
   int MPI_ERROR;
   int shim_bytes;     /* size of the matched message */
} MPI_Status;

typedef void (MPI_User_function)(void *invec, void *inoutvec, int *len, MPI_Datatype *datatype);

#define MPI_SUCCESS          0
#define MPI_ERR_OTHER        15
#define MPI_COMM_WORLD       ((MPI_Comm)0)
#define MPI_COMM_SELF        ((MPI_Comm)1)
#define MPI_COMM_NULL        ((MPI_Comm)-1)
#define MPI_GROUP_NULL       ((MPI_Group)-1)
#define MPI_REQUEST_NULL     ((MPI_Request)-1)
#define MPI_INFO_NULL        ((MPI_Info)0)
#define MPI_OP_NULL          ((MPI_Op)-1)
#define MPI_DATATYPE_NULL    ((MPI_Datatype)-1)
#define MPI_ANY_SOURCE       (-1)
#define MPI_ANY_TAG          (-1)
#define MPI_PROC_NULL        (-2)
#define MPI_UNDEFINED        (-32766)
#define MPI_BOTTOM           ((void *)0)
#define MPI_IN_PLACE         ((void *)-1)
#define MPI_STATUS_IGNORE    ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE  ((MPI_Status *)0)
#define MPI_COMM_TYPE_SHARED 1
#define MPI_MAX_PROCESSOR_NAME 64

/* basic datatypes (size table in mpishim.c) */
#define MPI_CHAR                  ((MPI_Datatype)1)
#define MPI_BYTE                  ((MPI_Datatype)2)
#define MPI_SHORT                 ((MPI_Datatype)3)
#define MPI_INT                   ((MPI_Datatype)4)
#define MPI_LONG                  ((MPI_Datatype)5)
#define MPI_LONG_LONG_INT         ((MPI_Datatype)6)
#define MPI_LONG_LONG             MPI_LONG_LONG_INT
#define MPI_UNSIGNED              ((MPI_Datatype)7)
#define MPI_UNSIGNED_LONG         ((MPI_Datatype)8)
#define MPI_FLOAT                 ((MPI_Datatype)9)
#define MPI_DOUBLE                ((MPI_Datatype)10)
#define MPI_LONG_DOUBLE           ((MPI_Datatype)11)
#define MPI_C_FLOAT_COMPLEX       ((MPI_Datatype)12)
#define MPI_C_DOUBLE_COMPLEX      ((MPI_Datatype)13)
#define MPI_C_LONG_DOUBLE_COMPLEX ((MPI_Datatype)14)
#define MPI_REAL                  MPI_FLOAT
#define MPI_COMPLEX               MPI_C_FLOAT_COMPLEX
#define MPI_UNSIGNED_CHAR         ((MPI_Datatype)15)

/* reduction operators; user operators start at 100 */
#define MPI_SUM   ((MPI_Op)1)
#define MPI_MIN   ((MPI_Op)2)
#define MPI_MAX   ((MPI_Op)3)
#define MPI_LOR   ((MPI_Op)4)
#define MPI_LAND  ((MPI_Op)5)
#define MPI_BOR   ((MPI_Op)6)
#define MPI_BAND  ((MPI_Op)7)
#define MPI_PROD  ((MPI_Op)8)

int MPI_Init(int *argc, char ***argv);
int MPI_Initialized(int *flag);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int errorcode);
double MPI_Wtime(void);
double MPI_Wtick(void);

int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_dup(MPI_Comm comm, MPI_Comm *newcomm);
int MPI_Comm_free(MPI_Comm *comm);
int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm *newcomm);
int MPI_Comm_split_type(MPI_Comm comm, int split_type, int key, MPI_Info info, MPI_Comm *newcomm);
int MPI_Comm_group(MPI_Comm comm, MPI_Group *group);
int MPI_Comm_create(MPI_Comm comm, MPI_Group group, MPI_Comm *newcomm);
MPI_Comm MPI_Comm_f2c(MPI_Fint comm);
int MPI_Group_incl(MPI_Group group, int n, const int *ranks, MPI_Group *newgroup);
int MPI_Group_free(MPI_Group *group);
int MPI_Info_create(MPI_Info *info);
int MPI_Info_free(MPI_Info *info);

int MPI_Send(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Isend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Irsend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Irecv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Send_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Recv_init(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Startall(int count, MPI_Request *reqs);
int MPI_Request_free(MPI_Request *req);
int MPI_Probe(int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Iprobe(int source, int tag, MPI_Comm comm, int *flag, MPI_Status *status);
int MPI_Test(MPI_Request *req, int *flag, MPI_Status *status);
int MPI_Testall(int count, MPI_Request *reqs, int *flag, MPI_Status *statuses);
int MPI_Wait(MPI_Request *req, MPI_Status *status);
int MPI_Waitall(int count, MPI_Request *reqs, MPI_Status *statuses);
int MPI_Waitany(int count, MPI_Request *reqs, int *index, MPI_Status *status);
int MPI_Get_count(const MPI_Status *status, MPI_Datatype dt, int *count);

int MPI_Barrier(MPI_Comm comm);
int MPI_Bcast(void *buf, int count, MPI_Datatype dt, int root, MPI_Comm comm);
int MPI_Reduce(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, int root, MPI_Comm comm);
int MPI_Allreduce(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm);
int MPI_Scan(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm);
int MPI_Gather(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Gatherv(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, const int *rcounts, const int *displs,
                MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Scatter(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Scatterv(const void *sbuf, const int *scounts, const int *displs, MPI_Datatype sdt, void *rbuf, int rcount,
                 MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Allgather(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, MPI_Comm comm);
int MPI_Allgatherv(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, const int *rcounts, const int *displs,
                   MPI_Datatype rdt, MPI_Comm comm);
int MPI_Alltoall(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, MPI_Comm comm);

int MPI_Op_create(MPI_User_function *fn, int commute, MPI_Op *op);
int MPI_Op_free(MPI_Op *op);
int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype *newtype);
int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype oldtype, MPI_Datatype *newtype);
int MPI_Type_hvector(int count, int blocklength, MPI_Aint stride, MPI_Datatype oldtype, MPI_Datatype *newtype);
int MPI_Type_create_hvector(int count, int blocklength, MPI_Aint stride, MPI_Datatype oldtype, MPI_Datatype *newtype);
int MPI_Type_struct(int count, int *blocklens, MPI_Aint *displs, MPI_Datatype *types, MPI_Datatype *newtype);
int MPI_Type_create_struct(int count, const int *blocklens, const MPI_Aint *displs, const MPI_Datatype *types, MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *dt);
int MPI_Type_free(MPI_Datatype *dt);
int MPI_Address(void *location, MPI_Aint *address);
int MPI_Get_address(const void *location, MPI_Aint *address);

#ifdef __cplusplus
}
#endif
#endif
