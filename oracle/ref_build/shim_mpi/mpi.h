/* The nine MPI calls the reference makes under -DDO_MPI (parallel.c:69-191, CoMD.c:108, timestep.c:318), implemented in
 * oracle/ref_build/mpi_shim.c over fork() + one anonymous shared mapping: MPI_Init forks COMD_SHIM_NRANKS - 1
 * children, so "mpirun -np R CoMD" becomes "COMD_SHIM_NRANKS=R CoMD-ref-mpi" on one node.  Test infrastructure only:
 * it exists so that the reference's CPU path can be timed as ONE decomposed job on the host cores (bench.py). */
#ifndef ORACLE_SHIM_MPI_FULL_H
#define ORACLE_SHIM_MPI_FULL_H
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int count; int MPI_SOURCE; int MPI_TAG; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_BYTE 1
#define MPI_INT 2
#define MPI_DOUBLE 3
#define MPI_FLOAT 4
#define MPI_CHAR 5
#define MPI_DOUBLE_INT 6
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MINLOC 3
#define MPI_MAXLOC 4
#define MPI_SUCCESS 0
int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm c, int* rank);
int MPI_Comm_size(MPI_Comm c, int* size);
int MPI_Barrier(MPI_Comm c);
int MPI_Sendrecv(const void* sendbuf, int sendcount, MPI_Datatype sendtype, int dest, int sendtag,
                 void* recvbuf, int recvcount, MPI_Datatype recvtype, int source, int recvtag, MPI_Comm c, MPI_Status* status);
int MPI_Get_count(const MPI_Status* status, MPI_Datatype t, int* count);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c);
int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm c);
#endif
