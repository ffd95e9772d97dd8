/* Type-only stand-in for <mpi.h>: the reference includes it unconditionally
 * (comm.h:3, gpu_utility.h:50) even in serial builds (no -DDO_MPI). */
#ifndef ORACLE_SHIM_MPI_H
#define ORACLE_SHIM_MPI_H
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int count; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_BYTE 1
#define MPI_INT 2
#define MPI_DOUBLE 3
#define MPI_FLOAT 4
#define MPI_CHAR 5
static inline int MPI_Comm_rank(MPI_Comm c, int* r) { (void)c; *r = 0; return 0; }
static inline int MPI_Comm_size(MPI_Comm c, int* n) { (void)c; *n = 1; return 0; }
static inline int MPI_Barrier(MPI_Comm c) { (void)c; return 0; }
static inline int MPI_Finalize(void) { return 0; }
#endif
