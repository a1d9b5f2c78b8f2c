/* Minimal stand-in for <mpi.h>: test infrastructure only (see oracle/README.md).
 *
 * The image has no MPI library, yet Microchannel3D includes <mpi.h> and calls
 * twelve MPI-1 functions.  This header declares exactly those; minimpi.cpp
 * implements them either for one rank (default) or, when the environment
 * variable LBM_MINIMPI_RANKS=P is set, for P forked processes that talk through
 * shared-memory mailboxes -- enough to run the reference's real z-slab
 * decomposition on P host cores.  Nothing here is product code. */
#ifndef LBM_ORACLE_SHIM_MPI_H
#define LBM_ORACLE_SHIM_MPI_H

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Request;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_INT 4
#define MPI_DOUBLE 8
#define MPI_MAX_PROCESSOR_NAME 64
#define MPI_SUCCESS 0

#ifdef __cplusplus
extern "C" {
#endif
int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Get_processor_name(char *name, int *len);
int MPI_Abort(MPI_Comm comm, int code);
int MPI_Barrier(MPI_Comm comm);
int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm comm, MPI_Status *st);
int MPI_Isend(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Wait(MPI_Request *req, MPI_Status *st);
int MPI_Sendrecv_replace(void *buf, int count, MPI_Datatype type, int dest, int sendtag,
                         int src, int recvtag, MPI_Comm comm, MPI_Status *st);
/* not MPI: lets the driver learn whether this process is a forked child */
int minimpi_rank(void);
int minimpi_size(void);
#ifdef __cplusplus
}
#endif
#endif
