/* mpi.h — minimal single-host MPI for building the UNMODIFIED reference.
 *
 * BASELINE TOOLING ONLY.  This image has no MPI (no mpi.h, libmpi or mpirun), and the
 * reference needs exactly nine MPI calls.  This header plus
 * minimpi.c provide those nine over POSIX shared memory with MPI's semantics where the
 * reference depends on them: non-blocking send/recv, per-(source,destination) FIFO
 * ordering ("non-overtaking"), tag matching, a barrier and a wall clock.  Nothing in
 * the product path includes or links this.
 */
#ifndef MINIMPI_H_
#define MINIMPI_H_
#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
typedef struct MPI_Request_s {
  int kind;      /* 0 = null/done, 1 = recv pending */
  void* buf;
  int bytes, peer, tag;
} MPI_Request;

#define MPI_COMM_WORLD 0
#define MPI_UNSIGNED_CHAR 1
#define MPI_CHAR 2
#define MPI_INT 3
#define MPI_SUCCESS 0

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Isend(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Wait(MPI_Request* req, MPI_Status* status);
int MPI_Barrier(MPI_Comm comm);
double MPI_Wtime(void);

#ifdef __cplusplus
}
#endif
#endif
