/* minimpi.c — the nine MPI calls of the reference over POSIX shared memory (baseline tooling). */
#define _GNU_SOURCE
#include "mpi.h"

#include <fcntl.h>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

#define MM_MAX_RANKS 16
#define MM_SLOTS 8
#define MM_SLOT_BYTES (128 * 1024)

typedef struct {
  volatile unsigned long head, tail; /* tail: next slot to fill; head: next slot to drain */
  char pad[48];
  struct {
    volatile int ready, bytes, tag, pad;
    char data[MM_SLOT_BYTES];
  } slot[MM_SLOTS];
} mm_queue;

typedef struct {
  volatile int barrier_count;
  volatile int barrier_sense;
  int size;
  char pad[52];
  mm_queue q[MM_MAX_RANKS][MM_MAX_RANKS]; /* [src][dst] */
} mm_shared;

static mm_shared* g_sh = NULL;
static int g_rank = 0, g_size = 1, g_local_sense = 0;

static int dt_bytes(MPI_Datatype dt) { return dt == MPI_INT ? 4 : 1; }

size_t minimpi_shared_bytes(void) { return sizeof(mm_shared); }

int MPI_Init(int* argc, char*** argv) {
  (void)argc; (void)argv;
  const char* r = getenv("MINIMPI_RANK");
  const char* s = getenv("MINIMPI_SIZE");
  const char* name = getenv("MINIMPI_SHM");
  g_rank = r ? atoi(r) : 0;
  g_size = s ? atoi(s) : 1;
  if (g_size > MM_MAX_RANKS) { fprintf(stderr, "minimpi: at most %d ranks\n", MM_MAX_RANKS); exit(1); }
  if (name) {
    int fd = shm_open(name, O_RDWR, 0600);
    if (fd < 0) { perror("minimpi: shm_open"); exit(1); }
    g_sh = (mm_shared*)mmap(NULL, sizeof(mm_shared), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
  } else { /* singleton: private zeroed region */
    g_sh = (mm_shared*)mmap(NULL, sizeof(mm_shared), PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
  }
  if (g_sh == MAP_FAILED) { perror("minimpi: mmap"); exit(1); }
  return MPI_SUCCESS;
}

int MPI_Finalize(void) { return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm c, int* rank) { (void)c; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int* size) { (void)c; *size = g_size; return MPI_SUCCESS; }

int MPI_Isend(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm c, MPI_Request* req) {
  (void)c;
  const int bytes = count * dt_bytes(dt);
  if (bytes > MM_SLOT_BYTES) { fprintf(stderr, "minimpi: message of %d bytes too large\n", bytes); exit(1); }
  mm_queue* q = &g_sh->q[g_rank][dest];
  while (q->tail - q->head >= MM_SLOTS) sched_yield(); /* eager protocol, bounded queue */
  const unsigned long t = q->tail;
  memcpy((void*)q->slot[t % MM_SLOTS].data, buf, (size_t)bytes);
  q->slot[t % MM_SLOTS].bytes = bytes;
  q->slot[t % MM_SLOTS].tag = tag;
  __atomic_store_n(&q->slot[t % MM_SLOTS].ready, 1, __ATOMIC_RELEASE);
  __atomic_store_n(&q->tail, t + 1, __ATOMIC_RELEASE);
  if (req) req->kind = 0;
  return MPI_SUCCESS;
}

int MPI_Irecv(void* buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm c, MPI_Request* req) {
  (void)c;
  req->kind = 1;
  req->buf = buf;
  req->bytes = count * dt_bytes(dt);
  req->peer = source;
  req->tag = tag;
  return MPI_SUCCESS;
}

int MPI_Wait(MPI_Request* req, MPI_Status* st) {
  if (!req || req->kind == 0) return MPI_SUCCESS;
  mm_queue* q = &g_sh->q[req->peer][g_rank];
  /* receives from one source complete in posting order against sends in posting order
     (non-overtaking); the reference waits on its receives in posting order */
  const unsigned long h = q->head;
  while (!__atomic_load_n(&q->slot[h % MM_SLOTS].ready, __ATOMIC_ACQUIRE)) sched_yield();
  int n = q->slot[h % MM_SLOTS].bytes;
  if (q->slot[h % MM_SLOTS].tag != req->tag) { fprintf(stderr, "minimpi: tag mismatch\n"); exit(1); }
  if (n > req->bytes) n = req->bytes;
  memcpy(req->buf, (const void*)q->slot[h % MM_SLOTS].data, (size_t)n);
  q->slot[h % MM_SLOTS].ready = 0;
  __atomic_store_n(&q->head, h + 1, __ATOMIC_RELEASE);
  if (st) { st->MPI_SOURCE = req->peer; st->MPI_TAG = req->tag; st->MPI_ERROR = 0; }
  req->kind = 0;
  return MPI_SUCCESS;
}

int MPI_Barrier(MPI_Comm c) {
  (void)c;
  if (g_size == 1) return MPI_SUCCESS;
  g_local_sense = !g_local_sense;
  if (__atomic_add_fetch(&g_sh->barrier_count, 1, __ATOMIC_ACQ_REL) == g_size) {
    g_sh->barrier_count = 0;
    __atomic_store_n(&g_sh->barrier_sense, g_local_sense, __ATOMIC_RELEASE);
  } else {
    while (__atomic_load_n(&g_sh->barrier_sense, __ATOMIC_ACQUIRE) != g_local_sense) sched_yield();
  }
  return MPI_SUCCESS;
}

double MPI_Wtime(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}
