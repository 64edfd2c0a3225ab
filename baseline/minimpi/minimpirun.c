/* minimpirun — `mpirun -n R prog args...` for minimpi: shared segment + fork/exec (baseline tooling). */
#define _GNU_SOURCE
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

size_t minimpi_shared_bytes(void);

int main(int argc, char** argv) {
  int n = 1, i = 1;
  while (i < argc && argv[i][0] == '-') {
    if ((!strcmp(argv[i], "-n") || !strcmp(argv[i], "-np")) && i + 1 < argc) { n = atoi(argv[i + 1]); i += 2; }
    else { fprintf(stderr, "minimpirun: unknown option %s\n", argv[i]); return 2; }
  }
  if (i >= argc || n < 1) { fprintf(stderr, "usage: minimpirun -n R prog [args]\n"); return 2; }
  char name[64];
  snprintf(name, sizeof name, "/minimpi_%d", (int)getpid());
  int fd = shm_open(name, O_CREAT | O_RDWR | O_EXCL, 0600);
  if (fd < 0) { perror("shm_open"); return 1; }
  if (ftruncate(fd, (off_t)minimpi_shared_bytes()) != 0) { perror("ftruncate"); shm_unlink(name); return 1; }
  close(fd);
  pid_t* pids = (pid_t*)calloc((size_t)n, sizeof(pid_t));
  for (int r = 0; r < n; ++r) {
    pid_t p = fork();
    if (p == 0) {
      char b[32];
      snprintf(b, sizeof b, "%d", r); setenv("MINIMPI_RANK", b, 1);
      snprintf(b, sizeof b, "%d", n); setenv("MINIMPI_SIZE", b, 1);
      setenv("MINIMPI_SHM", name, 1);
      execvp(argv[i], argv + i);
      perror("execvp");
      _exit(127);
    }
    pids[r] = p;
  }
  int rc = 0;
  for (int r = 0; r < n; ++r) {
    int st = 0;
    waitpid(pids[r], &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) rc = WIFEXITED(st) ? WEXITSTATUS(st) : 1;
  }
  shm_unlink(name);
  free(pids);
  return rc;
}
