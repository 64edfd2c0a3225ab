// cli_main.cpp — the `gol` command: same five positionals, messages, exit codes
// and output files as the reference driver (gol-main.c:30-146), on top of the
// B200 core.  The reference gets its rank count from an MPI launcher; this
// binary drives all ranks from one process (one rank per GPU, peer-to-peer over
// NVLink) and takes the count from --ranks (default: 1), or - like the
// reference under mpirun, gol-main.c:58-62 - runs one native process per rank
// with --spawn R: the ranks are forked before any CUDA call, swap their
// 256-byte export blobs (CUDA-IPC handle + offsets) through POSIX shared
// memory and map their two ring neighbours; no Python, no MPI.  The
// torchrun form of the same CLI is `torchrun ... -m gol_b200.cli`.
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../include/gol_b200.h"

static const char* kUsage =
    "GOL requires 5 arguments: pattern number, sq size of the world and the number of itterations, "
    "threads per block and output-on-off e.g. ./gol 0 32 2 512 0 \n";

// ---- --spawn: shared-memory rendezvous of the forked ranks -------------------------------
struct Rendezvous {
  int nranks;
  int arrived;
  int sense;
  int failed;
  float ms[64];
  unsigned long long bad[64];
  unsigned long long cells[64];
  gol_phase_counters_t phase[64];
  unsigned char blobs[64][GOL_EXPORT_BYTES];
};

static void rv_barrier(Rendezvous* rv, int* local_sense) {
  *local_sense = !*local_sense;
  if (__atomic_add_fetch(&rv->arrived, 1, __ATOMIC_ACQ_REL) == rv->nranks) {
    __atomic_store_n(&rv->arrived, 0, __ATOMIC_RELAXED);
    __atomic_store_n(&rv->sense, *local_sense, __ATOMIC_RELEASE);
  } else {
    while (__atomic_load_n(&rv->sense, __ATOMIC_ACQUIRE) != *local_sense) {
      if (__atomic_load_n(&rv->failed, __ATOMIC_ACQUIRE)) exit(255);   // a peer died: do not hang
      usleep(50);
    }
  }
}

static Rendezvous* g_rv = nullptr;

static void die(const char* what) {
  if (g_rv) __atomic_store_n(&g_rv->failed, 1, __ATOMIC_RELEASE);
  fprintf(stderr, "gol: %s: %s\n", what, gol_last_error());
  exit(255);
}

int main(int argc, char** argv) {
  std::vector<const char*> pos;
  int ranks = 1, mode = GOL_MODE_AUTO, k = 0, depth = 0, cpu = 0, check = 0, json = 0, ndev = 0, strip = 0, spawn = 0;
  long long rows = 0;
  double density = 0.5;
  unsigned long long seed = 1;
  const char* outdir = nullptr;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto val = [&](const char* name) -> const char* {
      if (i + 1 >= argc) {
        fprintf(stderr, "gol: %s needs a value\n", name);
        exit(255);
      }
      return argv[++i];
    };
    if (a == "--ranks") ranks = atoi(val("--ranks"));
    else if (a == "--spawn") spawn = atoi(val("--spawn"));
    else if (a == "--rows") rows = atoll(val("--rows"));
    else if (a == "--k") k = atoi(val("--k"));
    else if (a == "--depth") depth = atoi(val("--depth"));
    else if (a == "--devices") ndev = atoi(val("--devices"));
    else if (a == "--strip") strip = atoi(val("--strip"));
    else if (a == "--density") density = atof(val("--density"));
    else if (a == "--seed") seed = strtoull(val("--seed"), nullptr, 10);
    else if (a == "--outdir") outdir = val("--outdir");
    else if (a == "--cpu") cpu = 1;
    else if (a == "--check") check = 1;
    else if (a == "--json") json = 1;
    else if (a == "--mode") {
      const std::string m = val("--mode");
      if (m == "auto") mode = GOL_MODE_AUTO;
      else if (m == "generic") mode = GOL_MODE_GENERIC;
      else if (m == "fused") mode = GOL_MODE_FUSED;
      else if (m == "persistent") mode = GOL_MODE_PERSISTENT;
      else if (m == "tma") mode = GOL_MODE_TMA;
      else if (m == "resident") mode = GOL_MODE_RESIDENT;
      else {
        fprintf(stderr, "gol: unknown mode %s\n", m.c_str());
        return 255;
      }
    } else if (a.size() > 2 && a[0] == '-' && a[1] == '-') {
      fprintf(stderr, "gol: unknown option %s\n", a.c_str());
      return 255;
    } else {
      pos.push_back(argv[i]);
    }
  }
  if (pos.size() != 5) {  // gol-main.c:43-47
    printf("%s", kUsage);
    return 255;
  }
  const unsigned pattern = (unsigned)atoi(pos[0]);
  const long long N = atoll(pos[1]);
  const long long iters = atoll(pos[2]);
  const int threads = atoi(pos[3]);
  const int on_off = atoi(pos[4]);
  if (pattern >= 7) {  // gol-with-cuda.cu:324-326
    printf("Pattern %u has not been implemented \n", pattern);
    return 255;
  }
  if (N < 1 || iters < 0 || ranks < 1) {
    fprintf(stderr, "gol: world size, iterations and ranks must be positive\n");
    return 255;
  }
  if (spawn < 0 || spawn > 64 || (spawn && cpu)) {
    fprintf(stderr, "gol: --spawn takes 1..64 ranks and needs the CUDA backend\n");
    return 255;
  }

  // ---- process per rank: fork before anything touches CUDA --------------------------------
  int my_rank = 0, local_sense = 0;
  if (spawn) {
    ranks = spawn;
    Rendezvous* rv = (Rendezvous*)mmap(nullptr, sizeof(Rendezvous), PROT_READ | PROT_WRITE,
                                       MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (rv == MAP_FAILED) {
      perror("gol: mmap");
      return 255;
    }
    memset(rv, 0, sizeof *rv);
    rv->nranks = spawn;
    std::vector<pid_t> kids;
    bool child = false;
    for (int r = 0; r < spawn; ++r) {
      fflush(stdout);
      const pid_t pid = fork();
      if (pid < 0) {
        perror("gol: fork");
        __atomic_store_n(&rv->failed, 1, __ATOMIC_RELEASE);
        break;
      }
      if (pid == 0) {
        child = true;
        my_rank = r;
        g_rv = rv;
        break;
      }
      kids.push_back(pid);
    }
    if (!child) {
      int worst = 0;
      for (pid_t pid : kids) {
        int st = 0;
        waitpid(pid, &st, 0);
        const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 255;
        if (code) __atomic_store_n(&rv->failed, 1, __ATOMIC_RELEASE);
        worst = code > worst ? code : worst;
      }
      return worst;
    }
  }
  const int first_rank = spawn ? my_rank : 0;
  const int nlocal = spawn ? 1 : ranks;

  gol_config_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cfg.width = N;
  cfg.rows = rows > 0 ? rows : N;  // the reference gives every rank an N x N square (gol-main.c:76)
  cfg.nranks = ranks;
  cfg.first_rank = first_rank;
  cfg.nlocal = nlocal;
  cfg.backend = cpu ? GOL_BACKEND_CPU : GOL_BACKEND_CUDA;
  cfg.mode = mode;
  cfg.temporal_k = k;
  cfg.exchange_depth = depth;
  cfg.threads_per_block = threads;
  cfg.ndevices = ndev;
  cfg.device_base = spawn ? my_rank : 0;   // rank -> device like gol-with-cuda.cu:296
  cfg.reserved = strip;

  if (on_off == 1) {
    // like the reference, open every rank's output file before any work (gol-main.c:64-73)
    for (int r = first_rank; r < first_rank + nlocal; ++r) {
      char name[64];
      gol_rank_file_name(name, sizeof name, r, ranks);
      const std::string path = outdir ? std::string(outdir) + "/" + name : std::string(name);
      FILE* fp = fopen(path.c_str(), "w");
      if (!fp) {
        printf("ERROR IN RANK %d", r);
        return 255;
      }
      fclose(fp);
    }
  }

  if (spawn && gol_cuda_device_count() < ranks) {
    // two ranks on one GPU would be two kernels spinning on each other's flags with nothing to
    // guarantee that both are resident: a hang, not a slowdown
    if (my_rank == 0)
      fprintf(stderr, "gol: --spawn %d needs one GPU per rank, found %d (use --ranks %d --devices N to host "
                      "several ranks per GPU in one process)\n", ranks, gol_cuda_device_count(), ranks);
    __atomic_store_n(&g_rv->failed, 1, __ATOMIC_RELEASE);
    _exit(255);
  }
  gol_world_t* w = gol_world_create(&cfg);
  if (!w) die("cannot create world");
  if (spawn) {
    // every rank publishes its export blob, then maps its two ring neighbours
    if (gol_world_export(w, 0, g_rv->blobs[my_rank])) die("export");
    rv_barrier(g_rv, &local_sense);
    const int prev = (my_rank - 1 + ranks) % ranks, next = (my_rank + 1) % ranks;
    if (prev != my_rank && gol_world_import(w, prev, g_rv->blobs[prev])) die("import");
    if (next != my_rank && next != prev && gol_world_import(w, next, g_rv->blobs[next])) die("import");
  }
  if (gol_world_connect(w)) die("connect");
  if (gol_world_init_pattern(w, (int)pattern, seed, density)) die("init");
  if (gol_world_sync(w)) die("init sync");
  if (spawn) rv_barrier(g_rv, &local_sense);   // nobody steps before every neighbour is mapped and initialised

  std::vector<float> ms((size_t)nlocal, 0.f);
  if (gol_world_step_timed(w, iters, ms.data())) die("step");
  double seconds = *std::max_element(ms.begin(), ms.end()) * 1e-3;  // device time, max over ranks
  if (spawn) {
    g_rv->ms[my_rank] = ms[0];
    rv_barrier(g_rv, &local_sense);
    for (int r = 0; r < ranks; ++r) seconds = std::max(seconds, (double)g_rv->ms[r] * 1e-3);
  }
  const bool speaker = !spawn || my_rank == 0;   // rank 0 reports, like gol-main.c:121-133

  char report[512];
  gol_format_report(report, sizeof report, seconds, ranks, cfg.width, cfg.rows, iters);
  if (speaker) fputs(report, stdout);  // gol-main.c:124-125,132

  gol_plan_t plan;
  gol_world_plan(w, &plan);
  if (json) {
    const double updates = (double)ranks * (double)cfg.rows * (double)cfg.width * (double)iters;
    // per-phase device counters, summed over the local ranks (exposed halo wait = time edge warps
    // spent spinning on a neighbour's flag)
    gol_phase_counters_t pc, sum;
    memset(&sum, 0, sizeof sum);
    if (spawn) {
      if (gol_world_phase_counters(w, 0, &g_rv->phase[my_rank], 0)) die("phase counters");
      rv_barrier(g_rv, &local_sense);
    }
    for (int r = 0; r < ranks; ++r) {
      if (spawn) pc = g_rv->phase[r];
      else if (cpu || gol_world_phase_counters(w, r, &pc, 0)) break;
      for (int i = 0; i < 2; ++i) {
        sum.halo_wait_ns[i] += pc.halo_wait_ns[i];
        sum.halo_waits[i] += pc.halo_waits[i];
        sum.task_ns[i] += pc.task_ns[i];
        sum.tasks[i] += pc.tasks[i];
      }
      sum.dep_wait_ns += pc.dep_wait_ns;
    }
    if (speaker)
    printf("{\"gcups\": %.3f, \"seconds\": %.6f, \"mode\": %d, \"k\": %d, \"depth\": %d, "
           "\"threads\": %d, \"strip_rows\": %d, \"launches\": %llu, "
           "\"halo_wait_us\": [%.1f, %.1f], \"halo_waits\": [%llu, %llu], \"dep_wait_us\": %.1f, "
           "\"task_us\": [%.1f, %.1f], \"tasks\": [%llu, %llu]}\n",
           seconds > 0 ? updates / seconds * 1e-9 : 0.0, seconds, plan.mode, plan.temporal_k,
           plan.exchange_depth, plan.threads_per_block, plan.strip_rows,
           (unsigned long long)gol_launch_count(), sum.halo_wait_ns[0] * 1e-3, sum.halo_wait_ns[1] * 1e-3,
           (unsigned long long)sum.halo_waits[0], (unsigned long long)sum.halo_waits[1], sum.dep_wait_ns * 1e-3,
           sum.task_ns[0] * 1e-3, sum.task_ns[1] * 1e-3, (unsigned long long)sum.tasks[0],
           (unsigned long long)sum.tasks[1]);
  }

  int status = 0;
  if (check) {
    // whole-torus oracle comparison (small worlds only)
    const long long W = cfg.width, H = cfg.rows, GH = H * ranks;
    std::vector<uint8_t> init((size_t)(W * GH)), want((size_t)(W * GH)), got((size_t)(W * GH));
    for (int r = 0; r < ranks; ++r)
      gol_pattern_fill_bytes(init.data() + (size_t)r * W * H, (int)pattern, W, H, r, ranks, seed, density);
    gol_oracle_run(init.data(), want.data(), W, GH, iters);
    size_t bad = 0, total = 0;
    for (int r = first_rank; r < first_rank + nlocal; ++r) {
      uint8_t* mine = got.data() + (size_t)r * W * H;
      if (gol_world_read_bytes(w, r - first_rank, mine)) die("read");
      for (size_t i = 0; i < (size_t)(W * H); ++i) bad += mine[i] != want[(size_t)r * W * H + i];
      total += (size_t)(W * H);
    }
    if (spawn) {
      g_rv->bad[my_rank] = bad;
      g_rv->cells[my_rank] = total;
      rv_barrier(g_rv, &local_sense);
      bad = total = 0;
      for (int r = 0; r < ranks; ++r) {
        bad += g_rv->bad[r];
        total += g_rv->cells[r];
      }
    }
    if (speaker) printf("CHECK: %s (%zu of %zu cells differ from the oracle)\n", bad ? "FAIL" : "OK", bad, total);
    status = bad ? 1 : 0;
  }

  if (on_off == 1) {  // gol-main.c:135-139
    for (int r = 0; r < nlocal; ++r)
      if (gol_world_dump_text(w, r, outdir)) die("dump");
  }
  if (spawn) rv_barrier(g_rv, &local_sense);   // peers keep their slabs mapped until everyone is done
  gol_world_destroy(w);
  fflush(stdout);
  if (spawn) _exit(status);
  return status;
}
