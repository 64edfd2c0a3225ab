// cli_main.cpp — the `gol` command: same five positionals, messages, exit codes
// and output files as the reference driver (gol-main.c:30-146), on top of the
// B200 core.  The reference gets its rank count from an MPI launcher; this
// binary drives all ranks from one process (one rank per GPU, peer-to-peer over
// NVLink) and takes the count from --ranks (default: 1).  The process-per-GPU
// form of the same CLI is `torchrun ... -m gol_b200.cli`.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../include/gol_b200.h"

static const char* kUsage =
    "GOL requires 5 arguments: pattern number, sq size of the world and the number of itterations, "
    "threads per block and output-on-off e.g. ./gol 0 32 2 512 0 \n";

static void die(const char* what) {
  fprintf(stderr, "gol: %s: %s\n", what, gol_last_error());
  exit(255);
}

int main(int argc, char** argv) {
  std::vector<const char*> pos;
  int ranks = 1, mode = GOL_MODE_AUTO, k = 0, depth = 0, cpu = 0, check = 0, json = 0, ndev = 0, strip = 0;
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

  gol_config_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cfg.width = N;
  cfg.rows = rows > 0 ? rows : N;  // the reference gives every rank an N x N square (gol-main.c:76)
  cfg.nranks = ranks;
  cfg.first_rank = 0;
  cfg.nlocal = ranks;
  cfg.backend = cpu ? GOL_BACKEND_CPU : GOL_BACKEND_CUDA;
  cfg.mode = mode;
  cfg.temporal_k = k;
  cfg.exchange_depth = depth;
  cfg.threads_per_block = threads;
  cfg.ndevices = ndev;
  cfg.reserved = strip;

  if (on_off == 1) {
    // like the reference, open every rank's output file before any work (gol-main.c:64-73)
    for (int r = 0; r < ranks; ++r) {
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

  gol_world_t* w = gol_world_create(&cfg);
  if (!w) die("cannot create world");
  if (gol_world_connect(w)) die("connect");
  if (gol_world_init_pattern(w, (int)pattern, seed, density)) die("init");
  if (gol_world_sync(w)) die("init sync");

  std::vector<float> ms((size_t)ranks, 0.f);
  if (gol_world_step_timed(w, iters, ms.data())) die("step");
  const double seconds = *std::max_element(ms.begin(), ms.end()) * 1e-3;  // device time, max over ranks

  char report[512];
  gol_format_report(report, sizeof report, seconds, ranks, cfg.width, cfg.rows, iters);
  fputs(report, stdout);  // gol-main.c:124-125,132

  gol_plan_t plan;
  gol_world_plan(w, &plan);
  if (json) {
    const double updates = (double)ranks * (double)cfg.rows * (double)cfg.width * (double)iters;
    // per-phase device counters, summed over the local ranks (exposed halo wait = time edge warps
    // spent spinning on a neighbour's flag)
    gol_phase_counters_t pc, sum;
    memset(&sum, 0, sizeof sum);
    for (int r = 0; r < ranks; ++r) {
      if (cpu || gol_world_phase_counters(w, r, &pc, 0)) break;
      for (int i = 0; i < 2; ++i) {
        sum.halo_wait_ns[i] += pc.halo_wait_ns[i];
        sum.halo_waits[i] += pc.halo_waits[i];
        sum.task_ns[i] += pc.task_ns[i];
        sum.tasks[i] += pc.tasks[i];
      }
      sum.dep_wait_ns += pc.dep_wait_ns;
    }
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
    for (int r = 0; r < ranks; ++r)
      if (gol_world_read_bytes(w, r, got.data() + (size_t)r * W * H)) die("read");
    size_t bad = 0;
    for (size_t i = 0; i < got.size(); ++i) bad += got[i] != want[i];
    printf("CHECK: %s (%zu of %zu cells differ from the oracle)\n", bad ? "FAIL" : "OK", bad, got.size());
    status = bad ? 1 : 0;
  }

  if (on_off == 1) {  // gol-main.c:135-139
    for (int r = 0; r < ranks; ++r)
      if (gol_world_dump_text(w, r, outdir)) die("dump");
  }
  gol_world_destroy(w);
  return status;
}
