// world.cu — the slab/ring runtime behind gol_world_*.
//
// Replaces the reference's global-variable "API" (gol-main.c:11-13) and its
// host-side launch wrapper (gol-with-cuda.cu:264-284) with an explicit world
// object: per-rank slabs (double-buffered, bit-packed), recv buffers and flags
// the ring neighbours write into directly, a pass schedule that is identical
// on every rank, CUDA graphs for the launch-bound inner loop, and device-timed
// stepping.  A CPU backend shares the ring/ghost logic so that rank routing
// (R = 1, 2, >= 3) is testable without a GPU.
#include <cuda_runtime.h>
#include <errno.h>
#include <stddef.h>
#include <nvtx3/nvToolsExt.h>  // header-only; ranges show up in Nsight when a tool is attached
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "../include/gol_b200.h"
#include "internal.h"
#include "kernels.cuh"

namespace gol {
namespace {

#define CU_TRY(expr)                                                                     \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess)                                                              \
      return set_error(GOL_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
                       __FILE__, __LINE__);                                              \
  } while (0)

constexpr uint32_t kBlobMagic = 0x474F4C42u;  // "GOLB"
constexpr int kCanaryByte = 0xA5;

struct ExportBlob {
  uint32_t magic;
  int32_t rank, device, pid;
  int64_t W, H, stride;
  int32_t G, pad;
  uint64_t off_flags;
  uint64_t off_recv[2][2];
  uint64_t off_ll[2][2];
  uint64_t base_addr;
  cudaIpcMemHandle_t handle;
};
static_assert(sizeof(ExportBlob) <= GOL_EXPORT_BYTES, "export blob too large");

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Slab {
  int rank = 0, device = 0, group = 0, index_in_group = 0;
  char* base = nullptr;    // mailbox: flags + recv rows + flag-in-data rows (the only memory peers touch; IPC-exported)
  size_t bytes = 0;
  char* grid = nullptr;    // both slab buffers (never touched by a peer: kept out of the exported allocation)
  size_t grid_bytes = 0;
  std::vector<std::pair<char*, size_t>> canaries;   // guard rows between / around every buffer
  size_t off_flags = 0, off_recv[2][2] = {{0, 0}, {0, 0}}, off_ll[2][2] = {{0, 0}, {0, 0}}, off_buf[2] = {0, 0};
  SlabState h{};
  unsigned long long* d_partials = nullptr;
  unsigned long long* d_pop = nullptr;
  // end-to-end pipeline (gol_world_run_from_host_async): two slots of device staging
  uint64_t* stage_in[2] = {nullptr, nullptr};
  uint64_t* stage_out[2] = {nullptr, nullptr};
  cudaEvent_t ev_h2d[2] = {nullptr, nullptr};      // upload into stage_in[k] finished
  cudaEvent_t ev_in_used[2] = {nullptr, nullptr};  // stage_in[k] copied into the slab
  cudaEvent_t ev_out[2] = {nullptr, nullptr};      // result copied into stage_out[k]
  cudaEvent_t ev_done[2] = {nullptr, nullptr};     // stage_out[k] landed in host memory
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaEvent_t tr[2][6] = {{nullptr}};   // GOL_PIPE_TRACE: h2d begin/end, compute begin/end, d2h begin/end per slot
  // CPU backend
  std::vector<uint64_t> cbuf[2];
  std::vector<uint64_t> cghost[2];
  int ccur = 0;
  // CPU executor of the GPU pass schedule (GOL_FLAG_CPU_SCHEDULE)
  std::vector<uint64_t> ebuf[2];        // (H + 2G) rows with margins, like the device slab
  std::vector<uint64_t> erecv[2][2];    // [side][exchange parity], G rows each
  uint64_t epass = 0, eexch = 0;
};

struct PassItem {
  bool prime = false;
  bool generic = false;
  int K = 0;
  int depth = 0;  // prime depth
  PassParams p{};
};

// One launch thread per device: a 257-node graph launch costs the host ~0.2 ms, and eight of them
// issued one after the other from a single thread start the last GPU more than a millisecond
// after the first - its ring neighbours then wait for it in every pass (171 instead of 235 TCUPS
// on 8 GPUs from one process in round 1).  The reference gets the same effect from one MPI
// process per rank (gol-main.c:58-62).
struct Launcher {
  std::thread th;
  std::mutex m;
  std::condition_variable cv;
  std::function<cudaError_t()> job;
  bool has_job = false, quit = false;
  cudaError_t rc = cudaSuccess;

  void run() {
    std::unique_lock<std::mutex> lk(m);
    for (;;) {
      cv.wait(lk, [&] { return has_job || quit; });
      if (quit) return;
      std::function<cudaError_t()> j = job;
      lk.unlock();
      const cudaError_t r = j();
      lk.lock();
      rc = r;
      has_job = false;
      cv.notify_all();
    }
  }
  void post(std::function<cudaError_t()> j) {
    std::unique_lock<std::mutex> lk(m);
    if (!th.joinable()) th = std::thread([this] { run(); });
    cv.wait(lk, [&] { return !has_job; });
    job = std::move(j);
    has_job = true;
    cv.notify_all();
  }
  cudaError_t wait() {
    std::unique_lock<std::mutex> lk(m);
    cv.wait(lk, [&] { return !has_job; });
    return rc;
  }
  void stop() {
    {
      std::unique_lock<std::mutex> lk(m);
      if (!th.joinable()) return;
      cv.wait(lk, [&] { return !has_job; });
      quit = true;
      cv.notify_all();
    }
    th.join();
  }
};

struct Group {
  int device = 0;
  Launcher* launcher = nullptr;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  std::vector<int> slabs;
  SlabState* d_states = nullptr;
  unsigned int* d_barrier = nullptr;   // flow kernel control words: [0] ticket, [1] warps finished, [2] abort
  unsigned int* d_progress = nullptr;  // flow kernel: [slabs][flow_strips] tasks completed per strip
  uint64_t* d_mail = nullptr;          // resident kernel: CTA-to-CTA halo mailboxes
  CUtensorMap* d_tmaps = nullptr;  // [slabs][2], TMA mode only
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  std::map<int64_t, std::pair<cudaGraphExec_t, int>> graphs;  // gens -> (exec, kernels)
  void* l2_policy_stream = nullptr;  // stream the L2 access-policy window was last applied to
  bool l2_persist_held = false;      // this group holds a share of the device's persisting-L2 carve-out
};

struct Peer {
  bool imported = false;
  ExportBlob blob{};
  char* mapped = nullptr;
  bool ipc_opened = false;
  int opened_on = -1;
};

}  // namespace
}  // namespace gol

using namespace gol;

struct gol_world {
  gol_config_t cfg{};
  gol_plan_t plan{};
  int64_t W = 0, H = 0, ws = 0, stride = 0;
  int G = 0;
  bool cuda = false;
  bool connected = false;
  std::vector<Slab> slabs;
  std::vector<Group> groups;
  std::vector<Peer> peers;
  uint64_t pass_host = 0;  // passes enqueued so far (buffer parity of every slab)
  int resident_warps = 0, sms = 0;
  int flow_strips = 0;     // strips per slab of the flow kernel (fixed for the world)
  int flow_owner = 0;      // flow kernel: fixed owners (small slabs) instead of tickets
  bool auto_no_resident = false;  // AUTO tried the resident kernel and it did not fit: plan again without it
  int pipe_slot = 0;       // next slot of the end-to-end pipeline
};

namespace gol {
namespace {

int pick_strip_rows(int64_t rows_total, int col_groups, int K, int slots, int sms) {
  // One warp task = strip_rows output rows of one column group and costs strip_rows + 2K
  // row steps (pipeline fill).  With w tasks per scheduler the ALU pipe is shared w ways but
  // is also better utilised; measured on B200 (16384^2 and 4096^2 sweeps, profiles/) the time
  // per row step behaves like (w + 0.9), and more than `slots` tasks run in waves.
  if (slots < 1) slots = 1;
  if (sms < 1) sms = 1;
  const double wmax = std::max(1.0, (double)slots / (4.0 * sms));
  int best = 0;
  double best_cost = 1e300;
  const int lo = std::max(K, 4);
  const int hi = (int)std::min<int64_t>(rows_total, 2048);
  for (int rh = lo; rh <= std::max(lo, hi); ++rh) {
    const int64_t strips = (rows_total + rh - 1) / rh;
    const int64_t tasks = (strips + 2) * col_groups;   // + the two edge strips
    const double w = (double)tasks / (4.0 * sms);
    const double waves = std::max(1.0, ceil(w / wmax - 1e-9));
    const double cost = waves * ((double)rh + 2.0 * K + 2.0) * (w / waves + 0.9) * (1.0 + 0.02 * (waves - 1.0));
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = rh;
    }
  }
  return best > 0 ? best : lo;
}

int largest_k_le(int limit) {
  static const int ks[] = {8, 6, 4, 3, 2, 1};
  for (int k : ks)
    if (k <= limit) return k;
  return 1;
}

int make_plan(gol_world* w) {
  const gol_config_t& c = w->cfg;
  gol_plan_t& p = w->plan;
  memset(&p, 0, sizeof p);
  p.words_per_row = w->ws;
  p.row_stride = w->stride;
  int mode = c.mode;
  if (mode < GOL_MODE_AUTO || mode > GOL_MODE_RESIDENT)
    return set_error(GOL_ERR_ARG, "unknown mode %d", mode);
  int K = c.temporal_k > 0 ? c.temporal_k : 4;
  int depth = c.exchange_depth > 0 ? c.exchange_depth : K;
  const bool shape_ok = (w->W % 64) == 0;
  if (!w->cuda && !(c.flags & GOL_FLAG_CPU_SCHEDULE)) {
    p.mode = GOL_MODE_GENERIC;
    p.temporal_k = 1;
    p.exchange_depth = 1;
    p.threads_per_block = c.threads_per_block;
    p.slab_rows = w->H;
    return GOL_OK;
  }
  if (mode == GOL_MODE_AUTO) {
    if (shape_ok) {
      while (depth > w->H && depth > 1) depth /= 2;
      if (K > depth) K = largest_k_le(depth);
      mode = GOL_MODE_FUSED;
      // Small slabs (<= 8 MiB packed, tall enough to give every SM a strip) are bound by the
      // hand-off between passes, not by throughput: the shared-memory resident kernel steps a
      // 4096^2 slab in 4.9 instead of 8.5 ms per 4096 generations and an 8192^2 slab at 25 instead
      // of 16.6 TCUPS (profiles/r2/).  A slab whose strips do not fit shared memory falls back.
      if (w->cuda && !w->auto_no_resident && w->H >= 1024 && w->H * w->ws * 8 <= (8ll << 20) &&
          (c.exchange_depth <= 0 || c.exchange_depth == K))
        mode = GOL_MODE_RESIDENT;
    } else {
      mode = GOL_MODE_GENERIC;
    }
  }
  if (mode == GOL_MODE_GENERIC) {
    int t = c.threads_per_block > 0 ? c.threads_per_block : 256;
    if (t < 1 || t > 1024)
      return set_error(GOL_ERR_ARG, "threads per block must be in [1,1024], got %d", t);
    p.mode = mode;
    p.temporal_k = 1;
    p.exchange_depth = 1;
    p.threads_per_block = t;
    p.slab_rows = w->H + 2 * w->G;
    return GOL_OK;
  }
  // fused / persistent / resident
  if (!shape_ok)
    return set_error(GOL_ERR_UNSUPPORTED,
                     "fused kernels need a width that is a multiple of 64 (got %lld); use mode generic",
                     (long long)w->W);
  if (mode == GOL_MODE_RESIDENT) {
    // K = generations per halo exchange between strips (any 1..GOL_MAX_GHOST); the exchange depth is K
    if (K < 1 || K > GOL_MAX_GHOST || K > w->H)
      return set_error(GOL_ERR_ARG, "mode resident: temporal_k must be in [1, min(%d, rows)] (got %d)", GOL_MAX_GHOST, K);
    if (c.exchange_depth > 0 && c.exchange_depth != K)
      return set_error(GOL_ERR_ARG, "mode resident exchanges halos every temporal_k generations: exchange_depth must "
                       "be 0 or %d (got %d)", K, c.exchange_depth);
    int t = c.threads_per_block > 0 ? c.threads_per_block : 256;
    if (t < 1 || t > 1024)
      return set_error(GOL_ERR_ARG, "threads per block must be in [1,1024], got %d", t);
    t = std::min(512, std::max(32, (t / 32) * 32));
    p.mode = mode;
    p.temporal_k = K;
    p.exchange_depth = K;
    p.threads_per_block = t;
    p.col_groups = (int)((w->ws + 31) / 32);   // 32-word column blocks, no halo lanes
    p.slab_rows = w->H + 2 * w->G;
    return GOL_OK;
  }
  if (!fused_k_supported(K))
    return set_error(GOL_ERR_ARG, "temporal_k must be one of 1,2,3,4,6,8 (got %d)", K);
  if (depth % K != 0 || depth > GOL_MAX_GHOST || depth < K)
    return set_error(GOL_ERR_ARG, "exchange_depth must be a multiple of temporal_k and <= %d (got %d, k=%d)",
                     GOL_MAX_GHOST, depth, K);
  if (depth > w->H)
    return set_error(GOL_ERR_ARG, "exchange_depth %d exceeds the slab height %lld", depth, (long long)w->H);
  int t = c.threads_per_block > 0 ? c.threads_per_block : 256;
  if (t < 1 || t > 1024)
    return set_error(GOL_ERR_ARG, "threads per block must be in [1,1024], got %d", t);
  t = std::max(32, (t / 32) * 32);  // whole warps
  const int lim = fused_limits(K).max_threads;
  if (t > lim) t = lim;             // register budget of the K-generation pipeline
  if (mode == GOL_MODE_TMA) {
    if (!fused_tma_supported(K, w->ws))
      return set_error(GOL_ERR_UNSUPPORTED,
                       "mode tma needs temporal_k <= 4 and a width that is a multiple of 1024 and >= 9216 "
                       "(128-byte aligned 144-word row segments); got k=%d width=%lld", K, (long long)w->W);
    t = 32 * fused_tma_warps((int)((w->ws + 29) / 30));   // CTA tile = 3 or 4 column groups
  }
  p.mode = mode;
  p.temporal_k = K;
  p.exchange_depth = depth;
  p.threads_per_block = t;
  p.col_groups = (int)((w->ws + 29) / 30);
  p.slab_rows = w->H + 2 * w->G;
  return GOL_OK;
}

// Resident mode: how many strips (CTAs) per slab, in clusters of which size.
int finish_plan_resident(gol_world* w) {
  gol_plan_t& p = w->plan;
  const int dev = w->groups.empty() ? 0 : w->groups[0].device;
  CU_TRY(cudaSetDevice(dev));
  CU_TRY(cudaDeviceGetAttribute(&w->sms, cudaDevAttrMultiProcessorCount, dev));
  int max_smem = 0;
  CU_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  size_t slabs_per_dev = 1;
  for (auto& g : w->groups) slabs_per_dev = std::max(slabs_per_dev, g.slabs.size());
  const int K = p.temporal_k;
  int forced = 0;
  if (const char* e = getenv("GOL_RESIDENT_CLUSTER")) forced = atoi(e);
  // Cluster size: pairs by default.  148 SMs pack perfectly into clusters of 2, and the DSMEM
  // hand-off inside a pair is a little faster than the L2 mailbox (5.34 vs 5.43 ms per 4096
  // generations at 4096^2).  Clusters of 4 and 8 cannot cover every SM of the 16 / 18 / 20-SM GPCs
  // with one CTA each, so some SMs host two strips and the step takes 8 ms: GOL_RESIDENT_CLUSTER
  // selects them for experiments only.
  int best_c = 0, best_cs = 1;
  size_t smem = 0;
  int rows = 0;
  for (int cs : {8, 4, 2, 1}) {
    if (forced > 0 ? cs != forced : cs > 2) continue;
    // one CTA per SM: a second resident CTA would halve the strips and double the halo rows recomputed
    const int cap = std::min(resident_max_ctas(p.threads_per_block, 64 * 1024, cs, dev), w->sms / cs * cs);
    int64_t c = std::min<int64_t>(cap / (int64_t)slabs_per_dev, w->H / K);   // every strip holds at least K rows
    if (w->cfg.reserved > 0) c = std::min<int64_t>(c, (w->H + w->cfg.reserved - 1) / w->cfg.reserved);
    c = c / cs * cs;
    if (c < 1) continue;
    // the launch is cooperative: check co-residency with the shared memory the strips really need
    const int r = (int)((w->H + c - 1) / c);
    const size_t sm = resident_smem_bytes(r + 2 * K, w->ws, K);
    if (sm + 1024 > (size_t)max_smem) {
      if (best_c == 0) {
        rows = r;
        smem = sm;
      }
      continue;
    }
    if (resident_max_ctas(p.threads_per_block, sm, cs, dev) < c * (int64_t)slabs_per_dev) continue;
    if (c > best_c) {   // larger clusters are tried first and kept on ties
      best_c = (int)c;
      best_cs = cs;
      rows = r;
      smem = sm;
    }
  }
  if (best_c < 1) {
    if (smem)
      return set_error(GOL_ERR_UNSUPPORTED,
                       "mode resident keeps %d rows x %lld words per CTA in shared memory (%zu bytes, limit %d): the slab "
                       "is too large, use mode fused or persistent", rows + 2 * K, (long long)w->ws, smem, max_smem);
    return set_error(GOL_ERR_UNSUPPORTED, "mode resident: no co-resident launch shape for %zu slab(s) per device",
                     slabs_per_dev);
  }
  p.strip_rows = rows;
  p.ctas_per_slab = best_c;
  p.regs_hint = best_cs;   // cluster size
  for (auto& g : w->groups) {
    CU_TRY(cudaSetDevice(g.device));
    const size_t n = g.slabs.size() * (size_t)best_c * 4 * (size_t)(2 * K * w->ws) * 8;
    CU_TRY(cudaMalloc((void**)&g.d_mail, n));
    CU_TRY(cudaMemset(g.d_mail, 0, n));
    debug_bounds_register(g.d_mail, n);
  }
  return GOL_OK;
}

int finish_plan_cuda(gol_world* w) {
  gol_plan_t& p = w->plan;
  if (p.mode == GOL_MODE_RESIDENT) return finish_plan_resident(w);
  if (p.mode != GOL_MODE_FUSED && p.mode != GOL_MODE_PERSISTENT && p.mode != GOL_MODE_TMA) return GOL_OK;
  const int dev = w->groups.empty() ? 0 : w->groups[0].device;
  CU_TRY(cudaSetDevice(dev));
  CU_TRY(cudaDeviceGetAttribute(&w->sms, cudaDevAttrMultiProcessorCount, dev));
  int rw = fused_resident_warps_per_sm(p.temporal_k, p.threads_per_block, dev);
  if (rw < 1) rw = 8;
  w->resident_warps = rw;
  const int slabs_per_dev = (int)w->groups[0].slabs.size();
  int slots = w->sms * rw / std::max(1, slabs_per_dev);
  if (p.mode == GOL_MODE_PERSISTENT) {
    const int wpb = p.threads_per_block / 32;
    const int ctas = flow_max_ctas(p.temporal_k, p.threads_per_block, dev);
    if (ctas > 0) slots = ctas * wpb / std::max(1, slabs_per_dev);
  }
  // the two edge strips (exchange_depth rows each) are tasks of their own; the interior is tuned
  const int64_t interior = std::max<int64_t>(1, w->H - 2 * p.exchange_depth);
  // in TMA mode a CTA always carries a fixed number (3 or 4) of column groups
  const int tw = fused_tma_warps(p.col_groups);
  const int cg_eff = p.mode == GOL_MODE_TMA ? (p.col_groups + tw - 1) / tw * tw : p.col_groups;
  p.strip_rows = pick_strip_rows(interior, cg_eff, p.temporal_k, slots, w->sms);
  if (p.mode == GOL_MODE_PERSISTENT) {
    if (p.strip_rows < 8 * p.temporal_k && w->cfg.threads_per_block <= 0) {
      // Small slab (one wave of resident warps already means strips of a few dozen rows): the
      // run is bound by the per-pass hand-off between neighbouring strips, not by throughput.
      // One warp per scheduler that owns its strip for the whole launch did best on a 4096^2
      // slab (8.2 ms per 4096 generations; tickets 9.9 ms, graph of per-pass launches 9.7 ms).
      w->flow_owner = 1;
      p.threads_per_block = 128;
      const int64_t warps = (int64_t)w->sms * 4 / std::max(1, slabs_per_dev);
      const int64_t strips = std::max<int64_t>(1, warps / cg_eff - 2);
      p.strip_rows = (int)std::max<int64_t>(2 * p.temporal_k, (interior + strips - 1) / strips);
    } else {
      // The flow kernel hands tickets to free warps: it wants about three tasks per resident warp
      // and pass so that a free warp always finds one whose neighbours are done, but not strips so
      // short that the 2K pipeline-fill rows dominate (measured on B200: 36 rows at 16384^2, 64 at
      // 65536 x 8192, 128 at 131072 x 16384, 256 at 65536^2; profiles/flow_kernel.md).
      const int64_t per_warp = interior * cg_eff / (3 * (int64_t)std::max(1, slots));
      p.strip_rows = (int)std::max<int64_t>(std::max(9 * p.temporal_k, p.exchange_depth), (per_warp + 3) / 4 * 4);
    }
    if (p.strip_rows > interior) p.strip_rows = (int)interior;
  }
  if (w->cfg.reserved > 0) p.strip_rows = w->cfg.reserved;   // explicit strip height (tuning / experiments)
  const int wpb = p.threads_per_block / 32;
  const int64_t strips = strip_count(w->H, p.strip_rows, p.exchange_depth);
  p.ctas_per_slab = (int)((strips * p.col_groups + wpb - 1) / wpb);
  p.regs_hint = rw;
  if (p.mode == GOL_MODE_PERSISTENT) {
    // bookkeeping of the flow kernel: one progress counter per strip and slab
    w->flow_strips = flow_strip_count(w->H, p.exchange_depth, p.strip_rows);
    for (auto& g : w->groups) {
      CU_TRY(cudaSetDevice(g.device));
      const size_t n = g.slabs.size() * (size_t)w->flow_strips * sizeof(unsigned int);
      CU_TRY(cudaMalloc((void**)&g.d_progress, n));
      CU_TRY(cudaMemset(g.d_progress, 0, n));
      debug_bounds_register(g.d_progress, n);
    }
  }
  return GOL_OK;
}

// Tensor maps (TMA descriptors) over both buffers of every slab: uint64 elements, dims
// {ws, H + 2G}, row pitch = stride words, box = 144 words x 8 rows, no swizzle, zero fill
// outside the tensor.  cuTensorMapEncodeTiled is a driver entry point; it is looked up through
// the runtime so that the library does not link libcuda.
int make_tensor_maps(gol_world* w) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                               const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                               CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CU_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  if (!fn || q != cudaDriverEntryPointSuccess)
    return set_error(GOL_ERR_UNSUPPORTED, "driver has no cuTensorMapEncodeTiled");
  EncodeFn encode = (EncodeFn)fn;
  const int box_words = fused_tma_box_words(fused_tma_warps(w->plan.col_groups));
  if (w->H + 2 * w->G < fused_tma_box_rows())
    return set_error(GOL_ERR_UNSUPPORTED, "mode tma needs at least %d buffer rows", fused_tma_box_rows());
  for (auto& g : w->groups) {
    CU_TRY(cudaSetDevice(g.device));
    std::vector<CUtensorMap> maps(g.slabs.size() * 2);
    for (size_t i = 0; i < g.slabs.size(); ++i) {
      const Slab& sl = w->slabs[(size_t)g.slabs[i]];
      for (int b = 0; b < 2; ++b) {
        CUresult r;
        if (fused_tma_rank() == 3) {
          // each row as ws/16 groups of 16 words: the inner box extent is one 128-byte line
          const cuuint64_t dims[3] = {16, (cuuint64_t)(w->ws / 16), (cuuint64_t)(w->H + 2 * w->G)};
          const cuuint64_t strides[2] = {128, (cuuint64_t)w->stride * 8};
          const cuuint32_t box[3] = {16, (cuuint32_t)(box_words / 16), (cuuint32_t)fused_tma_box_rows()};
          const cuuint32_t estr[3] = {1, 1, 1};
          r = encode(&maps[i * 2 + (size_t)b], CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, (void*)sl.h.buf[b], dims, strides,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        } else {
          const cuuint64_t dims[2] = {(cuuint64_t)w->ws, (cuuint64_t)(w->H + 2 * w->G)};
          const cuuint64_t strides[1] = {(cuuint64_t)w->stride * 8};
          const cuuint32_t box[2] = {(cuuint32_t)box_words, (cuuint32_t)fused_tma_box_rows()};
          const cuuint32_t estr[2] = {1, 1};
          r = encode(&maps[i * 2 + (size_t)b], CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, (void*)sl.h.buf[b], dims, strides,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        }
        if (r != CUDA_SUCCESS) return set_error(GOL_ERR_CUDA, "cuTensorMapEncodeTiled failed with %d", (int)r);
      }
    }
    CU_TRY(cudaMalloc((void**)&g.d_tmaps, maps.size() * sizeof(CUtensorMap)));
    debug_bounds_register(g.d_tmaps, maps.size() * sizeof(CUtensorMap));
    CU_TRY(cudaMemcpy(g.d_tmaps, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
  }
  return GOL_OK;
}

void fill_strips(const gol_world* w, PassParams* pp) {
  const int ext_out = pp->ext_in - pp->k;
  const int64_t rows_total = w->H + 2 * ext_out;
  pp->strip_rows = w->plan.strip_rows;
  pp->col_groups = w->plan.col_groups;
  pp->edge_rows = w->plan.exchange_depth;
  pp->nstrips = strip_count(rows_total, pp->strip_rows, pp->edge_rows);
}

// The pass schedule for `gens` generations; identical on every rank.
std::vector<PassItem> build_schedule(const gol_world* w, int64_t gens) {
  std::vector<PassItem> items;
  if (gens <= 0) return items;
  const gol_plan_t& pl = w->plan;
  if (pl.mode == GOL_MODE_GENERIC) {
    PassItem pr;
    pr.prime = true;
    pr.depth = 1;
    items.push_back(pr);
    for (int64_t g = 0; g < gens; ++g) {
      PassItem it;
      it.generic = true;
      it.K = 1;
      it.p.ext_in = 1;
      it.p.k = 1;
      it.p.consume = 1;
      it.p.push = (g + 1 < gens) ? 1 : 0;
      items.push_back(it);
    }
    return items;
  }
  int64_t remaining = gens;
  bool first = true;
  while (remaining > 0) {
    const int d = (int)std::min<int64_t>(pl.exchange_depth, remaining);
    if (first) {
      PassItem pr;
      pr.prime = true;
      pr.depth = d;
      items.push_back(pr);
      first = false;
    }
    int ext = d;
    while (ext > 0) {
      const int k = largest_k_le(std::min(pl.temporal_k, ext));
      PassItem it;
      it.K = k;
      it.p.ext_in = ext;
      it.p.k = k;
      it.p.consume = (ext == d) ? 1 : 0;
      it.p.push = 0;
      ext -= k;
      if (ext == 0) {
        const int64_t left = remaining - d;
        it.p.push = left > 0 ? (int)std::min<int64_t>(pl.exchange_depth, left) : 0;
      }
      fill_strips(w, &it.p);
      items.push_back(it);
    }
    remaining -= d;
  }
  return items;
}

int launch_item(gol_world* w, Group& g, const PassItem& it) {
  const int n = (int)g.slabs.size();
  cudaError_t e;
  if (it.prime) {
    e = launch_prime(g.d_states, n, it.depth, w->ws, g.stream);
  } else if (it.generic) {
    e = launch_generic(g.d_states, n, it.p, w->plan.threads_per_block, w->H * w->ws, g.stream);
  } else if (w->plan.mode == GOL_MODE_TMA && fused_tma_supported(it.K, w->ws)) {
    e = launch_fused_tma(it.K, g.d_states, n, it.p, g.d_tmaps, g.stream);
  } else {
    e = launch_fused(it.K, g.d_states, n, it.p, w->plan.threads_per_block, g.stream);
  }
  if (e != cudaSuccess)
    return set_error(GOL_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
  return GOL_OK;
}

int count_passes(const std::vector<PassItem>& items) {
  int n = 0;
  for (auto& it : items) n += it.prime ? 0 : 1;
  return n;
}

int upload_states(gol_world* w) {
  for (auto& g : w->groups) {
    CU_TRY(cudaSetDevice(g.device));
    std::vector<SlabState> hs;
    for (int si : g.slabs) hs.push_back(w->slabs[(size_t)si].h);
    CU_TRY(cudaMemcpy(g.d_states, hs.data(), hs.size() * sizeof(SlabState), cudaMemcpyHostToDevice));
  }
  return GOL_OK;
}

uint64_t* cur_interior(gol_world* w, Slab& s) {
  return s.h.buf[w->pass_host & 1] + (int64_t)w->G * w->stride;
}

int check_local(gol_world* w, int li) {
  if (!w) return set_error(GOL_ERR_ARG, "null world");
  if (li < 0 || li >= (int)w->slabs.size())
    return set_error(GOL_ERR_ARG, "local index %d out of range (nlocal=%d)", li, (int)w->slabs.size());
  return GOL_OK;
}

// ---- CPU backend ------------------------------------------------------------
void cpu_exchange_local(gol_world* w) {
  // every rank local: ghost above rank r = last row of rank r-1; below = first row of r+1
  const int R = w->cfg.nranks;
  for (auto& s : w->slabs) {
    const Slab& prev = w->slabs[(size_t)((s.rank - 1 + R) % R)];
    const Slab& next = w->slabs[(size_t)((s.rank + 1) % R)];
    const uint64_t* pl = prev.cbuf[prev.ccur].data() + (w->H - 1) * w->stride;
    const uint64_t* nf = next.cbuf[next.ccur].data();
    std::copy(pl, pl + w->stride, s.cghost[0].begin());
    std::copy(nf, nf + w->stride, s.cghost[1].begin());
  }
}

void cpu_step_slab(gol_world* w, Slab& s) {
  const int64_t H = w->H, st = w->stride;
  std::vector<const uint64_t*> in_rows((size_t)H + 2);
  std::vector<uint64_t*> out_rows((size_t)H);
  const uint64_t* cur = s.cbuf[s.ccur].data();
  uint64_t* nxt = s.cbuf[s.ccur ^ 1].data();
  in_rows[0] = s.cghost[0].data();
  for (int64_t y = 0; y < H; ++y) in_rows[(size_t)y + 1] = cur + y * st;
  in_rows[(size_t)H + 1] = s.cghost[1].data();
  for (int64_t y = 0; y < H; ++y) out_rows[(size_t)y] = nxt + y * st;
  cpu_step_rows(in_rows.data(), out_rows.data(), H, w->W);
  s.ccur ^= 1;
}

// ---- CPU executor of the GPU pass schedule -----------------------------------------
// Runs exactly the launch sequence build_schedule() produces for the fused kernels - prime,
// passes of k generations on inputs with ext_in valid halo rows, pushes into the neighbours'
// double-buffered recv rows, in-slab margins for halos deeper than one pass - with the CPU
// row stepper in place of the kernels.  It exists so that the schedule / halo-depth logic is
// checked against the oracle without a GPU (tests/test_cpu_schedule.py).
void cpu_sched_prime(gol_world* w, int depth) {
  const int R = w->cfg.nranks;
  const int64_t st = w->stride, H = w->H;
  for (auto& s : w->slabs) {
    Slab& prev = w->slabs[(size_t)((s.rank - 1 + R) % R)];
    Slab& next = w->slabs[(size_t)((s.rank + 1) % R)];
    const uint64_t* in = s.ebuf[s.epass & 1].data() + (int64_t)w->G * st;
    // my first rows -> prev's bottom recv; my last rows -> next's top recv (parity of the next exchange)
    std::copy(in, in + depth * st, prev.erecv[1][prev.eexch & 1].begin());
    std::copy(in + (H - depth) * st, in + H * st, next.erecv[0][next.eexch & 1].begin());
  }
}

void cpu_sched_pass(gol_world* w, const PassParams& p) {
  const int R = w->cfg.nranks;
  const int64_t st = w->stride, H = w->H;
  const int ext_out = p.ext_in - p.k;
  const int64_t n0 = H + 2 * p.ext_in;  // input rows, y in [-ext_in, H + ext_in)
  std::vector<std::vector<uint64_t>> outs(w->slabs.size());
  for (size_t si = 0; si < w->slabs.size(); ++si) {
    Slab& s = w->slabs[si];
    const uint64_t* in = s.ebuf[s.epass & 1].data() + (int64_t)w->G * st;
    std::vector<uint64_t> a((size_t)(n0 * st)), b((size_t)(n0 * st));
    for (int64_t i = 0; i < n0; ++i) {
      const int64_t y = i - p.ext_in;
      const uint64_t* src = in + y * st;
      if (p.consume && y < 0) src = s.erecv[0][s.eexch & 1].data() + (y + p.ext_in) * st;
      if (p.consume && y >= H) src = s.erecv[1][s.eexch & 1].data() + (y - H) * st;
      std::copy(src, src + st, a.begin() + i * st);
    }
    // k generations on a shrinking window: after generation g rows [g, n0 - g) are valid
    int64_t lo = 0, hi = n0;
    for (int g = 0; g < p.k; ++g) {
      const int64_t n = hi - lo - 2;
      std::vector<const uint64_t*> in_rows((size_t)n + 2);
      std::vector<uint64_t*> out_rows((size_t)n);
      for (int64_t i = 0; i < n + 2; ++i) in_rows[(size_t)i] = a.data() + (lo + i) * st;
      for (int64_t i = 0; i < n; ++i) out_rows[(size_t)i] = b.data() + (lo + 1 + i) * st;
      cpu_step_rows(in_rows.data(), out_rows.data(), n, w->W);
      a.swap(b);
      ++lo;
      --hi;
    }
    outs[si].assign(a.begin() + lo * st, a.begin() + hi * st);  // rows y in [-ext_out, H + ext_out)
  }
  for (size_t si = 0; si < w->slabs.size(); ++si) {
    Slab& s = w->slabs[si];
    uint64_t* out = s.ebuf[(s.epass + 1) & 1].data() + ((int64_t)w->G - ext_out) * st;
    std::copy(outs[si].begin(), outs[si].end(), out);
    if (p.push) {
      Slab& prev = w->slabs[(size_t)((s.rank - 1 + R) % R)];
      Slab& next = w->slabs[(size_t)((s.rank + 1) % R)];
      const uint64_t* interior = outs[si].data() + (int64_t)ext_out * st;  // row 0
      const uint64_t pe_prev = prev.eexch + (p.consume ? 1 : 0), pe_next = next.eexch + (p.consume ? 1 : 0);
      std::copy(interior, interior + p.push * st, prev.erecv[1][pe_prev & 1].begin());
      std::copy(interior + (H - p.push) * st, interior + H * st, next.erecv[0][pe_next & 1].begin());
    }
  }
  for (auto& s : w->slabs) {
    s.epass += 1;
    if (p.consume) s.eexch += 1;
  }
}

int cpu_run_schedule(gol_world* w, int64_t gens) {
  const int64_t st = w->stride, H = w->H;
  for (auto& s : w->slabs) {  // bring the I/O buffer into the margin-carrying slab
    uint64_t* dst = s.ebuf[s.epass & 1].data() + (int64_t)w->G * st;
    std::copy(s.cbuf[s.ccur].begin(), s.cbuf[s.ccur].begin() + H * st, dst);
  }
  const std::vector<PassItem> items = build_schedule(w, gens);
  for (const PassItem& it : items) {
    if (it.prime) cpu_sched_prime(w, it.depth);
    else cpu_sched_pass(w, it.p);
  }
  for (auto& s : w->slabs) {
    const uint64_t* src = s.ebuf[s.epass & 1].data() + (int64_t)w->G * st;
    std::copy(src, src + H * st, s.cbuf[s.ccur].begin());
  }
  return GOL_OK;
}

}  // namespace
}  // namespace gol

extern "C" {

uint64_t gol_launch_count(void) { return launch_count_value(); }

int gol_debug_bounds_enabled(void) { return debug_bounds_enabled() ? 1 : 0; }

int gol_debug_bounds_report(uint64_t* violations, uint64_t* first_bad_address, int* source_line, int reset) {
  unsigned long long n = 0, bad = 0;
  int line = 0;
  if (debug_bounds_query(&n, &bad, &line, reset)) return set_error(GOL_ERR_CUDA, "cannot read the bounds registry");
  if (violations) *violations = n;
  if (first_bad_address) *first_bad_address = bad;
  if (source_line) *source_line = line;
  return GOL_OK;
}

int gol_debug_bounds_poke(const void* device_address) {
  return debug_bounds_poke(device_address) ? set_error(GOL_ERR_CUDA, "poke failed") : GOL_OK;
}

int gol_cuda_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

gol_world_t* gol_world_create(const gol_config_t* cfg) {
  if (!cfg) {
    set_error(GOL_ERR_ARG, "null config");
    return nullptr;
  }
  if (cfg->width < 1 || cfg->rows < 1 || cfg->nranks < 1 || cfg->nlocal < 1 || cfg->first_rank < 0 ||
      cfg->first_rank + cfg->nlocal > cfg->nranks) {
    set_error(GOL_ERR_ARG, "bad geometry: width=%lld rows=%lld nranks=%d first=%d nlocal=%d",
              (long long)cfg->width, (long long)cfg->rows, cfg->nranks, cfg->first_rank, cfg->nlocal);
    return nullptr;
  }
  gol_world* w = new gol_world();
  w->cfg = *cfg;
  w->W = cfg->width;
  w->H = cfg->rows;
  w->ws = words_per_row(w->W);
  w->stride = row_stride_words(w->W);
  w->cuda = cfg->backend == GOL_BACKEND_CUDA;
  w->G = (int)std::min<int64_t>(GOL_MAX_GHOST, w->H);
  w->peers.resize((size_t)cfg->nranks);
  if (make_plan(w) != GOL_OK) {
    delete w;
    return nullptr;
  }
  w->slabs.resize((size_t)cfg->nlocal);
  for (int i = 0; i < cfg->nlocal; ++i) w->slabs[(size_t)i].rank = cfg->first_rank + i;

  if (!w->cuda) {
    for (auto& s : w->slabs) {
      for (int b = 0; b < 2; ++b) s.cbuf[b].assign((size_t)(w->H * w->stride), 0);
      for (int b = 0; b < 2; ++b) s.cghost[b].assign((size_t)w->stride, 0);
      if (cfg->flags & GOL_FLAG_CPU_SCHEDULE) {
        for (int b = 0; b < 2; ++b) s.ebuf[b].assign((size_t)((w->H + 2 * w->G) * w->stride), 0);
        for (int side = 0; side < 2; ++side)
          for (int par = 0; par < 2; ++par) s.erecv[side][par].assign((size_t)(w->G * w->stride), 0);
      }
    }
    if ((cfg->flags & GOL_FLAG_CPU_SCHEDULE) && w->plan.strip_rows <= 0) w->plan.strip_rows = 64;
    w->connected = (cfg->nlocal == cfg->nranks);
    return w;
  }

  // ---- CUDA: devices, groups, allocations
  auto fail = [&](void) -> gol_world_t* {
    gol_world_destroy(w);
    return nullptr;
  };
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev < 1) {
    set_error(GOL_ERR_CUDA, "no CUDA device available: %s", cudaGetErrorString(e));
    return fail();
  }
  int use = cfg->ndevices > 0 ? std::min(cfg->ndevices, ndev) : ndev;
  for (int i = 0; i < cfg->nlocal; ++i) {
    Slab& s = w->slabs[(size_t)i];
    s.device = (cfg->device_base + i) % use;  // rank -> device like gol-with-cuda.cu:296
    int gi = -1;
    for (size_t k = 0; k < w->groups.size(); ++k)
      if (w->groups[k].device == s.device) gi = (int)k;
    if (gi < 0) {
      Group g;
      g.device = s.device;
      w->groups.push_back(g);
      gi = (int)w->groups.size() - 1;
    }
    s.group = gi;
    s.index_in_group = (int)w->groups[(size_t)gi].slabs.size();
    w->groups[(size_t)gi].slabs.push_back(i);
  }
  const size_t row_bytes = (size_t)w->stride * 8;
  const size_t recv_bytes = align_up((size_t)w->G * row_bytes, 256);
  const size_t buf_bytes = align_up((size_t)(w->H + 2 * w->G) * row_bytes, 256);
  for (auto& s : w->slabs) {
    if (cudaSetDevice(s.device) != cudaSuccess) {
      set_error(GOL_ERR_CUDA, "cudaSetDevice(%d) failed", s.device);
      return fail();
    }
    // Mailbox (the only memory ring neighbours write: flags, recv rows, flag-in-data rows; this
    // is what gets exported over CUDA IPC) and grid (the two slab buffers) are separate
    // allocations.  Every buffer is fenced by guard rows holding a fixed pattern, which
    // gol_world_sync verifies on request (GOL_CHECK_CANARIES=1, always in the bounds-checking build).
    const size_t guard = align_up(std::max<size_t>(row_bytes, 256), 256);
    std::vector<size_t> mail_guards, grid_guards;
    size_t off = 0;
    s.off_flags = off;
    off += 256;
    mail_guards.push_back(off);
    off += guard;
    for (int side = 0; side < 2; ++side)
      for (int p = 0; p < 2; ++p) {
        s.off_recv[side][p] = off;
        off += recv_bytes;
        mail_guards.push_back(off);
        off += guard;
      }
    for (int side = 0; side < 2; ++side)
      for (int p = 0; p < 2; ++p) {   // flag-in-data halo rows: two 8-byte words per halo word
        s.off_ll[side][p] = off;
        off += 2 * recv_bytes;
        mail_guards.push_back(off);
        off += guard;
      }
    s.bytes = off;
    s.grid_bytes = 0;
    grid_guards.push_back(s.grid_bytes);
    s.grid_bytes += guard;
    for (int b = 0; b < 2; ++b) {
      s.off_buf[b] = s.grid_bytes;
      s.grid_bytes += buf_bytes;
      grid_guards.push_back(s.grid_bytes);
      s.grid_bytes += guard;
    }
    e = cudaMalloc((void**)&s.base, s.bytes);
    if (e == cudaSuccess) e = cudaMemset(s.base, 0, s.bytes);
    if (e == cudaSuccess) e = cudaMalloc((void**)&s.grid, s.grid_bytes);
    if (e == cudaSuccess) e = cudaMemset(s.grid, 0, s.grid_bytes);
    for (size_t o : mail_guards) {
      if (e == cudaSuccess) e = cudaMemset(s.base + o, kCanaryByte, guard);
      s.canaries.emplace_back(s.base + o, guard);
    }
    for (size_t o : grid_guards) {
      if (e == cudaSuccess) e = cudaMemset(s.grid + o, kCanaryByte, guard);
      s.canaries.emplace_back(s.grid + o, guard);
    }
    if (e == cudaSuccess) e = cudaMalloc((void**)&s.d_partials, (kPopBlocks + 1) * sizeof(unsigned long long));
    if (e == cudaSuccess) {
      // bounds-checking build: each buffer is a region of its own, the guard rows belong to none
      debug_bounds_register(s.base + s.off_flags, 256);
      for (int side = 0; side < 2; ++side)
        for (int p = 0; p < 2; ++p) {
          debug_bounds_register(s.base + s.off_recv[side][p], recv_bytes);
          debug_bounds_register(s.base + s.off_ll[side][p], 2 * recv_bytes);
        }
      for (int b = 0; b < 2; ++b) debug_bounds_register(s.grid + s.off_buf[b], buf_bytes);
      debug_bounds_register(s.d_partials, (kPopBlocks + 1) * sizeof(unsigned long long));
    }
    if (e != cudaSuccess) {
      set_error(GOL_ERR_CUDA, "slab allocation of %zu bytes on device %d failed: %s", s.bytes + s.grid_bytes,
                s.device, cudaGetErrorString(e));
      return fail();
    }
    s.d_pop = s.d_partials + kPopBlocks;
    SlabState& h = s.h;
    memset(&h, 0, sizeof h);
    h.W = w->W;
    h.H = w->H;
    h.ws = w->ws;
    h.stride = w->stride;
    h.G = w->G;
    h.rank = s.rank;
    h.nranks = cfg->nranks;
    for (int b = 0; b < 2; ++b) h.buf[b] = (uint64_t*)(s.grid + s.off_buf[b]);
    for (int side = 0; side < 2; ++side)
      for (int p = 0; p < 2; ++p) {
        h.recv[side][p] = (uint64_t*)(s.base + s.off_recv[side][p]);
        h.ll_recv[side][p] = (uint64_t*)(s.base + s.off_ll[side][p]);
      }
    h.flag[0] = (unsigned long long*)(s.base + s.off_flags);
    h.flag[1] = (unsigned long long*)(s.base + s.off_flags + 128);
    {
      // how long a kernel waits for a neighbour's halo before it gives up (GOL_WATCHDOG_MS, default 10 s)
      const char* e = getenv("GOL_WATCHDOG_MS");
      const unsigned long long ms = e && atoll(e) > 0 ? (unsigned long long)atoll(e) : 10000ull;
      h.watchdog_ns = ms * 1000ull * 1000ull;
    }
  }
  for (auto& g : w->groups) {
    e = cudaSetDevice(g.device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking);
    g.own_stream = (e == cudaSuccess);
    if (e == cudaSuccess) e = cudaMalloc((void**)&g.d_states, g.slabs.size() * sizeof(SlabState));
    if (e == cudaSuccess) e = cudaMalloc((void**)&g.d_barrier, 256);
    if (e == cudaSuccess) e = cudaMemset(g.d_barrier, 0, 256);
    if (e == cudaSuccess) e = cudaEventCreate(&g.e0);
    if (e == cudaSuccess) e = cudaEventCreate(&g.e1);
    if (e == cudaSuccess) {
      debug_bounds_register(g.d_states, g.slabs.size() * sizeof(SlabState));
      debug_bounds_register(g.d_barrier, 256);
    }
    if (e != cudaSuccess) {
      set_error(GOL_ERR_CUDA, "device %d set-up failed: %s", g.device, cudaGetErrorString(e));
      return fail();
    }
  }
  if (finish_plan_cuda(w) != GOL_OK) {
    if (cfg->mode == GOL_MODE_AUTO && w->plan.mode == GOL_MODE_RESIDENT) {
      // e.g. several slabs per device: too few CTAs per slab for the strips to fit shared memory
      w->auto_no_resident = true;
      if (make_plan(w) != GOL_OK || finish_plan_cuda(w) != GOL_OK) return fail();
    } else {
      return fail();
    }
  }
  if (w->plan.mode == GOL_MODE_TMA && make_tensor_maps(w) != GOL_OK) return fail();
  return w;
}

static void persist_release(int device);

void gol_world_destroy(gol_world_t* w) {
  if (!w) return;
  if (w->cuda) {
    for (auto& g : w->groups) {
      cudaSetDevice(g.device);
      if (g.stream) cudaStreamSynchronize(g.stream);
      if (g.l2_persist_held) persist_release(g.device);
      if (g.launcher) {
        g.launcher->stop();
        delete g.launcher;
        g.launcher = nullptr;
      }
      for (auto& kv : g.graphs) cudaGraphExecDestroy(kv.second.first);
      if (g.e0) cudaEventDestroy(g.e0);
      if (g.e1) cudaEventDestroy(g.e1);
      debug_bounds_unregister(g.d_states);
      debug_bounds_unregister(g.d_barrier);
      debug_bounds_unregister(g.d_progress);
      debug_bounds_unregister(g.d_mail);
      if (g.d_mail) cudaFree(g.d_mail);
      if (g.d_states) cudaFree(g.d_states);
      if (g.d_barrier) cudaFree(g.d_barrier);
      if (g.d_progress) cudaFree(g.d_progress);
      if (g.d_tmaps) cudaFree(g.d_tmaps);
      if (g.own_stream && g.stream) cudaStreamDestroy(g.stream);
    }
    for (auto& p : w->peers)
      if (p.ipc_opened && p.mapped) {
        cudaSetDevice(p.opened_on);
        cudaIpcCloseMemHandle(p.mapped);
      }
    for (auto& s : w->slabs) {
      cudaSetDevice(s.device);
      if (debug_bounds_enabled() && s.base) {
        debug_bounds_unregister(s.base + s.off_flags);
        for (int side = 0; side < 2; ++side)
          for (int p = 0; p < 2; ++p) {
            debug_bounds_unregister(s.base + s.off_recv[side][p]);
            debug_bounds_unregister(s.base + s.off_ll[side][p]);
          }
        for (int b = 0; b < 2; ++b) debug_bounds_unregister(s.grid + s.off_buf[b]);
        debug_bounds_unregister(s.d_partials);
        for (int k = 0; k < 2; ++k) {
          debug_bounds_unregister(s.stage_in[k]);
          debug_bounds_unregister(s.stage_out[k]);
        }
      }
      if (s.base) cudaFree(s.base);
      if (s.grid) cudaFree(s.grid);
      if (s.d_partials) cudaFree(s.d_partials);
      for (int k = 0; k < 2; ++k) {
        if (s.stage_in[k]) cudaFree(s.stage_in[k]);
        if (s.stage_out[k]) cudaFree(s.stage_out[k]);
        if (s.ev_h2d[k]) cudaEventDestroy(s.ev_h2d[k]);
        if (s.ev_in_used[k]) cudaEventDestroy(s.ev_in_used[k]);
        if (s.ev_out[k]) cudaEventDestroy(s.ev_out[k]);
        if (s.ev_done[k]) cudaEventDestroy(s.ev_done[k]);
      }
      if (s.s_h2d) cudaStreamDestroy(s.s_h2d);
      if (s.s_d2h) cudaStreamDestroy(s.s_d2h);
    }
  }
  delete w;
}

int gol_schedule_describe(const gol_config_t* cfg, int64_t gens, gol_pass_desc_t* out, int max_out,
                          gol_plan_t* plan_out) {
  if (!cfg || gens < 0 || (max_out > 0 && !out)) return set_error(GOL_ERR_ARG, "gol_schedule_describe: bad argument");
  if (cfg->width < 1 || cfg->rows < 1) return set_error(GOL_ERR_ARG, "gol_schedule_describe: bad geometry");
  gol_world tmp;  // never touches a device: only the planner and the schedule builder run
  tmp.cfg = *cfg;
  tmp.cfg.backend = GOL_BACKEND_CUDA;
  tmp.W = cfg->width;
  tmp.H = cfg->rows;
  tmp.ws = words_per_row(tmp.W);
  tmp.stride = row_stride_words(tmp.W);
  tmp.cuda = true;
  tmp.G = (int)std::min<int64_t>(GOL_MAX_GHOST, tmp.H);
  int rc = make_plan(&tmp);
  if (rc) return rc;
  if (tmp.plan.strip_rows <= 0) tmp.plan.strip_rows = 64;  // the real value needs the device's occupancy
  if (plan_out) *plan_out = tmp.plan;
  const std::vector<PassItem> items = build_schedule(&tmp, gens);
  for (size_t i = 0; i < items.size() && (int)i < max_out; ++i) {
    gol_pass_desc_t& d = out[i];
    memset(&d, 0, sizeof d);
    d.prime = items[i].prime;
    d.generic = items[i].generic;
    d.prime_depth = items[i].depth;
    d.k = items[i].p.k;
    d.ext_in = items[i].p.ext_in;
    d.consume = items[i].p.consume;
    d.push = items[i].p.push;
  }
  return (int)items.size();
}

int gol_world_plan(const gol_world_t* w, gol_plan_t* out) {
  if (!w || !out) return set_error(GOL_ERR_ARG, "null argument");
  *out = w->plan;
  return GOL_OK;
}

int gol_world_export(gol_world_t* w, int li, void* blob) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!blob) return set_error(GOL_ERR_ARG, "null blob");
  if (!w->cuda) return set_error(GOL_ERR_UNSUPPORTED, "export is for the CUDA backend");
  Slab& s = w->slabs[(size_t)li];
  ExportBlob b;
  memset(&b, 0, sizeof b);
  b.magic = kBlobMagic;
  b.rank = s.rank;
  b.device = s.device;
  b.pid = (int32_t)getpid();
  b.W = w->W;
  b.H = w->H;
  b.stride = w->stride;
  b.G = w->G;
  b.off_flags = s.off_flags;
  for (int side = 0; side < 2; ++side)
    for (int p = 0; p < 2; ++p) {
      b.off_recv[side][p] = s.off_recv[side][p];
      b.off_ll[side][p] = s.off_ll[side][p];
    }
  b.base_addr = (uint64_t)(uintptr_t)s.base;
  CU_TRY(cudaSetDevice(s.device));
  CU_TRY(cudaIpcGetMemHandle(&b.handle, s.base));
  memset(blob, 0, GOL_EXPORT_BYTES);
  memcpy(blob, &b, sizeof b);
  return GOL_OK;
}

int gol_world_import(gol_world_t* w, int global_rank, const void* blob) {
  if (!w || !blob) return set_error(GOL_ERR_ARG, "null argument");
  if (global_rank < 0 || global_rank >= w->cfg.nranks)
    return set_error(GOL_ERR_ARG, "rank %d out of range", global_rank);
  ExportBlob b;
  memcpy(&b, blob, sizeof b);
  if (b.magic != kBlobMagic || b.rank != global_rank)
    return set_error(GOL_ERR_ARG, "export blob for rank %d is malformed", global_rank);
  if (b.W != w->W || b.H != w->H || b.stride != w->stride || b.G != w->G)
    return set_error(GOL_ERR_ARG, "rank %d has a different slab geometry", global_rank);
  Peer& p = w->peers[(size_t)global_rank];
  p.blob = b;
  p.imported = true;
  p.mapped = nullptr;
  return GOL_OK;
}

static int resolve_peer_base(gol_world* w, int rank, int for_device, char** base, ExportBlob* lay) {
  // local slab?
  for (auto& s : w->slabs)
    if (s.rank == rank) {
      if (s.device != for_device && !(w->cfg.flags & GOL_FLAG_NO_PEER_ACCESS)) {
        int can = 0;
        CU_TRY(cudaDeviceCanAccessPeer(&can, for_device, s.device));
        if (!can)
          return set_error(GOL_ERR_CUDA, "device %d cannot access device %d (no P2P path)", for_device, s.device);
        CU_TRY(cudaSetDevice(for_device));
        cudaError_t e = cudaDeviceEnablePeerAccess(s.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
          return set_error(GOL_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d->%d): %s", for_device, s.device,
                           cudaGetErrorString(e));
        cudaGetLastError();
      }
      *base = s.base;
      lay->off_flags = s.off_flags;
      for (int side = 0; side < 2; ++side)
        for (int p = 0; p < 2; ++p) {
          lay->off_recv[side][p] = s.off_recv[side][p];
          lay->off_ll[side][p] = s.off_ll[side][p];
        }
      return GOL_OK;
    }
  Peer& p = w->peers[(size_t)rank];
  if (!p.imported) return set_error(GOL_ERR_STATE, "rank %d is neither local nor imported", rank);
  if (!p.mapped) {
    if (p.blob.pid == (int32_t)getpid()) {
      // another world in this process: its pointer is usable as is on the same device, through
      // peer access on a different one
      if (p.blob.device != for_device) {
        if (w->cfg.flags & GOL_FLAG_NO_PEER_ACCESS)
          return set_error(GOL_ERR_CUDA, "rank %d lives on device %d and peer access is disabled", rank,
                           p.blob.device);
        int can = 0;
        CU_TRY(cudaDeviceCanAccessPeer(&can, for_device, p.blob.device));
        if (!can)
          return set_error(GOL_ERR_CUDA, "device %d cannot access device %d (no P2P path)", for_device,
                           p.blob.device);
        CU_TRY(cudaSetDevice(for_device));
        cudaError_t e = cudaDeviceEnablePeerAccess(p.blob.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
          return set_error(GOL_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d->%d): %s", for_device, p.blob.device,
                           cudaGetErrorString(e));
        cudaGetLastError();
      }
      p.mapped = (char*)(uintptr_t)p.blob.base_addr;
    } else {
      CU_TRY(cudaSetDevice(for_device));
      void* ptr = nullptr;
      CU_TRY(cudaIpcOpenMemHandle(&ptr, p.blob.handle, cudaIpcMemLazyEnablePeerAccess));
      p.mapped = (char*)ptr;
      p.ipc_opened = true;
      p.opened_on = for_device;
    }
  }
  *base = p.mapped;
  *lay = p.blob;
  return GOL_OK;
}

int gol_world_connect(gol_world_t* w) {
  if (!w) return set_error(GOL_ERR_ARG, "null world");
  const int R = w->cfg.nranks;
  if (!w->cuda) {
    w->connected = true;
    return GOL_OK;
  }
  for (auto& s : w->slabs) {
    const int prev = (s.rank - 1 + R) % R;  // ring, gol-main.c:86-87
    const int next = (s.rank + 1) % R;
    char* pb = nullptr;
    char* nb = nullptr;
    ExportBlob pl{}, nl{};
    int rc = resolve_peer_base(w, prev, s.device, &pb, &pl);
    if (rc) return rc;
    rc = resolve_peer_base(w, next, s.device, &nb, &nl);
    if (rc) return rc;
    if (debug_bounds_enabled()) {
      // the neighbours' mailboxes as seen from this device (same geometry, hence same sizes)
      CU_TRY(cudaSetDevice(s.device));
      const size_t recv_bytes = align_up((size_t)w->G * (size_t)w->stride * 8, 256);
      for (int side = 0; side < 2; ++side)
        for (int p = 0; p < 2; ++p) {
          debug_bounds_register(pb + pl.off_recv[side][p], recv_bytes);
          debug_bounds_register(pb + pl.off_ll[side][p], 2 * recv_bytes);
          debug_bounds_register(nb + nl.off_recv[side][p], recv_bytes);
          debug_bounds_register(nb + nl.off_ll[side][p], 2 * recv_bytes);
        }
      debug_bounds_register(pb + pl.off_flags, 256);
      debug_bounds_register(nb + nl.off_flags, 256);
    }
    for (int p = 0; p < 2; ++p) {
      s.h.peer_recv[0][p] = (uint64_t*)(pb + pl.off_recv[1][p]);  // prev's bottom recv
      s.h.peer_recv[1][p] = (uint64_t*)(nb + nl.off_recv[0][p]);  // next's top recv
      s.h.peer_ll[0][p] = (uint64_t*)(pb + pl.off_ll[1][p]);
      s.h.peer_ll[1][p] = (uint64_t*)(nb + nl.off_ll[0][p]);
    }
    s.h.peer_flag[0] = (unsigned long long*)(pb + pl.off_flags + 128);  // prev's bottom flag
    s.h.peer_flag[1] = (unsigned long long*)(nb + nl.off_flags);        // next's top flag
    // neighbours on this very device (R = 1, or several ranks per GPU) are reached at gpu scope
    s.h.local_ring = 0;
    for (auto& o : w->slabs) {
      if (o.rank == prev && o.device == s.device) s.h.local_ring |= 1;
      if (o.rank == next && o.device == s.device) s.h.local_ring |= 2;
    }
    // halo rows that cross devices use the flag-in-data form in the fused kernels (the generic
    // kernel reads whole ghost rows through pointers and keeps the flag protocol); GOL_NO_LL=1
    // falls back to flags + system-scope fences everywhere
    // (GOL_FORCE_LL=1 uses it for same-device neighbours too: how the single-GPU tests cover it)
    const int no_ll = getenv("GOL_NO_LL") ? atoi(getenv("GOL_NO_LL")) : 0;
    const int force_ll = getenv("GOL_FORCE_LL") ? atoi(getenv("GOL_FORCE_LL")) : 0;
    s.h.ll_ring = (w->plan.mode == GOL_MODE_GENERIC || no_ll) ? 0 : (force_ll ? 3 : (~s.h.local_ring & 3));
    if (w->plan.mode == GOL_MODE_RESIDENT) s.h.ll_ring = 3;   // every strip boundary, ring edges included, is flag-in-data
  }
  int rc = upload_states(w);
  if (rc) return rc;
  w->connected = true;
  return GOL_OK;
}

int gol_world_set_stream(gol_world_t* w, int li, void* stream) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!w->cuda) return GOL_OK;
  Group& g = w->groups[(size_t)w->slabs[(size_t)li].group];
  if ((cudaStream_t)stream == g.stream) return GOL_OK;
  CU_TRY(cudaSetDevice(g.device));
  if (g.stream) CU_TRY(cudaStreamSynchronize(g.stream));
  if (g.own_stream && g.stream) cudaStreamDestroy(g.stream);
  g.stream = (cudaStream_t)stream;
  g.own_stream = false;
  return GOL_OK;
}

int gol_world_init_pattern(gol_world_t* w, int pattern, uint64_t seed, double density) {
  if (!w) return set_error(GOL_ERR_ARG, "null world");
  for (auto& s : w->slabs) {
    PatternArgs a;
    int rc = make_pattern_args(&a, pattern, w->W, w->H, s.rank, w->cfg.nranks, seed, density);
    if (rc) return rc;
    if (!w->cuda) {
      uint64_t* dst = s.cbuf[s.ccur].data();
      for (int64_t y = 0; y < w->H; ++y)
        for (int64_t x = 0; x < w->stride; ++x)
          dst[y * w->stride + x] = x < w->ws ? pattern_word(a, y, x) : 0ull;
    } else {
      Group& g = w->groups[(size_t)s.group];
      CU_TRY(cudaSetDevice(s.device));
      CU_TRY(launch_pattern_init(cur_interior(w, s), w->stride, a, g.stream));
    }
  }
  return GOL_OK;
}

int gol_world_load_packed_async(gol_world_t* w, int li, const uint64_t* words) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!words) return set_error(GOL_ERR_ARG, "null buffer");
  Slab& s = w->slabs[(size_t)li];
  const size_t n = (size_t)(w->H * w->stride) * 8;
  if (!w->cuda) {
    memcpy(s.cbuf[s.ccur].data(), words, n);
    return GOL_OK;
  }
  Group& g = w->groups[(size_t)s.group];
  CU_TRY(cudaSetDevice(s.device));
  CU_TRY(cudaMemcpyAsync(cur_interior(w, s), words, n, cudaMemcpyHostToDevice, g.stream));
  return GOL_OK;
}

int gol_world_load_packed(gol_world_t* w, int li, const uint64_t* words) {
  int rc = gol_world_load_packed_async(w, li, words);
  if (rc) return rc;
  if (w->cuda) CU_TRY(cudaStreamSynchronize(w->groups[(size_t)w->slabs[(size_t)li].group].stream));
  return GOL_OK;
}

// Byte-cell I/O goes through a bounded device scratch (at most 64 MiB of byte cells at a time), so
// that a 2^32-cell slab does not need 4 GiB of extra device memory to be loaded or read.
static int64_t byte_io_chunk_rows(const gol_world* w) {
  const int64_t rows = (int64_t)(64ll << 20) / std::max<int64_t>(1, w->W);
  return std::max<int64_t>(1, std::min<int64_t>(rows, w->H));
}

int gol_world_load_bytes(gol_world_t* w, int li, const uint8_t* cells) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!cells) return set_error(GOL_ERR_ARG, "null buffer");
  Slab& s = w->slabs[(size_t)li];
  if (!w->cuda) return gol_pack_bytes(cells, s.cbuf[s.ccur].data(), w->W, w->H);
  Group& g = w->groups[(size_t)s.group];
  CU_TRY(cudaSetDevice(s.device));
  const int64_t chunk = byte_io_chunk_rows(w);
  uint8_t* d = nullptr;
  const size_t n = (size_t)w->W * (size_t)chunk;
  CU_TRY(cudaMalloc((void**)&d, n));
  debug_bounds_register(d, n);
  cudaError_t e = cudaSuccess;
  for (int64_t y0 = 0; y0 < w->H && e == cudaSuccess; y0 += chunk) {
    const int64_t rows = std::min(chunk, w->H - y0);
    e = cudaMemcpyAsync(d, cells + (size_t)y0 * (size_t)w->W, (size_t)rows * (size_t)w->W, cudaMemcpyHostToDevice,
                        g.stream);
    if (e == cudaSuccess) e = launch_pack(d, cur_interior(w, s) + y0 * w->stride, w->W, rows, w->stride, g.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g.stream);   // the scratch is reused by the next chunk
  }
  debug_bounds_unregister(d);
  cudaFree(d);
  if (e != cudaSuccess) return set_error(GOL_ERR_CUDA, "load_bytes: %s", cudaGetErrorString(e));
  return GOL_OK;
}

int gol_world_read_packed(gol_world_t* w, int li, uint64_t* words) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!words) return set_error(GOL_ERR_ARG, "null buffer");
  Slab& s = w->slabs[(size_t)li];
  const size_t n = (size_t)(w->H * w->stride) * 8;
  if (!w->cuda) {
    memcpy(words, s.cbuf[s.ccur].data(), n);
    return GOL_OK;
  }
  Group& g = w->groups[(size_t)s.group];
  CU_TRY(cudaSetDevice(s.device));
  CU_TRY(cudaMemcpyAsync(words, cur_interior(w, s), n, cudaMemcpyDeviceToHost, g.stream));
  CU_TRY(cudaStreamSynchronize(g.stream));
  return GOL_OK;
}

int gol_world_read_packed_async(gol_world_t* w, int li, uint64_t* pinned_words) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!pinned_words) return set_error(GOL_ERR_ARG, "null buffer");
  Slab& s = w->slabs[(size_t)li];
  const size_t n = (size_t)(w->H * w->stride) * 8;
  if (!w->cuda) {
    memcpy(pinned_words, s.cbuf[s.ccur].data(), n);
    return GOL_OK;
  }
  Group& g = w->groups[(size_t)s.group];
  CU_TRY(cudaSetDevice(s.device));
  CU_TRY(cudaMemcpyAsync(pinned_words, cur_interior(w, s), n, cudaMemcpyDeviceToHost, g.stream));
  return GOL_OK;
}

// ---- end-to-end pipeline -----------------------------------------------------------------
// One call = one user-visible step: upload this step's packed input from pinned host memory,
// advance `gens` generations, bring the packed result back to pinned host memory.  The upload
// and the download run on their own copy streams through two slots of device staging, so the
// transfer of step i+1's input and of step i's result overlap the compute of the neighbouring
// steps; the compute stream only sees two device-to-device copies (a few microseconds).
static int pipeline_prepare(gol_world* w, Slab& s) {
  if (s.s_h2d) return GOL_OK;
  const size_t n = (size_t)(w->H * w->stride) * 8;
  CU_TRY(cudaSetDevice(s.device));
  CU_TRY(cudaStreamCreateWithFlags(&s.s_h2d, cudaStreamNonBlocking));
  CU_TRY(cudaStreamCreateWithFlags(&s.s_d2h, cudaStreamNonBlocking));
  for (int k = 0; k < 2; ++k) {
    CU_TRY(cudaMalloc((void**)&s.stage_in[k], n));
    CU_TRY(cudaMalloc((void**)&s.stage_out[k], n));
    debug_bounds_register(s.stage_in[k], n);
    debug_bounds_register(s.stage_out[k], n);
    CU_TRY(cudaEventCreateWithFlags(&s.ev_h2d[k], cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&s.ev_in_used[k], cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&s.ev_out[k], cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&s.ev_done[k], cudaEventDisableTiming));
    // events start "completed" so that the first use of a slot does not wait
    CU_TRY(cudaEventRecord(s.ev_in_used[k], s.s_h2d));
    CU_TRY(cudaEventRecord(s.ev_done[k], s.s_d2h));
  }
  return GOL_OK;
}

int gol_world_run_from_host_async(gol_world_t* w, const uint64_t* const* pinned_in, int64_t gens,
                                  uint64_t* const* pinned_out) {
  if (!w || !pinned_in || !pinned_out) return set_error(GOL_ERR_ARG, "null argument");
  if (gens < 0) return set_error(GOL_ERR_ARG, "negative generation count");
  const size_t n = (size_t)(w->H * w->stride) * 8;
  if (!w->cuda) {
    for (size_t i = 0; i < w->slabs.size(); ++i) memcpy(w->slabs[i].cbuf[w->slabs[i].ccur].data(), pinned_in[i], n);
    int rc = gol_world_step(w, gens);
    if (rc) return rc;
    for (size_t i = 0; i < w->slabs.size(); ++i) memcpy(pinned_out[i], w->slabs[i].cbuf[w->slabs[i].ccur].data(), n);
    return 0;
  }
  const int k = w->pipe_slot;
  static const int trace = getenv("GOL_PIPE_TRACE") ? atoi(getenv("GOL_PIPE_TRACE")) : 0;
  static const int xflags = getenv("GOL_PIPE_FLAGS") ? atoi(getenv("GOL_PIPE_FLAGS")) : 0;  // experiments only
  for (size_t i = 0; i < w->slabs.size(); ++i) {
    Slab& s = w->slabs[i];
    if (!pinned_in[i] || !pinned_out[i]) return set_error(GOL_ERR_ARG, "null buffer for local slab %zu", i);
    int rc = pipeline_prepare(w, s);
    if (rc) return rc;
    Group& g = w->groups[(size_t)s.group];
    CU_TRY(cudaSetDevice(s.device));
    // upload: the slot's previous content must have been consumed by the compute stream
    CU_TRY(cudaStreamWaitEvent(s.s_h2d, s.ev_in_used[k], 0));
    if (trace) {
      for (int e = 0; e < 6; ++e)
        if (!s.tr[k][e]) CU_TRY(cudaEventCreate(&s.tr[k][e]));
      CU_TRY(cudaEventRecord(s.tr[k][0], s.s_h2d));
    }
    if (!(xflags & 1)) CU_TRY(cudaMemcpyAsync(s.stage_in[k], pinned_in[i], n, cudaMemcpyHostToDevice, s.s_h2d));
    if (trace) CU_TRY(cudaEventRecord(s.tr[k][1], s.s_h2d));
    CU_TRY(cudaEventRecord(s.ev_h2d[k], s.s_h2d));
    CU_TRY(cudaStreamWaitEvent(g.stream, s.ev_h2d[k], 0));
    if (!(xflags & 4)) CU_TRY(cudaMemcpyAsync(cur_interior(w, s), s.stage_in[k], n, cudaMemcpyDeviceToDevice, g.stream));
    CU_TRY(cudaEventRecord(s.ev_in_used[k], g.stream));
    if (trace) CU_TRY(cudaEventRecord(s.tr[k][2], g.stream));
  }
  int rc = gol_world_step(w, gens);
  if (rc) return rc;
  for (size_t i = 0; i < w->slabs.size(); ++i) {
    Slab& s = w->slabs[i];
    Group& g = w->groups[(size_t)s.group];
    CU_TRY(cudaSetDevice(s.device));
    if (trace) CU_TRY(cudaEventRecord(s.tr[k][3], g.stream));
    // result: wait until the slot's previous download has left the staging buffer
    CU_TRY(cudaStreamWaitEvent(g.stream, s.ev_done[k], 0));
    if (!(xflags & 4)) CU_TRY(cudaMemcpyAsync(s.stage_out[k], cur_interior(w, s), n, cudaMemcpyDeviceToDevice, g.stream));
    CU_TRY(cudaEventRecord(s.ev_out[k], g.stream));
    CU_TRY(cudaStreamWaitEvent(s.s_d2h, s.ev_out[k], 0));
    if (trace) CU_TRY(cudaEventRecord(s.tr[k][4], s.s_d2h));
    if (!(xflags & 2)) CU_TRY(cudaMemcpyAsync(pinned_out[i], s.stage_out[k], n, cudaMemcpyDeviceToHost, s.s_d2h));
    if (trace) CU_TRY(cudaEventRecord(s.tr[k][5], s.s_d2h));
    CU_TRY(cudaEventRecord(s.ev_done[k], s.s_d2h));
  }
  w->pipe_slot ^= 1;
  return k;
}

int gol_world_wait_result(gol_world_t* w, int slot) {
  if (!w || slot < 0 || slot > 1) return set_error(GOL_ERR_ARG, "bad slot");
  if (!w->cuda) return GOL_OK;
  for (auto& s : w->slabs) {
    if (!s.ev_done[slot]) return set_error(GOL_ERR_STATE, "no pipelined step has been issued");
    CU_TRY(cudaSetDevice(s.device));
    CU_TRY(cudaEventSynchronize(s.ev_done[slot]));
    if (s.tr[slot][5]) {   // GOL_PIPE_TRACE: where the step's time went, relative to its upload start
      float h = 0, c0 = 0, c = 0, d0 = 0, d = 0;
      cudaEventElapsedTime(&h, s.tr[slot][0], s.tr[slot][1]);
      cudaEventElapsedTime(&c0, s.tr[slot][0], s.tr[slot][2]);
      cudaEventElapsedTime(&c, s.tr[slot][2], s.tr[slot][3]);
      cudaEventElapsedTime(&d0, s.tr[slot][0], s.tr[slot][4]);
      cudaEventElapsedTime(&d, s.tr[slot][4], s.tr[slot][5]);
      fprintf(stderr, "pipe rank %d slot %d: h2d %.3f ms | compute starts +%.3f, runs %.3f | d2h starts +%.3f, runs %.3f\n",
              s.rank, slot, h, c0, c, d0, d);
    }
  }
  return GOL_OK;
}

int gol_world_read_bytes(gol_world_t* w, int li, uint8_t* cells) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!cells) return set_error(GOL_ERR_ARG, "null buffer");
  Slab& s = w->slabs[(size_t)li];
  if (!w->cuda) return gol_unpack_words(s.cbuf[s.ccur].data(), cells, w->W, w->H);
  Group& g = w->groups[(size_t)s.group];
  CU_TRY(cudaSetDevice(s.device));
  const int64_t chunk = byte_io_chunk_rows(w);
  uint8_t* d = nullptr;
  const size_t n = (size_t)w->W * (size_t)chunk;
  CU_TRY(cudaMalloc((void**)&d, n));
  debug_bounds_register(d, n);
  cudaError_t e = cudaSuccess;
  for (int64_t y0 = 0; y0 < w->H && e == cudaSuccess; y0 += chunk) {
    const int64_t rows = std::min(chunk, w->H - y0);
    e = launch_unpack(cur_interior(w, s) + y0 * w->stride, d, w->W, rows, w->stride, g.stream);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(cells + (size_t)y0 * (size_t)w->W, d, (size_t)rows * (size_t)w->W, cudaMemcpyDeviceToHost,
                          g.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g.stream);
  }
  debug_bounds_unregister(d);
  cudaFree(d);
  if (e != cudaSuccess) return set_error(GOL_ERR_CUDA, "read_bytes: %s", cudaGetErrorString(e));
  return GOL_OK;
}

int gol_world_population_async(gol_world_t* w, int li, uint64_t* out) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!out) return set_error(GOL_ERR_ARG, "null output");
  Slab& s = w->slabs[(size_t)li];
  if (!w->cuda) {
    uint64_t t = 0;
    const uint64_t* p = s.cbuf[s.ccur].data();
    for (int64_t y = 0; y < w->H; ++y)
      for (int64_t x = 0; x < w->ws; ++x) t += (uint64_t)__builtin_popcountll(p[y * w->stride + x]);
    *out = t;
    return GOL_OK;
  }
  Group& g = w->groups[(size_t)s.group];
  CU_TRY(cudaSetDevice(s.device));
  CU_TRY(launch_population(cur_interior(w, s), w->ws, w->stride, w->H, s.d_partials, s.d_pop, g.stream));
  CU_TRY(cudaMemcpyAsync(out, s.d_pop, sizeof(uint64_t), cudaMemcpyDeviceToHost, g.stream));
  return GOL_OK;
}

int gol_world_population(gol_world_t* w, int li, uint64_t* out) {
  int rc = gol_world_population_async(w, li, out);
  if (rc) return rc;
  if (w->cuda) CU_TRY(cudaStreamSynchronize(w->groups[(size_t)w->slabs[(size_t)li].group].stream));
  return GOL_OK;
}

// L2 residency control: mark the slab allocation of every slab on this device as "persisting" in
// L2 for kernels launched on the group's stream, so that streaming traffic (the end-to-end
// pipeline's PCIe uploads / downloads, which pass through L2) evicts itself instead of the
// double buffer the step kernels live in, and a step that starts on a cold or dirty L2 (another
// kernel ran in between) gets its working set resident within a pass or two: 8.05 instead of
// 8.7 ms per 1024 generations at 16384^2 after a 256 MiB write.  The carve-out is a device-wide
// limit: it is counted per device and given back when the last world using it goes away (left
// set, a later world whose slab does not fit ran at 31.6 instead of 40.5 TCUPS on 65536 x 32768).
// GOL_L2_PERSIST=0 disables it.
static int g_persist_users[64];
static void persist_release(int device) {
  if (device < 0 || device >= 64 || g_persist_users[device] <= 0) return;
  if (--g_persist_users[device] == 0) {
    cudaSetDevice(device);
    cudaCtxResetPersistingL2Cache();
    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, 0);
    cudaGetLastError();
  }
}
static int apply_l2_policy(gol_world* w, Group& g) {
  static const int enabled = getenv("GOL_L2_PERSIST") ? atoi(getenv("GOL_L2_PERSIST")) : 1;
  if (!enabled || g.l2_policy_stream == (void*)g.stream) return GOL_OK;
  g.l2_policy_stream = (void*)g.stream;
  int max_persist = 0, max_window = 0;
  cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, g.device);
  cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, g.device);
  if (max_persist <= 0 || max_window <= 0 || g.slabs.empty()) return GOL_OK;
  // one window per stream: cover the lowest..highest slab address on this device
  char* lo = nullptr;
  char* hi = nullptr;
  for (int si : g.slabs) {
    Slab& s = w->slabs[(size_t)si];
    if (!lo || s.grid < lo) lo = s.grid;
    if (!hi || s.grid + s.grid_bytes > hi) hi = s.grid + s.grid_bytes;
  }
  size_t span = (size_t)(hi - lo);
  // a window larger than the carve-out would demote part of the double buffer to "streaming":
  // measured 30.7 vs 34.8 TCUPS on a 16384 x 32768 slab (128 MiB), so only slabs that fit are pinned
  if (span > (size_t)max_window || span > (size_t)max_persist) return GOL_OK;
  if (g.device >= 64) return GOL_OK;
  if (!g.l2_persist_held) {
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)max_persist) != cudaSuccess) {
      cudaGetLastError();
      return GOL_OK;
    }
    g.l2_persist_held = true;
    ++g_persist_users[g.device];
  }
  cudaStreamAttrValue v;
  memset(&v, 0, sizeof v);
  v.accessPolicyWindow.base_ptr = lo;
  v.accessPolicyWindow.num_bytes = span;
  v.accessPolicyWindow.hitRatio = span <= (size_t)max_persist ? 1.0f : (float)max_persist / (float)span;
  v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
  v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  if (cudaStreamSetAttribute(g.stream, cudaStreamAttributeAccessPolicyWindow, &v) != cudaSuccess) cudaGetLastError();
  return GOL_OK;
}

static bool graph_capable(const Group& g) {
  // the legacy default stream cannot be captured
  return g.stream != nullptr && g.stream != cudaStreamLegacy && g.stream != cudaStreamPerThread;
}

// CUDA graph of the whole pass schedule for `gens` generations on one device (cached).
static int get_graph(gol_world* w, Group& g, int64_t gens, const std::vector<PassItem>& items,
                     std::pair<cudaGraphExec_t, int>** out) {
  auto it = g.graphs.find(gens);
  if (it == g.graphs.end()) {
    const uint64_t before = launch_count_value();
    cudaGraph_t graph = nullptr;
    CU_TRY(cudaStreamBeginCapture(g.stream, cudaStreamCaptureModeThreadLocal));
    int rc = GOL_OK;
    for (auto& item : items) {
      rc = launch_item(w, g, item);
      if (rc) break;
    }
    cudaError_t e = cudaStreamEndCapture(g.stream, &graph);
    count_launch(before - launch_count_value());  // capture is not execution
    if (rc) {
      if (graph) cudaGraphDestroy(graph);
      return rc;
    }
    if (e != cudaSuccess) return set_error(GOL_ERR_CUDA, "graph capture failed: %s", cudaGetErrorString(e));
    cudaGraphExec_t exec = nullptr;
    e = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) return set_error(GOL_ERR_CUDA, "graph instantiate failed: %s", cudaGetErrorString(e));
    if (g.graphs.size() > 16) {
      for (auto& kv : g.graphs) cudaGraphExecDestroy(kv.second.first);
      g.graphs.clear();
    }
    it = g.graphs.emplace(gens, std::make_pair(exec, (int)items.size())).first;
  }
  *out = &it->second;
  return GOL_OK;
}

// launch = false only builds (and caches) the CUDA graphs so that a timed step does not
// include capture / instantiation.
static int step_cuda(gol_world* w, int64_t gens, bool launch = true) {
  const gol_plan_t& pl = w->plan;
  for (auto& g : w->groups) {
    if (g.l2_policy_stream == (void*)g.stream) continue;
    CU_TRY(cudaSetDevice(g.device));
    apply_l2_policy(w, g);
    // kernel nodes capture the stream's access-policy window: graphs built before are stale
    for (auto& kv : g.graphs) cudaGraphExecDestroy(kv.second.first);
    g.graphs.clear();
  }
  const bool use_graph = !(w->cfg.flags & GOL_FLAG_NO_GRAPH);
  int64_t todo = gens;

  if (pl.mode == GOL_MODE_RESIDENT) {
    if (!launch) return GOL_OK;
    const int K = pl.temporal_k;
    for (auto& g : w->groups) {
      CU_TRY(cudaSetDevice(g.device));
      ResidentParams rp;
      memset(&rp, 0, sizeof rp);
      rp.K = K;
      rp.npasses = (int)((gens + K - 1) / K);
      rp.last_k = (int)(gens - (int64_t)(rp.npasses - 1) * K);
      rp.nslabs = (int)g.slabs.size();
      rp.ctas_per_slab = pl.ctas_per_slab;
      rp.cluster = pl.regs_hint;
      rp.rows_alloc = pl.strip_rows + 2 * K;
      rp.use_bulk = (w->stride == w->ws) && !getenv("GOL_RESIDENT_NO_BULK");
      rp.smem_words = (int)(resident_smem_bytes(rp.rows_alloc, w->ws, K) / 8);
      rp.mail = g.d_mail;
      rp.mail_slot = 2 * (int64_t)K * w->ws;
      rp.ctl = g.d_barrier;
      CU_TRY(launch_resident(g.d_states, rp, pl.threads_per_block, resident_smem_bytes(rp.rows_alloc, w->ws, K), g.stream));
    }
    w->pass_host += 1;   // one buffer flip per launch, whatever the generation count
    return GOL_OK;
  }
  if (pl.mode == GOL_MODE_PERSISTENT) {
    const int P = pl.exchange_depth / pl.temporal_k;
    const int64_t periods = todo / pl.exchange_depth;
    if (periods > 0 && !launch) {
      todo -= periods * pl.exchange_depth;
    } else if (periods > 0) {
      // One launch of the flow kernel per chunk of exchange periods; tickets are 32-bit.
      const int K = pl.temporal_k, E = pl.exchange_depth;
      const int nstrips = w->flow_strips;
      int dep_radius = 1;
      if (nstrips > 2 && (w->H - 2 * (int64_t)E) / (nstrips - 2) < K) dep_radius = nstrips;
      const int wpb = pl.threads_per_block / 32;
      int64_t max_tpp = 1;
      for (auto& g : w->groups)
        max_tpp = std::max<int64_t>(max_tpp, (int64_t)g.slabs.size() * nstrips * pl.col_groups);
      const int64_t max_periods = std::max<int64_t>(1, std::min<int64_t>(1 << 20, ((1ll << 31) - 1) / max_tpp / P));
      int64_t left = periods;
      while (left > 0) {
        const int64_t now = std::min(left, max_periods);
        for (auto& g : w->groups) {
          CU_TRY(cudaSetDevice(g.device));
          FlowParams fp;
          memset(&fp, 0, sizeof fp);
          fp.npasses = (int)(now * P);
          fp.passes_per_exchange = P;
          fp.depth = E;
          fp.nslabs = (int)g.slabs.size();
          fp.col_groups = pl.col_groups;
          fp.nstrips = nstrips;
          fp.edge_rows = E;
          fp.dep_radius = dep_radius;
          fp.max_strips = nstrips;
          fp.ctl = g.d_barrier;
          fp.progress = g.d_progress;
          fp.flags = w->flow_owner ? 4 : 0;
          if (const char* e = getenv("GOL_FLOW_FLAGS")) fp.flags = atoi(e);
          const int64_t tpp = (int64_t)fp.nslabs * nstrips * pl.col_groups;
          int nctas = flow_max_ctas(K, pl.threads_per_block, g.device);
          if (nctas < 1) return set_error(GOL_ERR_CUDA, "flow kernel does not fit an SM");
          // owner mode: one warp per task slot; ticket mode: more warps than two passes' worth of
          // tasks could only spin on dependencies
          const int64_t want_warps = (fp.flags & 4) ? tpp : 2 * tpp;
          nctas = (int)std::max<int64_t>(1, std::min<int64_t>(nctas, (want_warps + wpb - 1) / wpb));
          if (const char* e = getenv("GOL_FLOW_CTAS"))
            if (atoi(e) > 0) nctas = std::min(nctas, atoi(e));
          CU_TRY(launch_prime(g.d_states, fp.nslabs, E, w->ws, g.stream));
          CU_TRY(launch_flow(K, g.d_states, fp, pl.threads_per_block, nctas, g.stream));
        }
        w->pass_host += (uint64_t)(now * P);
        left -= now;
      }
      todo -= periods * pl.exchange_depth;
    }
    if (todo == 0) return GOL_OK;
  }

  const std::vector<PassItem> items = build_schedule(w, todo);
  if (items.empty()) return GOL_OK;
  const int npass = count_passes(items);
  if (use_graph && items.size() > 2) {
    for (auto& g : w->groups) {
      CU_TRY(cudaSetDevice(g.device));
      if (!graph_capable(g)) {
        if (launch)
          for (auto& item : items) {
            int rc = launch_item(w, g, item);
            if (rc) return rc;
          }
        continue;
      }
      std::pair<cudaGraphExec_t, int>* ge = nullptr;
      int rc = get_graph(w, g, todo, items, &ge);
      if (rc) return rc;
      if (!launch) continue;
      if (w->groups.size() > 1) {
        // all devices' graphs leave at the same time, each from its own thread
        if (!g.launcher) g.launcher = new Launcher();
        const int dev = g.device;
        cudaGraphExec_t exec = ge->first;
        cudaStream_t st = g.stream;
        g.launcher->post([dev, exec, st]() -> cudaError_t {
          cudaError_t e = cudaSetDevice(dev);
          return e != cudaSuccess ? e : cudaGraphLaunch(exec, st);
        });
      } else {
        CU_TRY(cudaGraphLaunch(ge->first, g.stream));
      }
      count_launch((uint64_t)ge->second);
    }
    if (!launch) return GOL_OK;
    if (w->groups.size() > 1)
      for (auto& g : w->groups)
        if (g.launcher) {
          const cudaError_t e = g.launcher->wait();
          if (e != cudaSuccess)
            return set_error(GOL_ERR_CUDA, "graph launch on device %d failed: %s", g.device, cudaGetErrorString(e));
        }
  } else {
    if (!launch) return GOL_OK;
    for (auto& item : items)
      for (auto& g : w->groups) {
        CU_TRY(cudaSetDevice(g.device));
        int rc = launch_item(w, g, item);
        if (rc) return rc;
      }
  }
  w->pass_host += (uint64_t)npass;
  return GOL_OK;
}

int gol_world_step(gol_world_t* w, int64_t gens) {
  if (!w) return set_error(GOL_ERR_ARG, "null world");
  if (gens < 0) return set_error(GOL_ERR_ARG, "negative generation count");
  if (gens == 0) return GOL_OK;
  if (!w->cuda) {
    if ((int)w->slabs.size() != w->cfg.nranks)
      return set_error(GOL_ERR_STATE,
                       "CPU world owns %d of %d ranks: drive it with cpu_get_edge / cpu_set_ghost / cpu_step_local",
                       (int)w->slabs.size(), w->cfg.nranks);
    if (w->cfg.flags & GOL_FLAG_CPU_SCHEDULE) return cpu_run_schedule(w, gens);
    for (int64_t g = 0; g < gens; ++g) {
      cpu_exchange_local(w);
      for (auto& s : w->slabs) cpu_step_slab(w, s);
    }
    return GOL_OK;
  }
  if (!w->connected) return set_error(GOL_ERR_STATE, "gol_world_connect has not been called");
  nvtxRangePushA("gol_world_step");
  // Long runs are cut into chunks of at most 4096 exchange periods so that a captured graph
  // stays small (a few thousand kernel nodes) however many generations are asked for; the
  // chunk graph is cached and replayed.
  const int64_t chunk = (int64_t)std::max(1, w->plan.exchange_depth) * 4096;
  int rc = GOL_OK;
  for (int64_t done = 0; done < gens && rc == GOL_OK; done += chunk)
    rc = step_cuda(w, std::min(chunk, gens - done));
  nvtxRangePop();
  return rc;
}

// Guard rows around every buffer must still hold their pattern (a kernel that ran past the end of
// a slab buffer, a recv buffer or a flag-in-data buffer lands in one first).
static int check_canaries(gol_world* w) {
  std::vector<unsigned char> h;
  for (auto& s : w->slabs) {
    CU_TRY(cudaSetDevice(s.device));
    for (size_t i = 0; i < s.canaries.size(); ++i) {
      h.resize(s.canaries[i].second);
      CU_TRY(cudaMemcpy(h.data(), s.canaries[i].first, h.size(), cudaMemcpyDeviceToHost));
      for (size_t b = 0; b < h.size(); ++b)
        if (h[b] != (unsigned char)kCanaryByte)
          return set_error(GOL_ERR_STATE, "rank %d: guard row %zu was overwritten at byte %zu (a kernel wrote outside "
                           "its buffer)", s.rank, i, b);
    }
  }
  return GOL_OK;
}

int gol_world_sync(gol_world_t* w) {
  if (!w) return set_error(GOL_ERR_ARG, "null world");
  if (!w->cuda) return GOL_OK;
  for (auto& g : w->groups) {
    CU_TRY(cudaSetDevice(g.device));
    CU_TRY(cudaStreamSynchronize(g.stream));
  }
  {
    const char* e = getenv("GOL_CHECK_CANARIES");
    if (debug_bounds_enabled() || (e && atoi(e))) {
      int rc = check_canaries(w);
      if (rc) return rc;
    }
  }
  if (debug_bounds_enabled())
    for (auto& g : w->groups) {
      CU_TRY(cudaSetDevice(g.device));
      unsigned long long n = 0, bad = 0;
      int line = 0;
      debug_bounds_query(&n, &bad, &line, 0);
      if (n)
        return set_error(GOL_ERR_STATE, "device %d: %llu out-of-bounds accesses, first at %#llx (csrc/kernels.cu:%d)",
                         g.device, n, bad, line);
    }
  for (auto& g : w->groups) {
    CU_TRY(cudaSetDevice(g.device));
    std::vector<SlabState> hs(g.slabs.size());
    CU_TRY(cudaMemcpy(hs.data(), g.d_states, hs.size() * sizeof(SlabState), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < hs.size(); ++i) {
      if (hs[i].error)
        return set_error(GOL_ERR_TIMEOUT,
                         "rank %d: halo from the %s neighbour did not arrive within the watchdog (device error %d)",
                         hs[i].rank, hs[i].error == 1 ? "previous" : "next", hs[i].error);
      if (w->connected && hs[i].pass != w->pass_host)
        return set_error(GOL_ERR_STATE, "rank %d: device pass counter %llu != host %llu", hs[i].rank,
                         hs[i].pass, (unsigned long long)w->pass_host);
    }
  }
  return GOL_OK;
}

int gol_world_step_timed(gol_world_t* w, int64_t gens, float* ms) {
  if (!w || !ms) return set_error(GOL_ERR_ARG, "null argument");
  if (!w->cuda) {
    timespec a, b;
    clock_gettime(CLOCK_MONOTONIC, &a);
    int rc = gol_world_step(w, gens);
    clock_gettime(CLOCK_MONOTONIC, &b);
    const float t = (float)((b.tv_sec - a.tv_sec) * 1e3 + (b.tv_nsec - a.tv_nsec) * 1e-6);
    for (size_t i = 0; i < w->slabs.size(); ++i) ms[i] = t;
    return rc;
  }
  if (!w->connected) return set_error(GOL_ERR_STATE, "gol_world_connect has not been called");
  const int64_t chunk = (int64_t)std::max(1, w->plan.exchange_depth) * 4096;
  int rc = step_cuda(w, std::min(chunk, gens), /*launch=*/false);  // graphs built outside the timed region
  if (rc == GOL_OK && gens > chunk && gens % chunk) rc = step_cuda(w, gens % chunk, /*launch=*/false);
  if (rc) return rc;
  for (auto& g : w->groups) {
    CU_TRY(cudaSetDevice(g.device));
    CU_TRY(cudaEventRecord(g.e0, g.stream));
  }
  rc = gol_world_step(w, gens);
  if (rc) return rc;
  for (auto& g : w->groups) {
    CU_TRY(cudaSetDevice(g.device));
    CU_TRY(cudaEventRecord(g.e1, g.stream));
  }
  rc = gol_world_sync(w);
  if (rc) return rc;
  for (auto& g : w->groups) {
    float t = 0;
    CU_TRY(cudaSetDevice(g.device));
    CU_TRY(cudaEventElapsedTime(&t, g.e0, g.e1));
    for (int si : g.slabs) ms[si] = t;
  }
  return GOL_OK;
}

int gol_world_phase_counters(gol_world_t* w, int li, gol_phase_counters_t* out, int reset) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (!out) return set_error(GOL_ERR_ARG, "null output");
  memset(out, 0, sizeof *out);
  if (!w->cuda) return GOL_OK;
  Slab& s = w->slabs[(size_t)li];
  Group& g = w->groups[(size_t)s.group];
  CU_TRY(cudaSetDevice(s.device));
  CU_TRY(cudaStreamSynchronize(g.stream));
  SlabState* d = g.d_states + s.index_in_group;
  SlabState h;
  CU_TRY(cudaMemcpy(&h, d, sizeof h, cudaMemcpyDeviceToHost));
  for (int i = 0; i < 2; ++i) {
    out->halo_wait_ns[i] = h.wait_ns[i];
    out->halo_waits[i] = h.wait_count[i];
    out->task_ns[i] = h.task_ns[i];
    out->tasks[i] = h.tasks[i];
  }
  out->dep_wait_ns = h.dep_wait_ns;
  if (reset) {
    const size_t off = offsetof(SlabState, wait_ns);
    CU_TRY(cudaMemset((char*)d + off, 0, sizeof(SlabState) - off));
  }
  return GOL_OK;
}

void* gol_world_device_ptr(gol_world_t* w, int li) {
  if (check_local(w, li) != GOL_OK) return nullptr;
  Slab& s = w->slabs[(size_t)li];
  if (!w->cuda) return s.cbuf[s.ccur].data();
  return cur_interior(w, s);
}

int gol_world_device_of(gol_world_t* w, int li) {
  if (check_local(w, li) != GOL_OK) return -1;
  return w->cuda ? w->slabs[(size_t)li].device : -1;
}

int gol_world_cpu_get_edge(gol_world_t* w, int li, int bottom, uint64_t* row) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (w->cuda || !row) return set_error(GOL_ERR_ARG, "cpu_get_edge needs a CPU world and a buffer");
  Slab& s = w->slabs[(size_t)li];
  const uint64_t* src = s.cbuf[s.ccur].data() + (bottom ? (w->H - 1) * w->stride : 0);
  memcpy(row, src, (size_t)w->stride * 8);
  return GOL_OK;
}

int gol_world_cpu_set_ghost(gol_world_t* w, int li, int bottom, const uint64_t* row) {
  int rc = check_local(w, li);
  if (rc) return rc;
  if (w->cuda || !row) return set_error(GOL_ERR_ARG, "cpu_set_ghost needs a CPU world and a buffer");
  memcpy(w->slabs[(size_t)li].cghost[bottom ? 1 : 0].data(), row, (size_t)w->stride * 8);
  return GOL_OK;
}

int gol_world_cpu_step_local(gol_world_t* w) {
  if (!w || w->cuda) return set_error(GOL_ERR_ARG, "cpu_step_local needs a CPU world");
  for (auto& s : w->slabs) cpu_step_slab(w, s);
  return GOL_OK;
}

int gol_world_dump_text(gol_world_t* w, int li, const char* dir) {
  int rc = check_local(w, li);
  if (rc) return rc;
  Slab& s = w->slabs[(size_t)li];
  char name[64];
  gol_rank_file_name(name, sizeof name, s.rank, w->cfg.nranks);
  std::string path = (dir && *dir) ? std::string(dir) + "/" + name : std::string(name);
  if (!w->cuda) {
    std::vector<uint8_t> cells((size_t)w->W * (size_t)w->H);
    rc = gol_unpack_words(s.cbuf[s.ccur].data(), cells.data(), w->W, w->H);
    if (rc) return rc;
    return gol_write_rank_file(path.c_str(), cells.data(), w->W, w->H, s.rank);
  }
  // device-formatted rows, streamed in chunks
  nvtxRangePushA("gol_world_dump_text");
  struct PopRange { ~PopRange() { nvtxRangePop(); } } pop_range;
  Group& g = w->groups[(size_t)s.group];
  CU_TRY(cudaSetDevice(s.device));
  const int64_t pitch = (int64_t)align_up((size_t)(kTextCellOff + 2 * w->W + 1), 16);
  // two slots of 32 MiB: chunk i+1 is formatted on the device and copied to the host while chunk i
  // is written to the file
  int64_t chunk_rows = std::max<int64_t>(1, (int64_t)(32ll << 20) / pitch);
  chunk_rows = std::min(chunk_rows, w->H);
  const size_t slot_bytes = (size_t)(chunk_rows * pitch);
  char* d_text[2] = {nullptr, nullptr};
  char* h_text[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  FILE* fp = nullptr;
  auto cleanup = [&]() {
    for (int k = 0; k < 2; ++k) {
      if (d_text[k]) {
        debug_bounds_unregister(d_text[k]);
        cudaFree(d_text[k]);
      }
      if (h_text[k]) cudaFreeHost(h_text[k]);
      if (ev[k]) cudaEventDestroy(ev[k]);
    }
  };
  cudaError_t e = cudaSuccess;
  for (int k = 0; k < 2 && e == cudaSuccess; ++k) {
    e = cudaMalloc((void**)&d_text[k], slot_bytes);
    if (e == cudaSuccess) debug_bounds_register(d_text[k], slot_bytes);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&h_text[k], slot_bytes);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    cleanup();
    return set_error(GOL_ERR_CUDA, "text staging allocation failed: %s", cudaGetErrorString(e));
  }
  fp = fopen(path.c_str(), "w");
  if (!fp) {
    cleanup();
    return set_error(GOL_ERR_IO, "cannot open %s: %s", path.c_str(), strerror(errno));
  }
  rc = write_rank_header(fp, s.rank);
  const uint64_t* rows = cur_interior(w, s);
  const int64_t nchunks = (w->H + chunk_rows - 1) / chunk_rows;
  auto issue = [&](int64_t c) -> cudaError_t {
    const int k = (int)(c & 1);
    const int64_t y0 = c * chunk_rows, n = std::min(chunk_rows, w->H - y0);
    cudaError_t ee = launch_format_rows(rows + y0 * w->stride, w->stride, w->W, n, w->H * s.rank + y0, d_text[k], pitch,
                                        g.stream);
    if (ee == cudaSuccess)
      ee = cudaMemcpyAsync(h_text[k], d_text[k], (size_t)(n * pitch), cudaMemcpyDeviceToHost, g.stream);
    if (ee == cudaSuccess) ee = cudaEventRecord(ev[k], g.stream);
    return ee;
  };
  e = issue(0);
  for (int64_t c = 0; c < nchunks && rc == GOL_OK && e == cudaSuccess; ++c) {
    if (c + 1 < nchunks) e = issue(c + 1);   // its slot was written out two chunks ago
    if (e == cudaSuccess) e = cudaEventSynchronize(ev[c & 1]);
    if (e != cudaSuccess) break;
    const int64_t y0 = c * chunk_rows, n = std::min(chunk_rows, w->H - y0);
    const int64_t first_global = w->H * s.rank + y0;
    for (int64_t r = 0; r < n; ++r) {
      char pre[40];
      const int plen = snprintf(pre, sizeof pre, "Row %2d: ", (int)(first_global + r));
      const char* start = h_text[c & 1] + r * pitch + kTextCellOff - plen;
      const size_t len = (size_t)plen + (size_t)(2 * w->W) + 1;
      if (fwrite(start, 1, len, fp) != len) {
        rc = set_error(GOL_ERR_IO, "short write to %s", path.c_str());
        break;
      }
    }
  }
  if (e != cudaSuccess) rc = set_error(GOL_ERR_CUDA, "format_rows: %s", cudaGetErrorString(e));
  cudaStreamSynchronize(g.stream);
  if (fclose(fp) != 0 && rc == GOL_OK) rc = set_error(GOL_ERR_IO, "close of %s failed", path.c_str());
  cleanup();
  return rc;
}

// ---- stateless device ops -------------------------------------------------------
int gol_cuda_pack(const uint8_t* d_cells, uint64_t* d_words, int64_t W, int64_t H, void* stream) {
  if (!d_cells || !d_words || W < 1 || H < 1) return set_error(GOL_ERR_ARG, "gol_cuda_pack: bad argument");
  debug_bounds_register(d_cells, (size_t)W * (size_t)H);
  debug_bounds_register(d_words, (size_t)H * (size_t)row_stride_words(W) * 8);
  CU_TRY(launch_pack(d_cells, d_words, W, H, row_stride_words(W), (cudaStream_t)stream));
  return GOL_OK;
}

int gol_cuda_unpack(const uint64_t* d_words, uint8_t* d_cells, int64_t W, int64_t H, void* stream) {
  if (!d_cells || !d_words || W < 1 || H < 1) return set_error(GOL_ERR_ARG, "gol_cuda_unpack: bad argument");
  debug_bounds_register(d_cells, (size_t)W * (size_t)H);
  debug_bounds_register(d_words, (size_t)H * (size_t)row_stride_words(W) * 8);
  CU_TRY(launch_unpack(d_words, d_cells, W, H, row_stride_words(W), (cudaStream_t)stream));
  return GOL_OK;
}

size_t gol_cuda_population_scratch_bytes(void) { return kPopBlocks * sizeof(unsigned long long); }

int gol_cuda_population(const uint64_t* d_words, int64_t W, int64_t H, void* d_scratch, uint64_t* d_out,
                        void* stream) {
  if (!d_words || !d_scratch || !d_out || W < 1 || H < 1)
    return set_error(GOL_ERR_ARG, "gol_cuda_population: bad argument");
  debug_bounds_register(d_words, (size_t)H * (size_t)row_stride_words(W) * 8);
  debug_bounds_register(d_scratch, gol_cuda_population_scratch_bytes());
  debug_bounds_register(d_out, 8);
  CU_TRY(launch_population(d_words, words_per_row(W), row_stride_words(W), H, (unsigned long long*)d_scratch,
                           (unsigned long long*)d_out, (cudaStream_t)stream));
  return GOL_OK;
}

int gol_cuda_step_torus(const uint64_t* d_in, uint64_t* d_out, int64_t W, int64_t H, int threads, void* stream) {
  if (!d_in || !d_out || d_in == d_out || W < 1 || H < 1)
    return set_error(GOL_ERR_ARG, "gol_cuda_step_torus: bad argument");
  if (threads <= 0) threads = 256;
  if (threads > 1024) return set_error(GOL_ERR_ARG, "threads per block must be <= 1024");
  const int64_t st = row_stride_words(W);
  debug_bounds_register(d_in, (size_t)H * (size_t)st * 8);
  debug_bounds_register(d_out, (size_t)H * (size_t)st * 8);
  CU_TRY(launch_step_rows(d_in, d_out, d_in + (H - 1) * st, d_in, W, H, st, threads, (cudaStream_t)stream));
  return GOL_OK;
}

int gol_cuda_step_rows(const uint64_t* d_in, uint64_t* d_out, const uint64_t* d_top, const uint64_t* d_bottom,
                       int64_t W, int64_t H, int threads, void* stream) {
  if (!d_in || !d_out || !d_top || !d_bottom || d_in == d_out || W < 1 || H < 1)
    return set_error(GOL_ERR_ARG, "gol_cuda_step_rows: bad argument");
  if (threads <= 0) threads = 256;
  if (threads > 1024) return set_error(GOL_ERR_ARG, "threads per block must be <= 1024");
  debug_bounds_register(d_in, (size_t)H * (size_t)row_stride_words(W) * 8);
  debug_bounds_register(d_out, (size_t)H * (size_t)row_stride_words(W) * 8);
  debug_bounds_register(d_top, (size_t)row_stride_words(W) * 8);
  debug_bounds_register(d_bottom, (size_t)row_stride_words(W) * 8);
  CU_TRY(launch_step_rows(d_in, d_out, d_top, d_bottom, W, H, row_stride_words(W), threads, (cudaStream_t)stream));
  return GOL_OK;
}

}  // extern "C"
