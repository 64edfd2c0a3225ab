// noalloc.cu — gol_cuda_run_torus_noalloc: stepping for callers that must not allocate.
//
// Who needs this: code that runs where device allocation is off the table — a plug-in
// living inside a host application that has handed every byte of device memory to its own
// pool, a real-time loop that may not take the driver's allocation lock, or a library
// called after the process has been told "no more cudaMalloc".  gol_world_create and the
// stateless helpers obtain their working memory (slab state, halo recv rows, flags) from
// cudaMalloc or from the caller; this entry point obtains it from nowhere: the working
// memory is a fixed array compiled into the library's device image.
//
// The kernels are the ordinary step kernels (life_fused<K> when the width is a multiple of
// 64 and the grid is tall enough, life_generic otherwise) launched through the same
// wrappers as everywhere else; only the home of their scratch differs.
#include <cuda_runtime.h>
#include <stdlib.h>

#include "../include/gol_b200.h"
#include "internal.h"
#include "kernels.cuh"

namespace gol {
namespace {

// Scratch held by every GPU that loads this library, from load until process exit.
// Layout: [SlabState | flag top | flag bottom | recv top p0 | recv top p1 | recv bottom p0 | recv bottom p1]
constexpr size_t kHeaderBytes = 1024;  // SlabState (< 512 B) + two flags 128 B apart
constexpr int kDepth = 4;              // halo rows per exchange = generations per fused pass
static __device__ __align__(256) unsigned char g_noalloc_scratch[GOL_NOALLOC_SCRATCH_BYTES];

static_assert(sizeof(SlabState) + 256 <= kHeaderBytes, "header too small");
static_assert(GOL_NOALLOC_MAX_WIDTH / 64 * 8 * 4 * kDepth + kHeaderBytes <= GOL_NOALLOC_SCRATCH_BYTES,
              "scratch array cannot hold the recv rows of the widest supported grid");

constexpr int kMaxDevices = 64;
unsigned char* g_scratch_addr[kMaxDevices];  // device address of g_noalloc_scratch, per device

// Writes the slab description into the scratch header and zeroes everything the step kernels
// expect to start at zero (pass / exchange counters, launch bookkeeping, both halo flags).
// The recv rows need no clearing: every row read in a call is written earlier in that call.
__global__ void noalloc_setup_kernel(SlabState s, unsigned char* scratch) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    *s.flag[0] = 0ull;
    *s.flag[1] = 0ull;
    *reinterpret_cast<SlabState*>(scratch) = s;
  }
}

int largest_fused_k(int64_t limit) {
  static const int ks[] = {4, 3, 2, 1};
  for (int k : ks)
    if (k <= limit) return k;
  return 1;
}

}  // namespace
}  // namespace gol

using namespace gol;

extern "C" int gol_cuda_run_torus_noalloc(uint64_t* d_a, uint64_t* d_b, int64_t W, int64_t H, int64_t gens,
                                          int threads_per_block, void* stream, int* result_in_b) {
  if (!d_a || !d_b || d_a == d_b || W < 1 || H < 1 || gens < 0)
    return set_error(GOL_ERR_ARG, "gol_cuda_run_torus_noalloc: bad argument");
  if (W > GOL_NOALLOC_MAX_WIDTH)
    return set_error(GOL_ERR_UNSUPPORTED,
                     "gol_cuda_run_torus_noalloc: width %lld exceeds the built-in scratch (at most %d columns); "
                     "use gol_world_* for wider grids",
                     (long long)W, GOL_NOALLOC_MAX_WIDTH);
  if (threads_per_block <= 0) threads_per_block = 128;
  if (threads_per_block > 1024) return set_error(GOL_ERR_ARG, "threads per block must be <= 1024");
  cudaStream_t s = (cudaStream_t)stream;

  // One-time set-up per device (first call on that device): resolve the device address of
  // the scratch array.  In host code the array's name is only a handle for the runtime, not
  // something a kernel can dereference.
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_error(GOL_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return set_error(GOL_ERR_UNSUPPORTED, "device index %d not supported", dev);
  unsigned char* scratch = __atomic_load_n(&g_scratch_addr[dev], __ATOMIC_ACQUIRE);
  if (!scratch) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return set_error(GOL_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
    scratch = static_cast<unsigned char*>(p);
    __atomic_store_n(&g_scratch_addr[dev], scratch, __ATOMIC_RELEASE);
  }
  if (result_in_b) *result_in_b = 0;
  if (gens == 0) return GOL_OK;

  const int64_t ws = words_per_row(W), stride = row_stride_words(W);
  const bool fused = (W % 64) == 0 && H >= kDepth;
  const size_t recv_bytes = (size_t)kDepth * (size_t)stride * 8;

  SlabState st;
  memset(&st, 0, sizeof st);
  st.W = W;
  st.H = H;
  st.ws = ws;
  st.stride = stride;
  st.G = 0;  // caller buffers have no margin rows: every pass takes its halo from recv[]
  st.rank = 0;
  st.nranks = 1;
  st.local_ring = 3;  // a one-rank ring: both neighbours are this slab itself
  st.buf[0] = d_a;
  st.buf[1] = d_b;
  st.flag[0] = reinterpret_cast<unsigned long long*>(scratch + 512);
  st.flag[1] = reinterpret_cast<unsigned long long*>(scratch + 512 + 128);
  unsigned char* r = scratch + kHeaderBytes;
  for (int side = 0; side < 2; ++side)
    for (int p = 0; p < 2; ++p) {
      st.recv[side][p] = reinterpret_cast<uint64_t*>(r);
      r += recv_bytes;
    }
  // a one-rank ring: my first rows are my own bottom halo, my last rows my own top halo
  for (int p = 0; p < 2; ++p) {
    st.peer_recv[0][p] = st.recv[1][p];
    st.peer_recv[1][p] = st.recv[0][p];
  }
  st.peer_flag[0] = st.flag[1];
  st.peer_flag[1] = st.flag[0];
  {
    // same knob as gol_world_create: how long a kernel waits for halo words before giving up
    const char* e = getenv("GOL_WATCHDOG_MS");
    const unsigned long long ms = e && atoll(e) > 0 ? (unsigned long long)atoll(e) : 10000ull;
    st.watchdog_ns = ms * 1000ull * 1000ull;
  }

  if (debug_bounds_enabled()) {   // bounds-checking build only: the registry is host-side bookkeeping
    debug_bounds_register(scratch, GOL_NOALLOC_SCRATCH_BYTES);
    debug_bounds_register(d_a, (size_t)H * (size_t)stride * 8);
    debug_bounds_register(d_b, (size_t)H * (size_t)stride * 8);
  }
  count_launch(1);
  noalloc_setup_kernel<<<1, 32, 0, s>>>(st, scratch);
  e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(GOL_ERR_CUDA, "noalloc set-up launch: %s", cudaGetErrorString(e));
  SlabState* d_state = reinterpret_cast<SlabState*>(scratch);

  int64_t passes = 0;
  if (!fused) {
    const int t = threads_per_block;
    e = launch_prime(d_state, 1, 1, ws, s);
    for (int64_t g = 0; g < gens && e == cudaSuccess; ++g) {
      PassParams p;
      memset(&p, 0, sizeof p);
      p.ext_in = 1;
      p.k = 1;
      p.consume = 1;
      p.push = (g + 1 < gens) ? 1 : 0;
      e = launch_generic(d_state, 1, p, t, H * ws, s);
      ++passes;
    }
  } else {
    int t = (threads_per_block / 32) * 32;
    if (t < 32) t = 32;
    if (t > 256) t = 256;
    int64_t left = gens;
    int k = largest_fused_k(left);
    e = launch_prime(d_state, 1, k, ws, s);
    while (left > 0 && e == cudaSuccess) {
      const int knext = left - k > 0 ? largest_fused_k(left - k) : 0;
      PassParams p;
      memset(&p, 0, sizeof p);
      p.ext_in = k;
      p.k = k;
      p.consume = 1;
      p.push = knext;
      p.strip_rows = 64;
      p.col_groups = (int)((ws + 29) / 30);
      p.nstrips = (int)((H + p.strip_rows - 1) / p.strip_rows);
      e = launch_fused(k, d_state, 1, p, t, s);
      left -= k;
      k = knext;
      ++passes;
    }
  }
  if (e != cudaSuccess) return set_error(GOL_ERR_CUDA, "noalloc step launch: %s", cudaGetErrorString(e));
  if (result_in_b) *result_in_b = (int)(passes & 1);
  return GOL_OK;
}
