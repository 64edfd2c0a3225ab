// kernels.cuh — host-callable launch wrappers of every device kernel.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "internal.h"

namespace gol {

// Kernel inventory ("new-framework kernel list"):
//   life_fused<K>      N-K1  register-streaming K-generation pass + P2P halo push + flags
//   life_generic       N-K2  one thread per word, any shape, 1 generation, same protocol
//   life_flow<K>       N-K3  whole step() in one launch: persistent warps pull (pass, strip) tickets and
//                            wait on per-strip progress counters instead of a grid barrier
//   life_resident      N-K3b whole step() with every CTA's strip resident in shared memory (small slabs):
//                            cp.async.bulk strip load / store, halo rows over DSMEM inside clusters
//   pattern_init       N-K4  patterns 0..6 straight into packed words
//   halo_prime         N-K5  push my edge rows into the neighbours' recv buffers
//   format_rows        N-K6  bits -> "0 1 0 ..." text rows on the device
//   pack / unpack      N-K7  byte cells <-> packed words
//   population         (new) popcount reduction

struct FusedLimits {
  int max_threads;   // __launch_bounds__ of the instantiation
};
FusedLimits fused_limits(int K);
bool fused_k_supported(int K);

// Resident warps per SM for the fused kernel at a block size (occupancy query).
int fused_resident_warps_per_sm(int K, int threads, int device);
int flow_max_ctas(int K, int threads, int device);

cudaError_t launch_fused(int K, SlabState* d_states, int nslabs, const PassParams& p, int threads,
                         cudaStream_t s);
// TMA-fed variant (K <= 4, ws % 16 == 0, ws >= 144): interior strips fed by 2-D tensor copies
// (cp.async.bulk.tensor) into an mbarrier ring.  d_tmaps: device array [nslabs][2] of tensor
// maps over buf[0] / buf[1] of each slab (box fused_tma_box_words(nw) x fused_tma_box_rows()).
bool fused_tma_supported(int K, int64_t ws);
int fused_tma_warps(int col_groups);   // compute warps per CTA (3 or 4)
int fused_tma_box_words(int nw);
int fused_tma_box_rows();
int fused_tma_rank();  // 2 or 3 (GOL_TMA_RANK), see make_tensor_maps
cudaError_t launch_fused_tma(int K, SlabState* d_states, int nslabs, const PassParams& p,
                             const CUtensorMap* d_tmaps, cudaStream_t s);
cudaError_t launch_generic(SlabState* d_states, int nslabs, const PassParams& p, int threads,
                           int64_t max_words, cudaStream_t s);
cudaError_t launch_prime(SlabState* d_states, int nslabs, int depth, int64_t max_ws, cudaStream_t s);

// Flow (persistent) kernel: `npasses` passes of K generations in one launch, dependency-driven
// (FlowParams in internal.h).  `nctas` CTAs of `threads` threads; ctl / progress must be zero before
// the first launch (the kernel leaves them zeroed).
cudaError_t launch_flow(int K, SlabState* d_states, const FlowParams& fp, int threads, int nctas,
                        cudaStream_t s);

// Resident kernel (small slabs, strips kept in shared memory, thread-block clusters; ResidentParams in
// internal.h).  smem = resident_smem_bytes(...); resident_max_ctas = CTAs that can be co-resident.
size_t resident_smem_bytes(int rows_alloc, int64_t ws, int K);
int resident_max_ctas(int threads, size_t smem, int cluster, int device);
cudaError_t launch_resident(SlabState* d_states, const ResidentParams& rp, int threads, size_t smem, cudaStream_t s);

cudaError_t launch_pattern_init(uint64_t* d_rows, int64_t stride, const PatternArgs& a,
                                cudaStream_t s);
cudaError_t launch_pack(const uint8_t* d_cells, uint64_t* d_words, int64_t W, int64_t H,
                        int64_t stride, cudaStream_t s);
cudaError_t launch_unpack(const uint64_t* d_words, uint8_t* d_cells, int64_t W, int64_t H,
                          int64_t stride, cudaStream_t s);
constexpr int kPopBlocks = 1184;  // fixed (it sizes the caller-visible scratch); rows are taken grid-stride, any SM count
constexpr int kPopThreads = 256;
cudaError_t launch_population(const uint64_t* d_words, int64_t ws, int64_t stride, int64_t H,
                              unsigned long long* d_partials /* kPopBlocks */,
                              unsigned long long* d_out, cudaStream_t s);
// Text rows: row r of the chunk starts at d_text + r * pitch; cells begin at
// byte `cell_off` (16-byte aligned), the prefix "Row %2d: " ends right before it.
constexpr int kTextCellOff = 32;
cudaError_t launch_format_rows(const uint64_t* d_rows, int64_t stride, int64_t W, int64_t nrows,
                               int64_t first_global_row, char* d_text, int64_t pitch,
                               cudaStream_t s);
cudaError_t launch_step_rows(const uint64_t* d_in, uint64_t* d_out, const uint64_t* d_top,
                             const uint64_t* d_bottom, int64_t W, int64_t H, int64_t stride,
                             int threads, cudaStream_t s);

// Bounds-checking build (-DGOL_DEBUG_BOUNDS=1, libgol_b200_debug.so): registry of the address ranges
// kernels may touch on the CURRENT device.  No-ops that report zero violations in the normal build.
bool debug_bounds_enabled();
int debug_bounds_register(const void* p, size_t n);
int debug_bounds_unregister(const void* p);
int debug_bounds_query(unsigned long long* violations, unsigned long long* first_bad, int* line, int reset);
int debug_bounds_poke(const void* p);   // one checked load of *p (tests that the detector fires)

void count_launch(uint64_t n);
uint64_t launch_count_value();

}  // namespace gol
