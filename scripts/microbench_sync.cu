// microbench_sync.cu - what a hand-off between two CTAs costs on this GPU (cycles and ns).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bin/microbench_sync scripts/microbench_sync.cu
//
// Measured because the design decisions around halo hand-offs rest on these numbers (DESIGN.md
// section 3): a cluster barrier, a flag-in-data word through distributed shared memory, a
// flag-in-data word through L2, and a release/acquire flag through L2.  Every ping-pong figure is
// a ROUND TRIP (A -> B -> A) averaged over kIters iterations; halve it for one direction.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

constexpr int kIters = 2000;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa(uint32_t a, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
  return r;
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// (a) barrier.cluster arrive + wait, all CTAs of the cluster, back to back
__global__ void k_cluster_barrier(long long* out) {
  cluster_sync();
  const long long t0 = clock64();
  for (int i = 0; i < kIters; ++i) cluster_sync();
  const long long t1 = clock64();
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = (t1 - t0) / kIters;
}

// (b) one 8-byte word with the iteration number in it, written into the partner CTA's shared memory
__global__ void k_dsmem_pingpong(long long* out) {
  __shared__ unsigned long long box;
  if (threadIdx.x == 0) box = 0;
  cluster_sync();
  const uint32_t me = cluster_rank(), other = me ^ 1u;
  const uint32_t remote = mapa(smem_u32(&box), other), local = smem_u32(&box);
  long long t0 = 0, t1 = 0;
  if (threadIdx.x == 0) {
    t0 = clock64();
    for (unsigned long long i = 1; i <= kIters; ++i) {
      unsigned long long v;
      if (me == 0) {
        asm volatile("st.shared::cluster.u64 [%0], %1;" ::"r"(remote), "l"(i) : "memory");
        do asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(v) : "r"(local) : "memory");
        while (v != i);
      } else {
        do asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(v) : "r"(local) : "memory");
        while (v != i);
        asm volatile("st.shared::cluster.u64 [%0], %1;" ::"r"(remote), "l"(i) : "memory");
      }
    }
    t1 = clock64();
  }
  cluster_sync();
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = (t1 - t0) / kIters;
}

// (c) the same word through global memory (L2), two CTAs on different SMs, no fence
// (d) a separate flag with release / acquire at gpu scope next to a data word
template <bool FENCED>
__global__ void k_l2_pingpong(unsigned long long* box, long long* out) {
  const int me = blockIdx.x;   // 0 or 1; box[0] is written by CTA 1 and read by CTA 0, box[16] the other way
  unsigned long long* mine = box + (me ? 16 : 0);
  unsigned long long* theirs = box + (me ? 0 : 16);
  long long t0 = 0, t1 = 0;
  if (threadIdx.x == 0) {
    t0 = clock64();
    for (unsigned long long i = 1; i <= kIters; ++i) {
      unsigned long long v;
      if (me == 0) {
        if (FENCED) asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(theirs), "l"(i) : "memory");
        else asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(theirs), "l"(i) : "memory");
        do {
          if (FENCED) asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
          else asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
        } while (v != i);
      } else {
        do {
          if (FENCED) asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
          else asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
        } while (v != i);
        if (FENCED) asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(theirs), "l"(i) : "memory");
        else asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(theirs), "l"(i) : "memory");
      }
    }
    t1 = clock64();
  }
  if (me == 0 && threadIdx.x == 0) out[0] = (t1 - t0) / kIters;
}

// (e) block barrier, 256 threads
__global__ void k_block_barrier(long long* out) {
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < kIters; ++i) __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[0] = (t1 - t0) / kIters;
}

template <class F>
static long long run_cluster(F kernel, int cluster, int threads, long long* d_out) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)cluster);
  cfg.blockDim = dim3((unsigned)threads);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cluster;
  attr[0].val.clusterDim.y = attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  void* args[] = {(void*)&d_out};
  long long best = -1;
  for (int rep = 0; rep < 5; ++rep) {
    if (cudaLaunchKernelExC(&cfg, (const void*)kernel, args) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
      printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError()));
      return -1;
    }
    long long v = 0;
    cudaMemcpy(&v, d_out, sizeof v, cudaMemcpyDeviceToHost);
    if (best < 0 || v < best) best = v;
  }
  return best;
}

int main() {
  int dev = 0, khz = 0;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  const double ns_per_cycle = 1e6 / (double)khz;
  printf("%s, %d SMs, SM clock %.0f MHz (clock64 cycles -> ns at that clock)\n", prop.name, prop.multiProcessorCount,
         khz * 1e-3);
  long long* d_out = nullptr;
  unsigned long long* d_box = nullptr;
  cudaMalloc(&d_out, 64);
  cudaMalloc(&d_box, 512);
  auto show = [&](const char* what, long long cyc) {
    if (cyc >= 0) printf("%-78s %7lld cycles  %8.1f ns\n", what, cyc, cyc * ns_per_cycle);
  };
  for (int cs : {2, 4, 8}) {
    char label[128];
    snprintf(label, sizeof label, "barrier.cluster arrive.release + wait.acquire, cluster of %d x 256 threads", cs);
    show(label, run_cluster(k_cluster_barrier, cs, 256, d_out));
  }
  show("DSMEM ping-pong round trip (st.shared::cluster.u64 + ld.volatile.shared poll), cluster of 2",
       run_cluster(k_dsmem_pingpong, 2, 32, d_out));
  long long best[2] = {-1, -1};
  for (int rep = 0; rep < 5; ++rep) {
    for (int f = 0; f < 2; ++f) {
      cudaMemset(d_box, 0, 512);
      // two CTAs of one launch: co-resident on any GPU with more than one SM
      if (f) k_l2_pingpong<true><<<2, 32>>>(d_box, d_out);
      else k_l2_pingpong<false><<<2, 32>>>(d_box, d_out);
      if (cudaDeviceSynchronize() != cudaSuccess) {
        printf("l2 ping-pong failed: %s\n", cudaGetErrorString(cudaGetLastError()));
        return 1;
      }
      long long v = 0;
      cudaMemcpy(&v, d_out, sizeof v, cudaMemcpyDeviceToHost);
      if (best[f] < 0 || v < best[f]) best[f] = v;
    }
  }
  show("L2 ping-pong round trip, flag in data (st.volatile / ld.volatile.global), 2 CTAs on 2 SMs", best[0]);
  show("L2 ping-pong round trip, st.release.gpu / ld.acquire.gpu, 2 CTAs on 2 SMs", best[1]);
  long long bb = -1;
  for (int rep = 0; rep < 5; ++rep) {
    k_block_barrier<<<1, 256>>>(d_out);
    cudaDeviceSynchronize();
    long long v = 0;
    cudaMemcpy(&v, d_out, sizeof v, cudaMemcpyDeviceToHost);
    if (bb < 0 || v < bb) bb = v;
  }
  show("__syncthreads, 256 threads", bb);
  cudaFree(d_out);
  cudaFree(d_box);
  return 0;
}
