// kernels.cu — every device kernel of the core, written for sm_100a.
//
// The hot kernel is life_fused<K>: one warp streams a strip of rows top to
// bottom through K pipelined generation stages held entirely in registers.
// Each lane owns one 64-cell word column; the two x-neighbour bits a stage
// needs come from the adjacent lanes by warp shuffle, lanes 0 and 31 are halo
// lanes whose columns belong to the neighbouring warp task.  There is no shared
// memory and no block barrier on the hot path: the kernel is bound by the
// integer (LOP3/SHF) pipes, which is the roofline for bit-sliced Life once the
// 16384^2 slab is L2-resident.
//
// The halo exchange of the reference (MPI_Irecv/Isend/Wait, gol-main.c:97-111)
// is fused into the same kernel: the warps that produce a slab's first / last
// rows also store them into the ring neighbour's recv buffer through a peer
// (NVLink P2P or CUDA-IPC) pointer and publish a word count with
// red.release.sys; the warps that need a fresh halo spin on ld.acquire.sys.
// Those warps are scheduled first (lowest task ids).
#include <stdlib.h>
#include <string.h>

#include "kernels.cuh"

#ifndef GOL_FUSED_COHERENT
#define GOL_FUSED_COHERENT 0   // 1: per-pass fused kernel also uses ld/st.global.cg for interior rows
#endif

#ifndef GOL_DEBUG_BOUNDS
#define GOL_DEBUG_BOUNDS 0     // 1: libgol_b200_debug.so - every global access of every kernel is checked
#endif

namespace gol {

namespace {

#if GOL_DEBUG_BOUNDS
// Memory-safety build (compute-sanitizer is closed on the GPU pool this was developed on): the
// host registers every allocation a kernel may touch - slab buffers, halo mailboxes (own and
// mapped neighbours'), text / byte scratch, partial sums, caller buffers of the stateless calls -
// and every global load, store and atomic of every kernel first looks its address range up.  A
// miss is counted, its address and source line are kept, and the access is skipped.
constexpr int kDbgMaxRegions = 96;
struct DebugRegions {
  unsigned long long lo[kDbgMaxRegions], hi[kDbgMaxRegions];
  int n;
  int first_line;
  unsigned long long violations;
  unsigned long long first_bad;
};
__device__ DebugRegions g_dbg;
__device__ __noinline__ bool dbg_ok(const void* p, long long bytes, int line) {
  const unsigned long long a = (unsigned long long)p, b = a + (unsigned long long)bytes;
  const int n = g_dbg.n;
  for (int i = 0; i < n; ++i)
    if (a >= g_dbg.lo[i] && b <= g_dbg.hi[i]) return true;
  if (atomicAdd(&g_dbg.violations, 1ull) == 0ull) {
    g_dbg.first_bad = a;
    g_dbg.first_line = line;
  }
  return false;
}
#define GOL_CHK(p, n) dbg_ok((const void*)(p), (long long)(n), __LINE__)
// shared-memory word offset `off` (n words) inside a dynamic shared-memory array of `limit` words
__device__ __noinline__ bool dbg_smem_bad(long long off, int line) {
  if (atomicAdd(&g_dbg.violations, 1ull) == 0ull) {
    g_dbg.first_bad = 0x5ed0000000000000ull | (unsigned long long)(off & 0xffffffffll);   // tag: shared-memory offset
    g_dbg.first_line = line;
  }
  return false;
}
#define GOL_SCHK(off, n, limit) \
  (((long long)(off) >= 0 && (long long)(off) + (long long)(n) <= (long long)(limit)) ? true : dbg_smem_bad((long long)(off), __LINE__))
#else
#define GOL_CHK(p, n) (true)
#define GOL_SCHK(off, n, limit) (true)
#endif

constexpr unsigned kFull = 0xffffffffu;
constexpr int kValidLanes = 30;  // lanes 1..30 own output columns

// ---- system-scope ordering primitives --------------------------------------
__device__ __forceinline__ void red_release_sys_add(unsigned long long* p, unsigned long long v) {
  if (!GOL_CHK(p, 8)) return;
  asm volatile("red.release.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_release_gpu_add(unsigned long long* p, unsigned long long v) {
  if (!GOL_CHK(p, 8)) return;
  asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Publish `v` more halo words to a ring neighbour: gpu scope when the neighbour's slab is on this
// device, system scope (NVLink peer) otherwise.
__device__ __forceinline__ void publish_words(const SlabState* st, int side, unsigned long long v) {
  if ((st->local_ring >> side) & 1) red_release_gpu_add(st->peer_flag[side], v);
  else red_release_sys_add(st->peer_flag[side], v);
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// ---- flag-in-data halo words (SlabState::ll_recv) ------------------------------------------------
__device__ __forceinline__ void ll_store(uint64_t* p, uint64_t v, uint32_t seq) {
  if (!GOL_CHK(p, 16)) return;
  const unsigned long long a = (v & 0xffffffffull) | ((unsigned long long)seq << 32);
  const unsigned long long b = (v >> 32) | ((unsigned long long)seq << 32);
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
// Warp-uniform: every lane polls its own word, the warp leaves together (a lane-divergent spin
// would leave the warp split for the shuffles that follow).
__device__ __noinline__ uint64_t ll_load(SlabState* st, const uint64_t* p, uint32_t seq, int side) {
  unsigned long long a, b;
  unsigned long long t0 = 0;
  unsigned ns = 0;
  if (!GOL_CHK(p, 16)) return 0;
  for (;;) {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
    const bool ok = (uint32_t)(a >> 32) == seq && (uint32_t)(b >> 32) == seq;
    if (__all_sync(0xffffffffu, ok)) break;
    if (!t0) t0 = globaltimer_ns();
    if (ns) __nanosleep(ns);
    ns = ns < 8 ? ns + 1 : (ns < 640 ? ns << 1 : ns);
    if (globaltimer_ns() - t0 > st->watchdog_ns || *(volatile int*)&st->error) {
      if ((threadIdx.x & 31) == 0) atomicCAS(&st->error, 0, 1 + side);
      break;
    }
  }
  if (t0 && (threadIdx.x & 31) == 0) {
    atomicAdd(&st->wait_ns[side], globaltimer_ns() - t0);
    atomicAdd(&st->wait_count[side], 1ull);
  }
  return (a & 0xffffffffull) | (b << 32);
}

// Polling loads are relaxed; one acquire fence follows the successful poll.  (An acquire load or
// fence at gpu / sys scope carries an L1 invalidate - SASS CCTL.IVALL - and polling with it slowed
// every other warp of the SM by up to 50 %: measured with the flow kernel, 16384^2, B200.)
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p) {
  if (!GOL_CHK(p, 8)) return ~0ull;
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int ld_relaxed_gpu_u32(const unsigned int* p) {
  if (!GOL_CHK(p, 4)) return ~0u;
  unsigned int v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ unsigned long long ld_relaxed_gpu(const unsigned long long* p) {
  if (!GOL_CHK(p, 8)) return ~0ull;
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
// Peer / streaming stores: the output of a pass is not re-read by this SM.
__device__ __forceinline__ void st_cg(uint64_t* p, uint64_t v) {
  if (!GOL_CHK(p, 8)) return;
  __stcg((unsigned long long*)p, v);
}
__device__ __forceinline__ uint64_t ld_cg(const uint64_t* p) {
  if (!GOL_CHK(p, 8)) return 0;
  return __ldcg((const unsigned long long*)p);
}

// One lane spins until the neighbour has delivered `target` words (monotonic
// counter).  A dead peer must not hang the GPU: after watchdog_ns the slab is
// marked failed and the pass continues on whatever is in the buffer.  Time spent
// spinning is accumulated per side (the "exposed halo wait" phase counter).
// Called by ALL lanes of the warp (they poll the same word: one request per warp).  A wait run
// by lane 0 alone, followed by __syncwarp(), leaves the warp split as far as the hardware's
// convergence tracking goes, and every later __shfl_sync of the task then takes the slow
// WARPSYNC.COLLECTIVE path (ncu: +57 % instructions executed, task time x2).
__device__ __noinline__ void wait_words(SlabState* st, int side, unsigned long long target) {
  const unsigned long long* f = st->flag[side];
  const bool lane0 = (threadIdx.x & 31) == 0;
  const bool local = (st->local_ring >> side) & 1;   // the writer is on this device: gpu scope
  if ((local ? ld_relaxed_gpu(f) : ld_relaxed_sys(f)) < target) {
    const unsigned long long t0 = globaltimer_ns();
    const unsigned long long limit = st->watchdog_ns;
    unsigned ns = 0;   // poll back to back first: the data is usually a few hundred ns away
    while ((local ? ld_relaxed_gpu(f) : ld_relaxed_sys(f)) < target) {
      if (ns) __nanosleep(ns);
      ns = ns < 8 ? ns + 1 : (ns < 640 ? ns << 1 : ns);
      if (globaltimer_ns() - t0 > limit || *(volatile int*)&st->error) {
        if (lane0) atomicCAS(&st->error, 0, 1 + side);
        break;
      }
    }
    if (lane0) {
      atomicAdd(&st->wait_ns[side], globaltimer_ns() - t0);
      atomicAdd(&st->wait_count[side], 1ull);
    }
  }
  // relaxed load + fence = acquire: the halo rows are visible from here on
  if (local) fence_acq_rel_gpu();
  else fence_acq_rel_sys();
}

// Everything a pass needs about one slab, resolved from SlabState counters.
struct PassCtx {
  const uint64_t* in;     // interior row 0 of the input buffer
  uint64_t* out;          // interior row 0 of the output buffer
  const uint64_t* rtop;   // recv top: row 0 <-> y = -ext_in
  const uint64_t* rbot;   // recv bottom: row 0 <-> y = H
  uint64_t* ptop;         // prev rank's bottom recv (my rows [0,push))
  uint64_t* pbot;         // next rank's top recv (my rows [H-push,H))
  unsigned long long target;
  int64_t H, ws, stride;
  int ext_in, consume, push;
  // flag-in-data halo path (neighbours on other devices)
  SlabState* st;
  const uint64_t* lltop;  // my top ll_recv of this exchange: halo row y (< 0) word w at ((y + ext_in) * stride + w) * 2
  const uint64_t* llbot;  // my bottom ll_recv: row y (>= H) at ((y - H) * stride + w) * 2
  uint64_t* pll_top;      // prev rank's bottom ll_recv for the exchange being pushed
  uint64_t* pll_bot;      // next rank's top ll_recv
  uint32_t seq_in, seq_out;
  int ll;                 // SlabState::ll_ring
};

__device__ __forceinline__ PassCtx make_ctx(SlabState* st, unsigned long long pass,
                                            unsigned long long exch, unsigned long long words_seen,
                                            int ext_in, int consume, int push) {
  PassCtx c;
  const int64_t stride = st->stride;
  c.in = st->buf[pass & 1] + (int64_t)st->G * stride;
  c.out = st->buf[(pass + 1) & 1] + (int64_t)st->G * stride;
  c.rtop = st->recv[0][exch & 1];
  c.rbot = st->recv[1][exch & 1];
  const unsigned long long pe = exch + (consume ? 1ull : 0ull);
  c.ptop = st->peer_recv[0][pe & 1];
  c.pbot = st->peer_recv[1][pe & 1];
  c.target = words_seen + (unsigned long long)ext_in * (unsigned long long)st->ws;
  c.H = st->H;
  c.ws = st->ws;
  c.stride = stride;
  c.ext_in = ext_in;
  c.consume = consume;
  c.push = push;
  c.st = st;
  c.ll = st->ll_ring;
  c.lltop = st->ll_recv[0][exch & 1];
  c.llbot = st->ll_recv[1][exch & 1];
  c.pll_top = st->peer_ll[0][pe & 1];
  c.pll_bot = st->peer_ll[1][pe & 1];
  c.seq_in = (uint32_t)(exch + 1ull);    // exchange numbers start at 1: zeroed memory never matches
  c.seq_out = (uint32_t)(pe + 1ull);
  return c;
}

// ---- the bit-sliced generation stage, spelled out in 32-bit LOP3s -----------------------------
// Each lane owns one 64-cell word as two 32-bit halves.  Per half and per generation:
//   1 SHF (funnel shift brings in the x-neighbour bit, across lanes via SHFL; the other
//   neighbour is the other half at the same bit index thanks to the even/odd layout)
//   2 LOP3 horizontal sum, 4 LOP3 vertical carry-save, 4 LOP3 rule decode  = 10 LOP3
// The LUTs are written out so that ptxas cannot pick a longer decomposition.
template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
  return r;
}

struct Half {
  uint32_t a0, a1;  // horizontal sum planes of row r-2
  uint32_t b0, b1;  // horizontal sum planes of row r-1
  uint32_t c;       // centre cells of row r-1
};
struct Stage {
  Half lo, hi;
};

// Feed the horizontal sum (h0,h1) and centre x of the newest row; returns the next
// generation of the previous row and rotates the window.
__device__ __forceinline__ uint32_t stage_half(Half& s, uint32_t h0, uint32_t h1, uint32_t x) {
  const uint32_t t0 = lop3<0x96>(s.a0, s.b0, h0);   // weight 1
  const uint32_t k0 = lop3<0xE8>(s.a0, s.b0, h0);   // weight 2
  const uint32_t t1 = lop3<0x96>(s.a1, s.b1, h1);   // weight 2
  const uint32_t k1 = lop3<0xE8>(s.a1, s.b1, h1);   // weight 4
  // 3x3 total (centre included) = t0 + 2(k0 + t1) + 4 k1;  next = total==3 | centre & total==4
  const uint32_t e3 = lop3<0x14>(k0, t1, k1);       // (k0^t1) & ~k1        -> total in {2,3}
  const uint32_t e4 = lop3<0x42>(k0, t1, k1);       // ~(k0^t1) & (k1^k0)   -> total in {4,5}
  const uint32_t v = lop3<0x08>(t0, s.c, e4);       // ~t0 & centre & e4    -> centre & total==4
  const uint32_t o = lop3<0xEA>(t0, e3, v);         // (t0 & e3) | v
  s.a0 = s.b0;
  s.a1 = s.b1;
  s.b0 = h0;
  s.b1 = h1;
  s.c = x;
  return o;
}

// Push one row through generation stages [0, nst).  Fed rows y_in+0, y_in+1, ... the value
// returned at step i (i >= 2K, nst == K) is generation +K of row y_in + i - K.  nst < K is used
// while the pipeline fills: stage t has no valid input before step 2t, so running it is waste.
//
// (V is kept as a template slot for kernel variants; only V = 0 exists.  A shift-by-multiply
// variant that moved the funnel shifts to the FMA pipe measured slower and was removed.)
template <int K, int V>
__device__ __forceinline__ uint64_t push_row(Stage (&s)[K], uint64_t x64, int nst) {
  uint32_t lo = (uint32_t)x64, hi = (uint32_t)(x64 >> 32);
#pragma unroll
  for (int t = 0; t < K; ++t) {
    if (t < nst) {
      // lo = even columns, hi = odd columns of my 64 cells (life_math.h layout).  The left
      // neighbour of an even column is the previous odd column (shift in the left lane's last
      // odd column); its right neighbour is the odd column at the SAME bit index - no shift.
      // Symmetrically for odd columns.  One funnel shift per half instead of two.
      const uint32_t lhi = __shfl_up_sync(kFull, hi, 1);    // left lane's odd columns: its bit 31 is my column -1
      const uint32_t rlo = __shfl_down_sync(kFull, lo, 1);  // right lane's even columns: its bit 0 is my column 64
      const uint32_t Llo = __funnelshift_l(lhi, hi, 1);     // (hi << 1) | (lhi >> 31)
      const uint32_t Rlo = hi;
      const uint32_t Lhi = lo;
      const uint32_t Rhi = __funnelshift_r(lo, rlo, 1);     // (lo >> 1) | (rlo << 31)
      const uint32_t h0lo = lop3<0x96>(Llo, lo, Rlo), h1lo = lop3<0xE8>(Llo, lo, Rlo);
      const uint32_t h0hi = lop3<0x96>(Lhi, hi, Rhi), h1hi = lop3<0xE8>(Lhi, hi, Rhi);
      const uint32_t olo = stage_half(s[t].lo, h0lo, h1lo, lo);
      const uint32_t ohi = stage_half(s[t].hi, h0hi, h1hi, hi);
      lo = olo;
      hi = ohi;
    }
  }
  return ((uint64_t)hi << 32) | lo;
}

// Geometry of one warp task, shared by the EDGE / interior instantiations of the row loop.
struct TaskGeom {
  int64_t y0, y1;   // output rows [y0, y1)
  int64_t wi;       // my word column
  bool owner;       // lanes 1..30 inside the row own output words
};

// Stream rows [y0-K, y1+K) through the K stages and store rows [y0, y1).  EDGE = the strip
// touches freshly exchanged halo rows (recv buffers) instead of in-slab rows.
template <int K, int V, bool EDGE>
__device__ __forceinline__ void stream_rows(const PassCtx& c, const TaskGeom& g) {
  constexpr int B = 4;  // rows per batch (software prefetch depth)
  const int64_t y_in = g.y0 - K;
  const int n_in = (int)(g.y1 - g.y0) + 2 * K;
  const int last = n_in - 1;
  const uint64_t* in_col = c.in + g.wi;
  const uint64_t* top_col = c.rtop + g.wi + (int64_t)c.ext_in * c.stride;  // row y (<0) at top_col + y*stride
  const uint64_t* bot_col = c.rbot + g.wi - c.H * c.stride;                // row y (>=H) at bot_col + y*stride
  uint64_t* out_col = c.out + g.wi;
  const bool fresh = c.consume != 0;
  // halo push straight from the registers that hold the new row (EDGE tasks only): the peer
  // store leaves as soon as the row exists and its NVLink latency overlaps the rest of the strip
  uint64_t* ptop_col = c.ptop + g.wi;                                // row y at ptop_col + y*stride
  uint64_t* pbot_col = c.pbot + g.wi - (c.H - c.push) * c.stride;    // row y at pbot_col + y*stride
  const int64_t push_lo = c.push, push_hi = c.H - c.push;

  auto load_row = [&](int i) -> uint64_t {
    i = i < last ? i : last;  // rows past the end are re-reads of the last row (never stored)
    const int64_t y = y_in + i;
    const uint64_t* p = in_col + y * c.stride;
    if (EDGE && fresh) {
      // y is warp-uniform, so these branches are too
      if (y < 0 && (c.ll & 1)) return ll_load(c.st, c.lltop + ((y + c.ext_in) * c.stride + g.wi) * 2, c.seq_in, 0);
      if (y >= c.H && (c.ll & 2)) return ll_load(c.st, c.llbot + ((y - c.H) * c.stride + g.wi) * 2, c.seq_in, 1);
      p = (y < 0) ? top_col + y * c.stride : p;
      p = (y >= c.H) ? bot_col + y * c.stride : p;
    }
    return ld_cg(p);
  };

  Stage s[K];
#pragma unroll
  for (int t = 0; t < K; ++t) {
    s[t].lo.a0 = s[t].lo.a1 = s[t].lo.b0 = s[t].lo.b1 = s[t].lo.c = 0;
    s[t].hi.a0 = s[t].hi.a1 = s[t].hi.b0 = s[t].hi.b1 = s[t].hi.c = 0;
  }

  // pipeline fill: the first 2K rows produce no output; row i only feeds stages 0..i/2
  if (K <= 4) {
    uint64_t pre[2 * K];
#pragma unroll
    for (int i = 0; i < 2 * K; ++i) pre[i] = load_row(i);
#pragma unroll
    for (int i = 0; i < 2 * K; ++i) push_row<K, V>(s, pre[i], (i / 2 + 1 < K) ? i / 2 + 1 : K);
  } else {
    for (int i = 0; i < 2 * K; ++i) push_row<K, V>(s, load_row(i), K);
  }

  // steady state: every row in yields one row out
  const int n_main = n_in - 2 * K;
  const int nb = (n_main + B - 1) / B;
  uint64_t nxt[B];
#pragma unroll
  for (int j = 0; j < B; ++j) nxt[j] = load_row(2 * K + j);
  for (int b = 0; b < nb; ++b) {
    uint64_t cur[B];
#pragma unroll
    for (int j = 0; j < B; ++j) cur[j] = nxt[j];
#pragma unroll
    for (int j = 0; j < B; ++j) nxt[j] = load_row(2 * K + (b + 1) * B + j);
    const int64_t yb = g.y0 + (int64_t)b * B;
#pragma unroll
    for (int j = 0; j < B; ++j) {
      const uint64_t o = push_row<K, V>(s, cur[j], K);
      const int64_t y = yb + j;
      if (g.owner && y < g.y1) {
        st_cg(out_col + y * c.stride, o);
        if (EDGE && c.push) {
          if (y >= 0 && y < push_lo) {
            if (c.ll & 1) ll_store(c.pll_top + (y * c.stride + g.wi) * 2, o, c.seq_out);
            else st_cg(ptop_col + y * c.stride, o);
          }
          if (y >= push_hi && y < c.H) {
            if (c.ll & 2) ll_store(c.pll_bot + ((y - push_hi) * c.stride + g.wi) * 2, o, c.seq_out);
            else st_cg(pbot_col + y * c.stride, o);
          }
        }
      }
    }
  }
}

// Interior strips (no recv rows, no halo push): the same K-stage pipeline with the per-row address
// arithmetic reduced to two pointer increments and the stores predicated instead of branched
// around, so that the row loop is straight-line code: 88 LOP3/SHF + 8 SHFL + 1 LDG + 1 STG + 2 adds
// per row at K = 4.  Nothing is read beyond row y1 + K - 1: caller-owned buffers (the no-allocation
// entry point) have no rows to spare.
// COH = true (flow kernel): rows written by other SMs earlier in the SAME launch are read, so loads
// and stores go to L2 (ld/st.global.cg).  COH = false (one launch per pass): the input is
// read-only for the whole launch and L1 is invalidated at every launch boundary, so plain
// accesses are enough.
template <bool COH>
__device__ __forceinline__ uint64_t ld_row(const uint64_t* p) {
  if (!COH && !GOL_CHK(p, 8)) return 0;
  return COH ? ld_cg(p) : *p;
}
template <bool COH>
__device__ __forceinline__ void st_row(uint64_t* p, uint64_t v) {
  if (COH) st_cg(p, v);
  else if (GOL_CHK(p, 8)) *p = v;
}

template <int K, int V, bool COH>
__device__ __forceinline__ void stream_interior(const PassCtx& c, const TaskGeom& g) {
  constexpr int B = 4;  // rows per batch (software prefetch depth)
  const int64_t stride = c.stride;
  const uint64_t* in = c.in + g.wi + (g.y0 - K) * stride;
  uint64_t* out = c.out + g.wi + g.y0 * stride;
  const bool owner = g.owner;
  const int n_out = (int)(g.y1 - g.y0);

  Stage s[K];
#pragma unroll
  for (int t = 0; t < K; ++t) {
    s[t].lo.a0 = s[t].lo.a1 = s[t].lo.b0 = s[t].lo.b1 = s[t].lo.c = 0;
    s[t].hi.a0 = s[t].hi.a1 = s[t].hi.b0 = s[t].hi.b1 = s[t].hi.c = 0;
  }
  // pipeline fill: the first 2K rows produce no output; row i only feeds stages 0..i/2
  {
    uint64_t pre[2 * K];
#pragma unroll
    for (int i = 0; i < 2 * K; ++i) {
      pre[i] = ld_row<COH>(in);
      in += stride;
    }
#pragma unroll
    for (int i = 0; i < 2 * K; ++i) push_row<K, V>(s, pre[i], (i / 2 + 1 < K) ? i / 2 + 1 : K);
  }
  // steady state: every row in yields one row out; whole batches first
  const int nfull = n_out / B, rem = n_out - nfull * B;
  uint64_t nxt[B];
  if (nfull > 0) {
#pragma unroll
    for (int j = 0; j < B; ++j) {
      nxt[j] = ld_row<COH>(in);
      in += stride;
    }
  }
#pragma unroll 1
  for (int b = 0; b < nfull; ++b) {
    uint64_t cur[B];
#pragma unroll
    for (int j = 0; j < B; ++j) cur[j] = nxt[j];
    if (b + 1 < nfull) {
#pragma unroll
      for (int j = 0; j < B; ++j) {
        nxt[j] = ld_row<COH>(in);
        in += stride;
      }
    }
#pragma unroll
    for (int j = 0; j < B; ++j) {
      const uint64_t o = push_row<K, V>(s, cur[j], K);
      if (owner) st_row<COH>(out, o);
      out += stride;
    }
  }
#pragma unroll 1
  for (int j = 0; j < rem; ++j) {
    const uint64_t x = ld_row<COH>(in);
    in += stride;
    const uint64_t o = push_row<K, V>(s, x, K);
    if (owner) st_row<COH>(out, o);
    out += stride;
  }
}

// One warp task of a fused pass: strip `strip` of column group `cg`.
__device__ __forceinline__ TaskGeom make_geom(const PassCtx& c, int64_t y0, int64_t y1, int cg, int lane) {
  TaskGeom g;
  g.y0 = y0;
  g.y1 = y1;
  // my word column (torus wrap in x, also when the row has < 32 words)
  int64_t wi = (int64_t)cg * kValidLanes - 1 + lane;
  wi %= c.ws;
  if (wi < 0) wi += c.ws;
  g.wi = wi;
  g.owner = lane >= 1 && lane <= kValidLanes && ((int64_t)cg * kValidLanes + lane - 1) < c.ws;
  return g;
}

// Rows of [y0, y1) that the ring neighbours need: [0, push) -> prev, [H-push, H) -> next.
__device__ __forceinline__ void push_ranges(const PassCtx& c, int64_t y0, int64_t y1, int* rows_top,
                                            int* rows_bot) {
  const int64_t t0 = y0 > 0 ? y0 : 0;
  const int64_t t1 = y1 < c.push ? y1 : c.push;
  const int64_t b0 = y0 > c.H - c.push ? y0 : c.H - c.push;
  const int64_t b1 = y1 < c.H ? y1 : c.H;
  *rows_top = (c.push && t1 > t0) ? (int)(t1 - t0) : 0;
  *rows_bot = (c.push && b1 > b0) ? (int)(b1 - b0) : 0;
}

// Halo publish (the completion half of the fused MPI_Isend, gol-main.c:104-107): the edge rows
// were stored into the neighbours' recv buffers as they were produced (stream_rows<EDGE>); here
// the word count is published.
__device__ __forceinline__ void publish_edges(SlabState* st, const PassCtx& c, const TaskGeom& g, int lane) {
  int rows_top, rows_bot;
  push_ranges(c, g.y0, g.y1, &rows_top, &rows_bot);
  if (c.ll & 1) rows_top = 0;   // flag-in-data sides need no publication
  if (c.ll & 2) rows_bot = 0;
  if (rows_top | rows_bot) {
    // No per-lane fence: the warp barrier orders every lane's stores before lane 0's
    // red.release.sys, and a release is cumulative over what happens-before it.  (A per-lane
    // __threadfence_system() is a fence.sc.sys; it kept the edge warp, and with it the kernel,
    // alive for an extra NVLink round trip or more: ~9 us per pass on the 4096^2 config.)
    __syncwarp();
    const unsigned owners = __popc(__ballot_sync(kFull, g.owner));
    if (lane == 0) {
      if (rows_top) publish_words(st, 0, (unsigned long long)owners * rows_top);
      if (rows_bot) publish_words(st, 1, (unsigned long long)owners * rows_bot);
    }
  }
}

// One warp task of a fused pass: output rows [y0, y1) of column group `cg`.
template <int K, int V, bool COH = true>
__device__ __forceinline__ void fused_task(SlabState* st, const PassCtx& c, int64_t y0, int64_t y1,
                                           int cg, int lane) {
  const TaskGeom g = make_geom(c, y0, y1, cg, lane);
  const bool reads_top = c.consume && (g.y0 - K < 0);
  const bool reads_bot = c.consume && (g.y1 + K > c.H);
  int rows_top, rows_bot;
  push_ranges(c, g.y0, g.y1, &rows_top, &rows_bot);
  if ((reads_top && !(c.ll & 1)) || (reads_bot && !(c.ll & 2))) {   // warp-uniform: every lane polls (see wait_words)
    if (reads_top && !(c.ll & 1)) wait_words(st, 0, c.target);
    if (reads_bot && !(c.ll & 2)) wait_words(st, 1, c.target);
    __syncwarp();
  }
  if (reads_top || reads_bot || rows_top || rows_bot) {
    stream_rows<K, V, true>(c, g);    // recv-row sources and inline halo push
    publish_edges(st, c, g, lane);
  } else {
    stream_interior<K, V, COH>(c, g);
  }
}

// Last CTA of a launch advances the slab's counters.
__device__ __forceinline__ void finish_pass(SlabState* st, const PassParams& p, unsigned nctas) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(&st->done, 1u);
    if (prev == nctas - 1) {
      st->done = 0;
      st->pass += 1;
      if (p.consume) {
        st->exch += 1;
        st->words_seen += (unsigned long long)p.ext_in * (unsigned long long)st->ws;
      }
      __threadfence();
    }
  }
}

// Programmatic dependent launch: when the host launches with the PDL attribute, this grid may be
// scheduled while its predecessor drains; nothing of the predecessor's output (not even the pass
// counters) is read before griddepcontrol.wait.  Without the attribute both are no-ops.
__device__ __forceinline__ void pdl_wait_and_release() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

template <int K, int MAXT, int MINB, int V>
__global__ void __launch_bounds__(MAXT, MINB) life_fused_kernel(SlabState* states, PassParams p) {
  pdl_wait_and_release();
  SlabState* st = states + blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const int task = blockIdx.x * wpb + (threadIdx.x >> 5);
  const int ntasks = p.nstrips * p.col_groups;
  if (task < ntasks) {
    const PassCtx c = make_ctx(st, st->pass, st->exch, st->words_seen, p.ext_in, p.consume, p.push);
    const int order = task / p.col_groups;
    const int cg = task - order * p.col_groups;
    const int ext_out = p.ext_in - K;
    int64_t y0, y1;
    strip_range(order, p.nstrips, -(int64_t)ext_out, c.H + ext_out, p.strip_rows, p.edge_rows, &y0, &y1);
    fused_task<K, V, GOL_FUSED_COHERENT != 0>(st, c, y0, y1, cg, lane);
  }
  finish_pass(st, p, gridDim.x);
}

// ---- N-K1 (TMA-fed): input rows arrive through the tensor memory accelerator ---------------------
// Same K-stage register pipeline, but the interior strips issue no LDGs: one elected thread per
// CTA streams 8-row x 144-word boxes global -> shared with cp.async.bulk.tensor.2d (SASS:
// UTMALDG) through a tensor map of the slab, into a 3-slot ring guarded by full/empty
// mbarriers; the four compute warps read their word with one conflict-free LDS.64 per row.
// The CTA tile is one strip x 4 adjacent column groups: words [120*c4 - 1, 120*c4 + 121), box
// start rounded down to 128 bytes.  TMA cannot wrap around the torus seam: columns outside
// [0, ws) arrive zero-filled, and the only two out-of-range words that matter (the halo word
// left of column 0 and right of column W-1) are fetched by their lane with a plain load.
// (A first version used one 1 KB cp.async.bulk per row: correct, but the per-SM copy engine
// retires roughly one descriptor per 200 cycles, which starved the pipeline: 9.5 vs 26.6 TCUPS.)
// Edge strips (the ones that read freshly exchanged halo rows) keep the LDG path: their data
// was just written by a peer GPU and is consumed through ld.acquire ordering.
namespace tma {
// NW compute warps per CTA (3 or 4, whichever wastes fewer column groups); box width in words:
// NW*30 + 2 needed, plus up to 15 because the box start is rounded down to 128 bytes
__host__ __device__ constexpr int seg_words(int nw) { return (nw * kValidLanes + 2 + 15 + 15) / 16 * 16; }  // 4 -> 144, 3 -> 112
constexpr int kChunk = 8;    // box height = rows per ring slot
constexpr int kSlots = 5;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Wait for a phase by polling test_wait.  Measured on B200 for this kernel (16384^2 / 65536-wide):
// suspending mbarrier.try_wait, with or without a suspend-time hint: 7-9 / 11 TCUPS; one polling
// lane + __syncwarp: 8 / 10; every thread polling test_wait: 15 / 29 with 3 ring slots and
// 17 / 30 with 5.  The LDG path does 26.8 / 33.0.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_test(bar, parity)) {
  }
}
__device__ __forceinline__ void tensor_g2s(void* dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tensor3_g2s(void* dst, const CUtensorMap* map, int x, int j, int y, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(x), "r"(j), "r"(y), "r"(smem_u32(bar))
      : "memory");
}
}  // namespace tma

template <int K, int NW>
__global__ void __launch_bounds__(NW * 32, 512 / (NW * 32))
    life_fused_tma_kernel(SlabState* states, PassParams p, const CUtensorMap* tmaps) {
  using namespace tma;
  constexpr int kWarps = NW;
  constexpr int kSeg = seg_words(NW);
  static_assert(2 * K <= kChunk, "pipeline fill must fit the first chunk");
  __shared__ __align__(128) uint64_t ring[kSlots][kChunk][kSeg];
  __shared__ __align__(8) uint64_t full_bar[kSlots];
  __shared__ __align__(8) uint64_t empty_bar[kSlots];

  pdl_wait_and_release();
  SlabState* st = states + blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ncg4 = (p.col_groups + kWarps - 1) / kWarps;
  const int order = blockIdx.x / ncg4;
  const int c4 = blockIdx.x - order * ncg4;
  const unsigned long long pass = st->pass;
  const PassCtx c = make_ctx(st, pass, st->exch, st->words_seen, p.ext_in, p.consume, p.push);
  const int ext_out = p.ext_in - K;
  int64_t y0, y1;
  strip_range(order, p.nstrips, -(int64_t)ext_out, c.H + ext_out, p.strip_rows, p.edge_rows, &y0, &y1);
  const int cg = c4 * kWarps + warp;
  int rows_top, rows_bot;
  push_ranges(c, y0, y1, &rows_top, &rows_bot);
  const bool edge = (c.consume && (y0 - K < 0 || y1 + K > c.H)) || rows_top || rows_bot;

  if (edge) {
    if (cg < p.col_groups) fused_task<K, 0>(st, c, y0, y1, cg, lane);
  } else {
    const int n_in = (int)(y1 - y0) + 2 * K;
    const int nchunks = (n_in + kChunk - 1) / kChunk;
    const int64_t y_in = y0 - K;
    const CUtensorMap* map = tmaps + 2 * blockIdx.y + (pass & 1);  // the input buffer's map
    if (threadIdx.x == 0) {
      for (int i = 0; i < kSlots; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], kWarps);
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int64_t need = (int64_t)c4 * (kWarps * kValidLanes) - 1;   // first word this CTA reads (halo)
    const int64_t seg_start = need & ~int64_t(15);                    // box x: 128-byte aligned, may be -16
    const int row0 = st->G + (int)y_in;                               // box y of chunk 0 in buffer rows
    auto issue_chunk = [&](int ch) {
      const int slot = ch % kSlots;
      mbar_expect_tx(&full_bar[slot], (uint32_t)(kChunk * kSeg * 8));
      if (p.tma_rank == 3)   // rows viewed as [ws/16][16 words]: 128-byte inner box
        tensor3_g2s(&ring[slot][0][0], map, 0, (int)(seg_start >> 4), row0 + ch * kChunk, &full_bar[slot]);
      else
        tensor_g2s(&ring[slot][0][0], map, (int)seg_start, row0 + ch * kChunk, &full_bar[slot]);
    };
    if (threadIdx.x == 0)
      for (int ch = 0; ch < kSlots && ch < nchunks; ++ch) issue_chunk(ch);

    const TaskGeom g = make_geom(c, y0, y1, cg, lane);
    uint64_t* out_col = c.out + g.wi;
    const int off = warp * kValidLanes + lane + (int)(need - seg_start);  // my word inside the box
    // the two words TMA cannot deliver: column -1 (= ws-1) and column ws (= 0)
    const int64_t gw = need + warp * kValidLanes + lane;
    const bool seam = (gw == -1) || (gw == c.ws);
    const uint64_t* seam_col = c.in + (gw < 0 ? c.ws - 1 : 0);
    Stage s[K];
#pragma unroll
    for (int t = 0; t < K; ++t) {
      s[t].lo.a0 = s[t].lo.a1 = s[t].lo.b0 = s[t].lo.b1 = s[t].lo.c = 0;
      s[t].hi.a0 = s[t].hi.a1 = s[t].hi.b0 = s[t].hi.b1 = s[t].hi.c = 0;
    }

    for (int ch = 0; ch < nchunks; ++ch) {
      const int slot = ch % kSlots;
      uint64_t rows[kChunk];
      if (seam) {  // issued before the barrier wait so that the latency overlaps it
#pragma unroll
        for (int r = 0; r < kChunk; ++r) {
          int i = ch * kChunk + r;
          i = i < n_in - 1 ? i : n_in - 1;
          rows[r] = ld_cg(seam_col + (y_in + i) * c.stride);
        }
      }
      mbar_wait(&full_bar[slot], (uint32_t)((ch / kSlots) & 1));
      if (!seam) {
#pragma unroll
        for (int r = 0; r < kChunk; ++r) rows[r] = ring[slot][r][off];
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[slot]);   // my warp no longer needs this slot
      if (threadIdx.x == 0 && ch + kSlots < nchunks) {
        mbar_wait(&empty_bar[slot], (uint32_t)((ch / kSlots) & 1));  // all four warps released it
        issue_chunk(ch + kSlots);
      }
      __syncwarp();  // reconverge after the single-lane sections: the stages below shuffle
      if (ch == 0) {
#pragma unroll
        for (int r = 0; r < kChunk; ++r) {
          if (r < 2 * K) {
            push_row<K, 0>(s, rows[r], (r / 2 + 1 < K) ? r / 2 + 1 : K);
          } else {
            const uint64_t o = push_row<K, 0>(s, rows[r], K);
            const int64_t y = y0 + (r - 2 * K);
            if (g.owner && y < y1) st_cg(out_col + y * c.stride, o);
          }
        }
      } else {
        const int64_t yb = y0 + (int64_t)ch * kChunk - 2 * K;
#pragma unroll
        for (int r = 0; r < kChunk; ++r) {
          const uint64_t o = push_row<K, 0>(s, rows[r], K);
          const int64_t y = yb + r;
          if (g.owner && y < y1) st_cg(out_col + y * c.stride, o);
        }
      }
    }
  }
  finish_pass(st, p, gridDim.x);
}

// ---- N-K2: generic kernel ----------------------------------------------------
// One thread per 64-cell word, one generation per pass, any W (masked tail
// word, bit-level x-wrap) and any H >= 1; blockDim is whatever the CLI asked
// for (gol-main.c:52).  Same recv/flag protocol with depth 1.
__global__ void life_generic_kernel(SlabState* states, PassParams p) {
  SlabState* st = states + blockIdx.y;
  const PassCtx c = make_ctx(st, st->pass, st->exch, st->words_seen, 1, p.consume, p.push);
  const int64_t total = c.H * c.ws;
  const int64_t first = (int64_t)blockIdx.x * blockDim.x;
  int64_t last = first + blockDim.x - 1;
  if (last >= total) last = total - 1;
  if (first < total && p.consume) {
    const bool top = first < c.ws;
    const bool bot = last >= (c.H - 1) * c.ws;
    if (threadIdx.x == 0) {
      if (top) wait_words(st, 0, c.target);
      if (bot) wait_words(st, 1, c.target);
    }
    if (top || bot) __syncthreads();
  }
  const int64_t idx = first + threadIdx.x;
  int sent_top = 0, sent_bot = 0;
  if (idx < total) {
    const int64_t y = idx / c.ws;
    const int64_t w = idx - y * c.ws;
    const uint64_t* cur = c.in + y * c.stride;
    const uint64_t* up = (y == 0 && p.consume) ? c.rtop : cur - c.stride;
    const uint64_t* down = (y == c.H - 1 && p.consume) ? c.rbot : cur + c.stride;
    uint64_t v = 0;
    if (GOL_CHK(up, c.ws * 8) && GOL_CHK(cur, c.ws * 8) && GOL_CHK(down, c.ws * 8))
      v = step_word(up, cur, down, w, c.ws, st->W);
    if (GOL_CHK(c.out + y * c.stride + w, 8)) c.out[y * c.stride + w] = v;
    if (p.push) {
      if (y == 0) { st_cg(c.ptop + w, v); sent_top = 1; }
      if (y == c.H - 1) { st_cg(c.pbot + w, v); sent_bot = 1; }
    }
  }
  if (p.push) {
    const int nt = __syncthreads_count(sent_top);   // block barrier + cumulative release below
    const int nb = __syncthreads_count(sent_bot);
    if (threadIdx.x == 0) {
      if (nt) publish_words(st, 0, (unsigned long long)nt);
      if (nb) publish_words(st, 1, (unsigned long long)nb);
    }
  }
  finish_pass(st, p, gridDim.x);
}

// ---- N-K5: halo prime ----------------------------------------------------------
// Replaces init_Ghost_rows (gol-with-cuda.cu:32-53): copies the current first
// and last `depth` rows into the neighbours' recv buffers and publishes them.
__global__ void halo_prime_kernel(SlabState* states, int depth) {
  SlabState* st = states + blockIdx.y;
  const int64_t ws = st->ws, stride = st->stride, H = st->H;
  const uint64_t* in = st->buf[st->pass & 1] + (int64_t)st->G * stride;
  uint64_t* ptop = st->peer_recv[0][st->exch & 1];
  uint64_t* pbot = st->peer_recv[1][st->exch & 1];
  uint64_t* lltop = st->peer_ll[0][st->exch & 1];
  uint64_t* llbot = st->peer_ll[1][st->exch & 1];
  const uint32_t seq = (uint32_t)(st->exch + 1ull);
  const int ll = st->ll_ring;
  const int64_t per_side = (int64_t)depth * ws;
  int sent_top = 0, sent_bot = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < 2 * per_side;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int side = i >= per_side;
    const int64_t j = side ? i - per_side : i;
    const int64_t r = j / ws, w = j - r * ws;
    if (!side) {
      const uint64_t v = GOL_CHK(in + r * stride + w, 8) ? in[r * stride + w] : 0;
      if (ll & 1) {
        ll_store(lltop + (r * stride + w) * 2, v, seq);   // flag-in-data: nothing to publish
      } else {
        st_cg(ptop + r * stride + w, v);
        ++sent_top;
      }
    } else {
      const uint64_t v = GOL_CHK(in + (H - depth + r) * stride + w, 8) ? in[(H - depth + r) * stride + w] : 0;
      if (ll & 2) {
        ll_store(llbot + (r * stride + w) * 2, v, seq);
      } else {
        st_cg(pbot + r * stride + w, v);
        ++sent_bot;
      }
    }
  }
  __shared__ unsigned long long acc[2];
  if (threadIdx.x == 0) acc[0] = acc[1] = 0;
  __syncthreads();
  if (sent_top) atomicAdd(&acc[0], (unsigned long long)sent_top);
  if (sent_bot) atomicAdd(&acc[1], (unsigned long long)sent_bot);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (acc[0]) publish_words(st, 0, acc[0]);
    if (acc[1]) publish_words(st, 1, acc[1]);
  }
}

// ---- N-K3: persistent "flow" kernel ---------------------------------------------------
// The whole generation loop of the reference (gol-main.c:94-116) in ONE ordinary launch, with no
// grid-wide barrier between passes.  Every (pass, strip, column group) is a ticket; warps pull
// tickets from a device counter in an order that is a topological order of the dependency graph
// (pass-major, inside a pass the two edge strips first and then the interior from the outside
// in).  A task of pass p waits - lane 0 with ld.acquire.gpu on per-strip progress counters - only
// for the three strips of pass p-1 whose rows it reads and overwrites; it publishes its own
// completion with red.release.gpu.  Passes therefore overlap: while the slow tail of pass p
// finishes, free warps are already deep in pass p+1, and the SMs never drain between passes (the
// per-pass launches of life_fused_kernel leave them empty for ~15 % of every 36 us pass).
// Cross-GPU ordering is unchanged: edge tasks wait for the ring neighbours' halo words
// (wait_words) and push their own edge rows as they produce them.
//
// Deadlock freedom: a warp waits only for tickets lower than its own, and a ticket is only ever
// held by a running warp, so the launch needs neither co-residency of all CTAs nor a cooperative
// launch.  Every wait is bounded by the slab's watchdog.
__device__ __forceinline__ void red_release_gpu_add_u32(unsigned int* p, unsigned int v) {
  if (!GOL_CHK(p, 4)) return;
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Called by all lanes of the warp (see wait_words).
__device__ __noinline__ void wait_strips(SlabState* st, unsigned int* ctl, const unsigned int* prog, int lo,
                                         int hi, unsigned int want, int fence) {
  const bool lane0 = (threadIdx.x & 31) == 0;
  unsigned long long t0 = 0;
  for (int jj = lo; jj <= hi; ++jj) {
    if (ld_relaxed_gpu_u32(prog + jj) >= want) continue;
    if (!t0) t0 = globaltimer_ns();
    unsigned ns = 0;
    while (ld_relaxed_gpu_u32(prog + jj) < want) {
      if (ns) __nanosleep(ns);
      ns = ns < 8 ? ns + 1 : (ns < 256 ? ns << 1 : ns);
      if (globaltimer_ns() - t0 > st->watchdog_ns || *(volatile unsigned int*)&ctl[2]) {
        if (lane0) {
          atomicCAS(&st->error, 0, 3);
          atomicExch(&ctl[2], 1u);
        }
        jj = hi;  // give up on the remaining strips too
        break;
      }
    }
  }
  if (fence) fence_acq_rel_gpu();
  if (t0 && lane0) atomicAdd(&st->dep_wait_ns, globaltimer_ns() - t0);
}

template <int K>
__device__ __forceinline__ void flow_task(SlabState* states, const FlowParams& fp, int pass, unsigned slot,
                                          unsigned percg, int lane) {
  const int q = (int)(slot / percg);
  const unsigned rem = slot - (unsigned)q * percg;
  const int sidx = (int)(rem / (unsigned)fp.col_groups);
  const int cg = (int)(rem - (unsigned)sidx * (unsigned)fp.col_groups);
  const int j = flow_order_to_strip(q, fp.nstrips);
  SlabState* st = states + sidx;

  const int pos = pass % fp.passes_per_exchange;
  const int ext_in = fp.depth - pos * K;
  const int ext_out = ext_in - K;
  const int consume = pos == 0;
  const int push = (ext_out == 0 && pass + 1 < fp.npasses) ? fp.depth : 0;
  const unsigned long long nex = (unsigned long long)(pass / fp.passes_per_exchange);
  const unsigned long long exch = st->exch + nex + (consume ? 0ull : 1ull);
  const unsigned long long seen = st->words_seen + nex * (unsigned long long)fp.depth * (unsigned long long)st->ws;
  const PassCtx c = make_ctx(st, st->pass + (unsigned long long)pass, exch, seen, ext_in, consume, push);
  int64_t y0, y1;
  flow_strip_range(j, fp.nstrips, c.H, fp.edge_rows, ext_out, &y0, &y1);

  unsigned int* prog = fp.progress + (size_t)sidx * (size_t)fp.max_strips;
  if (pass > 0 && !(fp.flags & 1)) {
    const int lo = j - fp.dep_radius, hi = j + fp.dep_radius;
    wait_strips(st, fp.ctl, prog, lo < 0 ? 0 : lo, hi > fp.nstrips - 1 ? fp.nstrips - 1 : hi,
                (unsigned)pass * (unsigned)fp.col_groups, !(fp.flags & 8));
    __syncwarp();
  }
  const unsigned long long t0 = globaltimer_ns();
  fused_task<K, 0>(st, c, y0, y1, cg, lane);
  // every lane's stores happen-before lane 0's release through the warp barrier
  __syncwarp();
  if (lane == 0) {
    red_release_gpu_add_u32(prog + j, 1u);
    const int edge = (j == 0 || j == fp.nstrips - 1) ? 0 : 1;
    atomicAdd(&st->task_ns[edge], globaltimer_ns() - t0);
    atomicAdd(&st->tasks[edge], 1ull);
  }
}

template <int K, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) life_flow_kernel(SlabState* states, FlowParams fp) {
  const int lane = threadIdx.x & 31;
  const unsigned percg = (unsigned)(fp.nslabs * fp.col_groups);
  const unsigned tpp = percg * (unsigned)fp.nstrips;          // tasks per pass

  if (!(fp.flags & 4)) {
    // ticket mode (default): any free warp takes the next task in topological order
    const unsigned total = tpp * (unsigned)fp.npasses;        // < 2^31 (bounded by the host)
    for (;;) {
      unsigned t = 0;
      if (lane == 0) t = atomicAdd(&fp.ctl[0], 1u);
      t = __shfl_sync(kFull, t, 0);
      if (t >= total) break;
      if (*(volatile unsigned int*)&fp.ctl[2]) continue;      // aborted: drain the tickets
      const int pass = (int)(t / tpp);
      flow_task<K>(states, fp, pass, t - (unsigned)pass * tpp, percg, lane);
    }
  } else {
    // owner mode (GOL_FLOW_FLAGS=4): warp w runs the same task slot(s) w, w + nwarps, ... in every pass, so in steady
    // state it only ever waits for the warps that own the neighbouring strips
    const unsigned gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    for (int pass = 0; pass < fp.npasses; ++pass) {
      if (*(volatile unsigned int*)&fp.ctl[2]) break;
      for (unsigned slot = gw; slot < tpp; slot += (unsigned)fp.total_warps)
        flow_task<K>(states, fp, pass, slot, percg, lane);
    }
  }

  // the last warp out advances the slab counters and leaves the bookkeeping zeroed for the next launch
  unsigned last = 0;
  if (lane == 0) {
    __threadfence();
    last = atomicAdd(&fp.ctl[1], 1u) == (unsigned)fp.total_warps - 1u;
  }
  last = __shfl_sync(kFull, last, 0);
  if (last) {
    __threadfence();
    const unsigned long long nex =
        (unsigned long long)((fp.npasses + fp.passes_per_exchange - 1) / fp.passes_per_exchange);
    for (int sidx = lane; sidx < fp.nslabs; sidx += 32) {
      SlabState* st = states + sidx;
      st->pass += (unsigned long long)fp.npasses;
      st->exch += nex;
      st->words_seen += nex * (unsigned long long)fp.depth * (unsigned long long)st->ws;
    }
    const int n = fp.nslabs * fp.max_strips;
    for (int i = lane; i < n; i += 32)
      if (GOL_CHK(fp.progress + i, 4)) fp.progress[i] = 0u;
    __syncwarp();
    if (lane == 0) {
      fp.ctl[0] = 0u;
      fp.ctl[1] = 0u;
      fp.ctl[2] = 0u;
      __threadfence();
    }
  }
}

// ---- N-K3b: resident kernel (small slabs) ---------------------------------------------------
// A slab that fits in shared memory (4096^2 cells = 2 MiB over 148 SMs) is not bound by
// throughput but by what happens between passes: a launch, or a release / acquire hand-off
// through L2, costs more than the pass itself.  Here every CTA keeps its strip of rows in shared
// memory for the WHOLE step: it is loaded once (cp.async.bulk, the TMA engine's 1-D copy) and
// written back once.  One generation is one sweep over the shared-memory strip by all warps
// (3-row register window per lane, x-neighbour words read from shared memory, so no halo lanes
// and no shuffles) followed by a block barrier.  Every K generations neighbouring CTAs swap K
// halo rows as flag-in-data words: straight into the neighbour's shared memory when it is in
// the same thread-block cluster (mapa + st.shared::cluster, DSMEM), through a mailbox in L2
// otherwise, and through the ring neighbour's ll_recv buffers at the two slab edges (NVLink).
// The reader polls the words themselves; there is no flag, fence or cluster barrier per pass.
// Replaces the whole loop of gol-main.c:94-116 for slabs of up to a few MiB.
namespace resident {
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cluster_map(uint32_t local_smem_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ void st_cluster_ll(uint32_t addr, uint64_t v, uint32_t seq) {
  const unsigned long long a = (v & 0xffffffffull) | ((unsigned long long)seq << 32);
  const unsigned long long b = (v >> 32) | ((unsigned long long)seq << 32);
  asm volatile("st.shared::cluster.v2.u64 [%0], {%1, %2};" ::"r"(addr), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_shared_ll(uint32_t addr, unsigned long long& a, unsigned long long& b) {
  asm volatile("ld.volatile.shared.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ld_global_ll(const uint64_t* p, unsigned long long& a, unsigned long long& b) {
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// cp.async.bulk global -> shared, completion on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_addr(src)), "r"(bytes)
               : "memory");
}
}  // namespace resident

namespace resident {
enum { kRing = 0, kCluster = 1, kMail = 2 };   // how a strip boundary's halo rows travel

// Where CTA `cta` of slab `sidx` receives the halo of side (0 = rows above, 1 = rows below), parity par.
__device__ __forceinline__ uint64_t* mail_slot(const ResidentParams& rp, int sidx, int cta, int side, int par) {
  return rp.mail + ((((int64_t)sidx * rp.ctas_per_slab + cta) * 2 + side) * 2 + par) * rp.mail_slot;
}

// One halo word, polled until both halves carry `seq` (flag-in-data); shared-memory or global slot.
// The watchdog is looked at every 256 polls only: its clock read and the error-flag load cost more
// than the poll itself.
__device__ __forceinline__ uint64_t poll_word(SlabState* st, bool in_smem, const uint64_t* slot, int64_t idx,
                                              uint32_t seq, int side) {
  unsigned long long a, b, t0 = 0;
  unsigned spins = 0;
  if (!in_smem && !GOL_CHK(slot + idx * 2, 16)) return 0;
  for (;;) {
    if (in_smem) ld_shared_ll(smem_addr(slot + idx * 2), a, b);
    else ld_global_ll(slot + idx * 2, a, b);
    if ((uint32_t)(a >> 32) == seq && (uint32_t)(b >> 32) == seq) break;
    if ((++spins & 255u) == 0) {
      if (!t0) t0 = globaltimer_ns();
      if (globaltimer_ns() - t0 > st->watchdog_ns || *(volatile int*)&st->error) {
        atomicCAS(&st->error, 0, 1 + side);
        break;
      }
    }
  }
  return (a & 0xffffffffull) | (b << 32);
}

// One generation of shared-memory rows [lo, hi) (row index = y - row_base) of the strip: words at
// sm[cur + row * ws + w] -> sm[nxt + ...].  Warp (cb, rs) sweeps its share of the rows of column
// block cb top to bottom; each lane keeps the horizontal sums of the two previous rows of its word in
// registers and has the three words of the next row in flight while it computes the current one.
// `sm` is the kernel's extern __shared__ array itself, so the accesses compile to LDS / STS.
__device__ __forceinline__ void sweep(uint64_t* sm, int cur, int nxt, int lo, int hi, int ws, int warp, int nwarps,
                                      int lane, int limit) {
  const int ncb = (ws + 31) >> 5;
  const int nrs = nwarps >= ncb ? nwarps / ncb : 1;          // warps that share a column block
  const int rs = nwarps >= ncb ? warp / ncb : 0;
  if (rs >= nrs) return;
  const int n = hi - lo;
  const int ya = lo + n * rs / nrs, yb = lo + n * (rs + 1) / nrs;
  if (ya >= yb) return;
  for (int cb = warp % ncb; cb < ncb; cb += (nwarps >= ncb ? ncb : nwarps)) {
    const int w = cb * 32 + lane;
    if (w >= ws) continue;
    const int wl = w ? w - 1 : ws - 1, wr = w + 1 < ws ? w + 1 : 0;   // torus in x
    Stage s;
    s.lo.a0 = s.lo.a1 = s.lo.b0 = s.lo.b1 = s.lo.c = 0;
    s.hi.a0 = s.hi.a1 = s.hi.b0 = s.hi.b1 = s.hi.c = 0;
    // feed one row (its word c and the neighbouring words l, r); returns generation + 1 of the row before
    auto feed = [&](uint64_t c, uint64_t l, uint64_t r) -> uint64_t {
      const uint32_t clo = (uint32_t)c, chi = (uint32_t)(c >> 32);
      // even columns (low half): left = previous odd column, right = odd column at the same index
      const uint32_t Llo = __funnelshift_l((uint32_t)(l >> 32), chi, 1);
      const uint32_t Rhi = __funnelshift_r(clo, (uint32_t)r, 1);
      const uint32_t h0lo = lop3<0x96>(Llo, clo, chi), h1lo = lop3<0xE8>(Llo, clo, chi);
      const uint32_t h0hi = lop3<0x96>(clo, chi, Rhi), h1hi = lop3<0xE8>(clo, chi, Rhi);
      const uint32_t olo = stage_half(s.lo, h0lo, h1lo, clo);
      const uint32_t ohi = stage_half(s.hi, h0hi, h1hi, chi);
      return ((uint64_t)ohi << 32) | olo;
    };
    int pc = cur + (ya - 1) * ws + w, pl = cur + (ya - 1) * ws + wl, pr = cur + (ya - 1) * ws + wr;
    int out = nxt + ya * ws + w;
    // bounds-checking build: the first and the last word this sweep touches (rows ya-1 .. yb+1 read,
    // rows ya .. yb-1 written) must lie inside the kernel's shared memory
    if (!GOL_SCHK(cur + (ya - 1) * ws, 1, limit) || !GOL_SCHK(cur + (yb + 1) * ws + ws - 1, 1, limit) ||
        !GOL_SCHK(out, 1, limit) || !GOL_SCHK(nxt + (yb - 1) * ws + w, 1, limit))
      continue;
    // rows ya-1 and ya only fill the window; from row ya+1 on every row fed yields the row above it.
    // The words of row y+1 are in flight while row y is computed; the last iteration reads one row
    // past yb, which is still inside the kernel's shared memory (the other buffer or the landing zones).
    uint64_t c0 = sm[pc], l0 = sm[pl], r0 = sm[pr];
    pc += ws; pl += ws; pr += ws;
    uint64_t c1 = sm[pc], l1 = sm[pl], r1 = sm[pr];
    pc += ws; pl += ws; pr += ws;
    feed(c0, l0, r0);
    c0 = sm[pc]; l0 = sm[pl]; r0 = sm[pr];
    pc += ws; pl += ws; pr += ws;
    feed(c1, l1, r1);
#pragma unroll 2
    for (int y = ya + 1; y <= yb; ++y) {
      c1 = sm[pc]; l1 = sm[pl]; r1 = sm[pr];
      pc += ws; pl += ws; pr += ws;
      sm[out] = feed(c0, l0, r0);
      out += ws;
      c0 = c1; l0 = l1; r0 = r1;
    }
  }
}
}  // namespace resident

__global__ void __launch_bounds__(512, 1) life_resident_kernel(SlabState* states, ResidentParams rp) {
  using namespace resident;
  extern __shared__ __align__(16) uint64_t rsmem[];
  __shared__ __align__(8) uint64_t load_bar;
  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarps = nthreads >> 5;
  const int C = rp.ctas_per_slab, K = rp.K;
  const int sidx = blockIdx.x / C, cta = blockIdx.x - sidx * C;
  SlabState* st = states + sidx;
  const int64_t H = st->H, stride = st->stride;
  const int ws = (int)st->ws;
  const int64_t r0 = H * cta / C, r1 = H * (cta + 1) / C;        // my rows [r0, r1)
  const int64_t row_base = r0 - K;                                // shared-memory row 0 <-> y = r0 - K
  const int nr = (int)(r1 - r0);                                  // my row count
  const int boff[2] = {0, rp.rows_alloc * ws};                    // word offsets of the two row buffers in rsmem
  uint64_t* buf[2] = {rsmem, rsmem + (int64_t)rp.rows_alloc * ws};
  uint64_t* land = rsmem + 2 * (int64_t)rp.rows_alloc * ws;       // [side][parity][K * ws * 2]: DSMEM landing zones
  const int64_t land_slot = 2 * (int64_t)K * ws;
  const int cs = rp.cluster, cr = cta % cs;
  const int up_kind = cta == 0 ? kRing : (cr != 0 ? kCluster : kMail);
  const int dn_kind = cta == C - 1 ? kRing : (cr != cs - 1 ? kCluster : kMail);

  // landing zones start with exchange number 0 (never valid); nobody may write into them before
  for (int64_t i = tid; i < 4 * land_slot; i += nthreads) land[i] = 0;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&load_bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (cs > 1) cluster_sync_all();

  // ---- load my strip (current generation) into buf[0]
  const uint64_t* gin = st->buf[st->pass & 1] + (int64_t)st->G * stride;
  const int64_t nrows = r1 - r0;
  if (rp.use_bulk) {
    // rows are contiguous (stride == ws): the TMA engine copies the strip in chunks of <= 64 KiB
    if (tid == 0) {
      const uint64_t total = (uint64_t)nrows * ws * 8;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&load_bar)),
                   "r"((uint32_t)total)
                   : "memory");
      for (uint64_t off = 0; off < total; off += 65536) {
        const uint32_t n = (uint32_t)(total - off < 65536 ? total - off : 65536);
        if (GOL_CHK((const char*)(gin + r0 * stride) + off, n))
          bulk_g2s((char*)(buf[0] + (int64_t)K * ws) + off, (const char*)(gin + r0 * stride) + off, n, &load_bar);
      }
    }
    uint32_t done = 0;
    while (!done)
      asm volatile(
          "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(smem_addr(&load_bar))
          : "memory");
  } else {
    for (int64_t i = tid; i < nrows * ws; i += nthreads) {
      const int64_t r = i / ws, w = i - r * ws;
      buf[0][(K + r) * ws + w] = ld_cg(gin + (r0 + r) * stride + w);
    }
  }
  __syncthreads();

  const unsigned long long exch0 = st->exch;
  unsigned long long wait_acc = 0;   // thread 0: time between "my rows are pushed" and "my halos are in"
  int cb = 0;
  for (int p = 0; p < rp.npasses; ++p) {
    const int kk = p == rp.npasses - 1 ? rp.last_k : K;
    const unsigned long long e = exch0 + (unsigned long long)p;
    const uint32_t seq = (uint32_t)(e + 1ull);
    const int par = (int)(e & 1ull);
    const uint64_t* cur = buf[cb];

    // ---- push my K top rows to the strip above (its bottom halo) and my K bottom rows to the strip below
    for (int i = tid; i < 2 * K * ws; i += nthreads) {
      const int side = i >= K * ws;                          // 0: my top rows go up, 1: my bottom rows go down
      const int j = side ? i - K * ws : i;
      const int r = j / ws, w = j - r * ws;
      const int64_t y = side ? r1 - K + r : r0 + r;
      const uint64_t v = GOL_SCHK((cur - rsmem) + (y - row_base) * ws + w, 1, rp.smem_words) ? cur[(y - row_base) * ws + w] : 0;
      const int kind = side ? dn_kind : up_kind;
      if (kind == kRing) {
        ll_store(st->peer_ll[side][par] + (r * stride + w) * 2, v, seq);
      } else if (kind == kCluster) {
        // the neighbour's landing zone for its opposite side, same offset inside its shared memory
        const int64_t loff = (land - rsmem) + ((1 - side) * 2 + par) * land_slot + (r * ws + w) * 2;
        const uint32_t mine = smem_addr(rsmem + loff);
        // same layout in every CTA of the cluster: the offset is valid there iff it is valid here
        if (GOL_SCHK(loff, 2, rp.smem_words) && (uint32_t)(side ? cr + 1 : cr - 1) < (uint32_t)cs)
          st_cluster_ll(cluster_map(mine, (uint32_t)(side ? cr + 1 : cr - 1)), v, seq);
      } else {
        ll_store(mail_slot(rp, sidx, side ? cta + 1 : cta - 1, 1 - side, par) + (r * ws + w) * 2, v, seq);
      }
    }
    // ---- receive the K rows above and below into the halo rows of the current buffer
    const unsigned long long tw0 = tid == 0 ? globaltimer_ns() : 0ull;
    uint64_t* curw = buf[cb];
    for (int i = tid; i < 2 * K * ws; i += nthreads) {
      const int side = i >= K * ws;
      const int j = side ? i - K * ws : i;
      const int r = j / ws, w = j - r * ws;
      const int kind = side ? dn_kind : up_kind;
      uint64_t v;
      if (kind == kRing) v = poll_word(st, false, st->ll_recv[side][par], r * stride + w, seq, side);
      else if (kind == kCluster) v = poll_word(st, true, land + (side * 2 + par) * land_slot, r * ws + w, seq, side);
      else v = poll_word(st, false, mail_slot(rp, sidx, cta, side, par), r * ws + w, seq, side);
      const int64_t y = side ? r1 + r : r0 - K + r;
      if (GOL_SCHK((curw - rsmem) + (y - row_base) * ws + w, 1, rp.smem_words)) curw[(y - row_base) * ws + w] = v;
    }
    __syncthreads();
    if (tid == 0) wait_acc += globaltimer_ns() - tw0;   // push is done, everything from here to the barrier is waiting

    // ---- kk generations on a shrinking window: after generation g rows [r0-(kk-1-g), r1+(kk-1-g)) are valid
    for (int g = 0; g < kk; ++g) {
      const int m = kk - 1 - g;
      sweep(rsmem, boff[cb], boff[cb ^ 1], K - m, K + nr + m, ws, warp, nwarps, lane, rp.smem_words);
      __syncthreads();
      cb ^= 1;
    }
  }

  // ---- write my strip back (next buffer of the slab)
  uint64_t* gout = st->buf[(st->pass + 1) & 1] + (int64_t)st->G * stride;
  if (rp.use_bulk) {
    if (tid == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> async proxy reads
      const uint64_t total = (uint64_t)nrows * ws * 8;
      for (uint64_t off = 0; off < total; off += 65536) {
        const uint32_t n = (uint32_t)(total - off < 65536 ? total - off : 65536);
        if (GOL_CHK((char*)(gout + r0 * stride) + off, n))
          bulk_s2g((char*)(gout + r0 * stride) + off, (const char*)(buf[cb] + (int64_t)K * ws) + off, n);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
  } else {
    for (int64_t i = tid; i < nrows * ws; i += nthreads) {
      const int64_t r = i / ws, w = i - r * ws;
      st_cg(gout + (r0 + r) * stride + w, buf[cb][(K + r) * ws + w]);
    }
  }
  __syncthreads();
  if (cs > 1) cluster_sync_all();   // nobody leaves while a cluster neighbour may still write into its shared memory

  // the last CTA of the launch advances the slab counters (one buffer flip per launch)
  if (tid == 0) {
    atomicAdd(&st->wait_ns[cta == 0 ? 0 : 1], wait_acc);   // [0]: the CTA at the ring edge, [1]: all the others
    atomicAdd(&st->wait_count[cta == 0 ? 0 : 1], (unsigned long long)rp.npasses);
    __threadfence();
    if (atomicAdd(&rp.ctl[1], 1u) == gridDim.x - 1) {
      __threadfence();
      for (int s2 = 0; s2 < rp.nslabs; ++s2) {
        states[s2].pass += 1ull;
        states[s2].exch += (unsigned long long)rp.npasses;
      }
      rp.ctl[1] = 0u;
      __threadfence();
    }
  }
}


// ---- N-K4: pattern initialisers ------------------------------------------------
__global__ void pattern_init_kernel(uint64_t* rows, int64_t stride, PatternArgs a) {
  const int64_t ws = words_per_row(a.W);
  const int64_t total = a.H * ws;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t y = i / ws, w = i - y * ws;
    uint64_t v;
    if (a.pattern == kAllOnes) {
      v = (w == ws - 1) ? tail_mask(a.W) : ~0ull;
    } else if (a.pattern == kRandom || a.pattern == kGlider) {
      v = pattern_word(a, y, w);
    } else {
      // sparse patterns: only a handful of words are non-zero
      const bool maybe = (y == 0 || y == a.H - 1) && (w <= 2 || w == ws - 1);
      v = maybe ? pattern_word(a, y, w) : 0ull;
    }
    if (GOL_CHK(rows + y * stride + w, 8)) rows[y * stride + w] = v;
  }
}

// ---- N-K7: pack / unpack ---------------------------------------------------------
__device__ __forceinline__ uint64_t gather8(uint64_t x) {
  // 8 bytes (any non-zero = alive) -> 8 bits
  x = (((x & 0x7f7f7f7f7f7f7f7full) + 0x7f7f7f7f7f7f7f7full) | x) & 0x8080808080808080ull;
  x >>= 7;
  return (x * 0x0102040810204080ull) >> 56;
}
__device__ __forceinline__ uint64_t spread8(uint64_t b) {
  // 8 bits -> 8 bytes of 0/1
  uint64_t x = (b * 0x0101010101010101ull) & 0x8040201008040201ull;
  return ((x + 0x7f7f7f7f7f7f7f7full) >> 7) & 0x0101010101010101ull;
}

__global__ void pack_kernel(const uint8_t* __restrict__ cells, uint64_t* __restrict__ words,
                            int64_t W, int64_t H, int64_t stride) {
  const int64_t ws = words_per_row(W);
  const int64_t total = H * ws;
  const bool aligned = (W % 8) == 0 && (((uintptr_t)cells) & 7) == 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t y = i / ws, w = i - y * ws;
    const uint8_t* src = cells + y * W + w * 64;
    const int n = (int)((W - w * 64) < 64 ? (W - w * 64) : 64);
    uint64_t v = 0;
    if (!GOL_CHK(src, n) || !GOL_CHK(words + y * stride + w, 8)) continue;
    if (aligned && n == 64) {
      const uint64_t* s8 = (const uint64_t*)src;
#pragma unroll
      for (int j = 0; j < 8; ++j) v |= gather8(s8[j]) << (8 * j);
    } else {
      for (int j = 0; j < n; ++j) v |= (uint64_t)(src[j] != 0) << j;
    }
    words[y * stride + w] = std_to_il(v);   // column order -> storage layout
  }
}

__global__ void unpack_kernel(const uint64_t* __restrict__ words, uint8_t* __restrict__ cells,
                              int64_t W, int64_t H, int64_t stride) {
  const int64_t ws = words_per_row(W);
  const int64_t total = H * ws;
  const bool aligned = (W % 8) == 0 && (((uintptr_t)cells) & 7) == 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t y = i / ws, w = i - y * ws;
    if (!GOL_CHK(words + y * stride + w, 8)) continue;
    const uint64_t v = il_to_std(words[y * stride + w]);   // storage layout -> column order
    uint8_t* dst = cells + y * W + w * 64;
    const int n = (int)((W - w * 64) < 64 ? (W - w * 64) : 64);
    if (!GOL_CHK(dst, n)) continue;
    if (aligned && n == 64) {
      uint64_t* d8 = (uint64_t*)dst;
#pragma unroll
      for (int j = 0; j < 8; ++j) d8[j] = spread8((v >> (8 * j)) & 0xffull);
    } else {
      for (int j = 0; j < n; ++j) dst[j] = (uint8_t)((v >> j) & 1ull);
    }
  }
}

// ---- population ---------------------------------------------------------------------
__global__ void __launch_bounds__(kPopThreads) population_partial_kernel(
    const uint64_t* __restrict__ words, int64_t ws, int64_t stride, int64_t H,
    unsigned long long* __restrict__ partials) {
  unsigned long long acc = 0;
  for (int64_t y = blockIdx.x; y < H; y += gridDim.x) {
    const uint64_t* row = words + y * stride;
    if (!GOL_CHK(row, ws * 8)) continue;
    for (int64_t w = threadIdx.x; w < ws; w += blockDim.x) acc += (unsigned)__popcll(row[w]);
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(kFull, acc, o);
  __shared__ unsigned long long wsum[kPopThreads / 32];
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int i = 0; i < kPopThreads / 32; ++i) t += wsum[i];
    if (GOL_CHK(partials + blockIdx.x, 8)) partials[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(kPopThreads) population_final_kernel(
    const unsigned long long* __restrict__ partials, int n, unsigned long long* __restrict__ out) {
  unsigned long long acc = 0;
  if (!GOL_CHK(partials, (long long)n * 8) || !GOL_CHK(out, 8)) return;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += partials[i];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(kFull, acc, o);
  __shared__ unsigned long long wsum[kPopThreads / 32];
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int i = 0; i < kPopThreads / 32; ++i) t += wsum[i];
    *out = t;
  }
}

// ---- N-K6: row text formatter ----------------------------------------------------------
// Layout of one output row (gol-main.c:17-28): "Row %2d: " then "%u " per cell
// then '\n'.  The cell text starts at a 16-byte aligned offset so a word of 64
// cells becomes eight 16-byte stores; the prefix is written right-aligned in
// front of it and the host knows its length.
__global__ void format_rows_kernel(const uint64_t* __restrict__ rows, int64_t stride, int64_t W,
                                   int64_t nrows, int64_t first_global_row, char* __restrict__ text,
                                   int64_t pitch) {
  const int64_t ws = words_per_row(W);
  const int64_t total = nrows * ws;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / ws, w = i - r * ws;
    if (!GOL_CHK(rows + r * stride + w, 8)) continue;
    const uint64_t v = il_to_std(rows[r * stride + w]);   // storage layout -> column order
    char* row_text = text + r * pitch;
    // everything this thread may write: its cells, the newline (last word), the prefix (first word)
    if (!GOL_CHK(row_text + kTextCellOff + w * 128, 2 * ((W - w * 64) < 64 ? (W - w * 64) : 64)) ||
        (w == ws - 1 && !GOL_CHK(row_text + kTextCellOff + 2 * W, 1)) || (w == 0 && !GOL_CHK(row_text, kTextCellOff)))
      continue;
    char* dst = row_text + kTextCellOff + w * 128;
    const int n = (int)((W - w * 64) < 64 ? (W - w * 64) : 64);
    if (n == 64) {
      uint4* d16 = (uint4*)dst;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t b = (uint32_t)(v >> (8 * j)) & 0xffu;
        uint4 q;
        q.x = 0x20302030u + (b & 1u) + (((b >> 1) & 1u) << 16);
        q.y = 0x20302030u + ((b >> 2) & 1u) + (((b >> 3) & 1u) << 16);
        q.z = 0x20302030u + ((b >> 4) & 1u) + (((b >> 5) & 1u) << 16);
        q.w = 0x20302030u + ((b >> 6) & 1u) + (((b >> 7) & 1u) << 16);
        d16[j] = q;
      }
    } else {
      for (int j = 0; j < n; ++j) {
        dst[2 * j] = (char)('0' + ((v >> j) & 1ull));
        dst[2 * j + 1] = ' ';
      }
    }
    if (w == ws - 1) row_text[kTextCellOff + 2 * W] = '\n';
    if (w == 0) {
      // "Row %2d: " right-aligned so that it ends at kTextCellOff
      int g = (int)(first_global_row + r);
      char* p = row_text + kTextCellOff;
      *--p = ' ';
      *--p = ':';
      int digits = 0;
      const bool neg = g < 0;
      unsigned u = neg ? 0u - (unsigned)g : (unsigned)g;
      do {
        *--p = (char)('0' + u % 10u);
        u /= 10u;
        ++digits;
      } while (u);
      if (neg) { *--p = '-'; ++digits; }
      while (digits < 2) { *--p = ' '; ++digits; }
      *--p = ' ';
      *--p = 'w';
      *--p = 'o';
      *--p = 'R';
    }
  }
}

// ---- stand-alone steps (no ring state) -----------------------------------------------------
// top / bottom are the rows above row 0 and below row H-1: the slab's own last /
// first row for a torus, or explicit ghost rows for the unfused NCCL baseline.
__global__ void step_rows_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ out,
                                 const uint64_t* __restrict__ top, const uint64_t* __restrict__ bottom,
                                 int64_t W, int64_t H, int64_t stride) {
  const int64_t ws = words_per_row(W);
  const int64_t total = H * ws;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t y = i / ws, w = i - y * ws;
    const uint64_t* cur = in + y * stride;
    const uint64_t* up = (y == 0) ? top : cur - stride;
    const uint64_t* down = (y == H - 1) ? bottom : cur + stride;
    if (GOL_CHK(up, ws * 8) && GOL_CHK(cur, ws * 8) && GOL_CHK(down, ws * 8) && GOL_CHK(out + y * stride + w, 8))
      out[y * stride + w] = step_word(up, cur, down, w, ws, W);
  }
}

// ---- instantiation table of the K-templated kernels -----------------------------------
// K      : generations per pass
// MAXT   : largest block the instantiation accepts (register budget)
template <int K> struct FusedCfg;
template <> struct FusedCfg<1> { static constexpr int kMaxT = 256, kMinB = 2; };
template <> struct FusedCfg<2> { static constexpr int kMaxT = 256, kMinB = 2; };
template <> struct FusedCfg<3> { static constexpr int kMaxT = 256, kMinB = 2; };
template <> struct FusedCfg<4> { static constexpr int kMaxT = 256, kMinB = 2; };
template <> struct FusedCfg<6> { static constexpr int kMaxT = 128, kMinB = 3; };
template <> struct FusedCfg<8> { static constexpr int kMaxT = 128, kMinB = 3; };

// Occupancy experiment (GOL_EXP_MINB=3|4, K = 2 and 3 only): the same kernel compiled for more resident blocks.
static int exp_minb() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GOL_EXP_MINB");
    v = e ? atoi(e) : 0;
  }
  return v;
}
template <int K, int V = 0> const void* fused_fn() {
  if (K == 2 || K == 3) {
    if (exp_minb() == 3) return (const void*)&life_fused_kernel<(K == 2 || K == 3) ? K : 2, 256, 3, V>;
    if (exp_minb() == 4) return (const void*)&life_fused_kernel<(K == 2 || K == 3) ? K : 2, 256, 4, V>;
  }
  return (const void*)&life_fused_kernel<K, FusedCfg<K>::kMaxT, FusedCfg<K>::kMinB, V>;
}
template <int K> const void* flow_fn() {
  return (const void*)&life_flow_kernel<K, FusedCfg<K>::kMaxT, FusedCfg<K>::kMinB>;
}

#define GOL_DISPATCH_K(K, expr)                       \
  switch (K) {                                        \
    case 1: { constexpr int KK = 1; expr; } break;    \
    case 2: { constexpr int KK = 2; expr; } break;    \
    case 3: { constexpr int KK = 3; expr; } break;    \
    case 4: { constexpr int KK = 4; expr; } break;    \
    case 6: { constexpr int KK = 6; expr; } break;    \
    case 8: { constexpr int KK = 8; expr; } break;    \
    default: break;                                   \
  }

}  // namespace

// ---- host side ---------------------------------------------------------------------------

// ---- bounds-checking build: host side of the region registry -----------------------------------
#if GOL_DEBUG_BOUNDS
bool debug_bounds_enabled() { return true; }
static int dbg_fetch(DebugRegions* h) { return cudaMemcpyFromSymbol(h, g_dbg, sizeof *h) == cudaSuccess ? 0 : -1; }
static int dbg_store(const DebugRegions* h) { return cudaMemcpyToSymbol(g_dbg, h, sizeof *h) == cudaSuccess ? 0 : -1; }
int debug_bounds_register(const void* p, size_t n) {
  if (!p || !n) return 0;
  cudaDeviceSynchronize();
  DebugRegions h;
  if (dbg_fetch(&h)) return -1;
  const unsigned long long lo = (unsigned long long)p, hi = lo + n;
  for (int i = 0; i < h.n; ++i)
    if (h.lo[i] == lo) {   // same base again: keep the larger extent
      if (hi > h.hi[i]) h.hi[i] = hi;
      return dbg_store(&h);
    }
  if (h.n >= kDbgMaxRegions) return -1;
  h.lo[h.n] = lo;
  h.hi[h.n] = hi;
  ++h.n;
  return dbg_store(&h);
}
int debug_bounds_unregister(const void* p) {
  if (!p) return 0;
  cudaDeviceSynchronize();
  DebugRegions h;
  if (dbg_fetch(&h)) return -1;
  for (int i = 0; i < h.n; ++i)
    if (h.lo[i] == (unsigned long long)p) {
      h.lo[i] = h.lo[h.n - 1];
      h.hi[i] = h.hi[h.n - 1];
      --h.n;
      break;
    }
  return dbg_store(&h);
}
int debug_bounds_query(unsigned long long* violations, unsigned long long* first_bad, int* line, int reset) {
  cudaDeviceSynchronize();
  DebugRegions h;
  if (dbg_fetch(&h)) return -1;
  if (violations) *violations = h.violations;
  if (first_bad) *first_bad = h.first_bad;
  if (line) *line = h.first_line;
  if (reset) {
    h.violations = 0;
    h.first_bad = 0;
    h.first_line = 0;
    return dbg_store(&h);
  }
  return 0;
}
namespace {
__global__ void debug_poke_kernel(const uint64_t* p, uint64_t* sink) {
  const uint64_t v = ld_cg(p);
  if (sink && GOL_CHK(sink, 8)) *sink = v;
}
}  // namespace
int debug_bounds_poke(const void* p) {
  debug_poke_kernel<<<1, 1>>>((const uint64_t*)p, nullptr);
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}
#else
bool debug_bounds_enabled() { return false; }
int debug_bounds_register(const void*, size_t) { return 0; }
int debug_bounds_unregister(const void*) { return 0; }
int debug_bounds_query(unsigned long long* violations, unsigned long long* first_bad, int* line, int) {
  if (violations) *violations = 0;
  if (first_bad) *first_bad = 0;
  if (line) *line = 0;
  return 0;
}
int debug_bounds_poke(const void*) { return 0; }
#endif

static uint64_t g_launches = 0;
void count_launch(uint64_t n) { __atomic_fetch_add(&g_launches, n, __ATOMIC_RELAXED); }
uint64_t launch_count_value() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

// GOL_PDL=1: launch fused passes with programmatic dependent launch.
static bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GOL_PDL");
    v = e ? atoi(e) : 0;
  }
  return v != 0;
}

bool fused_k_supported(int K) { return K == 1 || K == 2 || K == 3 || K == 4 || K == 6 || K == 8; }

FusedLimits fused_limits(int K) {
  FusedLimits l{0};
  GOL_DISPATCH_K(K, l.max_threads = FusedCfg<KK>::kMaxT);
  return l;
}

int fused_resident_warps_per_sm(int K, int threads, int device) {
  (void)device;
  const void* fn = nullptr;
  GOL_DISPATCH_K(K, fn = fused_fn<KK>());
  if (!fn) return 0;
  int nb = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, threads, 0) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return nb * (threads / 32);
}

int flow_max_ctas(int K, int threads, int device) {
  const void* fn = nullptr;
  GOL_DISPATCH_K(K, fn = flow_fn<KK>());
  if (!fn) return 0;
  int nb = 0, sms = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, threads, 0) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  return nb * sms;
}

cudaError_t launch_fused(int K, SlabState* d_states, int nslabs, const PassParams& p, int threads,
                         cudaStream_t s) {
  const void* fn = nullptr;
  GOL_DISPATCH_K(K, fn = fused_fn<KK>());
  if (!fn) return cudaErrorInvalidValue;
  const int wpb = threads / 32;
  const int ntasks = p.nstrips * p.col_groups;
  dim3 grid((unsigned)((ntasks + wpb - 1) / wpb), (unsigned)nslabs);
  PassParams pp = p;
  void* args[] = {(void*)&d_states, (void*)&pp};
  count_launch(1);
  if (pdl_enabled()) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3((unsigned)threads);
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelExC(&cfg, fn, args);
  }
  return cudaLaunchKernel(fn, grid, dim3((unsigned)threads), args, 0, s);
}

bool fused_tma_supported(int K, int64_t ws) {
  return K >= 1 && K <= 4 && (ws % 16) == 0 && ws >= tma::seg_words(4);
}
// 3 or 4 compute warps per CTA: the count that leaves fewer idle column groups (9 groups -> 3 x 3)
int fused_tma_warps(int col_groups) {
  const int w4 = (col_groups + 3) / 4 * 4, w3 = (col_groups + 2) / 3 * 3;
  return w3 < w4 ? 3 : 4;
}
int fused_tma_box_words(int nw) { return nw == 3 ? tma::seg_words(3) : tma::seg_words(4); }
int fused_tma_box_rows() { return tma::kChunk; }

int fused_tma_rank() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GOL_TMA_RANK");
    v = (e && atoi(e) == 2) ? 2 : 3;
  }
  return v;
}

cudaError_t launch_fused_tma(int K, SlabState* d_states, int nslabs, const PassParams& p0,
                             const CUtensorMap* d_tmaps, cudaStream_t s) {
  PassParams p = p0;
  p.tma_rank = fused_tma_rank();
  const int nw = fused_tma_warps(p.col_groups);
  const int ncg4 = (p.col_groups + nw - 1) / nw;
  dim3 grid((unsigned)(p.nstrips * ncg4), (unsigned)nslabs);
  dim3 block((unsigned)nw * 32);
  count_launch(1);
#define GOL_TMA_CASE(KK)                                                                      \
  case KK:                                                                                    \
    if (nw == 3) life_fused_tma_kernel<KK, 3><<<grid, block, 0, s>>>(d_states, p, d_tmaps);   \
    else life_fused_tma_kernel<KK, 4><<<grid, block, 0, s>>>(d_states, p, d_tmaps);           \
    break;
  switch (K) {
    GOL_TMA_CASE(1)
    GOL_TMA_CASE(2)
    GOL_TMA_CASE(3)
    GOL_TMA_CASE(4)
    default: return cudaErrorInvalidValue;
  }
#undef GOL_TMA_CASE
  return cudaGetLastError();
}

cudaError_t launch_generic(SlabState* d_states, int nslabs, const PassParams& p, int threads,
                           int64_t max_words, cudaStream_t s) {
  dim3 grid((unsigned)((max_words + threads - 1) / threads), (unsigned)nslabs);
  count_launch(1);
  life_generic_kernel<<<grid, threads, 0, s>>>(d_states, p);
  return cudaGetLastError();
}

cudaError_t launch_prime(SlabState* d_states, int nslabs, int depth, int64_t max_ws,
                         cudaStream_t s) {
  const int threads = 256;
  int64_t work = 2 * (int64_t)depth * max_ws;
  int blocks = (int)((work + threads - 1) / threads);
  if (blocks < 1) blocks = 1;
  if (blocks > 64) blocks = 64;
  count_launch(1);
  halo_prime_kernel<<<dim3((unsigned)blocks, (unsigned)nslabs), threads, 0, s>>>(d_states, depth);
  return cudaGetLastError();
}

cudaError_t launch_flow(int K, SlabState* d_states, const FlowParams& fp0, int threads, int nctas,
                        cudaStream_t s) {
  const void* fn = nullptr;
  GOL_DISPATCH_K(K, fn = flow_fn<KK>());
  if (!fn) return cudaErrorInvalidValue;
  FlowParams fp = fp0;
  fp.total_warps = nctas * (threads / 32);
  void* args[] = {(void*)&d_states, (void*)&fp};
  count_launch(1);
  return cudaLaunchKernel(fn, dim3((unsigned)nctas), dim3((unsigned)threads), args, 0, s);
}

// Grid of a grid-stride helper kernel: enough blocks for the data, at most 32 per SM of the
// current device.
// ---- resident kernel: launch with a cluster dimension, all CTAs co-resident (cooperative) ---------
static cudaError_t resident_config(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, int nctas, int threads,
                                   size_t smem, int cluster, cudaStream_t s) {
  cudaError_t e = cudaFuncSetAttribute((const void*)life_resident_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem);
  if (e != cudaSuccess) return e;
  // all of the SM's L1 / shared memory as shared memory: the kernel lives in it
  cudaFuncSetAttribute((const void*)life_resident_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                       cudaSharedmemCarveoutMaxShared);
  cudaGetLastError();
  *cfg = cudaLaunchConfig_t{};
  cfg->gridDim = dim3((unsigned)nctas);
  cfg->blockDim = dim3((unsigned)threads);
  cfg->dynamicSmemBytes = smem;
  cfg->stream = s;
  int n = 0;
  attr[n].id = cudaLaunchAttributeCooperative;   // CTAs poll each other's halo words: all must be resident
  attr[n].val.cooperative = 1;
  ++n;
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = (unsigned)cluster;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg->attrs = attr;
  cfg->numAttrs = (unsigned)n;
  return cudaSuccess;
}

size_t resident_smem_bytes(int rows_alloc, int64_t ws, int K) {
  return ((size_t)2 * (size_t)rows_alloc * (size_t)ws + (size_t)8 * (size_t)K * (size_t)ws) * 8;
}

// CTAs of `threads` threads and `smem` bytes that can be resident at once in clusters of `cluster`.
int resident_max_ctas(int threads, size_t smem, int cluster, int device) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (cluster <= 1) {
    int nb = 0;
    cudaFuncSetAttribute((const void*)life_resident_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                         cudaSharedmemCarveoutMaxShared);
    if (cudaFuncSetAttribute((const void*)life_resident_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
            cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, (const void*)life_resident_kernel, threads, smem) != cudaSuccess) {
      cudaGetLastError();
      return 0;
    }
    return nb > 0 ? sms : 0;   // one CTA per SM by design (__launch_bounds__(512, 1))
  }
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[2];
  if (resident_config(&cfg, attr, sms / cluster * cluster, threads, smem, cluster, nullptr) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  int nclusters = 0;
  if (cudaOccupancyMaxActiveClusters(&nclusters, (const void*)life_resident_kernel, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return nclusters * cluster;
}

cudaError_t launch_resident(SlabState* d_states, const ResidentParams& rp0, int threads, size_t smem, cudaStream_t s) {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[2];
  ResidentParams rp = rp0;
  cudaError_t e = resident_config(&cfg, attr, rp.nslabs * rp.ctas_per_slab, threads, smem, rp.cluster, s);
  if (e != cudaSuccess) return e;
  void* args[] = {(void*)&d_states, (void*)&rp};
  count_launch(1);
  return cudaLaunchKernelExC(&cfg, (const void*)life_resident_kernel, args);
}

static int grid_for(int64_t total, int threads) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
      sms < 1)
    sms = 128;
  int64_t b = (total + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > (int64_t)sms * 32) b = (int64_t)sms * 32;
  return (int)b;
}

cudaError_t launch_pattern_init(uint64_t* d_rows, int64_t stride, const PatternArgs& a,
                                cudaStream_t s) {
  const int64_t total = a.H * words_per_row(a.W);
  count_launch(1);
  pattern_init_kernel<<<grid_for(total, 256), 256, 0, s>>>(d_rows, stride, a);
  return cudaGetLastError();
}

cudaError_t launch_pack(const uint8_t* d_cells, uint64_t* d_words, int64_t W, int64_t H,
                        int64_t stride, cudaStream_t s) {
  count_launch(1);
  pack_kernel<<<grid_for(H * words_per_row(W), 256), 256, 0, s>>>(d_cells, d_words, W, H, stride);
  return cudaGetLastError();
}

cudaError_t launch_unpack(const uint64_t* d_words, uint8_t* d_cells, int64_t W, int64_t H,
                          int64_t stride, cudaStream_t s) {
  count_launch(1);
  unpack_kernel<<<grid_for(H * words_per_row(W), 256), 256, 0, s>>>(d_words, d_cells, W, H, stride);
  return cudaGetLastError();
}

cudaError_t launch_population(const uint64_t* d_words, int64_t ws, int64_t stride, int64_t H,
                              unsigned long long* d_partials, unsigned long long* d_out,
                              cudaStream_t s) {
  count_launch(2);
  population_partial_kernel<<<kPopBlocks, kPopThreads, 0, s>>>(d_words, ws, stride, H, d_partials);
  population_final_kernel<<<1, kPopThreads, 0, s>>>(d_partials, kPopBlocks, d_out);
  return cudaGetLastError();
}

cudaError_t launch_format_rows(const uint64_t* d_rows, int64_t stride, int64_t W, int64_t nrows,
                               int64_t first_global_row, char* d_text, int64_t pitch,
                               cudaStream_t s) {
  count_launch(1);
  format_rows_kernel<<<grid_for(nrows * words_per_row(W), 256), 256, 0, s>>>(
      d_rows, stride, W, nrows, first_global_row, d_text, pitch);
  return cudaGetLastError();
}

cudaError_t launch_step_rows(const uint64_t* d_in, uint64_t* d_out, const uint64_t* d_top,
                             const uint64_t* d_bottom, int64_t W, int64_t H, int64_t stride,
                             int threads, cudaStream_t s) {
  count_launch(1);
  step_rows_kernel<<<grid_for(H * words_per_row(W), threads), threads, 0, s>>>(d_in, d_out, d_top, d_bottom,
                                                                               W, H, stride);
  return cudaGetLastError();
}

}  // namespace gol
