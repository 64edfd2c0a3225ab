// internal.h — declarations shared between the translation units of the core.
#pragma once
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include <string>

#include "life_math.h"

namespace gol {

int set_error(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));

int make_pattern_args(PatternArgs* a, int pattern, int64_t W, int64_t H, int rank, int nranks,
                      uint64_t seed, double density);
int write_rank_header(FILE* fp, int rank);
void append_row_prefix(std::string* s, int64_t global_row);

// Packed CPU step of nrows rows; in_rows has nrows + 2 entries (ghost above /
// below at the ends), out_rows has nrows.
void cpu_step_rows(const uint64_t* const* in_rows, uint64_t* const* out_rows, int64_t nrows,
                   int64_t W);

// ---------------------------------------------------------------------------
// Device-resident description of one slab (one rank's share of the torus).
// Lives in device memory; the kernels take a pointer to an array of these, so a
// captured CUDA graph or a persistent kernel never needs host-side parameter
// patching: buffer parity, exchange parity and flag targets all come from the
// counters below, which the kernels themselves advance.
//
// Slab buffer: (H + 2*G) rows x stride words, rows [0,G) and [G+H, G+H+2G) are
// the in-slab margin used when a halo is deeper than one pass (the pass
// computes part of the neighbour's rows redundantly).  Freshly exchanged halo
// rows never land in the slab: they land in recv[side][exchange parity], which
// only the neighbour writes and only this slab reads.
// ---------------------------------------------------------------------------
struct SlabState {
  // geometry
  int64_t W, H;        // cells per row, interior rows
  int64_t ws;          // valid words per row
  int64_t stride;      // row stride in words
  int32_t G;           // margin rows per side (= GOL_MAX_GHOST capped by plan)
  int32_t ll_ring;     // bit 0 / 1: halo rows from / to the prev / next neighbour use the LL form
  int32_t rank, nranks;
  int32_t local_ring;  // bit 0 / 1: the prev / next ring neighbour lives on THIS device, so its flag
                       // and halo rows need gpu scope only.  System-scope fences travel to the host
                       // bridge: while a PCIe copy is in flight each one waits behind it, and the
                       // step kernels ran 25 % slower under a concurrent upload (B200, measured).
  // storage (all device pointers valid on the device that runs this slab)
  uint64_t* buf[2];              // buf[pass & 1] is the current generation
  uint64_t* recv[2][2];          // [side 0=top,1=bottom][exchange parity], G x stride each
  uint64_t* peer_recv[2][2];     // [0]: prev rank's bottom recv, [1]: next rank's top recv
  // Halo rows from a neighbour on ANOTHER device travel in flag-in-data ("LL") form: every 64-bit
  // halo word is sent as two 8-byte words {32 data bits | 32-bit exchange number}, each written
  // by one atomic 8-byte store, and the reader polls the data itself until both halves carry the
  // expected number.  No flag, no release / acquire fence on either side: a system-scope fence
  // waits for the host bridge as well, which costs microseconds and, while a PCIe copy is in
  // flight, far more (see local_ring).  Layout: word w of halo row r at ll[(r * stride + w) * 2].
  uint64_t* ll_recv[2][2];       // [side][exchange parity], 2 * G * stride words each
  uint64_t* peer_ll[2][2];       // [0]: prev rank's bottom ll_recv, [1]: next rank's top ll_recv
  unsigned long long* flag[2];       // words received so far on [top, bottom] (monotonic)
  unsigned long long* peer_flag[2];  // [0]: prev's bottom flag, [1]: next's top flag
  // counters advanced by the kernels
  unsigned long long pass;       // passes completed -> buffer parity
  unsigned long long exch;       // exchanges consumed -> recv parity
  unsigned long long words_seen; // flag value after the last consumed exchange
  unsigned int done;             // CTAs finished in the running launch
  int error;                     // != 0: device-side failure (1 / 2: halo from prev / next timed out,
                                 // 3: a strip dependency inside the flow kernel timed out)
  unsigned long long watchdog_ns;
  // per-phase counters (tracing; accumulated by the kernels, read by gol_world_phase_counters)
  unsigned long long wait_ns[2];   // time edge warps spent spinning on the [top, bottom] halo flag
  unsigned long long wait_count[2];// ... and how many waits actually had to spin
  unsigned long long dep_wait_ns;  // flow kernel: time warps spent waiting for a neighbouring strip
  unsigned long long task_ns[2];   // flow kernel: [edge, interior] task time, summed over warps
  unsigned long long tasks[2];     // flow kernel: [edge, interior] tasks executed
};

// Baked per launch (identical on every rank: the schedule is SPMD).
struct PassParams {
  int32_t ext_in;     // halo rows valid around the interior in the input
  int32_t k;          // generations this pass (== template K for the fused kernel)
  int32_t consume;    // 1: input halo comes from recv[] and must be waited for
  int32_t push;       // rows to push to each neighbour after this pass (0 = none)
  int32_t strip_rows; // output rows per warp task
  int32_t col_groups; // warp tasks per strip
  int32_t nstrips;
  int32_t edge_rows;  // > 0: the first / last edge_rows output rows are strips of their own (scheduled
                      // first and short, so the halo push leaves early and overlaps the interior)
  int32_t tma_rank;      // TMA mode: 2 = {ws, rows} map, 3 = {16, ws/16, rows} map (128-byte inner box)
};

// ---- strip decomposition of one pass (shared by host planner and kernels) ----------------
// Output rows [first, end) are cut into warp tasks.  With edge_rows > 0 the two edge strips
// come first in task order (order 0 = top, 1 = bottom), then the interior strips.
GOL_HD int strip_count(int64_t rows_total, int strip_rows, int edge_rows) {
  if (edge_rows > 0 && rows_total > 2 * (int64_t)edge_rows)
    return 2 + (int)((rows_total - 2 * edge_rows + strip_rows - 1) / strip_rows);
  return (int)((rows_total + strip_rows - 1) / strip_rows);
}

GOL_HD void strip_range(int order, int nstrips, int64_t first, int64_t end, int strip_rows, int edge_rows,
                        int64_t* y0, int64_t* y1) {
  if (edge_rows > 0 && end - first > 2 * (int64_t)edge_rows) {
    if (order == 0) {
      *y0 = first;
      *y1 = first + edge_rows;
    } else if (order == 1) {
      *y0 = end - edge_rows;
      *y1 = end;
    } else {
      *y0 = first + edge_rows + (int64_t)(order - 2) * strip_rows;
      *y1 = *y0 + strip_rows;
      if (*y1 > end - edge_rows) *y1 = end - edge_rows;
    }
  } else {
    // uniform strips, boundary ones first: order 0 -> strip 0, order 1 -> last strip
    const int strip = order == 0 ? 0 : (order == 1 ? nstrips - 1 : order - 1);
    *y0 = first + (int64_t)strip * strip_rows;
    *y1 = *y0 + strip_rows;
    if (*y1 > end) *y1 = end;
  }
}

// ---- strip decomposition of the flow (dependency-driven persistent) kernel ------------------
// Strips are anchored to the slab's interior rows so that they are the same in every pass, whatever
// the halo depth of the pass: strip 0 = top edge [-ext_out, E), strip n-1 = bottom edge
// [H-E, H+ext_out), the n-2 interior strips split [E, H-E) evenly.  Task (pass p, strip j) reads
// rows that strips j-1, j, j+1 wrote in pass p-1 and overwrites rows those strips read in pass p-1,
// so it waits for exactly those three (when every strip has at least K rows).
GOL_HD int flow_strip_count(int64_t H, int E, int strip_rows) {
  if (H <= 2 * (int64_t)E) return 1;
  const int64_t RI = H - 2 * (int64_t)E;
  const int64_t n = (RI + strip_rows - 1) / strip_rows;
  return 2 + (int)(n < 1 ? 1 : n);
}
GOL_HD void flow_strip_range(int j, int nstrips, int64_t H, int E, int ext_out, int64_t* y0, int64_t* y1) {
  if (nstrips == 1) {
    *y0 = -(int64_t)ext_out;
    *y1 = H + ext_out;
  } else if (j == 0) {
    *y0 = -(int64_t)ext_out;
    *y1 = E;
  } else if (j == nstrips - 1) {
    *y0 = H - E;
    *y1 = H + ext_out;
  } else {
    const int64_t RI = H - 2 * (int64_t)E;
    const int nint = nstrips - 2;
    *y0 = E + (RI * (j - 1)) / nint;
    *y1 = E + (RI * j) / nint;
  }
}
// Ticket order inside a pass: both edge strips first (their halo push leaves early), then the
// interior from the outside in, so that what a strip waits for was issued early in the pass before.
GOL_HD int flow_order_to_strip(int q, int nstrips) {
  if (q == 0) return 0;
  if (q == 1) return nstrips - 1;
  const int i = q - 2;
  return (i & 1) ? nstrips - 2 - (i >> 1) : 1 + (i >> 1);
}

// Baked per launch of the flow kernel: `npasses` passes of K generations, halo exchange of `depth`
// rows every passes_per_exchange passes (SPMD: identical on every rank).
struct FlowParams {
  int32_t npasses, passes_per_exchange, depth, nslabs;
  int32_t col_groups, nstrips, edge_rows;
  int32_t dep_radius;     // strips on each side a task waits for (1; nstrips when strips are shorter than K)
  int32_t max_strips;     // pitch of `progress` per slab
  int32_t total_warps;    // warps of the launch (the last one to finish resets the bookkeeping)
  int32_t flags;          // experiments (GOL_FLOW_FLAGS): 1 = skip dependency waits (wrong results, timing
                          // only), 4 = fixed owners (warp w always runs task slot w) instead of tickets,
                          // 8 = no acquire fence after a dependency poll
  unsigned int* ctl;      // [0] next ticket, [1] warps finished, [2] abort
  unsigned int* progress; // [nslabs][max_strips] column-group tasks completed per strip in this launch
};

// Baked per launch of the resident kernel (small slabs that fit in shared memory, one CTA per strip,
// thread-block clusters): the whole step() of `gens` generations in one launch.
struct ResidentParams {
  int32_t K;               // generations per halo exchange (halo depth)
  int32_t npasses;         // ceil(gens / K)
  int32_t last_k;          // generations of the last pass (1..K)
  int32_t nslabs, ctas_per_slab, cluster;
  int32_t rows_alloc;      // shared-memory rows per buffer: max strip rows + 2K
  int32_t use_bulk;        // strips are loaded / stored with cp.async.bulk (rows are contiguous: stride == ws)
  int32_t smem_words;      // 64-bit words of dynamic shared memory (bounds-checking build: every offset is checked)
  uint64_t* mail;          // CTA-to-CTA halo mailboxes in global memory (flag-in-data words)
  int64_t mail_slot;       // words per (cta, side, parity) slot = 2 * K * ws
  unsigned int* ctl;       // [1]: CTAs finished (the last one advances the slab counters)
};

}  // namespace gol
