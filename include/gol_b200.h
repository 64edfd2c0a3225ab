/* gol_b200.h — C ABI of the B200-native distributed Game-of-Life core.
 *
 * Everything a driver needs (the `gol` CLI, the Python/PyTorch layer, a third
 * party) goes through this header.  All functions return 0 on success and a
 * negative code on failure unless stated otherwise; gol_last_error() gives the
 * message of the last failure on the calling thread.
 *
 * Geometry (same as the reference, gol-main.c:22,76): rank r of R owns global
 * rows [r*H, (r+1)*H) of a W-wide, H*R-tall torus.  Cells are stored bit-packed,
 * 64 consecutive columns per uint64 word, even/odd interleaved: column 64*w + j of a row is
 * bit (j >> 1) + 32 * (j & 1) of word w (even columns in the low half, odd columns in the
 * high half; it saves one shift per 32 cells per generation).  gol_pack_bytes /
 * gol_unpack_words / gol_cuda_pack / gol_cuda_unpack convert from and to byte cells.
 */
#ifndef GOL_B200_H_
#define GOL_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GOL_OK 0
#define GOL_ERR_ARG (-1)
#define GOL_ERR_CUDA (-2)
#define GOL_ERR_STATE (-3)
#define GOL_ERR_IO (-4)
#define GOL_ERR_TIMEOUT (-5)
#define GOL_ERR_UNSUPPORTED (-6)

#define GOL_BACKEND_CPU 0
#define GOL_BACKEND_CUDA 1

/* step-kernel selection */
#define GOL_MODE_AUTO 0        /* fused where the shape allows, else generic   */
#define GOL_MODE_GENERIC 1     /* one thread per word, any W/H, 1 gen per pass  */
#define GOL_MODE_FUSED 2       /* register-streaming, K generations per pass    */
#define GOL_MODE_PERSISTENT 3  /* one launch for the whole step(): persistent warps,
                                  per-strip dependencies instead of grid barriers */
#define GOL_MODE_TMA 4         /* fused, input rows fed by TMA bulk copies      */
#define GOL_MODE_RESIDENT 5    /* small slabs: strips stay in shared memory for the whole
                                  step(), one CTA per strip, thread-block clusters,
                                  halo rows between CTAs over DSMEM / L2 mailboxes    */

#define GOL_MAX_GHOST 32       /* deepest exchange (rows per side)              */

const char* gol_last_error(void);
const char* gol_version(void);
/* Memory-safety build.  libgol_b200_debug.so is the same library compiled with
 * -DGOL_DEBUG_BOUNDS=1: every global load / store / atomic of every kernel first checks its address
 * range against a registry of the buffers the library allocated or was handed (each slab buffer,
 * recv buffer, flag ...; the guard rows between them belong to no region).  Misses are counted
 * and skipped; gol_world_sync fails when there are any, and also verifies the guard rows.
 * gol_debug_bounds_enabled(): 1 in that build, 0 in the normal one (where the report is always
 * zero).  The report is for the current CUDA device; _poke performs one checked 8-byte load of the
 * given device address (to test the detector).  The reference has no counterpart (SURVEY.md
 * section 5: no sanitizers, no error checks beyond device binding). */
int gol_debug_bounds_enabled(void);
int gol_debug_bounds_report(uint64_t* violations, uint64_t* first_bad_address, int* source_line, int reset);
int gol_debug_bounds_poke(const void* device_address);
/* CUDA devices visible to this process (0 when there is no driver / device). */
int gol_cuda_device_count(void);

/* ---- CPU oracle and helpers (no GPU needed) ------------------------------ */

/* True torus Life on a W x H byte grid (0/1 per cell), `gens` generations.
 * Restates the reference kernel's neighbourhood and rule
 * (gol-with-cuda.cu:207-257) for the whole global grid; this is the oracle. */
int gol_oracle_run(const uint8_t* in, uint8_t* out, int64_t W, int64_t H, int64_t gens);

/* Fill one rank's W x H byte slab with pattern 0..6 (gol-with-cuda.cu:56-171
 * for 0..4; 5 = glider, 6 = random density).  density in [0,1]. */
int gol_pattern_fill_bytes(uint8_t* cells, int pattern, int64_t W, int64_t H,
                           int rank, int nranks, uint64_t seed, double density);

/* bytes <-> packed words (row stride = gol_row_stride_words(W)). */
int64_t gol_words_per_row(int64_t W);
int64_t gol_row_stride_words(int64_t W);
int gol_pack_bytes(const uint8_t* cells, uint64_t* words, int64_t W, int64_t H);
int gol_unpack_words(const uint64_t* words, uint8_t* cells, int64_t W, int64_t H);

/* Bit-packed CPU torus stepper on the whole grid (timed CPU baseline). */
int gol_cpu_run_packed(uint64_t* words, int64_t W, int64_t H, int64_t gens);

/* Text of one rank's final world exactly as the reference writes it
 * (gol-main.c:17-28,135-139): header line, then "Row %2d: " + "%u " per cell.
 * gol_write_rank_file creates/truncates `path`. */
int gol_write_rank_file(const char* path, const uint8_t* cells, int64_t W, int64_t H,
                        int rank);
/* "Rank_<r>_of_<R>.txt" (gol-main.c:66). buf must hold >= 64 bytes. */
int gol_rank_file_name(char* buf, size_t n, int rank, int nranks);
/* The two stdout lines of rank 0 (gol-main.c:124-125,132) into buf. */
int gol_format_report(char* buf, size_t n, double seconds, int nranks, int64_t W, int64_t H,
                      int64_t iterations);

/* ---- world: slabs + ring + stepping --------------------------------------- */

typedef struct gol_world gol_world_t;

typedef struct gol_config {
  int64_t width;          /* W: columns                                         */
  int64_t rows;           /* H: rows per rank                                   */
  int32_t nranks;         /* R: ranks in the ring                               */
  int32_t first_rank;     /* this process owns ranks [first_rank, +nlocal)      */
  int32_t nlocal;
  int32_t backend;        /* GOL_BACKEND_*                                      */
  int32_t mode;           /* GOL_MODE_*                                         */
  int32_t temporal_k;     /* generations per pass (1,2,4,8); 0 = default        */
  int32_t exchange_depth; /* rows per halo exchange (multiple of k); 0 = k      */
  int32_t threads_per_block; /* CLI arg 4 (gol-main.c:52); 0 = default          */
  int32_t ndevices;       /* devices to spread local ranks over; 0 = all        */
  int32_t device_base;    /* local rank i -> device (device_base+i) % ndevices  */
  int32_t flags;          /* GOL_FLAG_*                                         */
  int32_t reserved;       /* > 0: force the strip height (rows per warp task) of the fused kernels */
} gol_config_t;

#define GOL_FLAG_NO_GRAPH 1      /* launch passes one by one, no CUDA graph     */
#define GOL_FLAG_NO_PEER_ACCESS 2 /* do not call cudaDeviceEnablePeerAccess     */
#define GOL_FLAG_CPU_SCHEDULE 4   /* CPU backend: execute the GPU pass schedule (k, exchange
                                     depth, margins, double-buffered recv rows) with the CPU
                                     stepper, to check that logic without a GPU             */

gol_world_t* gol_world_create(const gol_config_t* cfg);
void gol_world_destroy(gol_world_t* w);

/* What the planner chose (mode after AUTO resolution, k, depth, block size). */
typedef struct gol_plan {
  int32_t mode, temporal_k, exchange_depth, threads_per_block;
  int32_t strip_rows, col_groups, ctas_per_slab, regs_hint;
  int64_t words_per_row, row_stride, slab_rows;
} gol_plan_t;
int gol_world_plan(const gol_world_t* w, gol_plan_t* out);

/* The pass schedule step(gens) would run for a configuration, without touching a device
 * (planner + schedule builder only).  One entry per kernel launch, in order: a halo prime
 * (prime = 1, prime_depth rows pushed to each neighbour) or a pass of k generations whose
 * input carries ext_in valid halo rows (consume = 1: they come fresh from the neighbours and
 * are waited for) and which pushes `push` rows for the next exchange period.  Returns the
 * number of entries (may exceed max_out) or a negative error.  Identical on every rank. */
typedef struct gol_pass_desc {
  int32_t prime, generic, prime_depth, k, ext_in, consume, push, reserved;
} gol_pass_desc_t;
int gol_schedule_describe(const gol_config_t* cfg, int64_t gens, gol_pass_desc_t* out, int max_out,
                          gol_plan_t* plan_out);

/* Multi-process wiring.  Each process exports one blob per local rank, all
 * ranks gather them (torch.distributed / any transport) and import the ones
 * they do not own; gol_world_connect() then resolves every local slab's ring
 * neighbours (local pointer, P2P pointer or CUDA-IPC mapping).  With all ranks
 * local, call gol_world_connect() directly. */
#define GOL_EXPORT_BYTES 256
int gol_world_export(gol_world_t* w, int local_index, void* blob /* GOL_EXPORT_BYTES */);
int gol_world_import(gol_world_t* w, int global_rank, const void* blob);
int gol_world_connect(gol_world_t* w);

/* Streams: by default one non-blocking stream per device created by the world.
 * gol_world_set_stream lets PyTorch hand in its current stream. */
int gol_world_set_stream(gol_world_t* w, int local_index, void* cuda_stream);

/* State.  All of these enqueue on the slab's stream (CUDA) or run inline (CPU). */
int gol_world_init_pattern(gol_world_t* w, int pattern, uint64_t seed, double density);
int gol_world_load_bytes(gol_world_t* w, int local_index, const uint8_t* cells);      /* host, W*H */
int gol_world_load_packed(gol_world_t* w, int local_index, const uint64_t* words);    /* host, H*stride */
int gol_world_read_bytes(gol_world_t* w, int local_index, uint8_t* cells);            /* host out */
int gol_world_read_packed(gol_world_t* w, int local_index, uint64_t* words);          /* host out */
int gol_world_population(gol_world_t* w, int local_index, uint64_t* count);           /* syncs */
/* async pieces for end-to-end pipelines: result lands in *pinned_out after the stream drains */
int gol_world_load_packed_async(gol_world_t* w, int local_index, const uint64_t* pinned_words);
int gol_world_population_async(gol_world_t* w, int local_index, uint64_t* pinned_out);

int gol_world_read_packed_async(gol_world_t* w, int local_index, uint64_t* pinned_words);

/* End-to-end step pipeline.  One call uploads this step's packed input of every local slab from
 * pinned host memory (pinned_in[i], H * stride words), advances `gens` generations and brings
 * the packed result back into pinned_out[i].  Upload and download use their own copy streams and
 * two slots of device staging, so consecutive calls overlap: the input of step i+1 travels while
 * step i computes, and the result of step i travels while step i+1 computes.  Returns the slot
 * (0 or 1) to hand to gol_world_wait_result(), which blocks until that step's result is in host
 * memory; at most two steps may be outstanding.  Negative = error.  (The reference reaches its
 * data through managed memory and has no explicit copies, gol-with-cuda.cu:35-67.) */
int gol_world_run_from_host_async(gol_world_t* w, const uint64_t* const* pinned_in, int64_t gens,
                                  uint64_t* const* pinned_out);
int gol_world_wait_result(gol_world_t* w, int slot);

/* Advance every local slab by `gens` generations (halo exchange fused in).
 * Asynchronous on CUDA; gol_world_sync waits and reports device-side errors
 * (flag-wait watchdog, bad launch). */
int gol_world_step(gol_world_t* w, int64_t gens);
int gol_world_sync(gol_world_t* w);
/* Same, bracketed by CUDA events: ms[i] = device time of local slab i's stream. */
int gol_world_step_timed(gol_world_t* w, int64_t gens, float* ms /* nlocal */);

/* Per-phase counters accumulated on the device by the step kernels (tracing; the reference has
 * one host wall-clock interval, gol-main.c:82,122).  halo_wait_ns[side]: time edge warps spent
 * spinning on the [top, bottom] neighbour's halo flag, summed over the warps that had to wait
 * (halo_waits[side] of them): the exposed halo time.  dep_wait_ns: time warps of the persistent
 * kernel waited for a neighbouring strip of the previous pass.  task_ns / tasks: warp time and
 * count of [edge, interior] strip tasks (persistent kernel).  Synchronises the slab's stream;
 * reset != 0 zeroes the counters afterwards. */
typedef struct gol_phase_counters {
  uint64_t halo_wait_ns[2], halo_waits[2];
  uint64_t dep_wait_ns;
  uint64_t task_ns[2], tasks[2];
} gol_phase_counters_t;
int gol_world_phase_counters(gol_world_t* w, int local_index, gol_phase_counters_t* out, int reset);

/* Device pointer of the current packed slab interior (row 0, word 0) for
 * zero-copy views; stride in words via gol_world_plan(). */
void* gol_world_device_ptr(gol_world_t* w, int local_index);
int gol_world_device_of(gol_world_t* w, int local_index);

/* CPU-backend hooks for an external transport (gloo/MPI) when not all ranks
 * are local: read my edge rows, write my ghost rows, step once. */
int gol_world_cpu_get_edge(gol_world_t* w, int local_index, int bottom, uint64_t* row /* stride */);
int gol_world_cpu_set_ghost(gol_world_t* w, int local_index, int bottom, const uint64_t* row);
int gol_world_cpu_step_local(gol_world_t* w);

/* Write Rank_<r>_of_<R>.txt for a local slab into directory `dir` (NULL = cwd);
 * on CUDA the row text is formatted on the device (gol-main.c:17-28 layout). */
int gol_world_dump_text(gol_world_t* w, int local_index, const char* dir);

/* Kernels launched by this library in this process (graph replays counted per node). */
uint64_t gol_launch_count(void);

/* ---- stateless device ops on caller-owned buffers (stream = cudaStream_t) -- */
int gol_cuda_pack(const uint8_t* d_cells, uint64_t* d_words, int64_t W, int64_t H, void* stream);
int gol_cuda_unpack(const uint64_t* d_words, uint8_t* d_cells, int64_t W, int64_t H, void* stream);
/* Population of H rows of packed words.  d_scratch: >= gol_cuda_population_scratch_bytes()
 * bytes of device memory (per-block partial sums); d_out: one uint64. */
size_t gol_cuda_population_scratch_bytes(void);
int gol_cuda_population(const uint64_t* d_words, int64_t W, int64_t H, void* d_scratch,
                        uint64_t* d_out, void* stream);
/* One torus generation of a stand-alone packed W x H grid (no ring): out != in. */
int gol_cuda_step_torus(const uint64_t* d_in, uint64_t* d_out, int64_t W, int64_t H,
                        int threads_per_block, void* stream);
/* One generation of an H-row slab with explicit ghost rows above / below (each one
 * packed row).  This is the unfused stencil the NCCL-halo baseline uses after its
 * ncclSend/ncclRecv exchange; the product path fuses both (gol_world_step). */
int gol_cuda_step_rows(const uint64_t* d_in, uint64_t* d_out, const uint64_t* d_top,
                       const uint64_t* d_bottom, int64_t W, int64_t H, int threads_per_block,
                       void* stream);

/* ---- stepping without any allocation ---------------------------------------------
 * (No counterpart in the reference: it allocates everything with cudaMallocManaged.)
 *
 * For callers that are not allowed to allocate device memory at the point where they need
 * a step: a plug-in inside an application whose memory pool already owns the whole GPU, a
 * latency-critical loop that must not take the driver's allocation lock, a component that
 * runs after the host program has frozen its memory map.  Everything else in this header
 * gets working memory from cudaMalloc (gol_world_create) or from the caller; this call
 * uses a fixed scratch array of GOL_NOALLOC_SCRATCH_BYTES that is part of the library's
 * device image.  Every GPU that loads the library carries that array until the process
 * exits, whether or not the call is ever made.
 *
 * Advances a stand-alone packed W x H torus (row stride gol_row_stride_words(W), no ring
 * neighbours) by `gens` generations.  d_a holds the input and is overwritten; d_b is a
 * second buffer of the same size; *result_in_b tells which of the two holds the result
 * (0: d_a, 1: d_b).  Work is enqueued on `stream` and is asynchronous like the other
 * gol_cuda_* calls.  No cudaMalloc / cudaFree / allocating routine is called, now or later.
 *
 * Limits and rules:
 *  - W <= GOL_NOALLOC_MAX_WIDTH, otherwise GOL_ERR_UNSUPPORTED (the halo rows of a wider
 *    grid do not fit the scratch array); any H >= 1.
 *  - The first call on each device looks up the scratch array's device address
 *    (cudaGetSymbolAddress).  Make that first call outside any stream capture.
 *  - All calls on one device share the one scratch array, and it keeps whatever the last
 *    call left in it; each call re-initialises the counters and flags it needs.  Two calls
 *    whose GPU work can overlap (different streams, no dependency) corrupt each other:
 *    serialise them.
 *  - A device-side wait that runs into the watchdog (GOL_WATCHDOG_MS, default 10 s; in practice
 *    only when two overlapping calls trample each other's flags) is NOT reported by this call:
 *    it returns as soon as the work is enqueued, and the scratch header that records the
 *    failure is rewritten by the next call.  Callers that overlap calls get wrong cells, not an
 *    error code.
 */
#define GOL_NOALLOC_MAX_WIDTH 131072
#define GOL_NOALLOC_SCRATCH_BYTES (272 * 1024)
int gol_cuda_run_torus_noalloc(uint64_t* d_a, uint64_t* d_b, int64_t W, int64_t H, int64_t gens,
                               int threads_per_block, void* stream, int* result_in_b);

#ifdef __cplusplus
}
#endif
#endif /* GOL_B200_H_ */
