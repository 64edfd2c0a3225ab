// cpu_ref.cpp — CPU side of the core: the oracle, the packed CPU stepper, the
// byte<->bit converters, the pattern filler and the text writer.
//
// The oracle is deliberately the dumbest possible implementation (one byte per
// cell, explicit modulo wrap, integer neighbour count) so that it shares no
// code with the bit-sliced paths it is used to check.
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../include/gol_b200.h"
#include "internal.h"
#include "life_math.h"

namespace gol {

static thread_local std::string g_err;

int set_error(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

// ---- packed CPU stepping of a block of rows --------------------------------
// rows[-1] and rows[n] must be readable (ghost rows); all rows have `ws` valid
// words and zero bits beyond W.  Torus wrap in x is done here.
static void cpu_step_rows_serial(const uint64_t* const* in_rows, uint64_t* const* out_rows, int64_t nrows,
                                 int64_t W) {
  // in_rows has nrows + 2 entries: [0] = row above the first, [nrows+1] = below the last.
  const int64_t ws = words_per_row(W);
  const uint64_t tm = tail_mask(W);
  std::vector<HSum<uint64_t>> h[3];
  for (auto& v : h) v.resize((size_t)ws);
  auto fill = [&](int slot, const uint64_t* row) {
    for (int64_t w = 0; w < ws; ++w)
      h[slot][(size_t)w] = hsum3(left_of(row, w, W), row[w], right_of(row, w, ws, W));
  };
  fill(0, in_rows[0]);
  fill(1, in_rows[1]);
  for (int64_t y = 0; y < nrows; ++y) {
    const int a = (int)(y % 3), b = (int)((y + 1) % 3), d = (int)((y + 2) % 3);
    fill(d, in_rows[y + 2]);
    const uint64_t* c = in_rows[y + 1];
    uint64_t* o = out_rows[y];
    for (int64_t w = 0; w < ws; ++w) {
      uint64_t v = life_rule(h[a][(size_t)w], h[b][(size_t)w], h[d][(size_t)w], c[w]);
      if (w == ws - 1) v &= tm;
      o[w] = v;
    }
  }
}

// Row blocks are independent within one generation, so large grids are split over host
// threads (GOL_CPU_THREADS, default: hardware concurrency, capped at 16); each block reads
// one extra row above and below, exactly like a rank's ghost rows.
void cpu_step_rows(const uint64_t* const* in_rows, uint64_t* const* out_rows, int64_t nrows, int64_t W) {
  static const int max_threads = [] {
    const char* e = getenv("GOL_CPU_THREADS");
    int n = e ? atoi(e) : (int)std::thread::hardware_concurrency();
    return n < 1 ? 1 : (n > 16 ? 16 : n);
  }();
  const int64_t work = nrows * words_per_row(W);
  int nt = (int)std::min<int64_t>(max_threads, nrows / 16);
  if (work < (1 << 16) || nt < 2) {
    cpu_step_rows_serial(in_rows, out_rows, nrows, W);
    return;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < nt; ++t) {
    const int64_t y0 = nrows * t / nt, y1 = nrows * (t + 1) / nt;
    pool.emplace_back([=] { cpu_step_rows_serial(in_rows + y0, out_rows + y0, y1 - y0, W); });
  }
  for (auto& th : pool) th.join();
}

}  // namespace gol

using namespace gol;

extern "C" {

const char* gol_last_error(void) { return g_err.c_str(); }
const char* gol_version(void) { return "gol_b200 0.1 (sm_100a)"; }

int64_t gol_words_per_row(int64_t W) { return words_per_row(W); }
int64_t gol_row_stride_words(int64_t W) { return row_stride_words(W); }

int gol_oracle_run(const uint8_t* in, uint8_t* out, int64_t W, int64_t H, int64_t gens) {
  if (!in || !out || W < 1 || H < 1 || gens < 0)
    return set_error(GOL_ERR_ARG, "gol_oracle_run: bad argument (W=%lld H=%lld gens=%lld)",
                     (long long)W, (long long)H, (long long)gens);
  const size_t n = (size_t)W * (size_t)H;
  std::vector<uint8_t> a(in, in + n), b(n);
  for (int64_t g = 0; g < gens; ++g) {
    for (int64_t y = 0; y < H; ++y) {
      const int64_t ya = (y + H - 1) % H, yb = (y + 1) % H;
      for (int64_t x = 0; x < W; ++x) {
        const int64_t xa = (x + W - 1) % W, xb = (x + 1) % W;
        const int cnt = a[(size_t)(ya * W + xa)] + a[(size_t)(ya * W + x)] + a[(size_t)(ya * W + xb)] +
                        a[(size_t)(y * W + xa)] + a[(size_t)(y * W + xb)] +
                        a[(size_t)(yb * W + xa)] + a[(size_t)(yb * W + x)] + a[(size_t)(yb * W + xb)];
        const uint8_t alive = a[(size_t)(y * W + x)];
        b[(size_t)(y * W + x)] = alive ? (cnt == 2 || cnt == 3) : (cnt == 3);
      }
    }
    a.swap(b);
  }
  memcpy(out, a.data(), n);
  return GOL_OK;
}

int gol_pattern_fill_bytes(uint8_t* cells, int pattern, int64_t W, int64_t H, int rank,
                           int nranks, uint64_t seed, double density) {
  if (!cells || W < 1 || H < 1 || nranks < 1 || rank < 0 || rank >= nranks)
    return set_error(GOL_ERR_ARG, "gol_pattern_fill_bytes: bad argument");
  PatternArgs a;
  int rc = make_pattern_args(&a, pattern, W, H, rank, nranks, seed, density);
  if (rc) return rc;
  for (int64_t y = 0; y < H; ++y)
    for (int64_t x = 0; x < W; ++x) cells[(size_t)(y * W + x)] = pattern_cell(a, y, x) ? 1 : 0;
  return GOL_OK;
}

int gol_pack_bytes(const uint8_t* cells, uint64_t* words, int64_t W, int64_t H) {
  if (!cells || !words || W < 1 || H < 1) return set_error(GOL_ERR_ARG, "gol_pack_bytes: bad argument");
  const int64_t st = row_stride_words(W);
  for (int64_t y = 0; y < H; ++y) {
    uint64_t* row = words + y * st;
    for (int64_t w = 0; w < st; ++w) row[w] = 0;
    for (int64_t x = 0; x < W; ++x)
      if (cells[(size_t)(y * W + x)]) row[x >> 6] |= 1ull << bitpos((int)(x & 63));
  }
  return GOL_OK;
}

int gol_unpack_words(const uint64_t* words, uint8_t* cells, int64_t W, int64_t H) {
  if (!cells || !words || W < 1 || H < 1) return set_error(GOL_ERR_ARG, "gol_unpack_words: bad argument");
  const int64_t st = row_stride_words(W);
  for (int64_t y = 0; y < H; ++y)
    for (int64_t x = 0; x < W; ++x)
      cells[(size_t)(y * W + x)] = (uint8_t)col_bit(words + y * st, x);
  return GOL_OK;
}

int gol_cpu_run_packed(uint64_t* words, int64_t W, int64_t H, int64_t gens) {
  if (!words || W < 1 || H < 1 || gens < 0) return set_error(GOL_ERR_ARG, "gol_cpu_run_packed: bad argument");
  const int64_t st = row_stride_words(W);
  std::vector<uint64_t> other((size_t)(st * H), 0);
  uint64_t* cur = words;
  uint64_t* nxt = other.data();
  std::vector<const uint64_t*> in_rows((size_t)H + 2);
  std::vector<uint64_t*> out_rows((size_t)H);
  for (int64_t g = 0; g < gens; ++g) {
    in_rows[0] = cur + (H - 1) * st;
    for (int64_t y = 0; y < H; ++y) in_rows[(size_t)y + 1] = cur + y * st;
    in_rows[(size_t)H + 1] = cur;
    for (int64_t y = 0; y < H; ++y) out_rows[(size_t)y] = nxt + y * st;
    cpu_step_rows(in_rows.data(), out_rows.data(), H, W);
    uint64_t* t = cur; cur = nxt; nxt = t;
  }
  if (cur != words) memcpy(words, cur, (size_t)(st * H) * sizeof(uint64_t));
  return GOL_OK;
}

int gol_rank_file_name(char* buf, size_t n, int rank, int nranks) {
  if (!buf || n < 64) return set_error(GOL_ERR_ARG, "gol_rank_file_name: buffer too small");
  snprintf(buf, n, "Rank_%d_of_%d.txt", rank, nranks);
  return GOL_OK;
}

int gol_format_report(char* buf, size_t n, double seconds, int nranks, int64_t W, int64_t H,
                      int64_t iterations) {
  if (!buf || n < 256) return set_error(GOL_ERR_ARG, "gol_format_report: buffer too small");
  const long updates = (long)nranks * (long)H * (long)W * (long)iterations;
  snprintf(buf, n,
           "TOTAL DURATION : %.5lf, number of cell updates = %ld\n"
           "This is the Game of Life running in parallel on a GPU on multiple ranks.\n",
           seconds, updates);
  return GOL_OK;
}

int gol_write_rank_file(const char* path, const uint8_t* cells, int64_t W, int64_t H, int rank) {
  if (!path || !cells || W < 1 || H < 1) return set_error(GOL_ERR_ARG, "gol_write_rank_file: bad argument");
  FILE* fp = fopen(path, "w");
  if (!fp) return set_error(GOL_ERR_IO, "cannot open %s: %s", path, strerror(errno));
  int rc = write_rank_header(fp, rank);
  std::string line;
  line.reserve((size_t)(2 * W + 32));
  for (int64_t y = 0; y < H && rc == GOL_OK; ++y) {
    line.clear();
    append_row_prefix(&line, H * rank + y);
    for (int64_t x = 0; x < W; ++x) {
      line.push_back(cells[(size_t)(y * W + x)] ? '1' : '0');
      line.push_back(' ');
    }
    line.push_back('\n');
    if (fwrite(line.data(), 1, line.size(), fp) != line.size())
      rc = set_error(GOL_ERR_IO, "short write to %s", path);
  }
  if (fclose(fp) != 0 && rc == GOL_OK) rc = set_error(GOL_ERR_IO, "close of %s failed", path);
  return rc;
}

}  // extern "C"

namespace gol {

int make_pattern_args(PatternArgs* a, int pattern, int64_t W, int64_t H, int rank, int nranks,
                      uint64_t seed, double density) {
  if (pattern < 0 || pattern >= kNumPatterns)
    return set_error(GOL_ERR_ARG, "Pattern %u has not been implemented ", (unsigned)pattern);
  if (pattern == kTenOnLastRow && W < 137)
    return set_error(GOL_ERR_ARG,
                     "pattern 2 places cells in columns 127..136 and needs a world at least 137 wide (got %lld)",
                     (long long)W);
  if (pattern == kGlider && (W < 4 || H * nranks < 4))
    return set_error(GOL_ERR_ARG, "pattern 5 (glider) needs at least a 4x4 world");
  if (density < 0.0) density = 0.0;
  if (density > 1.0) density = 1.0;
  a->pattern = pattern;
  a->rank = rank;
  a->nranks = nranks;
  a->W = W;
  a->H = H;
  a->seed = seed;
  a->thresh24 = (uint32_t)(density * 16777216.0 + 0.5);
  return GOL_OK;
}

int write_rank_header(FILE* fp, int rank) {
  // 25 '#', the text, 31 '#'  (gol-main.c:136)
  if (fprintf(fp,
              "######################### FINAL WORLD IN RANK %d IS ###############################\n",
              rank) < 0)
    return set_error(GOL_ERR_IO, "header write failed");
  return GOL_OK;
}

void append_row_prefix(std::string* s, int64_t global_row) {
  char buf[40];
  snprintf(buf, sizeof buf, "Row %2d: ", (int)global_row);  // gol-main.c:22
  s->append(buf);
}

}  // namespace gol
