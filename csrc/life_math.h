// life_math.h — bit-sliced Conway B3/S23 arithmetic shared by host and device.
//
// One 64-bit word holds 64 consecutive cells of one row, EVEN/ODD INTERLEAVED: the low 32 bits
// are the even columns of the block, the high 32 bits the odd columns (column 64w + j lives at
// bit (j >> 1) + 32 * (j & 1)).  With this layout one of the two x-neighbours of every cell sits
// at the same bit index of the other half, so a generation needs one funnel shift per 32 cells
// instead of two (22 instead of 24 integer-pipe instructions per 64 cells).
// The same functions drive the CPU packed stepper, the CUDA kernels and the
// pattern initialisers, so a CPU test of this header is a test of the device
// math too.
//
// Behavioural parity: the rule is the one the reference spells out with
// if/else on a byte neighbour count (gol-with-cuda.cu:239-257) and the
// neighbourhood is its 8-cell Moore stencil (gol-with-cuda.cu:207-235).
// Nothing here is derived from that code; it is the textbook carry-save
// formulation on bit planes.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GOL_HD __host__ __device__ __forceinline__
#else
#define GOL_HD inline
#endif

namespace gol {

// ---- 3-input boolean primitives (each maps to one LOP3 on the GPU) --------
template <class T> GOL_HD T xor3(T a, T b, T c) { return a ^ b ^ c; }
template <class T> GOL_HD T maj3(T a, T b, T c) { return (a & b) | (c & (a | b)); }

// Horizontal 3-cell sum (left + centre + right) of one row as two bit planes.
template <class T> struct HSum { T s0, s1; };

template <class T> GOL_HD HSum<T> hsum3(T l, T c, T r) {
  HSum<T> h;
  h.s0 = xor3(l, c, r);
  h.s1 = maj3(l, c, r);
  return h;
}

// Next state of the middle row from the horizontal sums of the row above (a),
// the row itself (b, which includes the centre cell) and the row below (d).
// total = a + b + d counts the 3x3 block including the centre:
//   next = (total == 3) | (centre & total == 4)
template <class T> GOL_HD T life_rule(HSum<T> a, HSum<T> b, HSum<T> d, T centre) {
  const T t0 = xor3(a.s0, b.s0, d.s0);   // weight 1
  const T k0 = maj3(a.s0, b.s0, d.s0);   // weight 2
  const T t1 = xor3(a.s1, b.s1, d.s1);   // weight 2
  const T k1 = maj3(a.s1, b.s1, d.s1);   // weight 4
  // total = t0 + 2*(k0 + t1) + 4*k1
  const T eq3hi = (k0 ^ t1) & ~k1;             // bit1 = 1, bit2 = bit3 = 0
  const T eq4hi = ~(k0 ^ t1) & (k1 ^ k0);      // bit1 = 0, bit2 = 1, bit3 = 0
  return (t0 & eq3hi) | (~t0 & centre & eq4hi);
}

// ---- word layout -------------------------------------------------------------
// bit index of local column j (0..63) inside its word
GOL_HD int bitpos(int j) { return (j >> 1) + ((j & 1) << 5); }

// 32 bits -> the even bit positions of 64 bits
GOL_HD uint64_t spread32(uint64_t x) {
  x &= 0xffffffffull;
  x = (x | (x << 16)) & 0x0000ffff0000ffffull;
  x = (x | (x << 8)) & 0x00ff00ff00ff00ffull;
  x = (x | (x << 4)) & 0x0f0f0f0f0f0f0f0full;
  x = (x | (x << 2)) & 0x3333333333333333ull;
  x = (x | (x << 1)) & 0x5555555555555555ull;
  return x;
}
// the even bit positions of 64 bits -> 32 bits
GOL_HD uint64_t compact32(uint64_t x) {
  x &= 0x5555555555555555ull;
  x = (x | (x >> 1)) & 0x3333333333333333ull;
  x = (x | (x >> 2)) & 0x0f0f0f0f0f0f0f0full;
  x = (x | (x >> 4)) & 0x00ff00ff00ff00ffull;
  x = (x | (x >> 8)) & 0x0000ffff0000ffffull;
  x = (x | (x >> 16)) & 0x00000000ffffffffull;
  return x;
}
// column-ordered word (bit j = column j) <-> interleaved storage word
GOL_HD uint64_t std_to_il(uint64_t x) { return compact32(x) | (compact32(x >> 1) << 32); }
GOL_HD uint64_t il_to_std(uint64_t x) { return spread32(x) | (spread32(x >> 32) << 1); }

// ---- counter-based hash for the random-density pattern --------------------
GOL_HD uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// 24-bit uniform value for cell (global row gy, column x).
GOL_HD uint32_t cell_rand24(uint64_t seed, int64_t gy, int64_t x) {
  uint64_t h = mix64(seed ^ mix64((uint64_t)gy * 0xD1342543DE82EF95ull + (uint64_t)x));
  return (uint32_t)(h >> 40);
}

// ---- pattern predicate -----------------------------------------------------
// Patterns 0-4 place the same cells as the reference initialisers
// (gol-with-cuda.cu:56-171); 5 (glider) and 6 (random density) are new.
enum Pattern : int {
  kAllZeros = 0,        // gol-with-cuda.cu:56-69
  kAllOnes = 1,         // gol-with-cuda.cu:72-92
  kTenOnLastRow = 2,    // gol-with-cuda.cu:95-120  (last local row, cols 127..136, every rank)
  kCorners = 3,         // gol-with-cuda.cu:123-147 (rank 0 top corners, else-if last rank bottom corners)
  kSpinner = 4,         // gol-with-cuda.cu:150-171 (rank 0: row 0, cols 0, 1, W-1)
  kGlider = 5,          // new: glider straddling the rank-0 / rank-1 boundary
  kRandom = 6,          // new: Bernoulli(density) per cell, counter-based
  kNumPatterns = 7
};

struct PatternArgs {
  int pattern;
  int rank, nranks;
  int64_t W, H;          // per-rank slab: W columns, H rows
  uint64_t seed;
  uint32_t thresh24;     // alive iff cell_rand24 < thresh24
};

GOL_HD bool pattern_cell(const PatternArgs& a, int64_t y, int64_t x) {
  switch (a.pattern) {
    case kAllZeros: return false;
    case kAllOnes: return true;
    case kTenOnLastRow: return y == a.H - 1 && x >= 127 && x < 137;
    case kCorners:
      if (a.rank == 0) return y == 0 && (x == 0 || x == a.W - 1);
      if (a.rank == a.nranks - 1) return y == a.H - 1 && (x == 0 || x == a.W - 1);
      return false;
    case kSpinner: return a.rank == 0 && y == 0 && (x == 0 || x == 1 || x == a.W - 1);
    case kGlider: {
      // origin = global (H-2, 1): the glider starts across the first slab
      // boundary (or across the torus seam when there is one rank).
      const int64_t GH = a.H * a.nranks;
      const int64_t gy = a.H * a.rank + y;
      int64_t dy = gy - (a.H - 2);
      if (dy < 0) dy += GH;
      dy %= GH;
      const int64_t dx = x - 1;
      if (dy > 2 || dx < 0 || dx > 2) return false;
      // .O.
      // ..O
      // OOO
      return (dy == 0 && dx == 1) || (dy == 1 && dx == 2) || (dy == 2);
    }
    case kRandom: return cell_rand24(a.seed, a.H * a.rank + y, x) < a.thresh24;
    default: return false;
  }
}

// 64 cells of row y starting at column 64*w (bits beyond W are zero).
GOL_HD uint64_t pattern_word(const PatternArgs& a, int64_t y, int64_t w) {
  if (a.pattern == kAllZeros) return 0;
  uint64_t v = 0;
  const int64_t x0 = w * 64;
  for (int j = 0; j < 64; ++j) {
    const int64_t x = x0 + j;
    if (x < a.W && pattern_cell(a, y, x)) v |= (1ull << bitpos(j));
  }
  return v;
}

GOL_HD int64_t words_per_row(int64_t W) { return (W + 63) / 64; }
// Row stride in words: 16-byte aligned rows (TMA / 128-bit access friendly).
GOL_HD int64_t row_stride_words(int64_t W) { return (words_per_row(W) + 1) & ~int64_t(1); }
GOL_HD uint64_t tail_mask(int64_t W) {
  const int t = (int)(W & 63);  // valid columns in the last word
  if (!t) return ~0ull;
  const uint64_t even = (1ull << ((t + 1) >> 1)) - 1, odd = (1ull << (t >> 1)) - 1;
  return even | (odd << 32);
}
GOL_HD uint64_t col_bit(const uint64_t* row, int64_t x) {
  return (row[x >> 6] >> bitpos((int)(x & 63))) & 1ull;
}

// ---- x-neighbour words with torus wrap, any W (W % 64 may be non-zero) -----
// left_of: the word whose cell at column c is the cell at column c-1; right_of: column c+1.
// Rows must have zero bits beyond column W-1.  With A = even columns (low half) and B = odd
// columns (high half):  left = {(B << 1) | carry, A},  right = {B, (A >> 1) | carry << 31}.
GOL_HD uint64_t left_of(const uint64_t* row, int64_t w, int64_t W) {
  const uint64_t A = row[w] & 0xffffffffull, B = row[w] >> 32;
  // column 64w - 1: the previous word's last odd column (bit 63), or column W-1 across the seam
  const uint64_t carry = (w > 0) ? (row[w - 1] >> 63) : col_bit(row, W - 1);
  const uint64_t LA = ((B << 1) | carry) & 0xffffffffull;
  return LA | (A << 32);
}
GOL_HD uint64_t right_of(const uint64_t* row, int64_t w, int64_t ws, int64_t W) {
  const uint64_t A = row[w] & 0xffffffffull, B = row[w] >> 32;
  uint64_t RA = B, RB = A >> 1;
  if (w < ws - 1) {
    RB |= (row[w + 1] & 1ull) << 31;  // column 64(w+1): the next word's first even column
  } else {
    // last word: the cell right of column W-1 is column 0
    const int l = (int)((W - 1) & 63);
    const uint64_t c0 = row[0] & 1ull;
    if (l & 1) RB |= c0 << (l >> 1);
    else RA |= c0 << (l >> 1);
  }
  return RA | (RB << 32);
}

// One output word of a single generation given the three source rows.
GOL_HD uint64_t step_word(const uint64_t* up, const uint64_t* cur, const uint64_t* down, int64_t w,
                          int64_t ws, int64_t W) {
  const HSum<uint64_t> a = hsum3(left_of(up, w, W), up[w], right_of(up, w, ws, W));
  const HSum<uint64_t> b = hsum3(left_of(cur, w, W), cur[w], right_of(cur, w, ws, W));
  const HSum<uint64_t> d = hsum3(left_of(down, w, W), down[w], right_of(down, w, ws, W));
  uint64_t v = life_rule(a, b, d, cur[w]);
  if (w == ws - 1) v &= tail_mask(W);
  return v;
}

}  // namespace gol
