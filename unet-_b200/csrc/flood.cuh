// flood.cuh -- flood of a seed set through a candidate set on the bit planes of ONE frame held in shared memory.
// Used for the Canny hysteresis (8-connectivity: strong bits through candidate bits, cols_kernels.cuh) and for the exterior
// background of a mask (4-connectivity: the image frame through the background, annotate.cu).
//
// Layout: C (candidates) and S (flooded so far, S within C) as (H + 2) x (Wd + 2) words, one word / one row of padding all
// round (C = 0 there; S = 0 or all ones = "the outside is flooded").  A thread owns one word COLUMN over a SEGMENT of
// consecutive rows and visits them in order -- downwards in one sweep, upwards in the next -- updating S in place, so a sweep
// carries the flood along complete horizontal runs (Kogge-Stone fill inside a word, the neighbour words' edge bits across
// words) AND through a whole segment vertically; words are monotone (bits are only ever set), so the races between threads
// only make a sweep see more.  Sweeps repeat until one changes nothing (block-wide vote).
#pragma once
#include <stdint.h>

namespace fl {

// the runs of c that hold a bit of g (g within c)
__device__ __forceinline__ uint32_t fill_runs(uint32_t g, uint32_t c) {
  uint32_t gl = g, pl = c;
  gl |= pl & (gl << 1);  pl &= pl << 1;
  gl |= pl & (gl << 2);  pl &= pl << 2;
  gl |= pl & (gl << 4);  pl &= pl << 4;
  gl |= pl & (gl << 8);  pl &= pl << 8;
  gl |= pl & (gl << 16);
  uint32_t gr = g, pr = c;
  gr |= pr & (gr >> 1);  pr &= pr >> 1;
  gr |= pr & (gr >> 2);  pr &= pr >> 2;
  gr |= pr & (gr >> 4);  pr &= pr >> 4;
  gr |= pr & (gr >> 8);  pr &= pr >> 8;
  gr |= pr & (gr >> 16);
  return gl | gr;
}

// one visit of word idx; returns true when the word changed
template <bool CONN8>
__device__ __forceinline__ bool visit(const uint32_t* __restrict__ C, volatile uint32_t* S, int idx, int Wp) {
  const uint32_t c = C[idx];
  if (!c) return false;
  const uint32_t s = S[idx];
  if (s == c) return false;
  uint32_t seeds;
  if (CONN8) {
    const uint32_t m = S[idx - Wp] | S[idx + Wp] | s;
    const uint32_t l = (S[idx - Wp - 1] | S[idx - 1] | S[idx + Wp - 1]) >> 31;
    const uint32_t q = (S[idx - Wp + 1] | S[idx + 1] | S[idx + Wp + 1]) << 31;
    seeds = (m | (m << 1) | (m >> 1) | l | q) & c;
  } else {
    seeds = (S[idx - Wp] | S[idx + Wp] | s | (s << 1) | (s >> 1) | (S[idx - 1] >> 31) | (S[idx + 1] << 31)) & c;
  }
  if (seeds == s) return false;
  S[idx] = fill_runs(seeds, c);
  return true;
}

// floods until nothing changes; all threads of the block must call it (it synchronises)
template <bool CONN8>
__device__ __forceinline__ void flood(const uint32_t* __restrict__ C, volatile uint32_t* S, int H, int Wd) {
  const int Wp = Wd + 2, tid = threadIdx.x, nthr = blockDim.x;
  const int nseg = nthr >= Wd ? nthr / Wd : 1;                 // row segments per word column
  const int rpt = (H + nseg - 1) / nseg;                       // rows per segment
  int changed;
  do {
    changed = 0;
#pragma unroll 1
    for (int sweep = 0; sweep < 2; ++sweep) {
      for (int t = tid; t < Wd * nseg; t += nthr) {
        const int seg = t / Wd, w = t - seg * Wd;
        const int r0 = seg * rpt, r1 = min(r0 + rpt, H);
        if (sweep == 0) {
          for (int r = r0; r < r1; ++r) changed |= visit<CONN8>(C, S, (r + 1) * Wp + w + 1, Wp);
        } else {
          for (int r = r1 - 1; r >= r0; --r) changed |= visit<CONN8>(C, S, (r + 1) * Wp + w + 1, Wp);
        }
      }
    }
  } while (__syncthreads_or(changed));
}

}  // namespace fl
