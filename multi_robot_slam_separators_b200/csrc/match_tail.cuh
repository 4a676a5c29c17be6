// Shared tail of the two Hamming matchers (hamming_knn.cu: POPC pipe, hamming_mma.cu: tcgen05):
// NNDR decision, one-to-one uniqueness filter (R/src/myRegistrationVis.cpp:856-894) and the ascending
// compaction into the PnP correspondence list of util3d::estimateMotion3DTo2D.
#pragma once
#include "common.cuh"

namespace mtail {

// RTAB-Map compares in f32 and needs two neighbours: accept iff d1 <= nndr*d2 (global path,
// VWDictionary::addNewWords) or d1 < nndr*d2 (guess path, myRegistrationVis.cpp:584,744)
__device__ __forceinline__ int nndr_assign(int i1, int d1, int i2, int d2, float nndr, int strict) {
  if (i2 < 0) return -1;
  const float lhs = (float)d1, rhs = __fmul_rn(nndr, (float)d2);
  return (strict ? (lhs < rhs) : (lhs <= rhs)) ? i1 : -1;
}

// Block-wide; returns the list length.  Requires: s_assigned[0..nt) final and visible (caller synchronised), s_count[0..nf) zeroed,
// *s_dropped == 0.  s_scan holds THREADS/32 ints.  THREADS * 32 >= max features per frame.
template <int THREADS>
__device__ __forceinline__ int unique_and_compact(int tid, int nf, int nt, const float* __restrict__ p3,
                                                   const int32_t* s_assigned, uint32_t* s_count, uint16_t* s_owner,
                                                   int* s_scan, int* s_dropped, size_t obase, int pair,
                                                   int32_t* __restrict__ n_matches, uint16_t* __restrict__ match_from,
                                                   uint16_t* __restrict__ match_to, int32_t* __restrict__ n_unique_to) {
  // To ids used more than once are dropped (:882-894)
  for (int qi = tid; qi < nt; qi += THREADS) {
    const int a = s_assigned[qi];
    if (a >= 0) {
      atomicAdd(&s_count[a], 1u);
      s_owner[a] = (uint16_t)qi;  // only read back where count == 1 (single writer)
    }
  }
  __syncthreads();
  // ascending From index, drop non-finite From 3-D (util3d::isFinite) -> compact list
  const int chunk = (nf + THREADS - 1) / THREADS;
  const int lo = min(nf, tid * chunk), hi = min(nf, lo + chunk);
  int local = 0, dropped = 0;
  uint32_t keep_mask = 0;  // chunk <= 32
  for (int i = lo; i < hi; ++i) {
    const uint32_t c = s_count[i];
    if (c >= 2) dropped += (int)c;
    if (c == 1) {
      const float x = p3[3 * i], y = p3[3 * i + 1], z = p3[3 * i + 2];
      if (isfinite(x) && isfinite(y) && isfinite(z)) {
        keep_mask |= 1u << (i - lo);
        ++local;
      }
    }
  }
  if (dropped) atomicAdd(s_dropped, dropped);
  // block exclusive scan of `local`
  const int lane = tid & 31, warp = tid >> 5;
  int incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_scan[warp] = incl;
  __syncthreads();
  int base = 0, total = 0;
#pragma unroll
  for (int w = 0; w < THREADS / 32; ++w) {
    const int v = s_scan[w];
    if (w < warp) base += v;
    total += v;
  }
  int pos = base + incl - local;
  for (int i = lo; i < hi; ++i) {
    if (keep_mask & (1u << (i - lo))) {
      if (match_from) match_from[obase + pos] = (uint16_t)i;
      if (match_to) match_to[obase + pos] = s_owner[i];
      ++pos;
    }
  }
  if (tid == 0) {
    n_matches[pair] = total;
    if (n_unique_to) n_unique_to[pair] = nt - *s_dropped;
  }
  return total;   // the length of the correspondence list, in every thread
}

}  // namespace mtail
