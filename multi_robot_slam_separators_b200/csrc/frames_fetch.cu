// Lazy keypoints for host-resident keyframes (sepb_frames_bind_host_async): the descriptors of every frame travel to
// the device (K2 reads all of them), but key-point positions are only needed for the pairs that reach PnP -- a few
// percent of the candidates (R/src/myRegistrationVis.cpp:1069-1070: fewer than Vis/MinInliers correspondences => no
// motion estimate).  So the device PULLS them: a mark kernel flags the frames of the pairs with enough matches, a fetch
// kernel copies the flagged frames' kp2d / kp3d / octave rows straight out of the caller's pinned host arrays (mapped
// into the device address space, 128-bit loads over PCIe) into the store.  No host round trip, no host-side gather:
// the decision which rows cross the bus is taken where the match counts are.
#include "common.cuh"

namespace {

// thread per pair: need bits for the two frames of every pair that will run PnP (n_matches >= min_matches), or, for
// the second call of estimate_transformation, whose first call succeeded (status == 1)
__global__ void frames_mark_kernel(const int32_t* pairs, int n_pairs, const int32_t* n_matches, int min_matches,
                                   const int32_t* status, uint8_t* need_from, int from_size, uint8_t bits_from,
                                   uint8_t* need_to, int to_size, uint8_t bits_to) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pairs) return;
  if (status ? status[p] != 1 : n_matches[p] < min_matches) return;
  const int fi = pairs[2 * p], ti = pairs[2 * p + 1];
  if (fi < 0 || fi >= from_size || ti < 0 || ti >= to_size) return;
  // same-value byte stores race benignly; distinct bit sets for one frame only occur when From and To are one store
  if (need_from && bits_from) atomicOr(reinterpret_cast<unsigned int*>(need_from + (fi & ~3)), (unsigned)bits_from << (8 * (fi & 3)));
  if (need_to && bits_to) atomicOr(reinterpret_cast<unsigned int*>(need_to + (ti & ~3)), (unsigned)bits_to << (8 * (ti & 3)));
}

// CTA per slot of the binding: copy what is needed and not yet resident
__global__ void __launch_bounds__(128) frames_fetch_kernel(float* kp2d, float* kp3d, int32_t* octave, const int32_t* count,
                                                           int max_feat, int first, uint8_t* need, uint8_t* have,
                                                           const float* h_kp2d, const float* h_kp3d, const int32_t* h_octave,
                                                           unsigned long long* fetched) {
  const int rel = blockIdx.x, slot = first + rel;
  const uint8_t nd = need[slot];
  if (!nd) return;
  const uint8_t want = nd & ~have[slot];
  if (!want) {
    if (threadIdx.x == 0) need[slot] = 0;
    return;
  }
  const int n = min(max(count[slot], 0), max_feat);
  const size_t row = (size_t)slot * max_feat, hrow = (size_t)rel * max_feat;
  unsigned long long bytes = 0;
  {
    // pixels and 3-D points in one batch of loads (one PCIe round trip for both rows at F = 1000)
    const int b2 = ((want & 1) && h_kp2d) ? n * 8 : 0, b3 = ((want & 2) && h_kp3d) ? n * 12 : 0;
    if (b3)
      sepb_pull_rows(kp3d + row * 3, h_kp3d + hrow * 3, b3, kp2d + row * 2, h_kp2d ? h_kp2d + hrow * 2 : nullptr, b2, threadIdx.x,
                     blockDim.x);
    else if (b2)
      sepb_pull_row(kp2d + row * 2, h_kp2d + hrow * 2, b2);
    bytes += (unsigned long long)b2 + (unsigned long long)b3;
  }
  if ((want & 4) && h_octave) { sepb_pull_row(octave + row, h_octave + hrow, n * 4); bytes += (unsigned long long)n * 4; }
  __syncthreads();
  if (threadIdx.x == 0) {
    have[slot] |= want;
    need[slot] = 0;
    atomicAdd(fetched, bytes);
  }
}

}  // namespace

LazyPull sepb_lazy_pull_of(sepb_frames* from, sepb_frames* to, int min_matches) {
  LazyPull lp{};
  if (!(from->lazy || to->lazy) || min_matches <= 0) return lp;
  lp.min_matches = min_matches;
  if (from->lazy) {
    lp.from_kp3d = from->kp3d; lp.from_kp3d_host = from->h_kp3d; lp.from_have = from->have; lp.from_fetched = from->fetched;
    lp.from_first = from->lazy_first; lp.from_n = from->lazy_n;
  }
  if (to->lazy) {
    lp.to_kp2d = to->kp2d; lp.to_kp2d_host = to->h_kp2d; lp.to_have = to->have; lp.to_fetched = to->fetched;
    lp.to_first = to->lazy_first; lp.to_n = to->lazy_n;
  }
  return lp;
}

// flag + pull for one call.  bits: 1 kp2d, 2 kp3d, 4 octave.  No-op for stores that are not host-bound.
int sepb_launch_frames_fetch(sepb_ctx* ctx, sepb_frames* from, sepb_frames* to, const int32_t* d_pairs, int n_pairs,
                             const int32_t* d_n_matches, int min_matches, const int32_t* d_status, int bits_from, int bits_to) {
  const bool lf = from->lazy && bits_from, lt = to->lazy && bits_to;
  if ((!lf && !lt) || n_pairs == 0) return SEPB_OK;
  frames_mark_kernel<<<(n_pairs + 255) / 256, 256, 0, ctx->stream>>>(d_pairs, n_pairs, d_n_matches, min_matches, d_status,
                                                                     lf ? from->need : nullptr, from->size, (uint8_t)bits_from,
                                                                     lt ? to->need : nullptr, to->size, (uint8_t)bits_to);
  ctx->launches++;
  for (sepb_frames* f : {lf ? from : (sepb_frames*)nullptr, lt && !(lf && to == from) ? to : (sepb_frames*)nullptr}) {
    if (!f || f->lazy_n == 0) continue;
    frames_fetch_kernel<<<f->lazy_n, 128, 0, ctx->stream>>>(f->kp2d, f->kp3d, f->octave, f->count, f->max_feat, f->lazy_first,
                                                           f->need, f->have, f->h_kp2d, f->h_kp3d, f->h_octave, f->fetched);
    ctx->launches++;
  }
  SEPB_CUDA(ctx, cudaGetLastError());
  return SEPB_OK;
}
