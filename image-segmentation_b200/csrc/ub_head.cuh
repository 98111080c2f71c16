// ub_head.cuh — the fused head of the network, shared by the epilogues of conv_tc_kernel (kx-stacked / 9-tap) and
// conv_rs_kernel (row-streaming): per pixel, bias + ReLU on the 32 fp32 accumulators of dec4.conv.3, the 1x1 head
// (models/UNet.py:73), argmax(softmax(z)) with the first index winning ties (models/losses.py:48-50), the per-thread
// confusion counts and soft-Dice partial sums, and their reduction (warp shuffle -> shared-memory int64 histogram ->
// at most C*C global atomics per reference batch per epilogue group).
#pragma once
#include "ub_ptx.cuh"

namespace ub {

struct ConvParams;
__device__ __forceinline__ void epi_bar_sync(int grp);

// Per-thread state: counts of target class t live in pk[t], kMaxClasses 8-bit fields indexed by the prediction (4
// registers instead of 16, one shift and kMaxClasses predicated adds per pixel); they are reduced only when the
// reference batch changes, every kHeadFlushTiles tiles, and once at the end.
constexpr int kHeadMaxClasses = 4;
constexpr int kHeadFlushTiles = 250;  // < 255: the packed 8-bit count fields cannot overflow between flushes
static_assert(kHeadMaxClasses * 8 <= 32, "packed count fields");

struct HeadState {
  uint32_t pk[kHeadMaxClasses];
  float dsum[2 * kHeadMaxClasses];
  int cur_batch;
  int tiles_since_flush;
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int i = 0; i < kHeadMaxClasses; ++i) pk[i] = 0u;
#pragma unroll
    for (int i = 0; i < 2 * kHeadMaxClasses; ++i) dsum[i] = 0.0f;
    cur_batch = -1;
    tiles_since_flush = 0;
  }
};

// All 128 threads of epilogue group `grp` call this together (it contains the group's named barrier).
// s_cm / s_dice: the group's shared counters ([kHeadMaxClasses^2] int64, [2*kHeadMaxClasses] double), zero on entry and exit.
template <typename P>
__device__ __forceinline__ void head_flush(const P& p, HeadState& hs, int batch, int grp, int et, int lane,
                                           unsigned long long* s_cm, double* s_dice) {
  constexpr int kCells = kHeadMaxClasses * kHeadMaxClasses;
  const int C = p.n_classes;
#pragma unroll
  for (int i = 0; i < kCells; ++i) {
    uint32_t v = (hs.pk[i / kHeadMaxClasses] >> (8 * (i % kHeadMaxClasses))) & 255u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0 && v) atomicAdd(&s_cm[i], static_cast<unsigned long long>(v));
  }
#pragma unroll
  for (int i = 0; i < kHeadMaxClasses; ++i) hs.pk[i] = 0u;
  if (p.dice) {
#pragma unroll
    for (int i = 0; i < 2 * kHeadMaxClasses; ++i) {
      float v = hs.dsum[i];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) atomicAdd(&s_dice[i], static_cast<double>(v));
      hs.dsum[i] = 0.0f;
    }
  }
  epi_bar_sync(grp);  // all four warps of the group have added their share
  if (et < C * C) {
    const int tc = et / C, pc = et % C;
    const unsigned long long v = atomicExch(&s_cm[tc * kHeadMaxClasses + pc], 0ull);
    if (v) atomicAdd(&p.cm[static_cast<size_t>(batch) * C * C + et], v);
  } else if (p.dice && et >= 32 && et < 32 + 2 * C) {
    const int i = et - 32;
    const double v = s_dice[i];
    s_dice[i] = 0.0;
    atomicAdd(&p.dice[(static_cast<size_t>(batch) * C + (i >> 1)) * 3 + (i & 1)], v);
  }
  epi_bar_sync(grp);  // shared counters are zero again before the next batch is added
  hs.tiles_since_flush = 0;
}

// Called once per tile (row) by every thread of the group, with the reference batch of the tile's image: flushes when
// the batch changes or the 8-bit fields could overflow.
template <typename P>
__device__ __forceinline__ void head_begin_tile(const P& p, HeadState& hs, int batch, int grp, int et, int lane,
                                                unsigned long long* s_cm, double* s_dice) {
  if (batch != hs.cur_batch || hs.tiles_since_flush >= kHeadFlushTiles) {
    if (hs.cur_batch >= 0) head_flush(p, hs, hs.cur_batch, grp, et, lane, s_cm, s_dice);
    hs.cur_batch = batch;
  }
  ++hs.tiles_since_flush;
}

// One pixel: r[32] = this pixel's 32 fp32 accumulators of the last conv (before bias).  NC = 3: the usual class count
// known at compile time; NC = 0: p.n_classes at run time.
template <int NC, typename P>
__device__ __forceinline__ void head_pixel(const P& p, HeadState& hs, const uint32_t (&r)[32], bool inb, int b, int y,
                                           int x, int tgt_pre) {
  const int C = NC ? NC : p.n_classes;
  float z[kHeadMaxClasses];
#pragma unroll
  for (int c = 0; c < kHeadMaxClasses; ++c) z[c] = (c < C) ? p.head_b[c] : -INFINITY;
  // packed fp32x2 arithmetic (two IEEE operations per issue slot): even and odd input channels accumulate
  // separately and are added at the end
  float2 z2[kHeadMaxClasses];
#pragma unroll
  for (int c = 0; c < kHeadMaxClasses; ++c) z2[c] = make_float2(z[c], 0.0f);
#pragma unroll
  for (int j = 0; j < 32; j += 2) {
    float2 v = __fadd2_rn(make_float2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])),
                          make_float2(p.bias32[j], p.bias32[j + 1]));
    if (p.relu) { v.x = fmaxf(v.x, 0.0f); v.y = fmaxf(v.y, 0.0f); }
#pragma unroll
    for (int c = 0; c < kHeadMaxClasses; ++c)
      if (c < C) z2[c] = __ffma2_rn(make_float2(p.head_w[c * 32 + j], p.head_w[c * 32 + j + 1]), v, z2[c]);
  }
#pragma unroll
  for (int c = 0; c < kHeadMaxClasses; ++c)
    if (c < C) z[c] = z2[c].x + z2[c].y;
  // argmax(softmax(z)) with the first index winning ties (models/losses.py:48,50).  When the largest logit
  // leads the runner-up by more than 1e-4 the softmax cannot change the order: e_best = expf(0) = 1 exactly and
  // every other e_c <= expf(-1e-4) < 1 - 9e-5, far more than the rounding of the division by the common sum, so
  // pr_best > pr_c strictly and the prediction is the (unique) argmax of z.  Only near-ties, non-finite logits
  // and the soft-Dice sums (which need the probabilities) evaluate the softmax itself.
  float zmax = z[0];
#pragma unroll
  for (int c = 1; c < kHeadMaxClasses; ++c) zmax = fmaxf(zmax, z[c]);
  int pred = 0;
  float second = -INFINITY;
  if constexpr (NC == 3) {
    const float m12 = fmaxf(z[1], z[2]), m02 = fmaxf(z[0], z[2]), m01 = fmaxf(z[0], z[1]);
    pred = (z[0] == zmax) ? 0 : ((z[1] == zmax) ? 1 : 2);
    second = (pred == 0) ? m12 : ((pred == 1) ? m02 : m01);
  } else {
#pragma unroll
    for (int c = kHeadMaxClasses - 1; c >= 0; --c) if (c < C && z[c] == zmax) pred = c;
#pragma unroll
    for (int c = 0; c < kHeadMaxClasses; ++c) if (c < C && c != pred) second = fmaxf(second, z[c]);
  }
  const float gap = zmax - second;
  const bool decisive = (gap > 1e-4f) && (gap <= 3.0e38f) && !(p.dbg & 8);  // false for NaN / infinite logits
  float pr[kHeadMaxClasses];
  if (!decisive || p.dice) {
    float psum = 0.0f;
#pragma unroll
    for (int c = 0; c < kHeadMaxClasses; ++c) { pr[c] = (c < C) ? expf(z[c] - zmax) : 0.0f; psum += pr[c]; }
    int pred_s = 0;
    float best = -1.0f;
#pragma unroll
    for (int c = 0; c < kHeadMaxClasses; ++c) {
      pr[c] = pr[c] / psum;
      if (c < C && pr[c] > best) { best = pr[c]; pred_s = c; }
    }
    pred = pred_s;
  }
  const size_t pix = (static_cast<size_t>(b) * p.H + y) * p.W + x;
  if (inb) {
    if (p.pred_out) p.pred_out[pix] = static_cast<uint8_t>(pred);
    if (p.logits_out) {
#pragma unroll
      for (int c = 0; c < kHeadMaxClasses; ++c)
        if (c < C) p.logits_out[((static_cast<size_t>(b) * C + c) * p.H + y) * p.W + x] = z[c];
    }
  }
  if (p.targets) {
    const int tgt = (tgt_pre >= 0 && tgt_pre < C) ? tgt_pre : -1;
    const uint32_t inc = 1u << (8 * pred);
#pragma unroll
    for (int tc = 0; tc < kHeadMaxClasses; ++tc)
      if (tc < C) hs.pk[tc] += (tgt == tc) ? inc : 0u;
    if (p.dice) {
      // smp DiceLoss(from_logits=True) applies softmax again to the softmax output (losses.py:100)
      float q[kHeadMaxClasses], qmax = pr[0], qsum = 0.0f;
#pragma unroll
      for (int c = 1; c < kHeadMaxClasses; ++c) if (c < C) qmax = fmaxf(qmax, pr[c]);
#pragma unroll
      for (int c = 0; c < kHeadMaxClasses; ++c) { q[c] = (c < C) ? expf(pr[c] - qmax) : 0.0f; qsum += q[c]; }
#pragma unroll
      for (int c = 0; c < kHeadMaxClasses; ++c) {
        const float pc = (inb && c < C) ? q[c] / qsum : 0.0f;
        hs.dsum[2 * c + 0] += (tgt == c) ? pc : 0.0f;
        hs.dsum[2 * c + 1] += pc;
      }
    }
  }
}

// The target class of pixel (b, y, x), fetched before the accumulator is waited for so its latency hides behind the MMAs.
template <typename P>
__device__ __forceinline__ int head_load_target(const P& p, int b, int y, int x) {
  const size_t tp = (static_cast<size_t>(b) * p.H + y) * p.W + x;
  return p.target_fmt == 1 ? static_cast<int>(reinterpret_cast<const uint8_t*>(p.targets)[tp])
                           : static_cast<int>(reinterpret_cast<const long long*>(p.targets)[tp]);
}

}  // namespace ub
