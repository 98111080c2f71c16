// ub_vit.cuh — the CLIP ViT image tower around the tensor-core GEMMs (models/processing_blocks.py:156-233:
// CustomClipPreprocessor.forward = per-image torchvision Resize((224,224)) + Normalize, then
// CLIPModel.get_image_features).  The five linear maps of the tower (patch embedding, qkv, attention output, fc1, fc2)
// run on conv_tc_kernel as 1x1 "convolutions" over the token axis; everything else is here.
//
// Token tensors use the library's channel-blocked layout with the token index in place of the pixel index:
//   X[d/8][tp][8]   (tp = tokens rounded up to a multiple of 8 = the "image" is tp/8 rows of 8 tokens)
// bf16 for GEMM operands / results, fp32 for the residual stream.  Token t = image * ntok + position (position 0 = CLS).
//
// Programmatic dependent launch: every kernel here releases its dependents at once (the next GEMM's CTAs run their prologue —
// barrier init, TMEM allocation, descriptor prefetch — beside these small blocks) and touches global memory only after
// griddepcontrol.wait, i.e. after the preceding grid has completed and flushed: no read or write can overtake it.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/unet_b200.h"
#include "ub_ptx.cuh"

namespace ub {

constexpr int kVitHeadDim = 64;     // every OpenAI CLIP ViT: hidden / heads = 64
constexpr int kVitMaxTokens = 64;   // tokens per image the attention kernel holds in shared memory (ViT-B/32 @224: 50)
constexpr int kVitMaxDim = 1024;    // hidden size limit of the LayerNorm kernels (4 x 256)
constexpr int kClipMaxTaps = UB_CLIP_MAX_TAPS;  // antialiased-resize taps per axis (12: input up to ~5x the output size)

// One output coordinate of torchvision's Resize on a tensor = F.interpolate(mode="bilinear", antialias=True,
// align_corners=False) (ATen UpSampleKernel.cpp, _compute_indices_weights_aa): first input index, tap count, and the
// normalised triangle-filter weights.  Filled on the host in fp32 like ATen does.
struct ResizeTap {
  int start, count;
  float w[kClipMaxTaps];
};

// T = taps per axis evaluated (>= the largest count in the two tables; weights beyond a coordinate's count are zero and
// the indices are clamped into the image, so the loops unroll completely).  wx / xi: the column's weights and clamped indices.
template <int T>
__device__ __forceinline__ float clip_resized_t(const float* __restrict__ plane, int H, int W, int ys, const float* wy,
                                                const float (&wx)[T], const int (&xi)[T]) {
  float acc = 0.0f;
#pragma unroll
  for (int j = 0; j < T; ++j) {
    const float* row = plane + static_cast<size_t>(min(ys + j, H - 1)) * W;
    float r = 0.0f;
#pragma unroll
    for (int i = 0; i < T; ++i) r = fmaf(wx[i], row[xi[i]], r);
    acc = fmaf(wy[j], r, acc);
  }
  return acc;
}

// Resize + Normalize written as the patch matrix of the embedding conv (Conv2d(3, D, P, stride P, bias=False)):
//   A[kg][t][8],  k = c*P*P + py*P + px  (the conv weight's own flattening),  t = image * G*G + gy*G + gx.
// Block (gy, c, image): the P-row x S-column band of channel c that feeds patch row gy.  Phase 1 resamples it with lanes
// along the output columns (coalesced reads of the input rows) into shared memory as bf16; phase 2 writes it out as
// 16-byte items, consecutive threads on consecutive tokens of one k-group plane.
// Dynamic shared memory: bf16[P][S].
__host__ __device__ constexpr int clip_patch_smem_bytes(int P, int G) { return P * G * P * 2; }

// T = taps per axis the instantiation evaluates (>= the largest count in the two tables; the tables' weights beyond a
// coordinate's count are zero and the indices are clamped into the image, so the loops unroll completely: the first
// version with run-time tap counts spent 234 instructions per output pixel, 71 % SM-busy for 0.17 ms; this one 0.106 ms).
// SEP (UB_PATCH_SEP=1, experiment): separable — each input row the band needs is resampled along x into shared memory first
// (2.6x fewer global loads), then along y from there: 0.092 against 0.106 ms; the kernel is bound by its dependent
// load -> FMA chains at 8 warps per block, not by L2 -> L1 traffic.
template <int T, bool SEP = false>
__global__ void __launch_bounds__(256)
clip_patch_kernel(const float* __restrict__ x, int n, int H, int W, const ResizeTap* __restrict__ ry,
                  const ResizeTap* __restrict__ rx, int P, int G, float m0, float m1, float m2, float is0, float is1,
                  float is2, __nv_bfloat16* __restrict__ out, int tp, int nr_max = 0) {
  extern __shared__ __align__(16) unsigned char patch_sm[];
  const int S = G * P;
  __nv_bfloat16* band = reinterpret_cast<__nv_bfloat16*>(patch_sm);
  float* hrow = reinterpret_cast<float*>(patch_sm + static_cast<size_t>(P) * S * 2);  // SEP: [nr_max][S]
  const int gy = blockIdx.x, c = blockIdx.y, img = blockIdx.z, tid = threadIdx.x;
  (void)n;
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2), inv_std = c == 0 ? is0 : (c == 1 ? is1 : is2);
  const float* plane = x + (static_cast<size_t>(img) * 3 + c) * H * W;
  const ResizeTap* rows = ry + gy * P;  // (uniform per iteration: one broadcast load)
  for (int ox = tid; ox < S; ox += 256) {  // a thread keeps its column's taps in registers and walks down the band
    float wx[T];
    int xi[T];
#pragma unroll
    for (int i = 0; i < T; ++i) {
      wx[i] = rx[ox].w[i];
      xi[i] = min(rx[ox].start + i, W - 1);
    }
    if (SEP) {
      const int r0 = rows[0].start;
      const int nr = min(min(rows[P - 1].start + T, H) - r0, nr_max);
#pragma unroll 4
      for (int r = 0; r < nr; ++r) {
        const float* row = plane + static_cast<size_t>(r0 + r) * W;
        float h = 0.0f;
#pragma unroll
        for (int i = 0; i < T; ++i) h = fmaf(wx[i], row[xi[i]], h);
        hrow[r * S + ox] = h;
      }
#pragma unroll 4
      for (int py = 0; py < P; ++py) {  // (same thread, same column: no barrier between the passes)
        const int ys = rows[py].start - r0;
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < T; ++j) acc = fmaf(rows[py].w[j], hrow[min(ys + j, nr - 1) * S + ox], acc);
        band[py * S + ox] = __float2bfloat16((acc - mean) * inv_std);
      }
      continue;
    }
#pragma unroll 4
    for (int py = 0; py < P; ++py) {
      const float acc = clip_resized_t<T>(plane, H, W, rows[py].start, rows[py].w, wx, xi);
      band[py * S + ox] = __float2bfloat16((acc - mean) * inv_std);
    }
  }
  __syncthreads();
  const int pg = P / 8;
  for (int i = tid; i < P * pg * G; i += 256) {
    const int gx = i % G, px8 = (i / G) % pg, py = i / (G * pg);
    const int kg = c * (P * pg) + py * pg + px8;
    const int t = (img * G + gy) * G + gx;
    reinterpret_cast<uint4*>(out)[static_cast<size_t>(kg) * tp + t] =
        *reinterpret_cast<const uint4*>(band + py * S + gx * P + px8 * 8);
  }
}

// The same arithmetic (same instantiation as the patch kernel uses for this input size) as fp32 NCHW [n,3,S,S]: the test
// hook that pins the resize against torchvision.
template <int T>
__global__ void __launch_bounds__(256)
clip_preprocess_kernel(const float* __restrict__ x, int n, int H, int W, const ResizeTap* __restrict__ ry,
                       const ResizeTap* __restrict__ rx, int S, float m0, float m1, float m2, float is0, float is1, float is2,
                       float* __restrict__ out) {
  const size_t total = static_cast<size_t>(n) * 3 * S * S;
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int ox = static_cast<int>(idx % S);
  const int oy = static_cast<int>((idx / S) % S);
  const int c = static_cast<int>((idx / (static_cast<size_t>(S) * S)) % 3);
  const int img = static_cast<int>(idx / (static_cast<size_t>(S) * S * 3));
  const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2), inv_std = c == 0 ? is0 : (c == 1 ? is1 : is2);
  const float* plane = x + (static_cast<size_t>(img) * 3 + c) * H * W;
  float wx[T];
  int xi[T];
#pragma unroll
  for (int i = 0; i < T; ++i) {
    wx[i] = rx[ox].w[i];
    xi[i] = min(rx[ox].start + i, W - 1);
  }
  out[idx] = (clip_resized_t<T>(plane, H, W, ry[oy].start, ry[oy].w, wx, xi) - mean) * inv_std;
}

// ---- LayerNorm family.  A block owns 8 consecutive tokens (one row of the blocked layout); thread (tid & 7) = token in the
// row, (tid >> 3) = channel group within a stripe of 32 groups; a token's D values live in the registers of 32 threads.
// Two-pass statistics (mean, then centred sum of squares) like torch.nn.functional.layer_norm, in fp32.
struct LnReduce {
  float part[8][8];  // [warp][token in row]
};

__device__ __forceinline__ float ln_token_sum(float v, LnReduce& s, int tid) {
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 16);
  __syncthreads();  // (the previous round's readers are done)
  if ((tid & 31) < 8) s.part[tid >> 5][tid & 7] = v;
  __syncthreads();
  float t = 0.0f;
#pragma unroll
  for (int w = 0; w < 8; ++w) t += s.part[w][tid & 7];
  return t;
}

enum VitLnMode : int {
  VIT_LN_EMBED = 0,  // v = (CLS | patch embedding) + position embedding; resid = LN(v)           (CLIP's pre_layrnorm)
  VIT_LN_ADD = 1     // v = resid (+ branch);   resid = v (when a branch is given);   out = LN(v)  (layer_norm1 / layer_norm2)
};

// D = 256 * NS.  resid fp32 [D/8][tp][8]; branch / patches / out bf16 [D/8][tp*][8].
template <int NS, int MODE>
__global__ void __launch_bounds__(256)
vit_ln_kernel(float* __restrict__ resid, const __nv_bfloat16* __restrict__ branch, const float* __restrict__ gamma,
              const float* __restrict__ beta, float eps, __nv_bfloat16* __restrict__ out, int tokens, int tp,
              // VIT_LN_EMBED only
              const __nv_bfloat16* __restrict__ patches, int tp_patch, const float* __restrict__ cls,
              const float* __restrict__ pos, int ntok) {
  __shared__ LnReduce s;
  constexpr int D = 256 * NS;
  const int tid = threadIdx.x;
  const int t = blockIdx.x * 8 + (tid & 7);
  const bool live = t < tokens;
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  float v[NS][8];
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    const int g = (tid >> 3) + 32 * k;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[k][j] = 0.0f;
    if (!live) continue;
    if (MODE == VIT_LN_EMBED) {
      const int img = t / ntok, p = t % ntok;
      if (p == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[k][j] = cls[8 * g + j];
      } else {
        const uint4 q = reinterpret_cast<const uint4*>(patches)[static_cast<size_t>(g) * tp_patch + img * (ntok - 1) + p - 1];
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
        for (int j = 0; j < 4; ++j) { const float2 f = __bfloat1622float2(h[j]); v[k][2 * j] = f.x; v[k][2 * j + 1] = f.y; }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) v[k][j] += pos[static_cast<size_t>(p) * D + 8 * g + j];
    } else {
      const float4* r = reinterpret_cast<const float4*>(resid) + (static_cast<size_t>(g) * tp + t) * 2;
      const float4 a = r[0], b = r[1];
      v[k][0] = a.x; v[k][1] = a.y; v[k][2] = a.z; v[k][3] = a.w;
      v[k][4] = b.x; v[k][5] = b.y; v[k][6] = b.z; v[k][7] = b.w;
      if (branch) {
        const uint4 q = reinterpret_cast<const uint4*>(branch)[static_cast<size_t>(g) * tp + t];
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
        for (int j = 0; j < 4; ++j) { const float2 f = __bfloat1622float2(h[j]); v[k][2 * j] += f.x; v[k][2 * j + 1] += f.y; }
        float4* w = reinterpret_cast<float4*>(resid) + (static_cast<size_t>(g) * tp + t) * 2;
        w[0] = make_float4(v[k][0], v[k][1], v[k][2], v[k][3]);
        w[1] = make_float4(v[k][4], v[k][5], v[k][6], v[k][7]);
      }
    }
  }
  float sum = 0.0f;
#pragma unroll
  for (int k = 0; k < NS; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) sum += v[k][j];
  const float mean = ln_token_sum(sum, s, tid) * (1.0f / D);
  float sq = 0.0f;
#pragma unroll
  for (int k = 0; k < NS; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float d = v[k][j] - mean; sq = fmaf(d, d, sq); }
  const float rstd = rsqrtf(ln_token_sum(sq, s, tid) * (1.0f / D) + eps);
  if (!live) return;
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    const int g = (tid >> 3) + 32 * k;
    float y[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) y[j] = (v[k][j] - mean) * rstd * gamma[8 * g + j] + beta[8 * g + j];
    if (MODE == VIT_LN_EMBED) {
      float4* w = reinterpret_cast<float4*>(resid) + (static_cast<size_t>(g) * tp + t) * 2;
      w[0] = make_float4(y[0], y[1], y[2], y[3]);
      w[1] = make_float4(y[4], y[5], y[6], y[7]);
    } else {
      uint4 o;
      __nv_bfloat162* op = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) op[j] = __floats2bfloat162_rn(y[2 * j], y[2 * j + 1]);
      reinterpret_cast<uint4*>(out)[static_cast<size_t>(g) * tp + t] = o;
    }
  }
}

// ---- Multi-head self-attention of one (head, image): softmax(q k^T / sqrt(64)) v over ntok <= 64 tokens (CLIPAttention,
// eager path: scores scaled, softmax in fp32, no mask in the vision tower).  qkv bf16 [3D/8][tp][8] with q | k | v
// stacked along the channels; out bf16 [D/8][tp][8].
// 50 x 50 x 64 per head is far below the 128-row minimum of a tcgen05 tile, so the two small products run on the
// warp-level tensor-core path (mma.sync m16n8k16, bf16 x bf16 -> fp32): warp w owns query rows [16w, 16w+16), the scores
// stay in registers (the accumulator fragment of q k^T IS the A fragment of p v), softmax statistics in fp32 across the
// four lanes of a row.  Shared memory: q | k | v as [64][72] bf16 (144-byte rows: conflict-free ldmatrix), rows >= ntok zero.
constexpr int kAttnLd = 72;
constexpr int kAttnSmemBytes = 3 * kVitMaxTokens * kAttnLd * 2;

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], const void* smem_ptr) {
  const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(smem_ptr));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem_ptr) {
  const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(smem_ptr));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}

__global__ void __launch_bounds__(128)
vit_attention_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, int D, int ntok, int tp) {
  __shared__ __align__(16) __nv_bfloat16 sm[3 * kVitMaxTokens * kAttnLd];
  __nv_bfloat16* sq = sm;
  __nv_bfloat16* sk = sm + kVitMaxTokens * kAttnLd;
  __nv_bfloat16* sv = sm + 2 * kVitMaxTokens * kAttnLd;
  const int head = blockIdx.x, img = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int t0 = img * ntok;
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  // load: 3 matrices x 8 channel groups x 64 token rows of 16 bytes (rows >= ntok: zeros)
  for (int i = tid; i < 3 * 8 * kVitMaxTokens; i += 128) {
    const int tok = i % kVitMaxTokens, g = (i / kVitMaxTokens) % 8, m = i / (8 * kVitMaxTokens);
    uint4 q = make_uint4(0u, 0u, 0u, 0u);
    if (tok < ntok) q = reinterpret_cast<const uint4*>(qkv)[static_cast<size_t>(m * (D / 8) + head * 8 + g) * tp + t0 + tok];
    *reinterpret_cast<uint4*>(sm + (m * kVitMaxTokens + tok) * kAttnLd + 8 * g) = q;
  }
  __syncthreads();
  if (16 * warp >= ntok) return;  // (no barrier follows)
  // scores S[16 rows][64 keys] = q k^T: 4 k-steps (16 dims each) x 8 key tiles
  float sc[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt)
#pragma unroll
    for (int j = 0; j < 4; ++j) sc[nt][j] = 0.0f;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    uint32_t a[4];
    ldmatrix_x4(a, sq + (16 * warp + (lane & 15)) * kAttnLd + 16 * ks + 8 * (lane >> 4));
#pragma unroll
    for (int np = 0; np < 4; ++np) {  // two key tiles per ldmatrix: (keys 16np..+7 | +8..+15) x (dims 16ks..+7 | +8..+15)
      uint32_t b[4];
      ldmatrix_x4(b, sk + (16 * np + 8 * (lane >> 4) + (lane & 7)) * kAttnLd + 16 * ks + 8 * ((lane >> 3) & 1));
      mma_bf16_16816(sc[2 * np], a, b[0], b[1]);
      mma_bf16_16816(sc[2 * np + 1], a, b[2], b[3]);
    }
  }
  // softmax over the keys of rows r0 = lane/4 and r0 + 8 (this thread: columns 8nt + 2(lane%4) + {0,1})
  float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    const int c = 8 * nt + 2 * (lane & 3);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const bool ok = c + j < ntok;
      sc[nt][j] = ok ? sc[nt][j] * 0.125f : -INFINITY;
      sc[nt][2 + j] = ok ? sc[nt][2 + j] * 0.125f : -INFINITY;
      mx0 = fmaxf(mx0, sc[nt][j]);
      mx1 = fmaxf(mx1, sc[nt][2 + j]);
    }
  }
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
  float sum0 = 0.0f, sum1 = 0.0f;
  uint32_t pa[4][4];  // p as the A operand of p v: key block kb = score tiles 2kb, 2kb+1
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    const float e0 = expf(sc[nt][0] - mx0), e1 = expf(sc[nt][1] - mx0), e2 = expf(sc[nt][2] - mx1), e3 = expf(sc[nt][3] - mx1);
    sum0 += e0 + e1;
    sum1 += e2 + e3;
    pa[nt >> 1][(nt & 1) * 2 + 0] = pack2_bf16(e0, e1);
    pa[nt >> 1][(nt & 1) * 2 + 1] = pack2_bf16(e2, e3);
  }
  sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1); sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
  sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1); sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
  // out[16 rows][64 dims] = p v: 4 key blocks x 8 dim tiles
  float oc[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt)
#pragma unroll
    for (int j = 0; j < 4; ++j) oc[nt][j] = 0.0f;
#pragma unroll
  for (int kb = 0; kb < 4; ++kb) {
#pragma unroll
    for (int np = 0; np < 4; ++np) {  // (keys 16kb..+7 | +8..+15) x (dims 16np..+7 | +8..+15), transposed on load
      uint32_t b[4];
      ldmatrix_x4_trans(b, sv + (16 * kb + 8 * ((lane >> 3) & 1) + (lane & 7)) * kAttnLd + 16 * np + 8 * (lane >> 4));
      mma_bf16_16816(oc[2 * np], pa[kb], b[0], b[1]);
      mma_bf16_16816(oc[2 * np + 1], pa[kb], b[2], b[3]);
    }
  }
  const float inv0 = 1.0f / sum0, inv1 = 1.0f / sum1;
  const int r0 = 16 * warp + (lane >> 2), r1 = r0 + 8;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {  // dim tile nt = channel group nt of this head: a quad writes its 16 bytes
    uint32_t* dst = reinterpret_cast<uint32_t*>(out) + (static_cast<size_t>(head * 8 + nt) * tp + t0) * 4 + (lane & 3);
    if (r0 < ntok) dst[static_cast<size_t>(r0) * 4] = pack2_bf16(oc[nt][0] * inv0, oc[nt][1] * inv0);
    if (r1 < ntok) dst[static_cast<size_t>(r1) * 4] = pack2_bf16(oc[nt][2] * inv1, oc[nt][3] * inv1);
  }
}

// ---- pooled output: post_layernorm of the CLS token, then visual_projection (Linear(D, E, bias=False)), fp32 throughout.
// Block (image, 64-output slice): the normalised CLS vector is recomputed per slice (D values), each warp then owns 8
// outputs.  feats [n][E].  `branch`: the last encoder layer's fc2 output, not yet added to the residual stream.
__global__ void __launch_bounds__(256)
vit_pool_proj_kernel(const float* __restrict__ resid, const __nv_bfloat16* __restrict__ branch, int D, int ntok, int tp,
                     const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                     const float* __restrict__ w_proj, int E, float* __restrict__ feats) {
  __shared__ float pooled[kVitMaxDim];
  __shared__ float red[8];
  __shared__ float stat[2];
  const int img = blockIdx.x, e0 = blockIdx.y * 64, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int t = img * ntok;  // the CLS token
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  float loc = 0.0f;
  for (int d = tid; d < D; d += 256) {
    const size_t at = (static_cast<size_t>(d >> 3) * tp + t) * 8 + (d & 7);
    const float x = resid[at] + (branch ? __bfloat162float(branch[at]) : 0.0f);  // the last layer's MLP output joins here
    pooled[d] = x;
    loc += x;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) loc += __shfl_xor_sync(0xffffffffu, loc, o);
  if (lane == 0) red[warp] = loc;
  __syncthreads();
  if (tid == 0) { float s = 0.0f; for (int w = 0; w < 8; ++w) s += red[w]; stat[0] = s / D; }
  __syncthreads();
  const float mean = stat[0];
  loc = 0.0f;
  for (int d = tid; d < D; d += 256) { const float c = pooled[d] - mean; loc = fmaf(c, c, loc); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) loc += __shfl_xor_sync(0xffffffffu, loc, o);
  if (lane == 0) red[warp] = loc;
  __syncthreads();
  if (tid == 0) { float s = 0.0f; for (int w = 0; w < 8; ++w) s += red[w]; stat[1] = rsqrtf(s / D + eps); }
  __syncthreads();
  const float rstd = stat[1];
  for (int d = tid; d < D; d += 256) pooled[d] = (pooled[d] - mean) * rstd * gamma[d] + beta[d];
  __syncthreads();
  for (int e = e0 + warp; e < min(e0 + 64, E); e += 8) {
    const float* wr = w_proj + static_cast<size_t>(e) * D;
    float acc = 0.0f;
    for (int d = lane; d < D; d += 32) acc = fmaf(wr[d], pooled[d], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) feats[static_cast<size_t>(img) * E + e] = acc;
  }
}

}  // namespace ub
