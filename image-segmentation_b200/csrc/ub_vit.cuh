// ub_vit.cuh — the CLIP ViT image tower around the tensor-core GEMMs (models/processing_blocks.py:156-233:
// CustomClipPreprocessor.forward = per-image torchvision Resize((224,224)) + Normalize, then
// CLIPModel.get_image_features).  The five linear maps of the tower (patch embedding, qkv, attention output, fc1, fc2)
// run on conv_tc_kernel as 1x1 "convolutions" over the token axis; everything else is here.
//
// Token tensors use the library's channel-blocked layout with the token index in place of the pixel index:
//   X[d/8][tp][8]   (tp = tokens rounded up to a multiple of 8 = the "image" is tp/8 rows of 8 tokens)
// bf16 for GEMM operands / results, fp32 for the residual stream.  Token t = image * ntok + position (position 0 = CLS).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ub {

constexpr int kVitHeadDim = 64;     // every OpenAI CLIP ViT: hidden / heads = 64
constexpr int kVitMaxTokens = 64;   // tokens per image the attention kernel holds in shared memory (ViT-B/32 @224: 50)
constexpr int kVitMaxDim = 1024;    // hidden size limit of the LayerNorm kernels (4 x 256)
constexpr int kClipMaxTaps = 12;    // antialiased-resize taps per axis (input up to ~5x the output size)

// One output coordinate of torchvision's Resize on a tensor = F.interpolate(mode="bilinear", antialias=True,
// align_corners=False) (ATen UpSampleKernel.cpp, _compute_indices_weights_aa): first input index, tap count, and the
// normalised triangle-filter weights.  Filled on the host in fp32 like ATen does.
struct ResizeTap {
  int start, count;
  float w[kClipMaxTaps];
};

__device__ __forceinline__ float clip_resized(const float* __restrict__ plane, int W, const ResizeTap& ty, const ResizeTap& tx) {
  float acc = 0.0f;
  for (int j = 0; j < ty.count; ++j) {
    const float* row = plane + static_cast<size_t>(ty.start + j) * W + tx.start;
    float r = 0.0f;
    for (int i = 0; i < tx.count; ++i) r = fmaf(tx.w[i], row[i], r);
    acc = fmaf(ty.w[j], r, acc);
  }
  return acc;
}

// Resize + Normalize written as the patch matrix of the embedding conv (Conv2d(3, D, P, stride P, bias=False)):
//   A[kg][t][8],  k = c*P*P + py*P + px  (the conv weight's own flattening),  t = image * G*G + gy*G + gx.
// One thread per (k group, patch token): 8 horizontally adjacent output pixels -> one 16-byte store; consecutive threads
// are consecutive tokens of one plane (coalesced stores; the gathers hit L1/L2).
__global__ void __launch_bounds__(256)
clip_patch_kernel(const float* __restrict__ x, int n, int H, int W, const ResizeTap* __restrict__ ry,
                  const ResizeTap* __restrict__ rx, int P, int G, float m0, float m1, float m2, float is0, float is1,
                  float is2, __nv_bfloat16* __restrict__ out, int tp) {
  const int tokens = n * G * G;
  const int kgs_per_c = P * P / 8;
  const size_t total = static_cast<size_t>(3 * kgs_per_c) * tokens;
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int t = static_cast<int>(idx % tokens);
  const int kg = static_cast<int>(idx / tokens);
  const int c = kg / kgs_per_c, rem = kg % kgs_per_c;
  const int py = rem / (P / 8), px0 = (rem % (P / 8)) * 8;
  const int img = t / (G * G), g = t % (G * G);
  const int oy = (g / G) * P + py, ox0 = (g % G) * P + px0;
  const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2), inv_std = c == 0 ? is0 : (c == 1 ? is1 : is2);
  const float* plane = x + (static_cast<size_t>(img) * 3 + c) * H * W;
  const ResizeTap ty = ry[oy];
  uint4 o;
  __nv_bfloat162* op = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float a = (clip_resized(plane, W, ty, rx[ox0 + 2 * j]) - mean) * inv_std;
    const float b = (clip_resized(plane, W, ty, rx[ox0 + 2 * j + 1]) - mean) * inv_std;
    op[j] = __floats2bfloat162_rn(a, b);
  }
  reinterpret_cast<uint4*>(out)[static_cast<size_t>(kg) * tp + t] = o;
}

// The same arithmetic as fp32 NCHW [n,3,S,S] (test hook: pins the resize against torchvision).
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
  out[idx] = (clip_resized(plane, W, ry[oy], rx[ox]) - mean) * inv_std;
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

// ---- Multi-head self-attention of one (head, image): softmax(q k^T / sqrt(64)) v over ntok <= 64 tokens, fp32 in shared
// memory (CLIPAttention, eager path: scores scaled, softmax in fp32, no mask in the vision tower).
// qkv bf16 [3D/8][tp][8] with q | k | v stacked along the channels; out bf16 [D/8][tp][8].
// Dynamic shared memory: q[ntok][65] | k[ntok][65] | v[ntok][64] | s[ntok][ntok + 1] floats.
__host__ __device__ constexpr int vit_attn_smem_bytes(int ntok) { return 4 * (ntok * 65 * 2 + ntok * 64 + ntok * (ntok + 1)); }

__global__ void __launch_bounds__(128)
vit_attention_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, int D, int ntok, int tp) {
  extern __shared__ float vit_sm[];
  float* sq = vit_sm;
  float* sk = sq + ntok * 65;
  float* sv = sk + ntok * 65;
  float* ss = sv + ntok * 64;
  const int head = blockIdx.x, img = blockIdx.y, tid = threadIdx.x;
  const int t0 = img * ntok;
  const int ld_s = ntok + 1;
  // load: 3 matrices x 8 channel groups x ntok tokens of 16 bytes
  for (int i = tid; i < 3 * 8 * ntok; i += 128) {
    const int tok = i % ntok, g = (i / ntok) % 8, m = i / (8 * ntok);
    const uint4 q = reinterpret_cast<const uint4*>(qkv)[static_cast<size_t>(m * (D / 8) + head * 8 + g) * tp + t0 + tok];
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
    float* dst = (m == 0 ? sq + tok * 65 : (m == 1 ? sk + tok * 65 : sv + tok * 64)) + 8 * g;
#pragma unroll
    for (int j = 0; j < 4; ++j) { const float2 f = __bfloat1622float2(h[j]); dst[2 * j] = f.x; dst[2 * j + 1] = f.y; }
  }
  __syncthreads();
  // scores: 5 x 5 register tiles
  const int nt5 = (ntok + 4) / 5;
  for (int tile = tid; tile < nt5 * nt5; tile += 128) {
    const int i0 = (tile / nt5) * 5, j0 = (tile % nt5) * 5;
    float acc[5][5];
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
      for (int b = 0; b < 5; ++b) acc[a][b] = 0.0f;
    const float* qp[5];
    const float* kp[5];
#pragma unroll
    for (int a = 0; a < 5; ++a) {
      qp[a] = sq + min(i0 + a, ntok - 1) * 65;
      kp[a] = sk + min(j0 + a, ntok - 1) * 65;
    }
#pragma unroll 4
    for (int d = 0; d < 64; ++d) {
      float qa[5], kb[5];
#pragma unroll
      for (int a = 0; a < 5; ++a) { qa[a] = qp[a][d]; kb[a] = kp[a][d]; }
#pragma unroll
      for (int a = 0; a < 5; ++a)
#pragma unroll
        for (int b = 0; b < 5; ++b) acc[a][b] = fmaf(qa[a], kb[b], acc[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
      for (int b = 0; b < 5; ++b)
        if (i0 + a < ntok && j0 + b < ntok) ss[(i0 + a) * ld_s + j0 + b] = acc[a][b] * 0.125f;
  }
  __syncthreads();
  // softmax: one warp per row
  for (int i = tid >> 5; i < ntok; i += 4) {
    const int lane = tid & 31;
    float* row = ss + i * ld_s;
    const float a = lane < ntok ? row[lane] : -INFINITY, b = lane + 32 < ntok ? row[lane + 32] : -INFINITY;
    float mx = fmaxf(a, b);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float ea = lane < ntok ? expf(a - mx) : 0.0f, eb = lane + 32 < ntok ? expf(b - mx) : 0.0f;
    float sum = ea + eb;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane < ntok) row[lane] = ea / sum;
    if (lane + 32 < ntok) row[lane + 32] = eb / sum;
  }
  __syncthreads();
  // out = p v: 5 tokens x 8 channels (one channel group) per thread
  for (int tile = tid; tile < nt5 * 8; tile += 128) {
    const int i0 = (tile / 8) * 5, g = tile % 8;
    float acc[5][8];
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
      for (int b = 0; b < 8; ++b) acc[a][b] = 0.0f;
    const float* pp[5];
#pragma unroll
    for (int a = 0; a < 5; ++a) pp[a] = ss + min(i0 + a, ntok - 1) * ld_s;
    for (int j = 0; j < ntok; ++j) {
      const float4 v0 = *reinterpret_cast<const float4*>(sv + j * 64 + 8 * g);
      const float4 v1 = *reinterpret_cast<const float4*>(sv + j * 64 + 8 * g + 4);
#pragma unroll
      for (int a = 0; a < 5; ++a) {
        const float p = pp[a][j];
        acc[a][0] = fmaf(p, v0.x, acc[a][0]); acc[a][1] = fmaf(p, v0.y, acc[a][1]);
        acc[a][2] = fmaf(p, v0.z, acc[a][2]); acc[a][3] = fmaf(p, v0.w, acc[a][3]);
        acc[a][4] = fmaf(p, v1.x, acc[a][4]); acc[a][5] = fmaf(p, v1.y, acc[a][5]);
        acc[a][6] = fmaf(p, v1.z, acc[a][6]); acc[a][7] = fmaf(p, v1.w, acc[a][7]);
      }
    }
#pragma unroll
    for (int a = 0; a < 5; ++a) {
      if (i0 + a >= ntok) continue;
      uint4 o;
      __nv_bfloat162* op = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) op[j] = __floats2bfloat162_rn(acc[a][2 * j], acc[a][2 * j + 1]);
      reinterpret_cast<uint4*>(out)[static_cast<size_t>(head * 8 + g) * tp + t0 + i0 + a] = o;
    }
  }
}

// ---- pooled output: post_layernorm of the CLS token, then visual_projection (Linear(D, E, bias=False)); one block per
// image, fp32 throughout.  feats [n][E].  `branch`: the last encoder layer's fc2 output, not yet added to the stream.
__global__ void __launch_bounds__(256)
vit_pool_proj_kernel(const float* __restrict__ resid, const __nv_bfloat16* __restrict__ branch, int D, int ntok, int tp,
                     const float* __restrict__ gamma,
                     const float* __restrict__ beta, float eps, const float* __restrict__ w_proj, int E,
                     float* __restrict__ feats) {
  __shared__ float pooled[kVitMaxDim];
  __shared__ float red[8];
  __shared__ float stat[2];
  const int img = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int t = img * ntok;  // the CLS token
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
  for (int e = warp; e < E; e += 8) {
    const float* wr = w_proj + static_cast<size_t>(e) * D;
    float acc = 0.0f;
    for (int d = lane; d < D; d += 32) acc = fmaf(wr[d], pooled[d], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) feats[static_cast<size_t>(img) * E + e] = acc;
  }
}

}  // namespace ub
