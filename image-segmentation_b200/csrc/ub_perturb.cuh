// ub_perturb.cuh — perturbations (both reference families) + the 1x1 stem conv that consumes them.
//   uint8-space Dataset wrappers : customDatasets/perturbations.py  (exact integer arithmetic)
//   float-space nn.Modules       : models/processing_blocks.py:454-592 (fp32, no FMA contraction)
//   stem                         : models/UNet.py:39,62  nn.Conv2d(3,32,1)
// Pointwise families are evaluated inside the stem's input load (no intermediate image in HBM);
// the two blur families need neighbours and run as short ping-pong passes first.
#pragma once
#include <type_traits>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/unet_b200.h"
#include "ub_ptx.cuh"

namespace ub {

struct PerturbDev {
  int kind;
  int iparam;
  float fparam;
  float fparam2;
  unsigned long long seed;
  const void* aux;
  uint8_t lut[256];
};

// ---- counter-based RNG (Philox4x32-10) for throughput mode (aux == NULL)
__device__ __forceinline__ uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ uint4 rng4(unsigned long long seed, uint32_t a, uint32_t b, uint32_t c) {
  return philox4x32(make_uint4(a, b, c, 0x5eed5eedu),
                    make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32)));
}
__device__ __forceinline__ float u01(uint32_t r) { return (static_cast<float>(r >> 8) + 0.5f) * (1.0f / 16777216.0f); }
__device__ __forceinline__ float normal_from(uint32_t r0, uint32_t r1) {
  return sqrtf(-2.0f * logf(u01(r0))) * cospif(2.0f * u01(r1));
}

__device__ __forceinline__ float clamp01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }

// One pixel of the packed network input (ub_api.cu: pack_conv fold_at): (r, g, b, 1, r_lo, g_lo, b_lo, 0) in bf16, 16 bytes.
// Channels 4-6 carry what bf16 rounding dropped from r, g, b (same weights as channels 0-2): the layer sees the pixel to
// ~16 bits instead of 8.
__device__ __forceinline__ uint4 pack_pixel(float r, float g, float b) {
  const __nv_bfloat16 hr = __float2bfloat16(r), hg = __float2bfloat16(g), hb = __float2bfloat16(b);
  __nv_bfloat162 rg = __halves2bfloat162(hr, hg);
  __nv_bfloat162 b1 = __halves2bfloat162(hb, __float2bfloat16(1.0f));
  __nv_bfloat162 lo_rg = __floats2bfloat162_rn(r - __bfloat162float(hr), g - __bfloat162float(hg));
  __nv_bfloat162 lo_b0 = __floats2bfloat162_rn(b - __bfloat162float(hb), 0.0f);
  return make_uint4(*reinterpret_cast<uint32_t*>(&rg), *reinterpret_cast<uint32_t*>(&b1),
                    *reinterpret_cast<uint32_t*>(&lo_rg), *reinterpret_cast<uint32_t*>(&lo_b0));
}

// Raw pixel channel as the dataset would hand it over: fp32 in [0,1].
__device__ __forceinline__ float load_f(const void* img, int fmt, int b, int c, int y, int x, int H, int W) {
  if (fmt == UB_IMG_F32_NCHW)
    return reinterpret_cast<const float*>(img)[((static_cast<size_t>(b) * 3 + c) * H + y) * W + x];
  const uint8_t u = reinterpret_cast<const uint8_t*>(img)[((static_cast<size_t>(b) * H + y) * W + x) * 3 + c];
  return static_cast<float>(u) / 255.0f;
}
// (image*255).round().clamp(0,255).to(uint8)  — perturbations.py:34 etc.
__device__ __forceinline__ int load_u8(const void* img, int fmt, int b, int c, int y, int x, int H, int W) {
  if (fmt == UB_IMG_U8_NHWC)
    return reinterpret_cast<const uint8_t*>(img)[((static_cast<size_t>(b) * H + y) * W + x) * 3 + c];
  const float v = reinterpret_cast<const float*>(img)[((static_cast<size_t>(b) * 3 + c) * H + y) * W + x];
  return static_cast<int>(fminf(fmaxf(rintf(__fmul_rn(v, 255.0f)), 0.0f), 255.0f));
}

// One pixel (all 3 channels) of the perturbed network input, for every family except the blurs.
// `lut` is the 256-entry table of the LUT families: per-thread indexed, so callers stage it in shared memory
// (a data-dependent index into the constant bank would serialise the warp).
__device__ __forceinline__ void perturbed_pixel(const PerturbDev& P, const uint8_t* lut, const void* img, int fmt, int b,
                                                int y, int x, int H, int W, float (&out)[3]) {
  const size_t hw = static_cast<size_t>(H) * W;
  const size_t pix = static_cast<size_t>(y) * W + x;
  switch (P.kind) {
    default:
    case UB_P_NONE:
#pragma unroll
      for (int c = 0; c < 3; ++c) out[c] = load_f(img, fmt, b, c, y, x, H, W);
      break;
    case UB_P_U8_NOISE: {
      uint4 r0 = make_uint4(0, 0, 0, 0), r1 = r0;
      if (!P.aux) { r0 = rng4(P.seed, static_cast<uint32_t>(pix), b, 1); r1 = rng4(P.seed, static_cast<uint32_t>(pix), b, 2); }
      const uint32_t rr[6] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y};
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float nz = P.aux ? reinterpret_cast<const float*>(P.aux)[(static_cast<size_t>(b) * 3 + c) * hw + pix]
                               : __fmul_rn(P.fparam, normal_from(rr[2 * c], rr[2 * c + 1]));
        const float v = __fadd_rn(static_cast<float>(load_u8(img, fmt, b, c, y, x, H, W)), nz);
        out[c] = rintf(fminf(fmaxf(v, 0.0f), 255.0f)) / 255.0f;  // clamp then round, :45
      }
    } break;
    case UB_P_U8_LUT:
#pragma unroll
      for (int c = 0; c < 3; ++c) out[c] = static_cast<float>(lut[load_u8(img, fmt, b, c, y, x, H, W)]) / 255.0f;
      break;
    case UB_P_U8_OCCLUSION:
    case UB_P_F_OCCLUSION: {
      const int s = P.iparam;
      int y0, x0;
      if (P.aux) {
        y0 = reinterpret_cast<const int*>(P.aux)[2 * b];
        x0 = reinterpret_cast<const int*>(P.aux)[2 * b + 1];
      } else {
        const uint4 r = rng4(P.seed, 0u, b, 3);
        const bool draw = P.kind == UB_P_F_OCCLUSION ? true : (s > 0 && H - s > 0 && W - s > 0);
        y0 = draw ? (H > s ? static_cast<int>(r.x % static_cast<uint32_t>(H - s + 1)) : 0) : -1;
        x0 = draw ? (W > s ? static_cast<int>(r.y % static_cast<uint32_t>(W - s + 1)) : 0) : -1;
      }
      const bool hit = s > 0 && y0 >= 0 && x0 >= 0 && y >= y0 && y < y0 + s && x >= x0 && x < x0 + s;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float v = P.kind == UB_P_U8_OCCLUSION
                            ? static_cast<float>(load_u8(img, fmt, b, c, y, x, H, W)) / 255.0f
                            : load_f(img, fmt, b, c, y, x, H, W);
        out[c] = hit ? 0.0f : v;
      }
    } break;
    case UB_P_U8_SALT_PEPPER: {
      int code;
      if (P.aux) {
        code = reinterpret_cast<const uint8_t*>(P.aux)[static_cast<size_t>(b) * hw + pix];
      } else {
        // fparam2 = P(pixel hit at least once by the reference's n draws with replacement)
        const uint4 r = rng4(P.seed, static_cast<uint32_t>(pix), b, 4);
        code = (u01(r.x) < P.fparam2) ? ((r.y & 1u) ? 2 : 1) : 0;
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int u = code == 1 ? 0 : (code == 2 ? 255 : load_u8(img, fmt, b, c, y, x, H, W));
        out[c] = static_cast<float>(u) / 255.0f;
      }
    } break;
    case UB_P_F_NOISE: {
      uint4 r0 = make_uint4(0, 0, 0, 0), r1 = r0;
      if (!P.aux) { r0 = rng4(P.seed, static_cast<uint32_t>(pix), b, 5); r1 = rng4(P.seed, static_cast<uint32_t>(pix), b, 6); }
      const uint32_t rr[6] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y};
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float g = P.aux ? reinterpret_cast<const float*>(P.aux)[(static_cast<size_t>(b) * 3 + c) * hw + pix]
                              : normal_from(rr[2 * c], rr[2 * c + 1]);
        // randn * fl32(std/255.0), then img + noise (separate roundings, :472-473)
        out[c] = clamp01(__fadd_rn(load_f(img, fmt, b, c, y, x, H, W), __fmul_rn(g, P.fparam2)));
      }
    } break;
    case UB_P_F_CONTRAST:
#pragma unroll
      for (int c = 0; c < 3; ++c) out[c] = clamp01(__fmul_rn(load_f(img, fmt, b, c, y, x, H, W), P.fparam));
      break;
    case UB_P_F_BRIGHTNESS:
#pragma unroll
      for (int c = 0; c < 3; ++c) out[c] = clamp01(__fadd_rn(load_f(img, fmt, b, c, y, x, H, W), P.fparam));
      break;
    case UB_P_F_SALT_PEPPER: {
      // fparam = fl32(amount/2), fparam2 = fl32(1-amount/2); one draw per pixel, shared by channels
      const float r = P.aux ? reinterpret_cast<const float*>(P.aux)[static_cast<size_t>(b) * hw + pix]
                            : u01(rng4(P.seed, static_cast<uint32_t>(pix), b, 7).x);
      const float salt = r < P.fparam ? 1.0f : 0.0f;
      const float pepper = r > P.fparam2 ? 1.0f : 0.0f;
      const float mask = __fsub_rn(__fsub_rn(1.0f, salt), pepper);
#pragma unroll
      for (int c = 0; c < 3; ++c)
        out[c] = __fadd_rn(__fmul_rn(load_f(img, fmt, b, c, y, x, H, W), mask), salt);
    } break;
  }
}

// Perturbation alone -> fp32 NCHW (ub_perturb_images; also feeds perturbed_out).
__global__ void perturb_kernel(PerturbDev P, const void* img, int fmt, int n, int H, int W, float* out) {
  __shared__ uint8_t s_lut[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) s_lut[i] = P.lut[i];
  __syncthreads();
  const size_t hw = static_cast<size_t>(H) * W;
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= hw * n) return;
  const int b = static_cast<int>(i / hw);
  const int y = static_cast<int>((i % hw) / W), x = static_cast<int>(i % W);
  float v[3];
  perturbed_pixel(P, s_lut, img, fmt, b, y, x, H, W, v);
#pragma unroll
  for (int c = 0; c < 3; ++c) out[(static_cast<size_t>(b) * 3 + c) * hw + static_cast<size_t>(y) * W + x] = v[c];
}

// ---- blur passes
// uint8 family (perturbations.py:97-121): zero-padded 3x3 sum S -> floor((S+4)/9); planar u8 [n*3,H,W].
__global__ void blur_u8_quantise_kernel(const void* img, int fmt, int n, int H, int W, uint8_t* out) {
  const size_t hw = static_cast<size_t>(H) * W;
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= hw * n * 3) return;
  const int pl = static_cast<int>(i / hw);
  const int y = static_cast<int>((i % hw) / W), x = static_cast<int>(i % W);
  out[i] = static_cast<uint8_t>(load_u8(img, fmt, pl / 3, pl % 3, y, x, H, W));
}
__global__ void blur_u8_pass_kernel(const uint8_t* in, uint8_t* out, int planes, int H, int W) {
  const size_t hw = static_cast<size_t>(H) * W;
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= hw * planes) return;
  const uint8_t* pl = in + (i / hw) * hw;
  const int y = static_cast<int>((i % hw) / W), x = static_cast<int>(i % W);
  int s = 0;
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      const int yy = y + dy, xx = x + dx;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) s += pl[static_cast<size_t>(yy) * W + xx];
    }
  out[i] = static_cast<uint8_t>((s + 4) / 9);
}
__global__ void u8_planar_to_f32_kernel(const uint8_t* in, float* out, size_t count) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < count) out[i] = static_cast<float>(in[i]) / 255.0f;
}
// float family (processing_blocks.py:493-496 -> kornia box_blur, reflect border; fp32 planar).
__global__ void blur_f_load_kernel(const void* img, int fmt, int n, int H, int W, float* out) {
  const size_t hw = static_cast<size_t>(H) * W;
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= hw * n * 3) return;
  const int pl = static_cast<int>(i / hw);
  out[i] = load_f(img, fmt, pl / 3, pl % 3, static_cast<int>((i % hw) / W), static_cast<int>(i % W), H, W);
}
__global__ void blur_f_pass_kernel(const float* in, float* out, int planes, int H, int W) {
  const size_t hw = static_cast<size_t>(H) * W;
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= hw * planes) return;
  const float* pl = in + (i / hw) * hw;
  const int y = static_cast<int>((i % hw) / W), x = static_cast<int>(i % W);
  const float ninth = 1.0f / 9.0f;
  float acc = 0.0f;
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      int yy = y + dy, xx = x + dx;
      yy = yy < 0 ? -yy : (yy >= H ? 2 * H - 2 - yy : yy);
      xx = xx < 0 ? -xx : (xx >= W ? 2 * W - 2 - xx : xx);
      acc = __fadd_rn(acc, __fmul_rn(pl[static_cast<size_t>(yy) * W + xx], ninth));
    }
  out[i] = acc;
}

// ---- all blur passes of one tile in shared memory (both families, passes <= kBlurMaxPasses; more passes fall back to
// the pass-per-launch kernels above).  One CTA = one 32x64 tile of one IMAGE (its three channels one after the other,
// through the same two ping-pong regions) plus a halo of `passes` pixels; pass p recomputes the region shrunk by p, so
// after the last pass the tile itself is exact.  The image is read once (straight from its container: the quantise / load
// step is the tile load) and the result written once — as the PACKED network input (16 bytes per pixel into the level-0
// concat buffer, what image_pack_kernel writes for the other families: the blur levels need no fp32 staging image and
// no second launch) and/or as fp32 NCHW (ub_perturb_images, perturbed_out, the materialised-stem path).
//   uint8 family: positions outside the image hold 0 in every pass (conv2d zero padding, perturbations.py:108-115).
//   float family: neighbours outside the image are mirrored (kornia reflect border) when read; same summation order as
//   blur_f_pass_kernel.
constexpr int kBlurTH = 32, kBlurTW = 64, kBlurMaxPasses = 12;
__host__ __device__ constexpr int blur_region(int passes) { return (kBlurTH + 2 * passes) * (kBlurTW + 2 * passes); }
__host__ __device__ constexpr size_t blur_smem_bytes(int passes, bool is_float) {
  return (2 * static_cast<size_t>(blur_region(passes)) + 3 * kBlurTH * kBlurTW) * (is_float ? 4 : 1);
}
template <bool FLOAT>
__global__ void __launch_bounds__(256)
blur_fused_kernel(const void* __restrict__ img, int fmt, int H, int W, int passes, float* __restrict__ out,
                  __nv_bfloat16* __restrict__ packed, int cgs, int cg0) {
  using T = typename std::conditional<FLOAT, float, uint8_t>::type;
  extern __shared__ __align__(16) uint8_t blur_smem[];
  const int region = blur_region(passes);
  T* s_buf[2] = {reinterpret_cast<T*>(blur_smem), reinterpret_cast<T*>(blur_smem) + region};
  T* s_fin = reinterpret_cast<T*>(blur_smem) + 2 * region;  // [3][kBlurTH * kBlurTW]: the finished tile of each channel
  const int b = blockIdx.z;
  const int y0 = blockIdx.y * kBlurTH - passes, x0 = blockIdx.x * kBlurTW - passes;  // image coordinates of region (0,0)
  const int RH = kBlurTH + 2 * passes, RW = kBlurTW + 2 * passes;
  for (int c = 0; c < 3; ++c) {
    for (int i = threadIdx.x; i < RH * RW; i += blockDim.x) {
      const int y = y0 + i / RW, x = x0 + i % RW;
      const bool in = y >= 0 && y < H && x >= 0 && x < W;
      if constexpr (FLOAT) s_buf[0][i] = in ? load_f(img, fmt, b, c, y, x, H, W) : 0.0f;
      else s_buf[0][i] = in ? static_cast<uint8_t>(load_u8(img, fmt, b, c, y, x, H, W)) : static_cast<uint8_t>(0);
    }
    __syncthreads();
    int cur = 0;
    for (int p = 1; p <= passes; ++p) {
      const T* src = s_buf[cur];
      T* dst = s_buf[cur ^ 1];
      const int h = RH - 2 * p, w = RW - 2 * p;  // region still exact after this pass
      if constexpr (FLOAT) {
        for (int i = threadIdx.x; i < h * w; i += blockDim.x) {
          const int ry = p + i / w, rx = p + i % w;
          const int y = y0 + ry, x = x0 + rx;
          float acc = 0.0f;
          if (y >= 0 && y < H && x >= 0 && x < W) {
            const float ninth = 1.0f / 9.0f;
#pragma unroll
            for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
              for (int dx = -1; dx <= 1; ++dx) {
                int yy = y + dy, xx = x + dx;
                yy = yy < 0 ? -yy : (yy >= H ? 2 * H - 2 - yy : yy);
                xx = xx < 0 ? -xx : (xx >= W ? 2 * W - 2 - xx : xx);
                acc = __fadd_rn(acc, __fmul_rn(src[(yy - y0) * RW + (xx - x0)], ninth));
              }
          }
          dst[ry * RW + rx] = acc;
        }
      } else {
        // integer sums are order-free: a thread walks down one column of a row segment keeping the horizontal 3-sums of
        // the last two rows (3 shared-memory loads and one store per output instead of 9 loads)
        const int segs = static_cast<int>(blockDim.x) / w;  // 64 <= w <= 86: 2..4 row segments per column
        const int seg_rows = (h + segs - 1) / segs;
        const int col = static_cast<int>(threadIdx.x) % w, seg = static_cast<int>(threadIdx.x) / w;
        const int rx = p + col, r_begin = p + seg * seg_rows, r_end = min(p + h, r_begin + seg_rows);
        if (seg < segs && r_begin < r_end) {
          const int x = x0 + rx;
          const bool x_in = x >= 0 && x < W;
          auto hsum = [&](int ry) {
            const uint8_t* r = src + ry * RW + rx;
            return static_cast<int>(r[-1]) + static_cast<int>(r[0]) + static_cast<int>(r[1]);
          };
          int ha = hsum(r_begin - 1), hb = hsum(r_begin);
          for (int ry = r_begin; ry < r_end; ++ry) {
            const int hc = hsum(ry + 1);
            const int y = y0 + ry;
            const bool in = x_in && y >= 0 && y < H;
            dst[ry * RW + rx] = in ? static_cast<uint8_t>((ha + hb + hc + 4) / 9) : static_cast<uint8_t>(0);
            ha = hb;
            hb = hc;
          }
        }
      }
      __syncthreads();
      cur ^= 1;
    }
    const T* fin = s_buf[cur];
    for (int i = threadIdx.x; i < kBlurTH * kBlurTW; i += blockDim.x)
      s_fin[c * (kBlurTH * kBlurTW) + i] = fin[(i / kBlurTW + passes) * RW + i % kBlurTW + passes];
    __syncthreads();  // the regions are free for the next channel; s_fin[c] is complete
  }
  const size_t hw = static_cast<size_t>(H) * W;
  for (int i = threadIdx.x; i < kBlurTH * kBlurTW; i += blockDim.x) {
    const int y = blockIdx.y * kBlurTH + i / kBlurTW, x = blockIdx.x * kBlurTW + i % kBlurTW;
    if (y >= H || x >= W) continue;
    float v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const T t = s_fin[c * (kBlurTH * kBlurTW) + i];
      if constexpr (FLOAT) v[c] = t; else v[c] = static_cast<float>(t) / 255.0f;
    }
    const size_t pix = static_cast<size_t>(y) * W + x;
    if (out) {
#pragma unroll
      for (int c = 0; c < 3; ++c) out[(static_cast<size_t>(b) * 3 + c) * hw + pix] = v[c];
    }
    if (packed)  // consecutive threads = consecutive pixels of one row: a warp stores 512 contiguous bytes
      reinterpret_cast<uint4*>(packed)[(static_cast<size_t>(b) * cgs + cg0) * hw + pix] = pack_pixel(v[0], v[1], v[2]);
  }
}

// ---- stem: perturbed pixel -> Conv2d(3,32,1) -> bf16, written into channel groups [cg0, cg0+4) of the
// level-0 concat buffer (channel-blocked [n][cgs][H][W][8]).  One thread per pixel, 4 x 16-byte stores; a warp
// stores 512 contiguous bytes per instruction.
struct StemWeights {
  float w[32 * 3];
  float b[32];
};
constexpr int kStemPx = 4;  // pixels per thread: four independent load->FMA->store chains, 64 contiguous bytes per plane
__global__ void __launch_bounds__(256, 3)
stem_kernel(const __grid_constant__ PerturbDev P, const __grid_constant__ StemWeights sw, const void* __restrict__ img,
            int fmt, int n, int H, int W, __nv_bfloat16* __restrict__ out, int cgs, int cg0,
            float* __restrict__ perturbed_out) {
  __shared__ uint8_t s_lut[256];
  if (P.kind == UB_P_U8_LUT) {
    s_lut[threadIdx.x] = P.lut[threadIdx.x];  // blockDim.x == 256
    __syncthreads();
  }
  const size_t hw = static_cast<size_t>(H) * W;
  const size_t total = hw * n;
  static_assert(kStemPx == 4, "float4 / paired 32-byte stores below");
  const size_t i0 = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * kStemPx;
  if (i0 >= total) return;
  // W is a multiple of 8 (build_plan checks): the four pixels share image and row, and i0 + 3 < total
  const int b = static_cast<int>(i0 / hw);
  const size_t pix = i0 - static_cast<size_t>(b) * hw;
  const int y = static_cast<int>(pix / W), x = static_cast<int>(pix % W);
  float v[kStemPx][3];
#pragma unroll
  for (int k = 0; k < kStemPx; ++k) perturbed_pixel(P, s_lut, img, fmt, b, y, x + k, H, W, v[k]);
  if (perturbed_out) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      *reinterpret_cast<float4*>(perturbed_out + (static_cast<size_t>(b) * 3 + c) * hw + pix) =
          make_float4(v[0][c], v[1][c], v[2][c], v[3][c]);
  }
  // the 32x3 weights and 32 biases are kernel parameters: FFMA reads them straight from the constant bank.
  // One channel group (8 outputs = 16 B per pixel) at a time, so only 16 packed registers are live.
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint32_t packed[kStemPx][4];
#pragma unroll
    for (int k = 0; k < kStemPx; ++k) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int o = 8 * g + 2 * j;
        float a0 = sw.b[o], a1 = sw.b[o + 1];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          a0 = fmaf(sw.w[o * 3 + c], v[k][c], a0);
          a1 = fmaf(sw.w[(o + 1) * 3 + c], v[k][c], a1);
        }
        __nv_bfloat162 h = __floats2bfloat162_rn(a0, a1);
        packed[k][j] = *reinterpret_cast<uint32_t*>(&h);
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(out) + (static_cast<size_t>(b) * cgs + cg0 + g) * hw + pix;
#pragma unroll
    for (int k = 0; k < kStemPx; k += 2)
      st_global_256(dst + k, make_uint4(packed[k][0], packed[k][1], packed[k][2], packed[k][3]),
                    make_uint4(packed[k + 1][0], packed[k + 1][1], packed[k + 1][2], packed[k + 1][3]));
  }
}

// ---- stem folded into its readers (ub_api.cu: pack_conv fold_at): the perturbed pixel itself is the layer input.
// Writes channel group cg0 of the level-0 concat buffer with (r, g, b, 1, r_lo, g_lo, b_lo, 0) in bf16 — 16 bytes per pixel; the
// following group stays zero from the arena's memset (the layers read 16 channels).  Four pixels per thread, 64
// contiguous bytes per thread.
__global__ void __launch_bounds__(256)
image_pack_kernel(const __grid_constant__ PerturbDev P, const void* __restrict__ img, int fmt, int n, int H, int W,
                  __nv_bfloat16* __restrict__ out, int cgs, int cg0, float* __restrict__ perturbed_out) {
  __shared__ uint8_t s_lut[256];
  __shared__ float s_flut[256];
  const bool fast = fmt == UB_IMG_U8_NHWC && (P.kind == UB_P_NONE || P.kind == UB_P_U8_LUT);
  if (P.kind == UB_P_U8_LUT || fast) {
    const uint8_t u = P.kind == UB_P_U8_LUT ? P.lut[threadIdx.x] : static_cast<uint8_t>(threadIdx.x);  // blockDim.x == 256
    s_lut[threadIdx.x] = u;
    s_flut[threadIdx.x] = static_cast<float>(u) / 255.0f;
    __syncthreads();
  }
  const size_t hw = static_cast<size_t>(H) * W;
  const size_t total = hw * n;
  const size_t i0 = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * kStemPx;
  if (i0 >= total) return;
  const int b = static_cast<int>(i0 / hw);
  const size_t pix = i0 - static_cast<size_t>(b) * hw;
  const int y = static_cast<int>(pix / W), x = static_cast<int>(pix % W);
  float v[kStemPx][3];
  if (fast) {
    // uint8 container, value-wise family: the thread's 4 pixels are 12 contiguous bytes (three aligned words), and
    // lut[u] / 255 comes from one table (same IEEE division as perturbed_pixel, done once per table entry)
    const uint32_t* w = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(img) + i0 * 3);
    const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    const uint32_t bytes[12] = {w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u, w0 >> 24, w1 & 255u, (w1 >> 8) & 255u,
                                (w1 >> 16) & 255u, w1 >> 24, w2 & 255u, (w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24};
#pragma unroll
    for (int k = 0; k < kStemPx; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) v[k][c] = s_flut[bytes[3 * k + c]];
  } else {
#pragma unroll
    for (int k = 0; k < kStemPx; ++k) perturbed_pixel(P, s_lut, img, fmt, b, y, x + k, H, W, v[k]);
  }
  if (perturbed_out) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      *reinterpret_cast<float4*>(perturbed_out + (static_cast<size_t>(b) * 3 + c) * hw + pix) =
          make_float4(v[0][c], v[1][c], v[2][c], v[3][c]);
  }
  uint4* dst = reinterpret_cast<uint4*>(out) + (static_cast<size_t>(b) * cgs + cg0) * hw + pix;
  uint4 q[kStemPx];
#pragma unroll
  for (int k = 0; k < kStemPx; ++k) q[k] = pack_pixel(v[k][0], v[k][1], v[k][2]);
  st_global_256(dst, q[0], q[1]);
  st_global_256(dst + 2, q[2], q[3]);
}

}  // namespace ub
