// ub_api.cu — host side of libunetb200.so: context, weight folding/packing, the per-resolution
// execution plan (buffers + TMA descriptors) and the C ABI declared in include/unet_b200.h.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <vector>

#include "../../include/unet_b200.h"
#include "ub_aux.cuh"
#include "ub_conv_tc.cuh"
#include "ub_conv_rs.cuh"
#include "ub_conv_rs2.cuh"
#include "ub_perturb.cuh"
#include "ub_vit.cuh"

namespace {

using ub::ConvParams;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

std::string g_create_error;

struct HostTensor {
  std::vector<float> data;
  std::vector<int64_t> shape;
};

// One tensor-core GEMM layer (conv3x3+BN+ReLU, or ConvTranspose2d k2 s2).
struct GemmLayer {
  std::string key;  // state_dict prefix of the conv / up module
  int cin = 0, cout = 0, n_total = 0, ksize = 3, block_n = 0, bk = 64, relu = 1;
  int act = 0;      // ub::EpiAct of the 1x1 STORE epilogue (the ViT's fc1: QuickGELU)
  bool no_pair = false;  // never CTA pairs (the ViT's token GEMMs: one to four waves of tiles, single CTAs measured 9 % faster)
  int cin_src = 0;  // channels that exist in the source tensor when fewer than cin (the rest of the K range meets zero weights;
                    // the tensor map ends there, so TMA zero-fills those planes of the box without reading memory)
  bool is_convt = false;
  __nv_bfloat16* w = nullptr;  // [n_total][ksize*ksize*cin] bf16, BN folded
  __nv_bfloat16* w_kxs = nullptr;  // thin 3x3 layers: [(ky*3+kx)*cout + co][cin] for the kx-stacked kernel
  __nv_bfloat16* w_rs = nullptr;   // thin 3x3 layers: [(kx*3+(2-ky))*cout + co][cin] for the row-streaming kernel (ub_conv_rs.cuh)
  __nv_bfloat16* w_blend = nullptr;  // ConvTranspose: [4*co + 2a + b][cin], the four parities of a channel adjacent (EPI_CONVT_BLEND)
  float* bias = nullptr;       // [cout]
  std::vector<float> h_bias;   // host copy (the HEAD layer passes its 32 biases by value)
};

struct Step {
  enum Kind { GEMM, RESAMPLE, CLIP_FUSE, PROMPT_STEM, AE_OUT } kind = GEMM;  // CLIP_FUSE / PROMPT_STEM / AE_OUT reuse the RESAMPLE fields
  int layer = -1;
  ConvParams cp;   // GEMM: tensor maps + static params (B patched per launch)
  // RESAMPLE
  const __nv_bfloat16* rs_in = nullptr;
  __nv_bfloat16* rs_out = nullptr;
  int hi = 0, wi = 0, ho = 0, wo = 0, c = 0, out_cgs = 0, out_cg0 = 0;
};

struct VitState;  // the CLIP ViT image tower (ub_vit_host.inl)

struct Plan {
  int H = 0, W = 0;
  void* arena = nullptr;
  size_t arena_bytes = 0;
  __nv_bfloat16* stem_out = nullptr;  // level-0 concat buffer, channel-blocked
  int stem_cgs = 0, stem_cg0 = 0;
  float* pert_f32 = nullptr;   // [MB,3,H,W] perturbed image (blur families / perturbed_out staging)
  void* blur_a = nullptr;      // ping-pong planes (fp32-sized)
  void* blur_b = nullptr;
  std::vector<Step> steps;
  // the encoder's feature maps as the halves' entry points see them (ub_encoder_forward / ub_skip_decoder_forward):
  // buffer, planes per image, first plane, channels, level (resolution H >> level)
  struct Feat { __nv_bfloat16* buf; int cgs, cg0, c, level; };
  std::vector<Feat> feats;
};

}  // namespace

struct ub_ctx {
  int device = 0, arch = 0, in_ch = 3, n_classes = 3, max_mb = 16, num_sms = 148, n_enc = 3;
  std::string err;
  std::map<std::string, HostTensor> sd;
  std::vector<GemmLayer> layers;
  std::map<std::string, int> layer_by_key;
  ub::StemWeights stem;
  float* head_w = nullptr;  // device [C][32]
  // UB_ARCH_CLIP_UNET: cross_attention_fusion.cross_attn value / output projections, transposed ([in][out]), fp32
  float *fuse_wv_t = nullptr, *fuse_bv = nullptr, *fuse_wo_t = nullptr, *fuse_bo = nullptr;
  float* head_b = nullptr;
  std::vector<float> h_head_w, h_head_b;  // host copies, passed by value to the HEAD epilogue
  bool weights_ready = false;
  std::map<std::pair<int, int>, std::unique_ptr<Plan>> plans;
  EncodeTiledFn encode = nullptr;
  int64_t launches = 0;
  // optional per-launch timing (ub_profile_begin/end): event pairs around every GEMM launch
  bool force_stream = false;  // UB_FORCE_STREAM=1: never use the weight-stationary pipeline (tests)
  bool use_pdl = true;        // UB_NO_PDL=1: plain stream-ordered launches
  bool no_pairq = false;      // UB_NO_PAIRQ=1: the big ConvTranspose runs as single CTAs with streamed weights
  int dbg = 0;                // UB_DBG: ablation hooks of conv_tc_kernel (1 = drain-only epilogue, 2 = no MMA, 8 = head always evaluates the softmax)
  bool no_kxs = false;        // UB_NO_KXS=1: thin layers run the standard (9-tap) kernel
  bool stem_fold = true;      // UB_NO_STEM_FOLD=1: the 32-channel stem output is materialised (stem_kernel) instead of being folded into its readers' weights
  bool no_blur_fuse = false;  // UB_NO_BLUR_FUSE=1: blur families run one launch per pass (the path for > 12 passes)
  bool head_no_kxs = false;   // UB_HEAD_NO_KXS=1: only the head layer (dec4.conv.3) runs the 9-tap kernel
  bool no_cg2 = false;        // UB_NO_CG2=1: wide layers run one CTA per tile (no cta_group::2 pairs)
  bool fuse_enc1 = true;      // enc1.conv.0 -> conv.3 -> pool as one kernel (ub_conv_rs2.cuh); UB_NO_FUSE_ENC1=1 / UB_FUSE_ENC1=0: two kernels
  bool fuse_dec4 = true;      // the last decoder stage's conv.0 -> conv.3 -> head as one kernel; UB_NO_FUSE_DEC4=1 / UB_FUSE_DEC4=0: two
  bool pair_resident = false;    // UB_PAIR_RESIDENT=1 (experiment): CTA pairs for enc2.conv.0 as well
  bool rs2_no_spare = false;     // UB_RS2_NO_SPARE=1 (diagnostic): the fused last decoder stage with circular TMEM windows (split MMAs at the wrap)
  bool stem_zero_plane = false;  // UB_STEM_ZERO_PLANE=1 (diagnostic): the stem's readers fetch the zero plane after the packed image
  bool fuse_resize = true;    // the first decoder stage's ConvTranspose + align_corners resize as one GEMM epilogue (EPI_CONVT_BLEND);
                              // UB_NO_FUSE_RESIZE=1: ConvTranspose to a 2H x 2W buffer, then resample_bilinear_kernel
  bool no_rs = true;          // UB_RS=1 selects the row-streaming kernel (ub_conv_rs.cuh) for the thin layers from 32 rows up; default:
                              // kx-stacked (the two measure the same under the power cap)

  bool profiling = false;
  struct ProfRec { int layer; cudaEvent_t a, b; };
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> event_pool;
  std::vector<void*> owned;  // device allocations freed in ub_destroy
  // kernel instantiations whose dynamic shared-memory opt-in has been made ON THIS CONTEXT'S DEVICE (the attribute
  // is per device and per function, so a second context on another GPU of the same process makes its own)
  std::set<const void*> smem_opt_in;
  VitState* vit = nullptr;  // UB_ARCH_CLIP_VIT
  bool has_clip() const { return arch == UB_ARCH_CLIP_UNET || arch == UB_ARCH_CLIP_UNET_PROMPT; }
  bool is_half() const { return arch == UB_ARCH_ENCODER || arch == UB_ARCH_SKIP_DECODER; }
};

namespace {

int fail(ub_ctx* ctx, int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf; else g_create_error = buf;
  return code;
}

#define UB_CUDA(ctx, call)                                                                      \
  do {                                                                                          \
    cudaError_t e__ = (call);                                                                   \
    if (e__ != cudaSuccess)                                                                     \
      return fail(ctx, UB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),   \
                  __FILE__, __LINE__);                                                          \
  } while (0)

int dev_alloc(ub_ctx* ctx, void** p, size_t bytes) {
  UB_CUDA(ctx, cudaMalloc(p, bytes));
  ctx->owned.push_back(*p);
  return UB_OK;
}

const HostTensor* find(ub_ctx* ctx, const std::string& k) {
  auto it = ctx->sd.find(k);
  return it == ctx->sd.end() ? nullptr : &it->second;
}

int pick_block_n(int n_total) { return n_total >= 256 ? 256 : n_total; }

// conv3x3 + BatchNorm(eval) folded:  w' = w*g/sqrt(var+eps),  b' = (b-mean)*g/sqrt(var+eps)+beta
// (processing_blocks.py:43-44).  Packed K order = (ky*3+kx)*Cin + ci, K contiguous per output channel.
// fold_at >= 0: the 32 input channels [fold_at, fold_at + 32) are the stem output Conv2d(3,32,1)(x) (models/UNet.py:62,
// no activation before the next conv).  The layer is linear in them, so they are replaced by 16 channels
// (r, g, b, 1, r_lo, g_lo, b_lo, 0 ... 0) of the packed image with composed weights  w'[co][c][tap] = sum_ci w[co][ci][tap] * S[ci][c]  and
// w'[co][3][tap] = sum_ci w[co][ci][tap] * b_S[ci]  (the constant-one channel is zero outside the image, exactly like the
// zero padding of the stem output); fold_width (16 or 32) input channels are reserved for them, the rest with zero weights.
// Stored under "<conv_key>+stem".
int pack_conv(ub_ctx* ctx, const std::string& conv_key, const std::string& bn_key, int fold_at = -1, int fold_width = 16) {
  const HostTensor* w = find(ctx, conv_key + ".weight");
  const HostTensor* b = find(ctx, conv_key + ".bias");
  const HostTensor* g = find(ctx, bn_key + ".weight");
  const HostTensor* be = find(ctx, bn_key + ".bias");
  const HostTensor* mu = find(ctx, bn_key + ".running_mean");
  const HostTensor* var = find(ctx, bn_key + ".running_var");
  if (!w || !b || !g || !be || !mu || !var)
    return fail(ctx, UB_ERR_INVALID, "missing state_dict entries for %s / %s", conv_key.c_str(), bn_key.c_str());
  if (w->shape.size() != 4 || w->shape[2] != 3 || w->shape[3] != 3)
    return fail(ctx, UB_ERR_INVALID, "%s.weight is not [Cout,Cin,3,3]", conv_key.c_str());
  GemmLayer L;
  L.key = fold_at >= 0 ? conv_key + "+stem" : conv_key;
  L.cout = static_cast<int>(w->shape[0]);
  const int cin_w = static_cast<int>(w->shape[1]);
  if (fold_at >= 0 && (fold_at % 8 != 0 || fold_at + 32 != cin_w))
    return fail(ctx, UB_ERR_UNSUPPORTED, "%s: the stem channels must be the last 32 inputs", conv_key.c_str());
  // a layer with fewer than 32 input channels (the PromptEncoder's first conv, Cin = 1) reads a zero-padded
  // 32-channel operand: the extra weight columns are zero
  L.cin = fold_at >= 0 ? fold_at + fold_width : (cin_w < 32 ? 32 : cin_w);
  // the packed image is ONE channel group (r, g, b, 1, r_lo, g_lo, b_lo, 0); the other 8 (or 24) reserved input channels exist
  // only as zero weights and zero-filled operand planes, never as memory reads (UB_STEM_ZERO_PLANE=1: the all-zero second plane
  // of the buffer is read, as before)
  if (fold_at >= 0) L.cin_src = fold_at + (ctx->stem_zero_plane ? 16 : 8);
  L.n_total = L.cout;
  L.ksize = 3;
  L.block_n = pick_block_n(L.n_total);
  L.bk = (L.cin % 64 == 0) ? 64 : (L.cin % 32 == 0 ? 32 : 16);
  if (L.cin % 16 != 0 || L.cout % 32 != 0)
    return fail(ctx, UB_ERR_UNSUPPORTED, "%s: channel counts must be multiples of 32", conv_key.c_str());
  const size_t K = static_cast<size_t>(9) * L.cin;
  std::vector<__nv_bfloat16> packed(static_cast<size_t>(L.cout) * K, __float2bfloat16(0.0f));
  std::vector<float> bias(L.cout);
  for (int co = 0; co < L.cout; ++co) {
    const double s = static_cast<double>(g->data[co]) / std::sqrt(static_cast<double>(var->data[co]) + 1e-5);
    bias[co] = static_cast<float>((static_cast<double>(b->data[co]) - mu->data[co]) * s + be->data[co]);
    const int n_direct = fold_at >= 0 ? fold_at : cin_w;
    for (int ci = 0; ci < n_direct; ++ci)
      for (int t = 0; t < 9; ++t)
        packed[co * K + static_cast<size_t>(t) * L.cin + ci] =
            __float2bfloat16(static_cast<float>(w->data[(static_cast<size_t>(co) * cin_w + ci) * 9 + t] * s));
    if (fold_at >= 0) {
      for (int t = 0; t < 9; ++t) {
        double acc[4] = {0.0, 0.0, 0.0, 0.0};  // r, g, b, one
        for (int ci = 0; ci < 32; ++ci) {
          const double wv = w->data[(static_cast<size_t>(co) * cin_w + fold_at + ci) * 9 + t];
          for (int c = 0; c < 3; ++c) acc[c] += wv * ctx->stem.w[ci * 3 + c];
          acc[3] += wv * ctx->stem.b[ci];
        }
        for (int c = 0; c < 4; ++c)
          packed[co * K + static_cast<size_t>(t) * L.cin + fold_at + c] = __float2bfloat16(static_cast<float>(acc[c] * s));
        for (int c = 0; c < 3; ++c)  // channels 4-6: the bf16 rounding residue of r, g, b (image_pack_kernel), same weights
          packed[co * K + static_cast<size_t>(t) * L.cin + fold_at + 4 + c] = __float2bfloat16(static_cast<float>(acc[c] * s));
      }
    }
  }
  int rc;
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.w), packed.size() * 2))) return rc;
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.bias), bias.size() * 4))) return rc;
  UB_CUDA(ctx, cudaMemcpy(L.w, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
  UB_CUDA(ctx, cudaMemcpy(L.bias, bias.data(), bias.size() * 4, cudaMemcpyHostToDevice));
  L.h_bias = bias;
  if (L.cout <= 64) {
    // kx-stacked operand (ub_conv_tc.cuh): GEMM column n3 = kx*cout + co of the ky-th box, K = cin
    std::vector<__nv_bfloat16> ks(packed.size());
    for (int t = 0; t < 9; ++t)
      for (int co = 0; co < L.cout; ++co)
        for (int ci = 0; ci < L.cin; ++ci)
          ks[(static_cast<size_t>(t) * L.cout + co) * L.cin + ci] = packed[co * K + static_cast<size_t>(t) * L.cin + ci];
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.w_kxs), ks.size() * 2))) return rc;
    UB_CUDA(ctx, cudaMemcpy(L.w_kxs, ks.data(), ks.size() * 2, cudaMemcpyHostToDevice));
    // row-streaming operand: for each kx one box of 3*cout rows ordered (ky = 2, 1, 0) = the rows y'-1, y', y'+1 an
    // input row y' contributes to
    for (int ky = 0; ky < 3; ++ky)
      for (int kx = 0; kx < 3; ++kx)
        for (int co = 0; co < L.cout; ++co)
          for (int ci = 0; ci < L.cin; ++ci)
            ks[(static_cast<size_t>(kx * 3 + (2 - ky)) * L.cout + co) * L.cin + ci] =
                packed[co * K + static_cast<size_t>(ky * 3 + kx) * L.cin + ci];
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.w_rs), ks.size() * 2))) return rc;
    UB_CUDA(ctx, cudaMemcpy(L.w_rs, ks.data(), ks.size() * 2, cudaMemcpyHostToDevice));
  }
  ctx->layer_by_key[L.key] = static_cast<int>(ctx->layers.size());
  ctx->layers.push_back(L);
  return UB_OK;
}

// ConvTranspose2d(k2,s2) weight [Cin,Cout,2,2] (processing_blocks.py:102) as GEMM B operand
// W2[(a*2+b)*Cout + co][ci]: out[2i+a,2j+b,co] = sum_ci x[i,j,ci]*W[ci,co,a,b] + bias[co].
int pack_convt(ub_ctx* ctx, const std::string& key) {
  const HostTensor* w = find(ctx, key + ".weight");
  const HostTensor* b = find(ctx, key + ".bias");
  if (!w || !b) return fail(ctx, UB_ERR_INVALID, "missing state_dict entries for %s", key.c_str());
  if (w->shape.size() != 4 || w->shape[2] != 2 || w->shape[3] != 2)
    return fail(ctx, UB_ERR_INVALID, "%s.weight is not [Cin,Cout,2,2]", key.c_str());
  GemmLayer L;
  L.key = key;
  L.is_convt = true;
  L.cin = static_cast<int>(w->shape[0]);
  L.cout = static_cast<int>(w->shape[1]);
  L.n_total = 4 * L.cout;
  L.ksize = 1;
  L.relu = 0;
  L.block_n = pick_block_n(L.n_total);
  L.bk = (L.cin % 64 == 0) ? 64 : 32;
  if (L.cin % 32 != 0 || L.cout % 32 != 0)
    return fail(ctx, UB_ERR_UNSUPPORTED, "%s: channel counts must be multiples of 32", key.c_str());
  std::vector<__nv_bfloat16> packed(static_cast<size_t>(L.n_total) * L.cin);
  for (int ci = 0; ci < L.cin; ++ci)
    for (int co = 0; co < L.cout; ++co)
      for (int q = 0; q < 4; ++q)
        packed[(static_cast<size_t>(q) * L.cout + co) * L.cin + ci] =
            __float2bfloat16(w->data[(static_cast<size_t>(ci) * L.cout + co) * 4 + q]);
  int rc;
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.w), packed.size() * 2))) return rc;
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.bias), b->data.size() * 4))) return rc;
  UB_CUDA(ctx, cudaMemcpy(L.w, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
  UB_CUDA(ctx, cudaMemcpy(L.bias, b->data.data(), b->data.size() * 4, cudaMemcpyHostToDevice));
  // second order for the epilogue that blends the four parity products of a pixel (ConvTranspose + resize as one step)
  for (int ci = 0; ci < L.cin; ++ci)
    for (int co = 0; co < L.cout; ++co)
      for (int q = 0; q < 4; ++q)
        packed[(static_cast<size_t>(co) * 4 + q) * L.cin + ci] = __float2bfloat16(w->data[(static_cast<size_t>(ci) * L.cout + co) * 4 + q]);
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.w_blend), packed.size() * 2))) return rc;
  UB_CUDA(ctx, cudaMemcpy(L.w_blend, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
  ctx->layer_by_key[L.key] = static_cast<int>(ctx->layers.size());
  ctx->layers.push_back(L);
  return UB_OK;
}

// Plain Conv2d(k=1) with bias, no BatchNorm, no activation (prompt_segmentation.py:53 `prompt_fusion`): GEMM B operand
// W2[co][ci], stored by the STORE epilogue.
int pack_conv1x1(ub_ctx* ctx, const std::string& key) {
  const HostTensor* w = find(ctx, key + ".weight");
  const HostTensor* b = find(ctx, key + ".bias");
  if (!w || !b) return fail(ctx, UB_ERR_INVALID, "missing state_dict entries for %s", key.c_str());
  if (w->shape.size() != 4 || w->shape[2] != 1 || w->shape[3] != 1)
    return fail(ctx, UB_ERR_INVALID, "%s.weight is not [Cout,Cin,1,1]", key.c_str());
  GemmLayer L;
  L.key = key;
  L.cout = static_cast<int>(w->shape[0]);
  L.cin = static_cast<int>(w->shape[1]);
  L.n_total = L.cout;
  L.ksize = 1;
  L.relu = 0;
  L.block_n = pick_block_n(L.n_total);
  L.bk = (L.cin % 64 == 0) ? 64 : 32;
  if (L.cin % 32 != 0 || L.cout % 32 != 0)
    return fail(ctx, UB_ERR_UNSUPPORTED, "%s: channel counts must be multiples of 32", key.c_str());
  std::vector<__nv_bfloat16> packed(w->data.size());
  for (size_t i = 0; i < w->data.size(); ++i) packed[i] = __float2bfloat16(w->data[i]);
  int rc;
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.w), packed.size() * 2))) return rc;
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.bias), b->data.size() * 4))) return rc;
  UB_CUDA(ctx, cudaMemcpy(L.w, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
  UB_CUDA(ctx, cudaMemcpy(L.bias, b->data.data(), b->data.size() * 4, cudaMemcpyHostToDevice));
  L.h_bias = b->data;
  ctx->layer_by_key[L.key] = static_cast<int>(ctx->layers.size());
  ctx->layers.push_back(L);
  return UB_OK;
}

int pack_vit(ub_ctx* ctx);

int pack_all(ub_ctx* ctx) {
  int rc;
  if (ctx->arch == UB_ARCH_CLIP_VIT) return pack_vit(ctx);
  // the two halves of SegmentationModel hold one half of the weights each (models/autoencoder.py:5-26,
  // models/pre_trained.py:7-24); neither folds the stem (the decoder is handed the stem OUTPUT "x0")
  const bool enc_half = ctx->arch == UB_ARCH_ENCODER, dec_half = ctx->arch == UB_ARCH_SKIP_DECODER;
  if (enc_half || dec_half) ctx->stem_fold = false;
  const HostTensor* sw = find(ctx, "input.weight");
  const HostTensor* sb = find(ctx, "input.bias");
  const HostTensor* ow = find(ctx, "out.weight");
  const HostTensor* ob = find(ctx, "out.bias");
  if ((!dec_half && (!sw || !sb)) || (!enc_half && (!ow || !ob))) return fail(ctx, UB_ERR_INVALID, "missing input.* / out.* entries");
  const bool ae = ctx->arch == UB_ARCH_AUTOENCODER;  // Decoder.out is Conv2d(64, out_channels, 1) (autoencoder.py:34)
  if (!dec_half) {
    if (sw->data.size() != 32 * 3 || sb->data.size() != 32)
      return fail(ctx, UB_ERR_UNSUPPORTED, "stem must be Conv2d(3,32,1)");
    memcpy(ctx->stem.w, sw->data.data(), sizeof(ctx->stem.w));
    memcpy(ctx->stem.b, sb->data.data(), sizeof(ctx->stem.b));
  }
  if (!enc_half) {
    if (static_cast<int>(ow->shape[0]) != ctx->n_classes || ow->shape[1] != (ae ? 64 : 32))
      return fail(ctx, UB_ERR_INVALID, "out.weight must be [%d,%d,1,1]", ctx->n_classes, ae ? 64 : 32);
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&ctx->head_w), ow->data.size() * 4))) return rc;
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&ctx->head_b), ob->data.size() * 4))) return rc;
    UB_CUDA(ctx, cudaMemcpy(ctx->head_w, ow->data.data(), ow->data.size() * 4, cudaMemcpyHostToDevice));
    UB_CUDA(ctx, cudaMemcpy(ctx->head_b, ob->data.data(), ob->data.size() * 4, cudaMemcpyHostToDevice));
    ctx->h_head_w = ow->data;
    ctx->h_head_b = ob->data;
  }
  char k[128];
  for (int i = 1; i <= (dec_half ? 0 : ctx->n_enc); ++i) {
    snprintf(k, sizeof(k), "enc%d.block.0.conv", i);
    if ((rc = pack_conv(ctx, std::string(k) + ".0", std::string(k) + ".1"))) return rc;
    if ((rc = pack_conv(ctx, std::string(k) + ".3", std::string(k) + ".4"))) return rc;
  }
  if (ctx->arch == UB_ARCH_CLIP_UNET_PROMPT) {
    // PromptEncoder (prompt_segmentation.py:16-29): three ConvBlockDownsample stages and a ConvBlock, then prompt_fusion
    for (int i = 1; i <= 3; ++i) {
      snprintf(k, sizeof(k), "prompt_encoder.enc%d.block.0.conv", i);
      if ((rc = pack_conv(ctx, std::string(k) + ".0", std::string(k) + ".1"))) return rc;
      if ((rc = pack_conv(ctx, std::string(k) + ".3", std::string(k) + ".4"))) return rc;
    }
    if ((rc = pack_conv(ctx, "prompt_encoder.conv.conv.0", "prompt_encoder.conv.conv.1"))) return rc;
    if ((rc = pack_conv(ctx, "prompt_encoder.conv.conv.3", "prompt_encoder.conv.conv.4"))) return rc;
    if ((rc = pack_conv1x1(ctx, "prompt_fusion"))) return rc;
  }
  if (ctx->has_clip()) {
    // the bottleneck ConvBlock's output is dead in ClipUnet.forward (CLIP_models.py:124-126): its weights are not packed
    const int E = 2 * 32 * (1 << ctx->n_enc);  // bottleneck width = embed_dim of the fusion (512)
    const HostTensor* ipw = find(ctx, "cross_attention_fusion.cross_attn.in_proj_weight");
    const HostTensor* ipb = find(ctx, "cross_attention_fusion.cross_attn.in_proj_bias");
    const HostTensor* opw = find(ctx, "cross_attention_fusion.cross_attn.out_proj.weight");
    const HostTensor* opb = find(ctx, "cross_attention_fusion.cross_attn.out_proj.bias");
    if (!ipw || !ipb || !opw || !opb) return fail(ctx, UB_ERR_INVALID, "missing cross_attention_fusion.cross_attn.* entries");
    if (ipw->data.size() != static_cast<size_t>(3) * E * E || ipb->data.size() != static_cast<size_t>(3) * E ||
        opw->data.size() != static_cast<size_t>(E) * E || opb->data.size() != static_cast<size_t>(E))
      return fail(ctx, UB_ERR_UNSUPPORTED, "cross-attention fusion must have embed_dim %d (in_proj [3E,E])", E);
    // nn.MultiheadAttention packs q, k, v projections row-wise in in_proj_*: the value projection is rows [2E, 3E)
    std::vector<float> wv_t(static_cast<size_t>(E) * E), wo_t(static_cast<size_t>(E) * E);
    for (int o = 0; o < E; ++o)
      for (int i = 0; i < E; ++i) {
        wv_t[static_cast<size_t>(i) * E + o] = ipw->data[(static_cast<size_t>(2 * E + o)) * E + i];
        wo_t[static_cast<size_t>(i) * E + o] = opw->data[static_cast<size_t>(o) * E + i];
      }
    const size_t mb = static_cast<size_t>(E) * E * 4, vb = static_cast<size_t>(E) * 4;
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&ctx->fuse_wv_t), mb))) return rc;
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&ctx->fuse_wo_t), mb))) return rc;
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&ctx->fuse_bv), vb))) return rc;
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&ctx->fuse_bo), vb))) return rc;
    UB_CUDA(ctx, cudaMemcpy(ctx->fuse_wv_t, wv_t.data(), mb, cudaMemcpyHostToDevice));
    UB_CUDA(ctx, cudaMemcpy(ctx->fuse_wo_t, wo_t.data(), mb, cudaMemcpyHostToDevice));
    UB_CUDA(ctx, cudaMemcpy(ctx->fuse_bv, ipb->data.data() + 2 * E, vb, cudaMemcpyHostToDevice));
    UB_CUDA(ctx, cudaMemcpy(ctx->fuse_bo, opb->data.data(), vb, cudaMemcpyHostToDevice));
  } else if (!dec_half) {
    if ((rc = pack_conv(ctx, "bottleneck.conv.0", "bottleneck.conv.1"))) return rc;
    if ((rc = pack_conv(ctx, "bottleneck.conv.3", "bottleneck.conv.4"))) return rc;
  }
  if (enc_half) {
    UB_CUDA(ctx, cudaDeviceSynchronize());
    ctx->weights_ready = true;
    return UB_OK;
  }
  // the stem's two readers, once more with the stem folded into their weights (the plain versions stay for the
  // layer-wise debug entry and UB_NO_STEM_FOLD)
  if (!dec_half && (rc = pack_conv(ctx, "enc1.block.0.conv.0", "enc1.block.0.conv.1", 0))) return rc;
  if (!ae && !dec_half) {
    snprintf(k, sizeof(k), "dec%d.conv.conv", ctx->n_enc + 1);
    // (32 channels wide, the upper 16 with zero weights over planes nobody writes: 64 input channels keep the BK = 64
    // kx-stacked kernel — with 48 channels as three BK = 16 chunks the layer was slower, 22.2 against 20.7 ms)
    if ((rc = pack_conv(ctx, std::string(k) + ".0", std::string(k) + ".1", 32, 32))) return rc;
  }
  for (int j = 1; j <= ctx->n_enc + (ae ? 0 : 1); ++j) {
    snprintf(k, sizeof(k), "dec%d", j);
    if ((rc = pack_convt(ctx, std::string(k) + ".up"))) return rc;
    if ((rc = pack_conv(ctx, std::string(k) + ".conv.conv.0", std::string(k) + ".conv.conv.1"))) return rc;
    if ((rc = pack_conv(ctx, std::string(k) + ".conv.conv.3", std::string(k) + ".conv.conv.4"))) return rc;
  }
  // the uploads above are cudaMemcpy calls on the legacy default stream; the kernels run on the caller's (possibly
  // non-blocking) stream, so make the packed weights globally visible before the first launch can be issued
  UB_CUDA(ctx, cudaDeviceSynchronize());
  ctx->weights_ready = true;
  return UB_OK;
}

// ------------------------------------------------------------------ TMA descriptors
// Activations are channel-blocked: [B][cgs][H][W][8] bf16.  The map views `cin/8` consecutive planes starting at
// `base`: dims {8*W, H, cin/8, B}; one box = {8*(8+2*pad) elements, 16+2*pad rows, bk/8 planes} (ub_conv_tc.cuh).
int make_tm_act(ub_ctx* ctx, CUtensorMap* tm, const __nv_bfloat16* base, int cin, int cgs_total, int W, int H, int B,
                int ksize, int bk, int tw) {
  const int pady = ksize / 2, padx = ub::is_kxs(tw) ? 0 : pady;
  const cuuint64_t dims[4] = {static_cast<cuuint64_t>(8) * W, static_cast<cuuint64_t>(H), static_cast<cuuint64_t>(cin / 8),
                              static_cast<cuuint64_t>(B)};
  const cuuint64_t strides[3] = {static_cast<cuuint64_t>(W) * 16, static_cast<cuuint64_t>(H) * W * 16,
                                 static_cast<cuuint64_t>(cgs_total) * H * W * 16};
  const cuuint32_t box[4] = {static_cast<cuuint32_t>(8 * (tw + 2 * padx)), static_cast<cuuint32_t>(ub::tile_rows(tw) + 2 * pady),
                             static_cast<cuuint32_t>(bk / 8), 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = ctx->encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<__nv_bfloat16*>(base), dims,
                                 strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ctx, UB_ERR_CUDA, "cuTensorMapEncodeTiled(activation) failed: %d", static_cast<int>(r));
  return UB_OK;
}
int make_tm_w(ub_ctx* ctx, CUtensorMap* tm, const __nv_bfloat16* w, int k_total, int n_total, int block_n, int bk) {
  const cuuint64_t dims[2] = {static_cast<cuuint64_t>(k_total), static_cast<cuuint64_t>(n_total)};
  const cuuint64_t strides[1] = {static_cast<cuuint64_t>(k_total) * 2};
  const cuuint32_t box[2] = {static_cast<cuuint32_t>(bk), static_cast<cuuint32_t>(block_n)};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = ctx->encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(w), dims, strides,
                                 box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (bk == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B),
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ctx, UB_ERR_CUDA, "cuTensorMapEncodeTiled(weights) failed: %d", static_cast<int>(r));
  return UB_OK;
}

// Static part of a GEMM step: descriptors over (src view) and the layer's packed weights.  `src` points at the
// first channel-group plane the layer reads; `src_cgs` / `out_cgs` are the planes per image of the two buffers.
// Tile geometry of a layer: the thin 3x3 layers run kx-stacked (32-pixel-wide tiles, 16 when the pool is fused).
int pick_tw(const ub_ctx* ctx, const GemmLayer& L, int mode) {
  if (ctx->no_kxs || !L.w_kxs || L.ksize != 3 || mode == ub::EPI_CONVT || mode == ub::EPI_CONVT_BLEND) return 8;
  if (ctx->head_no_kxs && mode == ub::EPI_HEAD) return 8;
  if (L.bk == 32 && mode != ub::EPI_HEAD) return 8;  // Cin = 32: only the head layer gains (enc1.conv.0 is write-bound)
  if (L.bk == 16 && L.cout == 64) return 8;          // enc1.conv.0 with the stem folded in: 9 small MMAs, write-bound
  return mode == ub::EPI_POOL ? 16 : 32;
}

int fill_gemm(ub_ctx* ctx, ConvParams* cp, const GemmLayer& L, const __nv_bfloat16* src, int src_cgs,
              int B, int H, int W, int mode, __nv_bfloat16* out, int out_cgs, int out_cg0) {
  memset(cp, 0, sizeof(*cp));
  int rc;
  const bool convt = mode == ub::EPI_CONVT || mode == ub::EPI_CONVT_BLEND;
  if (convt != L.is_convt) return fail(ctx, UB_ERR_INVALID, "%s: epilogue mode %d does not fit the layer", L.key.c_str(), mode);
  const int tw = pick_tw(ctx, L, mode);
  cp->tw = tw;
  // the thin layers at the larger sizes: row-streaming kernel (a task is H/4 + 2 row tiles: worthwhile from 8 rows per
  // stream; below that, and with UB_NO_RS=1, the kx-stacked kernel)
  const bool rs = ub::is_kxs(tw) && !ctx->no_rs && L.w_rs && H % 8 == 0 && H >= 32 && W >= 32 && (L.cout == 32 || L.cout == 64) &&
                  ((L.cout == 32 && (L.bk == 32 || L.bk == 64)) || (L.cout == 64 && L.bk == 64));
  cp->rs = rs ? 1 : 0;
  cp->hq = H / 4;
  if (rs) {  // the row-streaming kernel reads the blocked tensor itself (cp.async); no activation tensor map
    cp->a_src = src;
    cp->a_cgs = src_cgs;
    cp->a_planes = (L.cin_src ? L.cin_src : L.cin) / 8;
  } else if ((rc = make_tm_act(ctx, &cp->tmA, src, L.cin_src ? L.cin_src : L.cin, src_cgs, W, H, B, L.ksize, L.bk, tw))) return rc;
  // CTA pairs (cta_group::2) for the wide layers whose weights cannot stay resident in shared memory
  const size_t w_tile_bytes = static_cast<size_t>(L.ksize) * L.ksize * L.cin * L.block_n * 2;
  const bool resident = w_tile_bytes + 2 * 24576 <= 220 * 1024;  // mirrors choose_pipeline's weight-stationary test
  cp->dbg = ctx->dbg;
  // (UB_PAIR_RESIDENT=1, experiment: also the one 3x3 layer whose whole n-tile fits a single CTA — enc2.conv.0, 64 -> 128 — runs as pairs)
  const bool pair_anyway = ctx->pair_resident && L.ksize == 3 && L.block_n == 128 && L.n_total == 128;
  cp->cg2 = (!ctx->no_cg2 && !L.no_pair && tw == 8 && L.bk == 64 && L.block_n >= 128 && mode != ub::EPI_HEAD && !convt &&
             (!resident || pair_anyway)) ? 1 : 0;
  // ConvTranspose whose n-tile does not fit one CTA but whose halves fit a pair (dec1.up: 512 -> 4 x 256): pairs, each
  // keeping its half of ONE n-tile resident, pair-granular tile order (ub_conv_tc.cuh) so the activations hit L2
  if (!ctx->no_cg2 && convt && L.ksize == 1 && L.bk == 64 && L.block_n == 256 && !resident &&
      w_tile_bytes / 2 + 2 * 16384 <= 200 * 1024 && !ctx->no_pairq) {
    cp->cg2 = 1;
    cp->pairq = 1;
  }
  if (rs) {
    if ((rc = make_tm_w(ctx, &cp->tmB, L.w_rs, L.cin, 9 * L.cout, 3 * L.cout, L.bk))) return rc;
  } else if (ub::is_kxs(tw)) {
    if ((rc = make_tm_w(ctx, &cp->tmB, L.w_kxs, L.cin, 9 * L.cout, 3 * L.cout, L.bk))) return rc;
  } else {
    const __nv_bfloat16* w = mode == ub::EPI_CONVT_BLEND ? L.w_blend : L.w;
    if (!w) return fail(ctx, UB_ERR_STATE, "%s: no weights packed for epilogue mode %d", L.key.c_str(), mode);
    if ((rc = make_tm_w(ctx, &cp->tmB, w, L.ksize * L.ksize * L.cin, L.n_total, cp->cg2 ? L.block_n / 2 : L.block_n, L.bk))) return rc;
  }
  // a single chunk whose tail is zero padding (the last decoder conv.0 with the stem folded in: 48 of 64 channels): the MMAs
  // over the padding are not issued
  cp->k_steps = (L.cin_src && L.cin == L.bk && L.cin_src < L.cin) ? (L.cin_src + 15) / 16 : 0;
  cp->B = B; cp->H = H; cp->W = W;
  cp->cin = L.cin; cp->ksize = L.ksize; cp->n_total = L.n_total;
  cp->mode = mode; cp->relu = L.relu; cp->act = L.act;
  cp->bias = L.bias; cp->n_bias = L.cout;
  cp->out = out; cp->out_cgs = out_cgs; cp->out_cg0 = out_cg0; cp->cout = L.cout;
  cp->n_classes = ctx->n_classes;
  cp->cout_shift = 0;
  while ((1 << cp->cout_shift) < L.cout) cp->cout_shift++;
  if (mode == ub::EPI_CONVT && (1 << cp->cout_shift) != L.cout)
    return fail(ctx, UB_ERR_UNSUPPORTED, "%s: ConvTranspose output channels must be a power of two", L.key.c_str());
  if (mode == ub::EPI_HEAD) {
    if (L.cout != 32 || ctx->h_head_w.size() != static_cast<size_t>(ctx->n_classes) * 32)
      return fail(ctx, UB_ERR_UNSUPPORTED, "head epilogue needs a 32-channel layer");
    memcpy(cp->head_w, ctx->h_head_w.data(), ctx->h_head_w.size() * 4);
    memcpy(cp->head_b, ctx->h_head_b.data(), ctx->h_head_b.size() * 4);
    memcpy(cp->bias32, L.h_bias.data(), 32 * 4);
  }
  cp->ref_batch = 1;
  return UB_OK;
}

// Pipeline shape for one launch: weight-stationary when every tap of an n-tile fits beside >= 2 A stages
// (and the persistent grid keeps each CTA on one n-tile), else a short A ring + as many B stages as fit.
template <int N, int BK, int KS, int TW, bool CG2 = false>
bool choose_pipeline(ConvParams* cp, int grid, bool allow_ws, int* smem_bytes) {
  using Cfg = ub::ConvCfg<N, BK, KS, TW, CG2>;
  const int n_tiles = Cfg::KXS ? 1 : cp->n_total / N;
  const int n_taps = (cp->cin / BK) * Cfg::TAPS;
  const int w_bytes = n_taps * Cfg::B_STAGE_BYTES;
  const int budget = Cfg::OPERAND_BUDGET;
  if (allow_ws && w_bytes + 2 * Cfg::A_STAGE_BYTES <= budget && grid % n_tiles == 0) {
    int sa = (budget - w_bytes) / Cfg::A_STAGE_BYTES;
    cp->sa = sa > ub::kMaxSA ? ub::kMaxSA : sa;
    cp->sb = 0;
    *smem_bytes = cp->sa * Cfg::A_STAGE_BYTES + w_bytes + Cfg::TAIL_BYTES + 1024;
    return true;
  }
  if (KS == 1) {
    // one tap: an A box feeds only BK/16 MMAs (512 cycles at N = 256), so two A stages cannot cover a TMA round trip —
    // the rings advance in step, equally deep (the ViT's token GEMMs: 4 + 4 stages; measured in profiles/r02_call34_*)
    int s = budget / (Cfg::A_STAGE_BYTES + Cfg::B_STAGE_BYTES);
    const int cap = ub::kMaxSA < ub::kMaxSB ? ub::kMaxSA : ub::kMaxSB;
    s = s > cap ? cap : (s < 2 ? 2 : s);
    cp->sa = cp->sb = s;
    while (cp->sa < ub::kMaxSA && (cp->sa + 1) * Cfg::A_STAGE_BYTES + cp->sb * Cfg::B_STAGE_BYTES <= budget) cp->sa++;
    *smem_bytes = cp->sa * Cfg::A_STAGE_BYTES + cp->sb * Cfg::B_STAGE_BYTES + Cfg::TAIL_BYTES + 1024;
    return false;
  }
  cp->sa = Cfg::A_STAGE_BYTES >= 16384 ? 2 : 3;
  if (N <= 128) cp->sa += 1;
  int sb = (budget - cp->sa * Cfg::A_STAGE_BYTES) / Cfg::B_STAGE_BYTES;
  cp->sb = sb > ub::kMaxSB ? ub::kMaxSB : sb;
  *smem_bytes = cp->sa * Cfg::A_STAGE_BYTES + cp->sb * Cfg::B_STAGE_BYTES + Cfg::TAIL_BYTES + 1024;
  return false;
}

template <int N, int BK, int KS, bool WS, int TW, bool CG2 = false>
int launch_kernel(ub_ctx* ctx, const ConvParams& cp, int grid, int smem, cudaStream_t st) {
  const void* fn = reinterpret_cast<const void*>(&ub::conv_tc_kernel<N, BK, KS, WS, TW, CG2>);
  if (!ctx->smem_opt_in.count(fn)) {
    UB_CUDA(ctx, cudaFuncSetAttribute(ub::conv_tc_kernel<N, BK, KS, WS, TW, CG2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      ub::kSmemLimit));
    ctx->smem_opt_in.insert(fn);
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(ub::conv_threads(N, TW, KS));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (ctx->use_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (CG2) {  // the two CTAs of a pair are one cluster (adjacent SMs of a TPC)
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  UB_CUDA(ctx, cudaLaunchKernelEx(&cfg, ub::conv_tc_kernel<N, BK, KS, WS, TW, CG2>, cp));
  ctx->launches++;
  return UB_OK;
}

template <int N, int BK, int KS, int TW = 8>
int launch_inst(ub_ctx* ctx, const ConvParams& cp_in, cudaStream_t st) {
  ConvParams cp = cp_in;
  constexpr int TH = ub::tile_rows(TW), XS = ub::tile_xstep(TW);
  const int tiles = cp.B * ((cp.H + TH - 1) / TH) * ((cp.W + XS - 1) / XS) * (ub::is_kxs(TW) ? 1 : cp.n_total / N);
  const int grid = tiles < ctx->num_sms ? tiles : ctx->num_sms;
  int smem = 0;
  // UB_FORCE_STREAM=1 (test hook): exercise the streaming path on layers that would be weight-stationary
  const bool ws = choose_pipeline<N, BK, KS, TW>(&cp, grid, !ctx->force_stream, &smem);
  return ws ? launch_kernel<N, BK, KS, true, TW>(ctx, cp, grid, smem, st)
            : launch_kernel<N, BK, KS, false, TW>(ctx, cp, grid, smem, st);
}

// CTA-pair launch: tiles are enumerated (column, row, image, n-tile) over an even, padded image count
// (ub_conv_tc.cuh), the grid is even and the weights always stream.
template <int N, int KS>
int launch_pair(ub_ctx* ctx, const ConvParams& cp_in, cudaStream_t st) {
  ConvParams cp = cp_in;
  int smem = 0;
  if constexpr (N == 256 && KS == 1) {
    if (cp.pairq) {  // weight-stationary pairs over several n-tiles: the pair count is a multiple of n_tiles
      const int n_tiles = cp.n_total / N;
      const int pairs_total = n_tiles * ((((cp.W + 7) / 8) + 1) / 2) * ((cp.H + 15) / 16) * cp.B;
      int pairs = ctx->num_sms / 2;
      if (pairs > pairs_total) pairs = pairs_total;
      const int pw = pairs - pairs % n_tiles;
      if (pw >= n_tiles && choose_pipeline<N, 64, KS, 8, true>(&cp, pw, !ctx->force_stream, &smem))
        return launch_kernel<N, 64, KS, true, 8, true>(ctx, cp, 2 * pw, smem, st);
      choose_pipeline<N, 64, KS, 8, true>(&cp, pairs, false, &smem);  // tiny launch or UB_FORCE_STREAM: same order, streamed
      return launch_kernel<N, 64, KS, false, 8, true>(ctx, cp, 2 * pairs, smem, st);
    }
  }
  const int per_img = ((cp.H + 15) / 16) * ((cp.W + 7) / 8);
  const int b_pad = cp.B + ((per_img * cp.B) & 1);
  const int tiles = per_img * b_pad * (cp.n_total / N);
  const int max_grid = ctx->num_sms & ~1;
  const int grid = tiles < max_grid ? tiles : max_grid;
  if constexpr (N == 128 && KS == 3) {  // one n-tile and the pair's halves fit: each CTA keeps its half of the weights resident
    if (choose_pipeline<N, 64, KS, 8, true>(&cp, grid, cp.n_total == N && !ctx->force_stream, &smem))
      return launch_kernel<N, 64, KS, true, 8, true>(ctx, cp, grid, smem, st);
  } else {
    choose_pipeline<N, 64, KS, 8, true>(&cp, grid, false, &smem);
  }
  return launch_kernel<N, 64, KS, false, 8, true>(ctx, cp, grid, smem, st);
}

// Row-streaming launch: a task = (image, 30-column band); persistent CTAs walk their tasks row by row.
template <int COUT, int BK>
int launch_rs(ub_ctx* ctx, const ConvParams& cp_in, cudaStream_t st) {
  using Cfg = ub::RsCfg<COUT, BK>;
  ConvParams cp = cp_in;
  const int bands = (cp.W + ub::kRsOut - 1) / ub::kRsOut;
  const int tasks = cp.B * bands;
  const int grid = tasks < ctx->num_sms ? tasks : ctx->num_sms;
  const int w_bytes = 3 * (cp.cin / BK) * Cfg::B_STAGE_BYTES;
  int sa = (Cfg::OPERAND_BUDGET - w_bytes) / Cfg::A_STAGE_BYTES;
  if (sa > ub::kRsMaxSA) sa = ub::kRsMaxSA;
  { const char* e = getenv("UB_RS_SA"); if (e && atoi(e) >= 2 && atoi(e) < sa) sa = atoi(e); }  // (experiment: ring depth)
  if (sa < 2) return fail(ctx, UB_ERR_UNSUPPORTED, "row-streaming kernel: the layer's weights (%d bytes) leave no room for 2 A stages", w_bytes);
  cp.sa = sa;
  const int smem = Cfg::TAIL_BYTES + sa * Cfg::A_STAGE_BYTES + w_bytes + 1024;
  const void* fn = reinterpret_cast<const void*>(&ub::conv_rs_kernel<COUT, BK>);
  if (!ctx->smem_opt_in.count(fn)) {
    UB_CUDA(ctx, cudaFuncSetAttribute(ub::conv_rs_kernel<COUT, BK>, cudaFuncAttributeMaxDynamicSharedMemorySize, ub::kSmemLimit));
    ctx->smem_opt_in.insert(fn);
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(ub::rs_threads(COUT));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  int na = 0;
  if (ctx->use_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  UB_CUDA(ctx, cudaLaunchKernelEx(&cfg, ub::conv_rs_kernel<COUT, BK>, cp));
  ctx->launches++;
  return UB_OK;
}

// Fused conv pairs (ub_conv_rs2.cuh): a task = (image, 28-column band); persistent CTAs.
template <int C1, int C2, int BK1, bool HEAD, bool SPARE = false>
int launch_rs2_inst(ub_ctx* ctx, const ConvParams& cp_in, cudaStream_t st) {
  using Cfg = ub::Rs2Cfg<C1, C2, BK1, HEAD, SPARE>;
  ConvParams cp = cp_in;
  const int bands = (cp.W + ub::kRs2Out - 1) / ub::kRs2Out;
  const int tasks = cp.B * bands;
  const int grid = tasks < ctx->num_sms ? tasks : ctx->num_sms;
  const int fixed = Cfg::B1_BYTES + Cfg::B2_BYTES;
  cp.sa = Cfg::A1_STAGE <= 4096 ? 8 : 6;                                               // A1 stages (4 or 16 KB; 6 against 4: -2 % on dec4)
  { const char* e = getenv("UB_RS2_SA"); if (e && atoi(e) >= 2 && atoi(e) <= ub::kRs2MaxSA1) cp.sa = atoi(e); }  // (experiment)
  int s2 = (Cfg::OPERAND_BUDGET - fixed - cp.sa * Cfg::A1_STAGE) / Cfg::A2_STAGE;      // A2 stages (16 or 8 KB)
  if (s2 > ub::kRs2MaxSA2) s2 = ub::kRs2MaxSA2;
  if (s2 < 3) return fail(ctx, UB_ERR_UNSUPPORTED, "fused conv pair: no room for 3 stage-1 row buffers");
  cp.sb = s2;
  const int smem = Cfg::TAIL_BYTES + cp.sa * Cfg::A1_STAGE + cp.sb * Cfg::A2_STAGE + fixed + 1024;
  const void* fn = reinterpret_cast<const void*>(&ub::conv_rs2_kernel<C1, C2, BK1, HEAD, SPARE>);
  if (!ctx->smem_opt_in.count(fn)) {
    UB_CUDA(ctx, cudaFuncSetAttribute(ub::conv_rs2_kernel<C1, C2, BK1, HEAD, SPARE>, cudaFuncAttributeMaxDynamicSharedMemorySize, ub::kSmemLimit));
    ctx->smem_opt_in.insert(fn);
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(Cfg::THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  int na = 0;
  if (ctx->use_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  UB_CUDA(ctx, cudaLaunchKernelEx(&cfg, ub::conv_rs2_kernel<C1, C2, BK1, HEAD, SPARE>, cp));
  ctx->launches++;
  return UB_OK;
}

int launch_rs2(ub_ctx* ctx, const ConvParams& cp, cudaStream_t st) {
  if (cp.mode == ub::EPI_HEAD)                                                           // dec4.conv.0 -> conv.3 -> head
    return ctx->rs2_no_spare ? launch_rs2_inst<32, 32, 64, true, false>(ctx, cp, st)     // (UB_RS2_NO_SPARE=1: circular TMEM windows)
                             : launch_rs2_inst<32, 32, 64, true, true>(ctx, cp, st);
  return launch_rs2_inst<64, 64, 16, false>(ctx, cp, st);                              // enc1.conv.0 -> conv.3 -> pool
}

template <int N, int BK>
int launch_ks(ub_ctx* ctx, const ConvParams& cp, cudaStream_t st) {
  if (cp.ksize == 3) return launch_inst<N, BK, 3>(ctx, cp, st);
  if (cp.ksize == 1) return launch_inst<N, BK, 1>(ctx, cp, st);
  return fail(ctx, UB_ERR_UNSUPPORTED, "kernel size %d", cp.ksize);
}

int launch_gemm_raw(ub_ctx* ctx, const ConvParams& cp, int block_n, int bk, cudaStream_t st);

cudaEvent_t take_event(ub_ctx* ctx) {
  if (!ctx->event_pool.empty()) { cudaEvent_t e = ctx->event_pool.back(); ctx->event_pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}

int launch_gemm(ub_ctx* ctx, const ConvParams& cp, int block_n, int bk, cudaStream_t st, int layer = -1) {
  if (!ctx->profiling) return launch_gemm_raw(ctx, cp, block_n, bk, st);
  ub_ctx::ProfRec r{layer, take_event(ctx), take_event(ctx)};
  cudaEventRecord(r.a, st);
  const int rc = launch_gemm_raw(ctx, cp, block_n, bk, st);
  cudaEventRecord(r.b, st);
  ctx->prof.push_back(r);
  return rc;
}

int launch_gemm_raw(ub_ctx* ctx, const ConvParams& cp, int block_n, int bk, cudaStream_t st) {
  if (cp.mode == ub::EPI_HEAD && cp.cout != 32) return fail(ctx, UB_ERR_UNSUPPORTED, "head epilogue needs 32 channels");
  if (cp.rs == 2) return launch_rs2(ctx, cp, st);
  if (cp.rs) {  // thin 3x3 layers at the larger sizes: ky-stacked, row-streaming
    if (cp.cout == 64 && bk == 64) return launch_rs<64, 64>(ctx, cp, st);
    if (cp.cout == 32 && bk == 64) return launch_rs<32, 64>(ctx, cp, st);
    if (cp.cout == 32 && bk == 32) return launch_rs<32, 32>(ctx, cp, st);
    return fail(ctx, UB_ERR_UNSUPPORTED, "no row-streaming kernel for cout=%d BK=%d", cp.cout, bk);
  }
  if (ub::is_kxs(cp.tw)) {  // thin 3x3 layers, N = 3 * cout
    if (bk == 32 && cp.tw == 32 && cp.cout == 64) return launch_inst<192, 32, 3, 32>(ctx, cp, st);
    if (bk == 32 && cp.tw == 32 && cp.cout == 32) return launch_inst<96, 32, 3, 32>(ctx, cp, st);
    if (bk == 16 && cp.tw == 32 && cp.cout == 32) return launch_inst<96, 16, 3, 32>(ctx, cp, st);
    if (bk != 64) return fail(ctx, UB_ERR_UNSUPPORTED, "no kx-stacked kernel for BK=%d tw=%d", bk, cp.tw);
    if (cp.tw == 32 && cp.cout == 64) return launch_inst<192, 64, 3, 32>(ctx, cp, st);
    if (cp.tw == 32 && cp.cout == 32) return launch_inst<96, 64, 3, 32>(ctx, cp, st);
    if (cp.tw == 16 && cp.cout == 64) return launch_inst<192, 64, 3, 16>(ctx, cp, st);
    if (cp.tw == 16 && cp.cout == 32) return launch_inst<96, 64, 3, 16>(ctx, cp, st);
    return fail(ctx, UB_ERR_UNSUPPORTED, "no kx-stacked kernel for cout=%d tw=%d", cp.cout, cp.tw);
  }
  if (cp.cg2) {
    if (block_n == 256 && cp.ksize == 3) return launch_pair<256, 3>(ctx, cp, st);
    if (block_n == 256 && cp.ksize == 1) return launch_pair<256, 1>(ctx, cp, st);
    if (block_n == 128 && cp.ksize == 3) return launch_pair<128, 3>(ctx, cp, st);
    if (block_n == 128 && cp.ksize == 1) return launch_pair<128, 1>(ctx, cp, st);
    return fail(ctx, UB_ERR_UNSUPPORTED, "no CTA-pair kernel for BLOCK_N=%d ksize=%d", block_n, cp.ksize);
  }
  if (bk == 64) {
    switch (block_n) {
      case 32: return launch_ks<32, 64>(ctx, cp, st);
      case 64: return launch_ks<64, 64>(ctx, cp, st);
      case 128: return launch_ks<128, 64>(ctx, cp, st);
      case 256: return launch_ks<256, 64>(ctx, cp, st);
    }
  } else if (bk == 32 && cp.ksize == 3) {  // the two Cin = 32 layers (enc1.conv.0, dec4.conv.3)
    switch (block_n) {
      case 32: return launch_inst<32, 32, 3>(ctx, cp, st);
      case 64: return launch_inst<64, 32, 3>(ctx, cp, st);
    }
  } else if (bk == 16 && cp.ksize == 3) {  // the stem's two readers with the stem folded in (16 / 48 input channels)
    switch (block_n) {
      case 32: return launch_inst<32, 16, 3>(ctx, cp, st);
      case 64: return launch_inst<64, 16, 3>(ctx, cp, st);
    }
  }
  return fail(ctx, UB_ERR_UNSUPPORTED, "no kernel for BLOCK_N=%d BK=%d ksize=%d", block_n, bk, cp.ksize);
}

// ------------------------------------------------------------------ execution plan for one resolution
int skip_width(int level) { return level == 0 ? 32 : (32 << level); }

// True when F.interpolate(align_corners=True) from 2n back to n samples reads, for every output sample o, only the two
// children of ConvTranspose input sample o (ub_conv_tc.cuh: convt_resample_weights, evaluated here with the kernel's own
// fp32 arithmetic): then ConvTranspose + resize is a per-pixel blend and needs no 2n-sized tensor.
bool convt_resize_blends(int n) {
  for (int o = 0; o < n; ++o) {
    float w0, w1;
    if (!ub::convt_resample_weights(o, n, &w0, &w1)) return false;
  }
  return true;
}

int build_plan(ub_ctx* ctx, int H, int W, cudaStream_t st, Plan** out) {
  auto key = std::make_pair(H, W);
  auto it = ctx->plans.find(key);
  if (it != ctx->plans.end()) { *out = it->second.get(); return UB_OK; }
  const int n = ctx->n_enc, MB = ctx->max_mb;
  if (H % (1 << n) || W % (1 << n) || H < (2 << n) || W < (2 << n))
    return fail(ctx, UB_ERR_UNSUPPORTED, "H and W must be multiples of %d and at least %d", 1 << n, 2 << n);
  std::unique_ptr<Plan> pl(new Plan());
  pl->H = H; pl->W = W;
  std::vector<size_t> sizes;
  auto want = [&](size_t bytes) { sizes.push_back((bytes + 1023) & ~static_cast<size_t>(1023)); return sizes.size() - 1; };
  auto hl = [&](int l) { return H >> l; };
  auto wl = [&](int l) { return W >> l; };
  std::vector<size_t> i_cat(n + 1), i_enc(n + 1), i_da(n + 2), i_db(n + 2);
  for (int l = 0; l <= n; ++l) i_cat[l] = want(static_cast<size_t>(MB) * hl(l) * wl(l) * 2 * skip_width(l) * 2);
  for (int i = 1; i <= n; ++i) i_enc[i] = want(static_cast<size_t>(MB) * hl(i - 1) * wl(i - 1) * skip_width(i) * 2);
  const int cb = 2 * skip_width(n);
  const size_t i_ba = want(static_cast<size_t>(MB) * hl(n) * wl(n) * cb * 2);
  const size_t i_bb = want(static_cast<size_t>(MB) * hl(n) * wl(n) * cb * 2);
  const size_t i_up = want(static_cast<size_t>(MB) * 4 * hl(n) * wl(n) * skip_width(n) * 2);
  for (int j = 1; j <= n + 1; ++j) {
    const int l = n + 1 - j;
    i_da[j] = want(static_cast<size_t>(MB) * hl(l) * wl(l) * skip_width(l) * 2);
    i_db[j] = want(static_cast<size_t>(MB) * hl(l) * wl(l) * skip_width(l) * 2);
  }
  // ClipUnetPrompt: the PromptEncoder's activations (32 -> 32 -> 64 -> 128 channels at H .. H/8) and the 1024-channel
  // concat [attention output | prompt embedding] that prompt_fusion reads
  const bool prompt = ctx->arch == UB_ARCH_CLIP_UNET_PROMPT;
  size_t i_p0 = 0, i_pa[4] = {0, 0, 0, 0}, i_pb[4] = {0, 0, 0, 0}, i_pcat = 0;
  if (prompt) {
    if (n != 3) return fail(ctx, UB_ERR_UNSUPPORTED, "ClipUnetPrompt has three encoder stages");
    i_p0 = want(static_cast<size_t>(MB) * H * W * 32 * 2);
    for (int i = 1; i <= 3; ++i) {
      i_pa[i] = want(static_cast<size_t>(MB) * hl(i - 1) * wl(i - 1) * (16 << i) * 2);  // conv.0 output at the input resolution
      i_pb[i] = want(static_cast<size_t>(MB) * hl(i) * wl(i) * (16 << i) * 2);          // pooled
    }
    i_pcat = want(static_cast<size_t>(MB) * hl(n) * wl(n) * 2 * cb * 2);
  }
  // Autoencoder: decoder stage j (no skip) works at level n - j with 512 >> j channels: ConvT output, conv.0, conv.3
  const bool ae = ctx->arch == UB_ARCH_AUTOENCODER;
  std::vector<size_t> i_au(n + 1), i_aa(n + 1), i_ab(n + 1);
  if (ae)
    for (int j = 1; j <= n; ++j) {
      const size_t bytes = static_cast<size_t>(MB) * hl(n - j) * wl(n - j) * (cb >> j) * 2;
      i_au[j] = want(bytes); i_aa[j] = want(bytes); i_ab[j] = want(bytes);
    }
  const size_t img_bytes = static_cast<size_t>(MB) * 3 * H * W * 4;
  const size_t i_pert = want(img_bytes), i_bla = want(img_bytes), i_blb = want(img_bytes);
  size_t total = 0;
  std::vector<size_t> offs(sizes.size());
  for (size_t i = 0; i < sizes.size(); ++i) { offs[i] = total; total += sizes[i]; }
  int rc;
  if ((rc = dev_alloc(ctx, &pl->arena, total))) return rc;
  pl->arena_bytes = total;
  // zeroed ON THE CALLER'S STREAM: the kernels that rely on the zero planes (plane 1 of the packed image, the prompt
  // stem's padding) are launched on `st`, which need not synchronise with the legacy default stream
  UB_CUDA(ctx, cudaMemsetAsync(pl->arena, 0, total, st));
  auto bufp = [&](size_t idx) { return reinterpret_cast<__nv_bfloat16*>(static_cast<uint8_t*>(pl->arena) + offs[idx]); };
  pl->pert_f32 = reinterpret_cast<float*>(bufp(i_pert));
  pl->blur_a = bufp(i_bla);
  pl->blur_b = bufp(i_blb);
  // every activation buffer is channel-blocked [MB][cgs][h][w][8]; plane(l) = elements of one channel group
  auto plane = [&](int l) { return static_cast<size_t>(hl(l)) * wl(l) * 8; };
  auto skip_half = [&](int l) { return bufp(i_cat[l]) + static_cast<size_t>(skip_width(l) / 8) * plane(l); };
  pl->stem_out = bufp(i_cat[0]);
  pl->stem_cgs = 2 * skip_width(0) / 8;
  pl->stem_cg0 = skip_width(0) / 8;

  auto add_gemm = [&](const std::string& lkey, const __nv_bfloat16* src, int src_cgs, int h, int w, int mode,
                      __nv_bfloat16* dst, int dst_cgs, int dst_cg0) -> int {
    auto li = ctx->layer_by_key.find(lkey);
    if (li == ctx->layer_by_key.end()) return fail(ctx, UB_ERR_STATE, "layer %s not packed", lkey.c_str());
    Step s;
    s.kind = Step::GEMM;
    s.layer = li->second;
    int r2 = fill_gemm(ctx, &s.cp, ctx->layers[s.layer], src, src_cgs, MB, h, w, mode, dst, dst_cgs, dst_cg0);
    if (r2) return r2;
    pl->steps.push_back(s);
    return UB_OK;
  };
  char k[128];
  // the feature maps of Encoder.forward's dict (models/autoencoder.py:17-25): x0 and the pooled stage outputs live in
  // the skip planes of the concat buffers, the bottleneck in its own buffer
  for (int l = 0; l <= n; ++l)
    pl->feats.push_back({bufp(i_cat[l]), 2 * skip_width(l) / 8, skip_width(l) / 8, skip_width(l), l});
  pl->feats.push_back({bufp(i_bb), cb / 8, 0, cb, n});
  const bool enc_half = ctx->arch == UB_ARCH_ENCODER, dec_half = ctx->arch == UB_ARCH_SKIP_DECODER;
  // encoder (models/UNet.py:63-65): conv.0 -> conv.3 + MaxPool into the skip planes of the next concat buffer
  for (int i = 1; i <= (dec_half ? 0 : n); ++i) {
    snprintf(k, sizeof(k), "enc%d.block.0.conv", i);
    const bool fold = (i == 1 && ctx->stem_fold);  // reads the packed image (16 channels) through composed weights
    if (fold && ctx->fuse_enc1 && hl(0) % 8 == 0 && hl(0) >= 32 && wl(0) >= 32) {
      // one kernel for the whole stage (ub_conv_rs2.cuh): no conv.0 output buffer is touched
      auto l0 = ctx->layer_by_key.find(std::string(k) + ".0+stem");
      auto l3 = ctx->layer_by_key.find(std::string(k) + ".3");
      if (l0 == ctx->layer_by_key.end() || l3 == ctx->layer_by_key.end()) return fail(ctx, UB_ERR_STATE, "enc1 layers not packed");
      const GemmLayer& L0 = ctx->layers[l0->second];
      const GemmLayer& L3 = ctx->layers[l3->second];
      if (L0.cin == 16 && L0.cout == 64 && L3.cin == 64 && L3.cout == 64 && L0.w_rs && L3.w_rs) {
        Step s;
        s.kind = Step::GEMM;
        s.layer = l3->second;
        ConvParams& cp = s.cp;
        memset(&cp, 0, sizeof(cp));
        if ((rc = make_tm_w(ctx, &cp.tmA, L0.w_rs, L0.cin, 9 * L0.cout, 3 * L0.cout, L0.bk))) return rc;
        if ((rc = make_tm_w(ctx, &cp.tmB, L3.w_rs, L3.cin, 9 * L3.cout, 3 * L3.cout, L3.bk))) return rc;
        cp.rs = 2; cp.tw = 32; cp.hq = hl(0) / 4;
        cp.a_src = skip_half(0); cp.a_cgs = 2 * skip_width(0) / 8; cp.a_planes = (L0.cin_src ? L0.cin_src : L0.cin) / 8;
        cp.B = MB; cp.H = hl(0); cp.W = wl(0);
        cp.cin = L0.cin; cp.ksize = 3; cp.n_total = L3.n_total; cp.mode = ub::EPI_POOL; cp.relu = 1;
        cp.bias1 = L0.bias; cp.bias = L3.bias; cp.n_bias = L3.cout;
        memcpy(cp.bias64[0], L0.h_bias.data(), 64 * sizeof(float));
        memcpy(cp.bias64[1], L3.h_bias.data(), 64 * sizeof(float));
        cp.out = bufp(i_cat[1]); cp.out_cgs = 2 * skip_width(1) / 8; cp.out_cg0 = skip_width(1) / 8; cp.cout = L3.cout;
        cp.dbg = ctx->dbg; cp.ref_batch = 1; cp.n_classes = ctx->n_classes;
        pl->steps.push_back(s);
        continue;
      }
    }
    if ((rc = add_gemm(std::string(k) + (fold ? ".0+stem" : ".0"), skip_half(i - 1), 2 * skip_width(i - 1) / 8, hl(i - 1), wl(i - 1),
                       ub::EPI_STORE, bufp(i_enc[i]), skip_width(i) / 8, 0))) return rc;
    if ((rc = add_gemm(std::string(k) + ".3", bufp(i_enc[i]), skip_width(i) / 8, hl(i - 1), wl(i - 1), ub::EPI_POOL,
                       bufp(i_cat[i]), 2 * skip_width(i) / 8, skip_width(i) / 8))) return rc;
  }
  // bottleneck (UNet.py:67), not pooled
  if (ctx->arch == UB_ARCH_CLIP_UNET) {
    // ClipUnet: dec1's input is the cross-attention fusion of the CLIP embedding, broadcast over the grid
    Step s;
    s.kind = Step::CLIP_FUSE;
    s.rs_out = bufp(i_bb); s.ho = hl(n); s.wo = wl(n); s.c = cb; s.out_cgs = cb / 8;
    pl->steps.push_back(s);
  } else if (prompt) {
    // ClipUnetPrompt (prompt_segmentation.py:76-86): PromptEncoder over the heat-map, cross-attention fusion of the CLIP
    // embedding into planes [0, cb/8) and the prompt embedding into planes [cb/8, 2cb/8) of one buffer
    // (= cat([attention_output, prompt_embedding], dim=1)), then the 1x1 prompt_fusion -> dec1's input
    Step s0;
    s0.kind = Step::PROMPT_STEM;
    s0.rs_out = bufp(i_p0); s0.ho = H; s0.wo = W; s0.out_cgs = 4;
    pl->steps.push_back(s0);
    const __nv_bfloat16* px = bufp(i_p0);
    int pc = 32;
    for (int i = 1; i <= 3; ++i) {
      const int co = 16 << i;
      snprintf(k, sizeof(k), "prompt_encoder.enc%d.block.0.conv", i);
      if ((rc = add_gemm(std::string(k) + ".0", px, pc / 8, hl(i - 1), wl(i - 1), ub::EPI_STORE, bufp(i_pa[i]), co / 8, 0))) return rc;
      if ((rc = add_gemm(std::string(k) + ".3", bufp(i_pa[i]), co / 8, hl(i - 1), wl(i - 1), ub::EPI_POOL, bufp(i_pb[i]), co / 8, 0))) return rc;
      px = bufp(i_pb[i]); pc = co;
    }
    if ((rc = add_gemm("prompt_encoder.conv.conv.0", px, pc / 8, hl(n), wl(n), ub::EPI_STORE, bufp(i_ba), cb / 8, 0))) return rc;
    if ((rc = add_gemm("prompt_encoder.conv.conv.3", bufp(i_ba), cb / 8, hl(n), wl(n), ub::EPI_STORE, bufp(i_pcat), 2 * cb / 8, cb / 8))) return rc;
    Step s;
    s.kind = Step::CLIP_FUSE;
    s.rs_out = bufp(i_pcat); s.ho = hl(n); s.wo = wl(n); s.c = cb; s.out_cgs = 2 * cb / 8;
    pl->steps.push_back(s);
    if ((rc = add_gemm("prompt_fusion", bufp(i_pcat), 2 * cb / 8, hl(n), wl(n), ub::EPI_STORE, bufp(i_bb), cb / 8, 0))) return rc;
  } else if (!dec_half) {
    if ((rc = add_gemm("bottleneck.conv.0", skip_half(n), 2 * skip_width(n) / 8, hl(n), wl(n), ub::EPI_STORE,
                       bufp(i_ba), cb / 8, 0))) return rc;
    if ((rc = add_gemm("bottleneck.conv.3", bufp(i_ba), cb / 8, hl(n), wl(n), ub::EPI_STORE, bufp(i_bb), cb / 8, 0))) return rc;
  }
  if (enc_half) {  // Encoder on its own: no decoder
    *out = pl.get();
    ctx->plans[key] = std::move(pl);
    return UB_OK;
  }
  // decoder (UNet.py:69-72, processing_blocks.py:105-109)
  const __nv_bfloat16* x = bufp(i_bb);
  int xc = cb, xh = hl(n), xw = wl(n);
  if (ae) {
    // Autoencoder decoder (autoencoder.py:36-41, processing_blocks.py:131-133): ConvT -> ConvBlock, no skip, no resize
    for (int j = 1; j <= n; ++j) {
      const int l = n - j, c = cb >> j;
      snprintf(k, sizeof(k), "dec%d", j);
      if ((rc = add_gemm(std::string(k) + ".up", x, xc / 8, xh, xw, ub::EPI_CONVT, bufp(i_au[j]), c / 8, 0))) return rc;
      if ((rc = add_gemm(std::string(k) + ".conv.conv.0", bufp(i_au[j]), c / 8, hl(l), wl(l), ub::EPI_STORE, bufp(i_aa[j]), c / 8, 0))) return rc;
      if ((rc = add_gemm(std::string(k) + ".conv.conv.3", bufp(i_aa[j]), c / 8, hl(l), wl(l), ub::EPI_STORE, bufp(i_ab[j]), c / 8, 0))) return rc;
      x = bufp(i_ab[j]); xc = c; xh = hl(l); xw = wl(l);
    }
    Step s;
    s.kind = Step::AE_OUT;
    s.rs_in = x; s.c = xc; s.ho = xh; s.wo = xw;
    pl->steps.push_back(s);
    *out = pl.get();
    ctx->plans[key] = std::move(pl);
    return UB_OK;
  }
  for (int j = 1; j <= n + 1; ++j) {
    const int l = n + 1 - j;
    snprintf(k, sizeof(k), "dec%d", j);
    const int sw = skip_width(l);
    if (2 * xh == hl(l) && 2 * xw == wl(l)) {
      // F.interpolate is the identity: ConvT scatters straight into planes [0, sw/8) of the concat buffer
      if ((rc = add_gemm(std::string(k) + ".up", x, xc / 8, xh, xw, ub::EPI_CONVT, bufp(i_cat[l]), 2 * sw / 8, 0))) return rc;
    } else if (ctx->fuse_resize && xh == hl(l) && xw == wl(l) && convt_resize_blends(xh) && convt_resize_blends(xw)) {
      // the resize undoes the doubling (dec1: the bottleneck is not pooled): every resized pixel is a blend of the four
      // parity products of ONE ConvTranspose input pixel -> done in the GEMM's epilogue, straight into the concat buffer
      if ((rc = add_gemm(std::string(k) + ".up", x, xc / 8, xh, xw, ub::EPI_CONVT_BLEND, bufp(i_cat[l]), 2 * sw / 8, 0))) return rc;
    } else {
      if ((rc = add_gemm(std::string(k) + ".up", x, xc / 8, xh, xw, ub::EPI_CONVT, bufp(i_up), sw / 8, 0))) return rc;
      Step s;
      s.kind = Step::RESAMPLE;
      s.rs_in = bufp(i_up); s.rs_out = bufp(i_cat[l]);
      s.hi = 2 * xh; s.wi = 2 * xw; s.ho = hl(l); s.wo = wl(l); s.c = sw; s.out_cgs = 2 * sw / 8; s.out_cg0 = 0;
      pl->steps.push_back(s);
    }
    const bool last = (j == n + 1);
    if (last && ctx->stem_fold && ctx->fuse_dec4 && hl(l) % 8 == 0 && hl(l) >= 32 && wl(l) >= 32) {
      auto l0 = ctx->layer_by_key.find(std::string(k) + ".conv.conv.0+stem");
      auto l3 = ctx->layer_by_key.find(std::string(k) + ".conv.conv.3");
      if (l0 == ctx->layer_by_key.end() || l3 == ctx->layer_by_key.end()) return fail(ctx, UB_ERR_STATE, "last decoder stage not packed");
      const GemmLayer& L0 = ctx->layers[l0->second];
      const GemmLayer& L3 = ctx->layers[l3->second];
      if (L0.cin == 64 && L0.cout == 32 && L3.cin == 32 && L3.cout == 32 && L0.w_rs && L3.w_rs &&
          ctx->h_head_w.size() == static_cast<size_t>(ctx->n_classes) * 32) {
        Step s;
        s.kind = Step::GEMM;
        s.layer = l3->second;
        ConvParams& cp = s.cp;
        memset(&cp, 0, sizeof(cp));
        if ((rc = make_tm_w(ctx, &cp.tmA, L0.w_rs, L0.cin, 9 * L0.cout, 3 * L0.cout, L0.bk))) return rc;
        if ((rc = make_tm_w(ctx, &cp.tmB, L3.w_rs, L3.cin, 9 * L3.cout, 3 * L3.cout, L3.bk))) return rc;
        cp.rs = 2; cp.tw = 32; cp.hq = hl(l) / 4;
        cp.a_src = bufp(i_cat[l]); cp.a_cgs = 2 * sw / 8; cp.a_planes = (L0.cin_src ? L0.cin_src : L0.cin) / 8;
        cp.k_steps = (L0.cin_src && L0.cin_src < L0.cin) ? (L0.cin_src + 15) / 16 : 0;
        cp.B = MB; cp.H = hl(l); cp.W = wl(l);
        cp.cin = L0.cin; cp.ksize = 3; cp.n_total = L3.n_total; cp.mode = ub::EPI_HEAD; cp.relu = 1;
        cp.bias1 = L0.bias; cp.bias = L3.bias; cp.n_bias = L3.cout; cp.cout = L3.cout;
        memcpy(cp.bias64[0], L0.h_bias.data(), 32 * sizeof(float));
        memcpy(cp.bias64[1], L3.h_bias.data(), 32 * sizeof(float));
        memcpy(cp.bias32, L3.h_bias.data(), 32 * sizeof(float));
        memcpy(cp.head_w, ctx->h_head_w.data(), ctx->h_head_w.size() * 4);
        memcpy(cp.head_b, ctx->h_head_b.data(), ctx->h_head_b.size() * 4);
        cp.dbg = ctx->dbg; cp.ref_batch = 1; cp.n_classes = ctx->n_classes;
        pl->steps.push_back(s);
        x = nullptr; xc = sw; xh = hl(l); xw = wl(l);   // (nothing follows the head)
        continue;
      }
    }
    if ((rc = add_gemm(std::string(k) + ((last && ctx->stem_fold) ? ".conv.conv.0+stem" : ".conv.conv.0"), bufp(i_cat[l]), 2 * sw / 8,
                       hl(l), wl(l), ub::EPI_STORE, bufp(i_da[j]), sw / 8, 0))) return rc;
    if ((rc = add_gemm(std::string(k) + ".conv.conv.3", bufp(i_da[j]), sw / 8, hl(l), wl(l), last ? ub::EPI_HEAD : ub::EPI_STORE,
                       bufp(i_db[j]), sw / 8, 0))) return rc;
    x = bufp(i_db[j]); xc = sw; xh = hl(l); xw = wl(l);
  }
  *out = pl.get();
  ctx->plans[key] = std::move(pl);
  return UB_OK;
}
}  // namespace
namespace {

int to_dev_perturb(ub_ctx* ctx, const ub_perturb* p, const void* aux, ub::PerturbDev* d) {
  memset(d, 0, sizeof(*d));
  if (!p) { d->kind = UB_P_NONE; return UB_OK; }
  if (p->kind < UB_P_NONE || p->kind > UB_P_F_SALT_PEPPER) return fail(ctx, UB_ERR_INVALID, "unknown perturbation kind %d", p->kind);
  if ((p->kind == UB_P_U8_BLUR || p->kind == UB_P_F_BLUR) && (p->iparam < 0 || p->iparam > 64))
    return fail(ctx, UB_ERR_INVALID, "blur passes out of range: %d", p->iparam);
  d->kind = p->kind; d->iparam = p->iparam; d->fparam = p->fparam; d->fparam2 = p->fparam2;
  d->seed = p->seed; d->aux = aux;
  memcpy(d->lut, p->lut, 256);
  return UB_OK;
}

inline unsigned blocks_for(size_t n, int bs) { return static_cast<unsigned>((n + bs - 1) / bs); }

// Blur families.  Up to kBlurMaxPasses passes: one launch of blur_fused_kernel, which writes the packed network input
// (`packed`, may be null) and/or the fp32 NCHW image (`dst`, may be null).  More passes (or UB_NO_BLUR_FUSE): quantise/load
// -> one launch per pass (ping-pong planes) -> fp32 NCHW in `dst`; the caller packs it.  Returns through *did_pack
// whether `packed` has been written.
int run_blur(ub_ctx* ctx, Plan* pl, const ub::PerturbDev& P, const void* img, int fmt, int n, float* dst,
             __nv_bfloat16* packed, int cgs, int cg0, bool* did_pack, cudaStream_t st) {
  const int H = pl->H, W = pl->W;
  const size_t count = static_cast<size_t>(n) * 3 * H * W;
  const unsigned g = blocks_for(count, 256);
  if (did_pack) *did_pack = false;
  if (P.iparam <= ub::kBlurMaxPasses && !ctx->no_blur_fuse) {
    // every pass of a tile in shared memory: one launch, the image read once, the result written once
    const bool is_f = P.kind != UB_P_U8_BLUR;
    const size_t smem = ub::blur_smem_bytes(P.iparam, is_f);
    if (smem > 48 * 1024) {  // the float family with many passes: opt in (per device, like the conv kernels)
      const void* fn = reinterpret_cast<const void*>(&ub::blur_fused_kernel<true>);
      if (!ctx->smem_opt_in.count(fn)) {
        UB_CUDA(ctx, cudaFuncSetAttribute(ub::blur_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          static_cast<int>(ub::blur_smem_bytes(ub::kBlurMaxPasses, true))));
        ctx->smem_opt_in.insert(fn);
      }
    }
    const dim3 grid((W + ub::kBlurTW - 1) / ub::kBlurTW, (H + ub::kBlurTH - 1) / ub::kBlurTH, n);
    if (!is_f) ub::blur_fused_kernel<false><<<grid, 256, smem, st>>>(img, fmt, H, W, P.iparam, dst, packed, cgs, cg0);
    else ub::blur_fused_kernel<true><<<grid, 256, smem, st>>>(img, fmt, H, W, P.iparam, dst, packed, cgs, cg0);
    ctx->launches++;
    UB_CUDA(ctx, cudaGetLastError());
    if (did_pack) *did_pack = packed != nullptr;
    return UB_OK;
  }
  if (!dst) return fail(ctx, UB_ERR_STATE, "the pass-per-launch blur path needs an fp32 destination");
  if (P.kind == UB_P_U8_BLUR) {
    uint8_t* a = static_cast<uint8_t*>(pl->blur_a);
    uint8_t* b = static_cast<uint8_t*>(pl->blur_b);
    ub::blur_u8_quantise_kernel<<<g, 256, 0, st>>>(img, fmt, n, H, W, a);
    ctx->launches++;
    for (int i = 0; i < P.iparam; ++i) {
      ub::blur_u8_pass_kernel<<<g, 256, 0, st>>>(a, b, n * 3, H, W);
      ctx->launches++;
      std::swap(a, b);
    }
    ub::u8_planar_to_f32_kernel<<<g, 256, 0, st>>>(a, dst, count);
    ctx->launches++;
  } else {
    float* a = static_cast<float*>(pl->blur_a);
    float* b = static_cast<float*>(pl->blur_b);
    if (P.iparam == 0) a = dst;
    ub::blur_f_load_kernel<<<g, 256, 0, st>>>(img, fmt, n, H, W, a);
    ctx->launches++;
    for (int i = 0; i < P.iparam; ++i) {
      float* o = (i == P.iparam - 1) ? dst : b;
      ub::blur_f_pass_kernel<<<g, 256, 0, st>>>(a, o, n * 3, H, W);
      ctx->launches++;
      b = a; a = o;
    }
  }
  UB_CUDA(ctx, cudaGetLastError());
  return UB_OK;
}

struct HeadArgs {
  const void* targets = nullptr; int target_fmt = 0; int ref_batch = 1; int img_base = 0;
  unsigned long long* cm = nullptr; double* dice = nullptr; uint8_t* pred = nullptr; float* logits = nullptr;
  const float* clip = nullptr;  // UB_ARCH_CLIP_UNET(_PROMPT): this micro-batch's CLIP embeddings [nb, 512]
  const float* prompt = nullptr;  // UB_ARCH_CLIP_UNET_PROMPT: this micro-batch's prompt heat-maps [nb, 1, H, W]
};

// One micro-batch: perturb+stem, then every planned step.
int run_microbatch(ub_ctx* ctx, Plan* pl, const ub::PerturbDev& P, const void* img, int fmt, int nb,
                   const HeadArgs& ha, float* perturbed_out, cudaStream_t st) {
  const int H = pl->H, W = pl->W;
  const size_t px = static_cast<size_t>(nb) * H * W;
  int rc;
  if (!img) {
    // SegmentationDecoderSkip on its own: the feature maps are already in the plan's buffers (ub_skip_decoder_forward)
  } else if (P.kind == UB_P_U8_BLUR || P.kind == UB_P_F_BLUR) {
    // fused passes + folded stem: the blur kernel writes the packed image itself (no fp32 staging image, no second
    // launch); perturbed_out (tests) is written by the same launch when asked for
    const bool direct = ctx->stem_fold && P.iparam <= ub::kBlurMaxPasses && !ctx->no_blur_fuse;
    float* dst = perturbed_out ? perturbed_out : (direct ? nullptr : pl->pert_f32);
    bool packed_done = false;
    if ((rc = run_blur(ctx, pl, P, img, fmt, nb, dst, direct ? pl->stem_out : nullptr, pl->stem_cgs, pl->stem_cg0, &packed_done, st)))
      return rc;
    if (packed_done) {
      // (the launch below is not needed)
    } else {
    ub::PerturbDev none;
    memset(&none, 0, sizeof(none));
    if (ctx->stem_fold)
      ub::image_pack_kernel<<<blocks_for(blocks_for(px, ub::kStemPx), 256), 256, 0, st>>>(none, dst, UB_IMG_F32_NCHW, nb, H, W, pl->stem_out,
                                                                pl->stem_cgs, pl->stem_cg0, nullptr);
    else
      ub::stem_kernel<<<blocks_for(blocks_for(px, ub::kStemPx), 256), 256, 0, st>>>(none, ctx->stem, dst, UB_IMG_F32_NCHW, nb, H, W, pl->stem_out,
                                                          pl->stem_cgs, pl->stem_cg0, nullptr);
    UB_CUDA(ctx, cudaGetLastError());
    ctx->launches++;
    }
  } else if (ctx->stem_fold) {
    ub::image_pack_kernel<<<blocks_for(blocks_for(px, ub::kStemPx), 256), 256, 0, st>>>(P, img, fmt, nb, H, W, pl->stem_out, pl->stem_cgs,
                                                              pl->stem_cg0, perturbed_out);
    UB_CUDA(ctx, cudaGetLastError());
    ctx->launches++;
  } else {
    ub::stem_kernel<<<blocks_for(blocks_for(px, ub::kStemPx), 256), 256, 0, st>>>(P, ctx->stem, img, fmt, nb, H, W, pl->stem_out, pl->stem_cgs,
                                                        pl->stem_cg0, perturbed_out);
    UB_CUDA(ctx, cudaGetLastError());
    ctx->launches++;
  }
  for (Step& s : pl->steps) {
    if (s.kind == Step::CLIP_FUSE) {
      if (!ha.clip) return fail(ctx, UB_ERR_INVALID, "this model needs CLIP embeddings: call ub_eval_batches_clip");
      ub::clip_fusion_broadcast_kernel<<<nb, 512, 2 * s.c * sizeof(float), st>>>(
          ha.clip, ctx->fuse_wv_t, ctx->fuse_bv, ctx->fuse_wo_t, ctx->fuse_bo, s.rs_out, s.c, s.ho * s.wo, s.out_cgs);
      UB_CUDA(ctx, cudaGetLastError());
      ctx->launches++;
      continue;
    }
    if (s.kind == Step::AE_OUT) {
      if (!ha.logits) return fail(ctx, UB_ERR_INVALID, "the autoencoder writes its reconstruction: call ub_forward");
      const size_t hw1 = static_cast<size_t>(s.ho) * s.wo;
      const int C = ctx->n_classes;
      ub::out1x1_sigmoid_kernel<ub::kMaxClasses><<<blocks_for(hw1 * nb, 256), 256, (C * s.c + C) * sizeof(float), st>>>(
          s.rs_in, ctx->head_w, ctx->head_b, ha.logits, nb, hw1, s.c, C, 1);
      UB_CUDA(ctx, cudaGetLastError());
      ctx->launches++;
      continue;
    }
    if (s.kind == Step::PROMPT_STEM) {
      if (!ha.prompt) return fail(ctx, UB_ERR_INVALID, "this model needs a prompt heat-map: call ub_forward_prompt");
      const size_t hw1 = static_cast<size_t>(s.ho) * s.wo;
      ub::prompt_stem_kernel<<<blocks_for(hw1 * nb, 256), 256, 0, st>>>(ha.prompt, hw1, nb, s.rs_out, s.out_cgs);
      UB_CUDA(ctx, cudaGetLastError());
      ctx->launches++;
      continue;
    }
    if (s.kind == Step::RESAMPLE) {
      const size_t total = static_cast<size_t>(nb) * s.ho * s.wo * (s.c / 8);
      ub::resample_bilinear_kernel<<<blocks_for(total, 256), 256, 0, st>>>(s.rs_in, nb, s.hi, s.wi, s.c, s.rs_out, s.ho, s.wo,
                                                                           s.out_cgs, s.out_cg0);
      UB_CUDA(ctx, cudaGetLastError());
      ctx->launches++;
      continue;
    }
    ConvParams cp = s.cp;
    cp.B = nb;
    if (cp.mode == ub::EPI_HEAD) {
      cp.targets = ha.targets; cp.target_fmt = ha.target_fmt; cp.ref_batch = ha.ref_batch; cp.img_base = ha.img_base;
      cp.cm = ha.cm; cp.dice = ha.dice; cp.pred_out = ha.pred; cp.logits_out = ha.logits;
    }
    const GemmLayer& L = ctx->layers[s.layer];
    if ((rc = launch_gemm(ctx, cp, L.block_n, L.bk, st, s.layer))) return rc;
  }
  return UB_OK;
}

int check_ready(ub_ctx* ctx) {
  if (!ctx) return UB_ERR_INVALID;
  if (!ctx->weights_ready) return fail(ctx, UB_ERR_STATE, "ub_load_weights has not been called");
  UB_CUDA(ctx, cudaSetDevice(ctx->device));
  return UB_OK;
}

#include "ub_vit_host.inl"

}  // namespace

extern "C" {

int ub_abi_version(void) { return UB_ABI_VERSION; }

int ub_make_lut(ub_perturb* p, int lut_kind, double param) {
  if (!p) return UB_ERR_INVALID;
  const float f = static_cast<float>(param);
  for (int u = 0; u < 256; ++u) {
    float v;
    const float uf = static_cast<float>(u);
    switch (lut_kind) {
      case UB_LUT_CONTRAST: v = fminf(fmaxf(rintf(uf * f), 0.0f), 255.0f); break;
      case UB_LUT_BRIGHTNESS_INCREASE: v = rintf(fminf(fmaxf(uf + f, 0.0f), 255.0f)); break;
      case UB_LUT_BRIGHTNESS_DECREASE: v = fminf(fmaxf(rintf(uf - f), 0.0f), 255.0f); break;
      default: return UB_ERR_INVALID;
    }
    p->lut[u] = static_cast<uint8_t>(v);
  }
  p->kind = UB_P_U8_LUT;
  return UB_OK;
}

int ub_create(ub_ctx** out, int device, int arch, int in_channels, int n_classes, int max_microbatch) {
  if (!out) return UB_ERR_INVALID;
  *out = nullptr;
  if (arch < UB_ARCH_UNET || arch > UB_ARCH_CLIP_VIT)
    return fail(nullptr, UB_ERR_INVALID, "unknown arch %d", arch);
  if (in_channels != 3) return fail(nullptr, UB_ERR_UNSUPPORTED, "only in_channels=3 is supported");
  if (n_classes < 1 || n_classes > ub::kMaxClasses) return fail(nullptr, UB_ERR_UNSUPPORTED, "n_classes must be 1..%d", ub::kMaxClasses);
  if (max_microbatch < 1 || max_microbatch > 4096) return fail(nullptr, UB_ERR_INVALID, "max_microbatch out of range");
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) return fail(nullptr, UB_ERR_CUDA, "cudaGetDeviceProperties(%d): %s", device, cudaGetErrorString(e));
  if (prop.major != 10)
    return fail(nullptr, UB_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library contains sm_100a code only (no fallback)",
                device, prop.major, prop.minor);
  if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(nullptr, UB_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
  std::unique_ptr<ub_ctx> ctx(new ub_ctx());
  ctx->device = device; ctx->arch = arch; ctx->in_ch = in_channels; ctx->n_classes = n_classes;
  ctx->max_mb = max_microbatch; ctx->num_sms = prop.multiProcessorCount;
  ctx->n_enc = arch == UB_ARCH_LARGE_UNET ? 4 : 3;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
    return fail(nullptr, UB_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  ctx->encode = reinterpret_cast<EncodeTiledFn>(fn);
  const char* fs = getenv("UB_FORCE_STREAM");
  ctx->force_stream = fs && fs[0] == '1';
  const char* np = getenv("UB_NO_PDL");
  ctx->use_pdl = !(np && np[0] == '1');
  const char* nk = getenv("UB_NO_KXS");
  ctx->no_kxs = nk && nk[0] == '1';
  { const char* e = getenv("UB_NO_PAIRQ"); ctx->no_pairq = e && e[0] == '1'; }
  { const char* e = getenv("UB_HEAD_NO_KXS"); ctx->head_no_kxs = e && e[0] == '1'; }
  { const char* e = getenv("UB_NO_BLUR_FUSE"); ctx->no_blur_fuse = e && e[0] == '1'; }
  { const char* e = getenv("UB_NO_STEM_FOLD"); ctx->stem_fold = !(e && e[0] == '1'); }
  { const char* e = getenv("UB_DBG"); ctx->dbg = e ? atoi(e) : 0; }
  { const char* e = getenv("UB_RS"); ctx->no_rs = !(e && e[0] == '1'); }
  { const char* e = getenv("UB_FUSE_ENC1"); const char* n = getenv("UB_NO_FUSE_ENC1");   // default on
    ctx->fuse_enc1 = !(e && e[0] == '0') && !(n && n[0] == '1'); }
  { const char* e = getenv("UB_FUSE_DEC4"); const char* n = getenv("UB_NO_FUSE_DEC4");   // default on
    ctx->fuse_dec4 = !(e && e[0] == '0') && !(n && n[0] == '1'); }
  { const char* e = getenv("UB_NO_FUSE_RESIZE"); ctx->fuse_resize = !(e && e[0] == '1'); }
  { const char* e = getenv("UB_STEM_ZERO_PLANE"); ctx->stem_zero_plane = e && e[0] == '1'; }
  { const char* e = getenv("UB_RS2_NO_SPARE"); ctx->rs2_no_spare = e && e[0] == '1'; }
  { const char* e = getenv("UB_PAIR_RESIDENT"); ctx->pair_resident = e && e[0] == '1'; }
  const char* nc = getenv("UB_NO_CG2");
  ctx->no_cg2 = nc && nc[0] == '1';
  *out = ctx.release();
  return UB_OK;
}

void ub_destroy(ub_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  for (void* p : ctx->owned) cudaFree(p);
  for (auto& r : ctx->prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  for (cudaEvent_t e : ctx->event_pool) cudaEventDestroy(e);
  delete ctx->vit;
  delete ctx;
}

const char* ub_last_error(const ub_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }
int64_t ub_launch_count(const ub_ctx* ctx) { return ctx ? ctx->launches : 0; }

int ub_profile_begin(ub_ctx* ctx) {
  if (!ctx) return UB_ERR_INVALID;
  for (auto& r : ctx->prof) { ctx->event_pool.push_back(r.a); ctx->event_pool.push_back(r.b); }
  ctx->prof.clear();
  ctx->profiling = true;
  return UB_OK;
}

int ub_profile_end(ub_ctx* ctx, double* gemm_ms_total, int64_t* gemm_launches, char* report, int report_len) {
  if (!ctx) return UB_ERR_INVALID;
  ctx->profiling = false;
  UB_CUDA(ctx, cudaDeviceSynchronize());
  std::vector<double> ms(ctx->layers.size(), 0.0);
  std::vector<int64_t> cnt(ctx->layers.size(), 0);
  double total = 0.0;
  for (auto& r : ctx->prof) {
    float t = 0.f;
    UB_CUDA(ctx, cudaEventElapsedTime(&t, r.a, r.b));
    total += t;
    if (r.layer >= 0) { ms[r.layer] += t; cnt[r.layer]++; }
  }
  if (gemm_ms_total) *gemm_ms_total = total;
  if (gemm_launches) *gemm_launches = static_cast<int64_t>(ctx->prof.size());
  if (report && report_len > 0) {
    int off = 0;
    report[0] = 0;
    for (size_t i = 0; i < ctx->layers.size() && off < report_len - 96; ++i)
      off += snprintf(report + off, report_len - off, "%s %d %d %d %lld %.4f\n", ctx->layers[i].key.c_str(), ctx->layers[i].cin,
                      ctx->layers[i].n_total, ctx->layers[i].ksize, static_cast<long long>(cnt[i]), ms[i]);
  }
  for (auto& r : ctx->prof) { ctx->event_pool.push_back(r.a); ctx->event_pool.push_back(r.b); }
  ctx->prof.clear();
  return UB_OK;
}

int ub_load_weights(ub_ctx* ctx, int n, const char* const* names, const void* const* host_ptrs,
                    const int64_t* const* shapes, const int* ndims) {
  if (!ctx || n <= 0 || !names || !host_ptrs || !shapes || !ndims) return fail(ctx, UB_ERR_INVALID, "bad arguments");
  if (ctx->weights_ready) return fail(ctx, UB_ERR_STATE, "weights already loaded; create a new context");
  UB_CUDA(ctx, cudaSetDevice(ctx->device));
  for (int i = 0; i < n; ++i) {
    std::string k(names[i]);
    const std::string pre = "_orig_mod.";  // torch.compile prefix stripped by every reference loader
    if (k.compare(0, pre.size(), pre) == 0) k = k.substr(pre.size());
    if (k.size() > 19 && k.compare(k.size() - 19, 19, "num_batches_tracked") == 0) continue;
    HostTensor t;
    size_t cnt = 1;
    for (int d = 0; d < ndims[i]; ++d) { t.shape.push_back(shapes[i][d]); cnt *= static_cast<size_t>(shapes[i][d]); }
    t.data.assign(static_cast<const float*>(host_ptrs[i]), static_cast<const float*>(host_ptrs[i]) + cnt);
    ctx->sd[k] = std::move(t);
  }
  const int rc = pack_all(ctx);
  ctx->sd.clear();
  return rc;
}

}  // extern "C"  (the shared body of the two evaluation entries follows)
namespace {
int eval_impl(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format, const void* targets,
              int target_format, const void* aux, const float* clip_feats, int n, int h, int w, int ref_batch,
              int64_t* cm, double* dice_sums, uint8_t* pred_out, float* logits_out, float* perturbed_out,
              void* stream, const float* prompt = nullptr) {
  int rc;
  if ((rc = check_ready(ctx))) return rc;
  if (ctx->has_clip() && !clip_feats)
    return fail(ctx, UB_ERR_INVALID, "this model needs the CLIP embeddings: call ub_eval_batches_clip / ub_forward_prompt");
  if (!ctx->has_clip() && clip_feats)
    return fail(ctx, UB_ERR_INVALID, "clip_feats given to a model without a CLIP fusion");
  if ((ctx->arch == UB_ARCH_CLIP_UNET_PROMPT) != (prompt != nullptr))
    return fail(ctx, UB_ERR_INVALID, "a prompt heat-map goes with UB_ARCH_CLIP_UNET_PROMPT (ub_forward_prompt) and only with it");
  if (ctx->is_half())
    return fail(ctx, UB_ERR_INVALID, "UB_ARCH_ENCODER / UB_ARCH_SKIP_DECODER compute through ub_encoder_forward / ub_skip_decoder_forward");
  if (!images || n <= 0 || ref_batch <= 0) return fail(ctx, UB_ERR_INVALID, "bad arguments");
  if (ctx->arch == UB_ARCH_AUTOENCODER && (targets || cm || dice_sums || pred_out))
    return fail(ctx, UB_ERR_INVALID, "UB_ARCH_AUTOENCODER reconstructs images (ub_forward); it has no argmax / confusion counts");
  if (image_format != UB_IMG_F32_NCHW && image_format != UB_IMG_U8_NHWC) return fail(ctx, UB_ERR_INVALID, "bad image_format");
  if (targets && !cm) return fail(ctx, UB_ERR_INVALID, "targets given without a confusion-matrix buffer");
  if (dice_sums && !targets) return fail(ctx, UB_ERR_INVALID, "dice_sums needs targets");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Plan* pl;
  if ((rc = build_plan(ctx, h, w, st, &pl))) return rc;
  ub::PerturbDev P;
  if ((rc = to_dev_perturb(ctx, p, aux, &P))) return rc;
  const int C = ctx->n_classes;
  const int n_batches = (n + ref_batch - 1) / ref_batch;
  if (cm) UB_CUDA(ctx, cudaMemsetAsync(cm, 0, sizeof(int64_t) * n_batches * C * C, st));
  if (dice_sums) UB_CUDA(ctx, cudaMemsetAsync(dice_sums, 0, sizeof(double) * n_batches * C * 3, st));
  const size_t hw = static_cast<size_t>(h) * w;
  const size_t img_stride = image_format == UB_IMG_U8_NHWC ? hw * 3 : hw * 3 * 4;
  const size_t tgt_stride = target_format == UB_TGT_U8 ? hw : hw * 8;
  // Micro-batches.  A layer's launch is a whole number of waves only when (images x tiles per image) is a multiple of the
  // SM count; the coarsest level has the fewest tiles per image (16 at 256x256: 2 x 4 tiles x 2 n-tiles), so its quantum
  // q = SMs / gcd(SMs, tiles) = 37 images decides.  Sizes are multiples of q wherever the image count allows and the
  // remainder rides on one micro-batch (a rank's share of a sharded set — 460 images on 8 GPUs — is 111+111+111+127
  // instead of 4 x 115: the 32x32 layers then run 12 + 12 + 12 + 14 full waves instead of 4 x 12.4 -> 13).
  std::vector<int> mb_sizes;
  {
    const int lv = ctx->n_enc;
    const int tiles = (((h >> lv) + 15) / 16) * (((w >> lv) + 7) / 8) * ((64 << lv) / 256 > 0 ? (64 << lv) / 256 : 1);
    int a = ctx->num_sms, b = tiles;
    while (b) { const int t = a % b; a = b; b = t; }
    const int q = ctx->num_sms / a;
    const int mb_count = (n + ctx->max_mb - 1) / ctx->max_mb;
    const int base = q > 1 && q <= ctx->max_mb ? (n / mb_count) / q * q : 0;
    if (base == 0) {  // balanced: ceil(n / max_mb) launches of (almost) equal size instead of a short last one
      const int sz = (n + mb_count - 1) / mb_count;
      for (int i0 = 0; i0 < n; i0 += sz) mb_sizes.push_back((n - i0) < sz ? (n - i0) : sz);
    } else {
      mb_sizes.assign(mb_count, base);
      int r = n - base * mb_count;
      for (bool grew = true; grew && r >= q;) {
        grew = false;
        for (int i = 0; i < mb_count && r >= q; ++i)
          if (mb_sizes[i] + q <= ctx->max_mb) { mb_sizes[i] += q; r -= q; grew = true; }
      }
      for (int i = mb_count - 1; i >= 0 && r > 0; --i) {
        const int add = (ctx->max_mb - mb_sizes[i]) < r ? (ctx->max_mb - mb_sizes[i]) : r;
        mb_sizes[i] += add; r -= add;
      }
      if (r > 0) mb_sizes.push_back(r);
    }
  }
  int i0 = 0;
  for (size_t mbi = 0; mbi < mb_sizes.size(); i0 += mb_sizes[mbi], ++mbi) {
    const int nb = mb_sizes[mbi];
    ub::PerturbDev Pm = P;
    if (aux) {
      size_t per_img = 0;
      switch (P.kind) {
        case UB_P_U8_NOISE: case UB_P_F_NOISE: per_img = hw * 3 * 4; break;
        case UB_P_U8_OCCLUSION: case UB_P_F_OCCLUSION: per_img = 8; break;
        case UB_P_U8_SALT_PEPPER: per_img = hw; break;
        case UB_P_F_SALT_PEPPER: per_img = hw * 4; break;
        default: break;
      }
      Pm.aux = static_cast<const uint8_t*>(aux) + per_img * i0;
    }
    // device RNG: decorrelate micro-batches
    Pm.seed = P.seed + 0x9E3779B97F4A7C15ull * static_cast<unsigned long long>(i0);
    HeadArgs ha;
    ha.targets = targets ? static_cast<const uint8_t*>(targets) + tgt_stride * i0 : nullptr;
    ha.target_fmt = target_format; ha.ref_batch = ref_batch; ha.img_base = i0;
    ha.cm = reinterpret_cast<unsigned long long*>(cm);
    ha.dice = dice_sums;
    ha.pred = pred_out ? pred_out + hw * i0 : nullptr;
    ha.logits = logits_out ? logits_out + hw * C * i0 : nullptr;
    ha.clip = clip_feats ? clip_feats + static_cast<size_t>(2 * 32 * (1 << ctx->n_enc)) * i0 : nullptr;
    ha.prompt = prompt ? prompt + hw * i0 : nullptr;
    if ((rc = run_microbatch(ctx, pl, Pm, static_cast<const uint8_t*>(images) + img_stride * i0, image_format, nb, ha,
                             perturbed_out ? perturbed_out + hw * 3 * i0 : nullptr, st)))
      return rc;
  }
  if (dice_sums) {
    ub::dice_tsum_kernel<<<blocks_for(static_cast<size_t>(n_batches) * C, 128), 128, 0, st>>>(
        reinterpret_cast<const unsigned long long*>(cm), dice_sums, n_batches, C);
    UB_CUDA(ctx, cudaGetLastError());
    ctx->launches++;
  }
  return UB_OK;
}
}  // namespace
extern "C" {

int ub_eval_batches(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format, const void* targets,
                    int target_format, const void* aux, int n, int h, int w, int ref_batch, int64_t* cm,
                    double* dice_sums, uint8_t* pred_out, float* logits_out, float* perturbed_out, void* stream) {
  return eval_impl(ctx, p, images, image_format, targets, target_format, aux, nullptr, n, h, w, ref_batch, cm, dice_sums,
                   pred_out, logits_out, perturbed_out, stream);
}

int ub_eval_batches_clip(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format, const void* targets,
                         int target_format, const void* aux, const float* clip_feats, int n, int h, int w,
                         int ref_batch, int64_t* cm, double* dice_sums, uint8_t* pred_out, float* logits_out,
                         float* perturbed_out, void* stream) {
  if (!clip_feats) return fail(ctx, UB_ERR_INVALID, "clip_feats is null");
  return eval_impl(ctx, p, images, image_format, targets, target_format, aux, clip_feats, n, h, w, ref_batch, cm,
                   dice_sums, pred_out, logits_out, perturbed_out, stream);
}

int ub_forward_prompt(ub_ctx* ctx, const float* x_nchw, const float* prompt_heatmap, const float* clip_feats,
                      float* logits_nchw, int n, int h, int w, void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  if (!logits_nchw || !prompt_heatmap || !clip_feats) return fail(ctx, UB_ERR_INVALID, "null argument");
  return eval_impl(ctx, nullptr, x_nchw, UB_IMG_F32_NCHW, nullptr, UB_TGT_I64, nullptr, clip_feats, n, h, w, 1, nullptr,
                   nullptr, nullptr, logits_nchw, nullptr, stream, prompt_heatmap);
}

int ub_forward(ub_ctx* ctx, const float* x_nchw, float* logits_nchw, int n, int h, int w, void* stream) {
  if (!logits_nchw) return fail(ctx, UB_ERR_INVALID, "logits buffer is null");
  return ub_eval_batches(ctx, nullptr, x_nchw, UB_IMG_F32_NCHW, nullptr, UB_TGT_I64, nullptr, n, h, w, 1, nullptr, nullptr,
                         nullptr, logits_nchw, nullptr, stream);
}

// One micro-batch loop shared by the two halves: `encode` runs stem + encoder steps and copies the features out,
// otherwise the features are copied in and the decoder steps run.
static int run_half(ub_ctx* ctx, bool encode, const float* x_nchw, float* const* feats_out, const float* const* feats_in,
                    int n_feats, float* logits, int n, int h, int w, void* stream) {
  int rc;
  if ((rc = check_ready(ctx))) return rc;
  if (ctx->arch != (encode ? UB_ARCH_ENCODER : UB_ARCH_SKIP_DECODER))
    return fail(ctx, UB_ERR_INVALID, encode ? "ub_encoder_forward needs a UB_ARCH_ENCODER context"
                                            : "ub_skip_decoder_forward needs a UB_ARCH_SKIP_DECODER context");
  if (n <= 0 || n_feats != ctx->n_enc + 2 || (encode ? (!x_nchw || !feats_out) : (!feats_in || !logits)))
    return fail(ctx, UB_ERR_INVALID, "bad arguments (n_feats must be %d)", ctx->n_enc + 2);
  for (int f = 0; f < n_feats; ++f)
    if (!(encode ? static_cast<const void*>(feats_out[f]) : static_cast<const void*>(feats_in[f])))
      return fail(ctx, UB_ERR_INVALID, "feature map %d is null", f);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Plan* pl;
  if ((rc = build_plan(ctx, h, w, st, &pl))) return rc;
  const size_t hw = static_cast<size_t>(h) * w;
  ub::PerturbDev none;
  memset(&none, 0, sizeof(none));
  for (int i0 = 0; i0 < n; i0 += ctx->max_mb) {
    const int nb = (n - i0) < ctx->max_mb ? (n - i0) : ctx->max_mb;
    HeadArgs ha;
    ha.logits = logits ? logits + hw * ctx->n_classes * i0 : nullptr;
    if (!encode) {
      for (int f = 0; f < n_feats; ++f) {
        const Plan::Feat& F = pl->feats[f];
        const size_t fhw = hw >> (2 * F.level);
        ub::block_from_nchw_kernel<<<blocks_for(static_cast<size_t>(nb) * (F.c / 8) * fhw, 256), 256, 0, st>>>(
            feats_in[f] + static_cast<size_t>(i0) * F.c * fhw, F.c, fhw, nb, F.buf, F.cgs, F.cg0);
        ctx->launches++;
      }
      UB_CUDA(ctx, cudaGetLastError());
    }
    if ((rc = run_microbatch(ctx, pl, none, encode ? x_nchw + hw * 3 * i0 : nullptr, UB_IMG_F32_NCHW, nb, ha, nullptr, st)))
      return rc;
    if (encode) {
      for (int f = 0; f < n_feats; ++f) {
        const Plan::Feat& F = pl->feats[f];
        const size_t fhw = hw >> (2 * F.level);
        ub::unblock_to_nchw_kernel<<<blocks_for(static_cast<size_t>(nb) * (F.c / 8) * fhw, 256), 256, 0, st>>>(
            F.buf, F.cgs, F.cg0, F.c, fhw, nb, feats_out[f] + static_cast<size_t>(i0) * F.c * fhw);
        ctx->launches++;
      }
      UB_CUDA(ctx, cudaGetLastError());
    }
  }
  return UB_OK;
}

int ub_encoder_forward(ub_ctx* ctx, const float* x_nchw, int n, int h, int w, float* const* feats, int n_feats, void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  return run_half(ctx, true, x_nchw, feats, nullptr, n_feats, nullptr, n, h, w, stream);
}

int ub_skip_decoder_forward(ub_ctx* ctx, const float* const* feats, int n_feats, float* logits_nchw, int n, int h, int w,
                            void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  return run_half(ctx, false, nullptr, nullptr, feats, n_feats, logits_nchw, n, h, w, stream);
}

int ub_perturb_images(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format, const void* aux, int n,
                      int h, int w, float* out_nchw, void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  if (!images || !out_nchw || n <= 0) return fail(ctx, UB_ERR_INVALID, "bad arguments");
  UB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ub::PerturbDev P;
  int rc;
  if ((rc = to_dev_perturb(ctx, p, aux, &P))) return rc;
  const size_t hw = static_cast<size_t>(h) * w;
  if (P.kind == UB_P_U8_BLUR || P.kind == UB_P_F_BLUR) {
    Plan tmp;
    tmp.H = h; tmp.W = w;
    if (P.iparam <= ub::kBlurMaxPasses && !ctx->no_blur_fuse)  // every pass in shared memory: no planes needed
      return run_blur(ctx, &tmp, P, images, image_format, n, out_nchw, nullptr, 0, 0, nullptr, st);
    // pass-per-launch path: ping-pong planes, allocated per call on the stream (not on the evaluation hot path)
    void *a = nullptr, *b = nullptr;
    const size_t bytes = static_cast<size_t>(n) * 3 * hw * 4;
    UB_CUDA(ctx, cudaMallocAsync(&a, bytes, st));
    UB_CUDA(ctx, cudaMallocAsync(&b, bytes, st));
    tmp.blur_a = a; tmp.blur_b = b;
    rc = run_blur(ctx, &tmp, P, images, image_format, n, out_nchw, nullptr, 0, 0, nullptr, st);
    cudaFreeAsync(a, st);
    cudaFreeAsync(b, st);
    return rc;
  }
  ub::perturb_kernel<<<blocks_for(hw * n, 256), 256, 0, st>>>(P, images, image_format, n, h, w, out_nchw);
  UB_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return UB_OK;
}

}  // extern "C"

// ---- ub_blur_images_static: the pass-per-launch blur with its ping-pong planes in a file-scope device array -----------
// For callers that may not allocate (CUDA graph capture, tight serving loops).  No counterpart in the reference.
// Two planes of UB_STATIC_BLUR_MAX_ELEMS fp32-sized elements each; every GPU that loads the library holds them (8 MiB)
// until the process exits.  Every call on a device shares them: calls that overlap on the GPU race.  Nothing in them has
// to start at zero — the load / quantise kernel writes all n*3*h*w elements of plane A before any pass reads them and
// every pass writes all of its destination — so a call never sees what the previous one left.
namespace {
static __device__ __align__(256) unsigned char g_static_blur_planes[2 * sizeof(float) * UB_STATIC_BLUR_MAX_ELEMS];
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_blur_mu;
void* g_static_blur_dev_ptr[kStaticMaxDevices] = {};  // device address of the array, per device (the symbol itself is
                                                      // a host-side stand-in and must never reach a kernel)

// Once per device, with that device current: the array's device address.
int static_blur_planes(ub_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= kStaticMaxDevices) return fail(ctx, UB_ERR_INVALID, "device index %d out of range", ctx->device);
  std::lock_guard<std::mutex> lock(g_static_blur_mu);
  void*& slot = g_static_blur_dev_ptr[ctx->device];
  if (!slot) {
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_blur_planes);
    if (e != cudaSuccess || !ptr)
      return fail(ctx, UB_ERR_CUDA, "cudaGetSymbolAddress(static blur planes) on device %d: %s", ctx->device, cudaGetErrorString(e));
    slot = ptr;
  }
  *out = static_cast<unsigned char*>(slot);
  return UB_OK;
}
}  // namespace

extern "C" {

int ub_blur_images_static(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format, int n, int h, int w,
                          float* out_nchw, void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  if (!p || !images || !out_nchw || n <= 0 || h <= 0 || w <= 0) return fail(ctx, UB_ERR_INVALID, "bad arguments");
  if (p->kind != UB_P_U8_BLUR && p->kind != UB_P_F_BLUR)
    return fail(ctx, UB_ERR_INVALID, "ub_blur_images_static takes the blur families only (kind %d)", p->kind);
  const unsigned long long count = 3ull * static_cast<unsigned long long>(n) * static_cast<unsigned long long>(h) *
                                   static_cast<unsigned long long>(w);
  if (count > UB_STATIC_BLUR_MAX_ELEMS)
    return fail(ctx, UB_ERR_INVALID,
                "ub_blur_images_static: n*3*h*w = %llu elements exceeds the static planes' %llu (UB_STATIC_BLUR_MAX_ELEMS); "
                "split the call or use ub_perturb_images",
                count, static_cast<unsigned long long>(UB_STATIC_BLUR_MAX_ELEMS));
  UB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ub::PerturbDev P;
  int rc;
  if ((rc = to_dev_perturb(ctx, p, nullptr, &P))) return rc;
  unsigned char* planes = nullptr;
  if ((rc = static_blur_planes(ctx, &planes))) return rc;
  unsigned char* plane_a = planes;
  unsigned char* plane_b = planes + sizeof(float) * UB_STATIC_BLUR_MAX_ELEMS;
  const unsigned g = blocks_for(static_cast<size_t>(count), 256);
  if (P.kind == UB_P_U8_BLUR) {
    uint8_t* a = plane_a;
    uint8_t* b = plane_b;
    ub::blur_u8_quantise_kernel<<<g, 256, 0, st>>>(images, image_format, n, h, w, a);
    ctx->launches++;
    for (int i = 0; i < P.iparam; ++i) {
      ub::blur_u8_pass_kernel<<<g, 256, 0, st>>>(a, b, n * 3, h, w);
      ctx->launches++;
      std::swap(a, b);
    }
    ub::u8_planar_to_f32_kernel<<<g, 256, 0, st>>>(a, out_nchw, static_cast<size_t>(count));
    ctx->launches++;
  } else {
    float* a = reinterpret_cast<float*>(plane_a);
    float* b = reinterpret_cast<float*>(plane_b);
    if (P.iparam == 0) a = out_nchw;
    ub::blur_f_load_kernel<<<g, 256, 0, st>>>(images, image_format, n, h, w, a);
    ctx->launches++;
    for (int i = 0; i < P.iparam; ++i) {
      float* o = (i == P.iparam - 1) ? out_nchw : b;
      ub::blur_f_pass_kernel<<<g, 256, 0, st>>>(a, o, n * 3, h, w);
      ctx->launches++;
      b = a; a = o;
    }
  }
  UB_CUDA(ctx, cudaGetLastError());
  return UB_OK;
}

int ub_clip_image_features(ub_ctx* ctx, const float* x_nchw, int n, int h, int w, float* feats, void* stream) {
  int rc;
  if ((rc = check_ready(ctx))) return rc;
  if (ctx->arch != UB_ARCH_CLIP_VIT || !ctx->vit) return fail(ctx, UB_ERR_STATE, "ub_clip_image_features needs a UB_ARCH_CLIP_VIT context");
  if (n == 0) return UB_OK;
  if (!x_nchw || !feats || n < 0 || h <= 0 || w <= 0) return fail(ctx, UB_ERR_INVALID, "bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t img = static_cast<size_t>(3) * h * w;
  const int chunk = ctx->vit->chunk;
  for (int i0 = 0; i0 < n; i0 += chunk) {
    const int nb = (n - i0) < chunk ? (n - i0) : chunk;
    if ((rc = vit_forward_chunk(ctx, x_nchw + img * i0, nb, h, w, feats + static_cast<size_t>(ctx->vit->E) * i0, st))) return rc;
  }
  return UB_OK;
}

int ub_clip_resize_taps(int in_size, int out_size, int32_t* start, int32_t* count, float* weights) {
  if (in_size <= 0 || out_size <= 0 || !start || !count || !weights) return UB_ERR_INVALID;
  std::vector<ub::ResizeTap> taps;
  const int rc = make_resize_taps(nullptr, in_size, out_size, &taps);
  if (rc) return rc;
  for (int i = 0; i < out_size; ++i) {
    start[i] = taps[i].start;
    count[i] = taps[i].count;
    memcpy(weights + static_cast<size_t>(i) * UB_CLIP_MAX_TAPS, taps[i].w, sizeof(float) * UB_CLIP_MAX_TAPS);
  }
  return UB_OK;
}

int ub_clip_dims(const ub_ctx* ctx, int* image_size, int* embed_dim, int* chunk) {
  if (!ctx || !ctx->vit) return UB_ERR_STATE;
  if (image_size) *image_size = ctx->vit->S;
  if (embed_dim) *embed_dim = ctx->vit->E;
  if (chunk) *chunk = ctx->vit->chunk;
  return UB_OK;
}

int ub_debug_clip_preprocess(ub_ctx* ctx, const float* x_nchw, int n, int h, int w, float* out_nchw, void* stream) {
  int rc;
  if ((rc = check_ready(ctx))) return rc;
  if (ctx->arch != UB_ARCH_CLIP_VIT || !ctx->vit) return fail(ctx, UB_ERR_STATE, "ub_debug_clip_preprocess needs a UB_ARCH_CLIP_VIT context");
  if (!x_nchw || !out_nchw || n <= 0 || h <= 0 || w <= 0) return fail(ctx, UB_ERR_INVALID, "bad arguments");
  ub::ResizeTap *ry = nullptr, *rx = nullptr;
  int max_taps = ub::kClipMaxTaps;
  if ((rc = vit_taps(ctx, h, w, &ry, &rx, &max_taps))) return rc;
  const int S = ctx->vit->S;
  const unsigned blocks = blocks_for(static_cast<size_t>(n) * 3 * S * S, 256);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define UB_PRE(T)                                                                                                              \
  ub::clip_preprocess_kernel<T><<<blocks, 256, 0, st>>>(x_nchw, n, h, w, ry, rx, S, kClipMean[0], kClipMean[1], kClipMean[2], \
                                                        1.0f / kClipStd[0], 1.0f / kClipStd[1], 1.0f / kClipStd[2], out_nchw)
  if (max_taps <= 2) UB_PRE(2);  // (the same dispatch as the patch kernel: vit_forward_chunk)
  else if (max_taps <= 3) UB_PRE(3);
  else if (max_taps <= 4) UB_PRE(4);
  else if (max_taps <= 6) UB_PRE(6);
  else if (max_taps <= 8) UB_PRE(8);
  else UB_PRE(ub::kClipMaxTaps);
#undef UB_PRE
  ctx->launches++;
  UB_CUDA(ctx, cudaGetLastError());
  return UB_OK;
}

int ub_decode_masks(ub_ctx* ctx, const uint8_t* raw_masks, uint8_t* masks_out, int n, int64_t pixels_per_image,
                    void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  if (n == 0) return UB_OK;  // empty set: nothing to read, pointers may be NULL
  if (!raw_masks || !masks_out || n < 0 || pixels_per_image <= 0) return fail(ctx, UB_ERR_INVALID, "bad arguments");
  UB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ub::decode_masks_kernel<<<n, 1024, 0, st>>>(raw_masks, masks_out, static_cast<size_t>(pixels_per_image));
  UB_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return UB_OK;
}

int ub_pack_dataset(ub_ctx* ctx, const float* images_nchw, const int64_t* masks_i64, int n, int h, int w,
                    uint8_t* images_u8_nhwc, uint8_t* masks_u8, void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  if (n == 0) return UB_OK;
  if (n < 0 || h <= 0 || w <= 0 || (images_nchw && !images_u8_nhwc) || (masks_i64 && !masks_u8) ||
      (!images_nchw && !masks_i64))
    return fail(ctx, UB_ERR_INVALID, "bad arguments");
  UB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t hw = static_cast<size_t>(h) * w;
  ub::pack_dataset_kernel<<<blocks_for(hw * n, 256), 256, 0, st>>>(
      images_nchw, reinterpret_cast<const long long*>(masks_i64), n, hw, images_u8_nhwc, masks_u8);
  UB_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return UB_OK;
}

int ub_metrics(ub_ctx* ctx, const float* logits_nchw, const void* targets, int target_format, int n, int n_classes,
               int h, int w, int ref_batch, int64_t* cm, double* dice_sums, void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  if (!logits_nchw || !targets || !cm || n <= 0 || ref_batch <= 0 || n_classes < 1 || n_classes > ub::kMaxClasses)
    return fail(ctx, UB_ERR_INVALID, "bad arguments");
  UB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int n_batches = (n + ref_batch - 1) / ref_batch;
  UB_CUDA(ctx, cudaMemsetAsync(cm, 0, sizeof(int64_t) * n_batches * n_classes * n_classes, st));
  if (dice_sums) UB_CUDA(ctx, cudaMemsetAsync(dice_sums, 0, sizeof(double) * n_batches * n_classes * 3, st));
  dim3 grid(blocks_for(static_cast<size_t>(h) * w, 256), n);
  ub::metrics_kernel<<<grid, 256, 0, st>>>(logits_nchw, targets, target_format, n_classes, h, w, ref_batch,
                                           reinterpret_cast<unsigned long long*>(cm), dice_sums);
  UB_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  if (dice_sums) {
    ub::dice_tsum_kernel<<<blocks_for(static_cast<size_t>(n_batches) * n_classes, 128), 128, 0, st>>>(
        reinterpret_cast<const unsigned long long*>(cm), dice_sums, n_batches, n_classes);
    UB_CUDA(ctx, cudaGetLastError());
    ctx->launches++;
  }
  return UB_OK;
}

static int debug_layer(ub_ctx* ctx, const char* key, const void* in, void* out, int n, int h, int w, int mode, void* stream) {
  int rc;
  if ((rc = check_ready(ctx))) return rc;
  if (!key || !in || !out || n <= 0) return fail(ctx, UB_ERR_INVALID, "bad arguments");
  auto it = ctx->layer_by_key.find(key);
  if (it == ctx->layer_by_key.end()) return fail(ctx, UB_ERR_INVALID, "no such layer: %s", key);
  const GemmLayer& L = ctx->layers[it->second];
  ConvParams cp;
  if ((rc = fill_gemm(ctx, &cp, L, static_cast<const __nv_bfloat16*>(in), L.cin / 8, n, h, w, mode,
                      static_cast<__nv_bfloat16*>(out), L.cout / 8, 0)))
    return rc;
  return launch_gemm(ctx, cp, L.block_n, L.bk, static_cast<cudaStream_t>(stream), it->second);
}

int ub_debug_conv3x3(ub_ctx* ctx, const char* conv_key, const void* in_bf16_cg8, void* out_bf16_cg8, int n, int h,
                     int w, int pool, void* stream) {
  if (ctx && conv_key) {
    auto it = ctx->layer_by_key.find(conv_key);
    if (it != ctx->layer_by_key.end() && ctx->layers[it->second].is_convt) return fail(ctx, UB_ERR_INVALID, "%s is a ConvTranspose", conv_key);
  }
  return debug_layer(ctx, conv_key, in_bf16_cg8, out_bf16_cg8, n, h, w, pool ? ub::EPI_POOL : ub::EPI_STORE, stream);
}

int ub_debug_convt2x2(ub_ctx* ctx, const char* up_key, const void* in_bf16_cg8, void* out_bf16_cg8, int n, int h,
                      int w, void* stream) {
  if (ctx && up_key) {
    auto it = ctx->layer_by_key.find(up_key);
    if (it != ctx->layer_by_key.end() && !ctx->layers[it->second].is_convt) return fail(ctx, UB_ERR_INVALID, "%s is not a ConvTranspose", up_key);
  }
  return debug_layer(ctx, up_key, in_bf16_cg8, out_bf16_cg8, n, h, w, ub::EPI_CONVT, stream);
}

int ub_debug_resample(ub_ctx* ctx, const void* in_bf16_cg8, void* out_bf16_cg8, int n, int c, int hi, int wi, int ho,
                      int wo, void* stream) {
  if (!ctx) return UB_ERR_INVALID;
  if (!in_bf16_cg8 || !out_bf16_cg8 || n <= 0 || c <= 0 || c % 8 || hi <= 0 || wi <= 0 || ho <= 0 || wo <= 0)
    return fail(ctx, UB_ERR_INVALID, "bad arguments");
  UB_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t total = static_cast<size_t>(n) * ho * wo * (c / 8);
  ub::resample_bilinear_kernel<<<blocks_for(total, 256), 256, 0, st>>>(
      static_cast<const __nv_bfloat16*>(in_bf16_cg8), n, hi, wi, c, static_cast<__nv_bfloat16*>(out_bf16_cg8), ho, wo,
      c / 8, 0);
  UB_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return UB_OK;
}

}  // extern "C"
