/* unet_b200.h — C ABI of libunetb200.so, the sm_100a (B200) implementation of the robustness-
 * evaluation hot path of MattiDeBeer/image-segmentation:
 *
 *   perturbation -> U-Net forward -> per-pixel argmax -> per-reference-batch confusion counts
 *
 * Nothing like this interface exists in the reference (it is pure Python); each entry point names
 * the reference code it replaces (paths relative to the reference root).  Plain pointers and sizes
 * only: no torch / C++ types cross this boundary.  All device pointers are owned by the caller; the
 * library owns the packed weights and its workspace inside ub_ctx.  Every call is asynchronous on
 * the passed stream (a cudaStream_t / CUstream handle passed as void*), never synchronises, and
 * returns 0 or a negative ub_status; ub_last_error() gives the message.  One ctx per device per
 * process; a ctx is not thread-safe (the reference's callers are single-threaded).
 * There is no CPU fallback: ub_create fails on a device that is not compute capability 10.x.
 */
#ifndef UNET_B200_H_
#define UNET_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UB_ABI_VERSION 1

typedef struct ub_ctx ub_ctx;

enum ub_status {
  UB_OK = 0,
  UB_ERR_INVALID = -1,     /* bad argument / shape / missing weight */
  UB_ERR_CUDA = -2,        /* a CUDA call failed; see ub_last_error */
  UB_ERR_UNSUPPORTED = -3, /* wrong GPU architecture, unsupported size */
  UB_ERR_STATE = -4        /* weights not loaded yet */
};

/* models/UNet.py:7-76 (UNet) and :78-148 (LargeUNet) */
enum ub_arch {
  UB_ARCH_UNET = 0,       /* models/UNet.py:7-76 */
  UB_ARCH_LARGE_UNET = 1, /* models/UNet.py:78-148 */
  UB_ARCH_CLIP_UNET = 2,  /* models/CLIP_models.py:89-134: UNet encoder/decoder, bottleneck = cross-attention fusion of the
                             CLIP image embedding (the embedding itself is computed by the caller) */
  UB_ARCH_CLIP_UNET_PROMPT = 3, /* models/prompt_segmentation.py:31-95: ClipUnet plus the PromptEncoder (:16-29) over a
                                  1-channel prompt heat-map and the 1x1 prompt_fusion of the two 512-channel maps */
  UB_ARCH_AUTOENCODER = 4, /* models/autoencoder.py:5-58: Encoder -> Decoder (three ConvBlockUpsample stages without skips,
                              1x1 conv 64 -> n_classes channels) -> sigmoid; ub_forward returns the reconstruction
                              (after the sigmoid), the evaluation entries do not apply */
  UB_ARCH_ENCODER = 5,      /* models/autoencoder.py:5-26 `Encoder` on its own (weights input.*, enc1-3.*, bottleneck.*):
                              ub_encoder_forward returns its dict of feature maps */
  UB_ARCH_SKIP_DECODER = 6, /* models/pre_trained.py:7-24 `SegmentationDecoderSkip` on its own (weights dec1-4.*, out.*):
                              ub_skip_decoder_forward takes that dict and returns the logits */
  UB_ARCH_CLIP_VIT = 7      /* models/processing_blocks.py:156-233 `ClipFeatureExtractor`: CustomClipPreprocessor (Resize +
                              Normalize) and the CLIP ViT image tower + visual projection of HuggingFace CLIPModel
                              (weights: its state_dict keys "vision_model.*", "visual_projection.weight");
                              ub_clip_image_features.  ub_create's n_classes is ignored (pass 1) */
};

/* Input image container. */
enum ub_image_format {
  UB_IMG_F32_NCHW = 0, /* fp32 [n,3,H,W] in [0,1] — what customDatasets/datasets.py:157-158 yields */
  UB_IMG_U8_NHWC = 1   /* uint8 [n,H,W,3] — the decoded bytes before datasets.py:115 divides by 255 */
};
enum ub_target_format { UB_TGT_I64 = 0 /* int64 [n,H,W], datasets.py:118-131 */, UB_TGT_U8 = 1 };

/* Perturbation families.
 * UB_P_U8_*: uint8-space Dataset wrappers, customDatasets/perturbations.py (image is quantised with
 *            rint(x*255) first, result is k/255.0f) — exact integer arithmetic.
 * UB_P_F_* : float-space nn.Modules, models/processing_blocks.py:454-592. */
enum ub_perturb_kind {
  UB_P_NONE = 0,
  UB_P_U8_NOISE = 1,       /* perturbations.py:31-50;   fparam = std; aux = fp32 [n,3,H,W] N(0,std) or NULL */
  UB_P_U8_BLUR = 2,        /* perturbations.py:85-124;  iparam = passes (0..64; up to 12 run fused in shared memory) */
  UB_P_U8_LUT = 3,         /* perturbations.py:140-156,182-198,222-235,262-275; lut[] filled by ub_make_lut */
  UB_P_U8_OCCLUSION = 4,   /* perturbations.py:301-327; iparam = size; aux = int32 [n,2] (y0,x0), <0 = none, or NULL */
  UB_P_U8_SALT_PEPPER = 5, /* perturbations.py:354-382; fparam = amount; aux = uint8 [n,H,W] 0 keep/1 pepper/2 salt, or NULL */
  UB_P_F_NOISE = 6,        /* processing_blocks.py:471-474; fparam = std; aux = fp32 [n,3,H,W] randn or NULL */
  UB_P_F_BLUR = 7,         /* processing_blocks.py:493-496 (kornia box_blur, reflect); iparam = times (0..64) */
  UB_P_F_CONTRAST = 8,     /* processing_blocks.py:516-518; fparam = factor */
  UB_P_F_BRIGHTNESS = 9,   /* processing_blocks.py:535-539; fparam = fl32(offset/255.0) computed by the caller in double */
  UB_P_F_OCCLUSION = 10,   /* processing_blocks.py:557-563; iparam = size; aux = int32 [n,2] (y,x) or NULL */
  UB_P_F_SALT_PEPPER = 11  /* processing_blocks.py:586-592; fparam = amount; aux = fp32 [n,1,H,W] rand or NULL */
};

typedef struct ub_perturb {
  int32_t kind;      /* ub_perturb_kind */
  int32_t iparam;
  float fparam;
  float fparam2;     /* reserved */
  uint64_t seed;     /* device RNG seed, used only when the family is random and aux == NULL */
  uint8_t lut[256];  /* UB_P_U8_LUT */
} ub_perturb;

enum ub_lut_kind {
  UB_LUT_CONTRAST = 0,            /* clamp(rint(fl32(u*fl32(param))),0,255)   perturbations.py:146-150 */
  UB_LUT_BRIGHTNESS_INCREASE = 1, /* rint(clamp(u+param,0,255))               perturbations.py:228-230 */
  UB_LUT_BRIGHTNESS_DECREASE = 2  /* clamp(rint(u-param),0,255)               perturbations.py:268-270 */
};

/* Host-only helper: fills p->lut and sets p->kind = UB_P_U8_LUT. */
int ub_make_lut(ub_perturb* p, int lut_kind, double param);

int ub_abi_version(void);

/* Replaces nn.Module construction + .to(device) (models/model_wrappers.py:335-338).
 * max_microbatch: images processed per internal pass (workspace ~45 MB per image at 256x256). */
int ub_create(ub_ctx** out, int device, int arch, int in_channels, int n_classes, int max_microbatch);
void ub_destroy(ub_ctx* ctx);
const char* ub_last_error(const ub_ctx* ctx); /* ctx may be NULL: returns the last creation error */

/* Replaces load_state_dict (models/model_wrappers.py:323-332, scripts/robustness_evaluation.py:39-46).
 * names[i] are reference state_dict keys ("enc1.block.0.conv.0.weight", ...), host_ptrs[i] fp32 host
 * tensors in PyTorch layout, shapes[i]/ndims[i] their shapes.  BatchNorm (eval) is folded into the
 * conv weights here; "num_batches_tracked" and unknown keys are ignored; a missing key is an error. */
int ub_load_weights(ub_ctx* ctx, int n, const char* const* names, const void* const* host_ptrs,
                    const int64_t* const* shapes, const int* ndims);

/* Replaces UNet.forward (models/UNet.py:60-76): x fp32 NCHW device -> logits fp32 NCHW device. */
int ub_forward(ub_ctx* ctx, const float* x_nchw, float* logits_nchw, int n, int h, int w, void* stream);

/* Replaces the body of the evaluation hot loop, models/model_wrappers.py:440-463 and
 * scripts/robustness_evaluation.py:116-122, for n images at once: perturb -> forward -> argmax
 * (first index on ties) -> confusion counts.  Image i belongs to reference batch i / ref_batch.
 *   cm        int64  [ceil(n/ref_batch), C, C]   rows = target, cols = prediction; zeroed by the callee
 *   dice_sums double [ceil(n/ref_batch), C, 3]   (sum p*1[t=c], sum p, sum 1[t=c]), p = softmax(softmax(z))
 *                                                — the sums smp DiceLoss needs (models/losses.py:98-100); may be NULL
 *   pred_out  uint8  [n,H,W] argmax map, may be NULL;  logits_out fp32 [n,C,H,W], may be NULL
 *   perturbed_out fp32 [n,3,H,W] the perturbed network input, may be NULL
 * Precondition: every target value is a class id in [0, C).  A foreign label (e.g. 255 "ignore") is dropped from every
 * cell, whereas the reference's IoU (models/losses.py:52-60) still counts the prediction made on that pixel in the
 * union; the reference's dataset only yields {0,1,2} (datasets.py:118-131).  The Python host checks this once when a
 * dataset is made resident (evaluator.py). */
int ub_eval_batches(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format,
                    const void* targets, int target_format, const void* aux, int n, int h, int w,
                    int ref_batch, int64_t* cm, double* dice_sums, uint8_t* pred_out,
                    float* logits_out, float* perturbed_out, void* stream);

/* ub_eval_batches for UB_ARCH_CLIP_UNET: clip_feats fp32 [n, 512] (device) = `ClipFeatureExtractor.forward`'s output
 * for the same n images (models/processing_blocks.py:216-233).  The library applies `CrossAttentionFusion`
 * (processing_blocks.py:310-322; weights `cross_attention_fusion.cross_attn.*` from ub_load_weights) and feeds the
 * result to dec1 with the encoder skips.  The reference's bottleneck ConvBlock output is never used by its forward
 * (CLIP_models.py:124-126) and is not computed. */
int ub_eval_batches_clip(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format,
                         const void* targets, int target_format, const void* aux, const float* clip_feats, int n,
                         int h, int w, int ref_batch, int64_t* cm, double* dice_sums, uint8_t* pred_out,
                         float* logits_out, float* perturbed_out, void* stream);

/* Replaces ClipUnetPrompt.forward (models/prompt_segmentation.py:63-95) for UB_ARCH_CLIP_UNET_PROMPT: x fp32 [n,3,H,W],
 * prompt_heatmap fp32 [n,1,H,W], clip_feats fp32 [n,512] (all device) -> logits fp32 [n,C,H,W] (before `activation`).
 * The library runs stem + enc1-3 on x, the PromptEncoder on the heat-map, the cross-attention fusion of clip_feats,
 * prompt_fusion over their channel concat, dec1-4 and the 1x1 head. */
int ub_forward_prompt(ub_ctx* ctx, const float* x_nchw, const float* prompt_heatmap, const float* clip_feats,
                      float* logits_nchw, int n, int h, int w, void* stream);

/* The two halves of SegmentationModel as separate modules (models/pre_trained.py:47-56: `decoder(encoder(x))`), for
 * callers that take the encoder's features (the reference's pre-training scripts do).  All tensors fp32 NCHW on the
 * device, in the order of Encoder.forward's dict (models/autoencoder.py:17-25):
 *   feats[0] "x0" [n,32,H,W], feats[1] "enc1" [n,64,H/2,W/2], feats[2] "enc2" [n,128,H/4,W/4],
 *   feats[3] "enc3" [n,256,H/8,W/8], feats[4] "bottleneck" [n,512,H/8,W/8].
 * The maps are the library's bf16 activations widened to fp32 (each within one bf16 rounding of the fp32 reference's
 * layer output); handing them back to ub_skip_decoder_forward loses nothing (bf16 -> fp32 -> bf16 is exact). */
int ub_encoder_forward(ub_ctx* ctx, const float* x_nchw, int n, int h, int w, float* const* feats, int n_feats,
                       void* stream);
int ub_skip_decoder_forward(ub_ctx* ctx, const float* const* feats, int n_feats, float* logits_nchw, int n, int h,
                            int w, void* stream);

/* Replaces ClipFeatureExtractor.forward (models/processing_blocks.py:219-233) for UB_ARCH_CLIP_VIT: x fp32 [n,3,H,W]
 * (device; any H, W up to ~5x the CLIP input size) -> feats fp32 [n,E] (device), E = rows of visual_projection.weight.
 *   CustomClipPreprocessor (:156-170): per image torchvision Resize((S,S)) on a tensor = bilinear with antialiasing,
 *   align_corners=False, then Normalize with CLIP's mean / std — evaluated while the patch matrix is written;
 *   CLIPModel.get_image_features (:228-231): patch embedding (Conv2d(3,D,P,stride P), no bias) + class token + position
 *   embedding, pre_layrnorm, L pre-norm encoder layers (LayerNorm -> multi-head self-attention (head width 64) -> +,
 *   LayerNorm -> fc1 -> QuickGELU -> fc2 -> +), post_layernorm of the class token, visual_projection (no bias).
 * The linear maps run on the tensor cores in bf16 with fp32 accumulation; residual stream, LayerNorm statistics and
 * softmax are fp32.  Supported towers: hidden size a multiple of 256 up to 1024 with heads = hidden / 64, MLP width a
 * multiple of 256, patch size a multiple of 8, at most 64 tokens per image (ViT-B/32 at 224x224: 768 / 12 / 3072 / 32 / 50),
 * hidden_act quick_gelu, layer_norm_eps 1e-5.  ub_clip_dims reports S, E and the number of images the library embeds per
 * internal pass (chosen <= max_microbatch so that the GEMM tile counts fill whole waves of the SMs; callers that batch
 * should hand over multiples of it) once the weights are loaded. */
int ub_clip_image_features(ub_ctx* ctx, const float* x_nchw, int n, int h, int w, float* feats, void* stream);
int ub_clip_dims(const ub_ctx* ctx, int* image_size, int* embed_dim, int* chunk);
/* Host-only helper (no GPU needed): the per-axis taps of that Resize for in_size -> out_size, as ATen's
 * _compute_indices_weights_aa builds them for bilinear + antialias (torchvision Resize on a tensor): output coordinate i
 * reads inputs [start[i], start[i] + count[i]) with weights[i*UB_CLIP_MAX_TAPS + j].  Arrays of out_size entries
 * (weights: out_size * UB_CLIP_MAX_TAPS).  UB_ERR_UNSUPPORTED when an output needs more than UB_CLIP_MAX_TAPS inputs. */
#define UB_CLIP_MAX_TAPS 12
int ub_clip_resize_taps(int in_size, int out_size, int32_t* start, int32_t* count, float* weights);
/* Test hook: CustomClipPreprocessor.forward alone (models/processing_blocks.py:163-170): out fp32 [n,3,S,S]. */
int ub_debug_clip_preprocess(ub_ctx* ctx, const float* x_nchw, int n, int h, int w, float* out_nchw, void* stream);

/* Perturbation alone (the Dataset wrappers' __getitem__ / the nn.Modules' forward): out fp32 [n,3,H,W]. */
int ub_perturb_images(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format,
                      const void* aux, int n, int h, int w, float* out_nchw, void* stream);

/* The blur families of ub_perturb_images (UB_P_U8_BLUR / UB_P_F_BLUR, 0..64 passes) for callers that may not allocate
 * device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE.  It runs the pass-per-launch blur (load /
 * quantise, one launch per pass, result as fp32 [n,3,H,W] in out_nchw — bit-equal to ub_perturb_images) with its two
 * ping-pong planes in a file-scope `static __device__` array instead of stream-ordered allocations: the call performs
 * no cudaMalloc / cudaFree / cudaMallocAsync and calls no allocating library routine.
 *   Limit: n*3*h*w <= UB_STATIC_BLUR_MAX_ELEMS (1 Mi elements: 5 images at 256x256, 1 at 512x512); larger calls
 *   return UB_ERR_INVALID with a message naming the limit — split the call or use ub_perturb_images.
 *   Cost: 2 x 4 MiB of device memory on every GPU that loads the library, held until the process exits.
 *   Sharing: every call on a device uses the same planes.  Calls in one stream are ordered and safe; calls that overlap
 *   on the GPU (different streams, no dependency) race.  The planes keep the last call's contents; nothing needs to
 *   start at zero because each call writes every element it later reads. */
#define UB_STATIC_BLUR_MAX_ELEMS (1u << 20)
int ub_blur_images_static(ub_ctx* ctx, const ub_perturb* p, const void* images, int image_format, int n, int h,
                          int w, float* out_nchw, void* stream);

/* Dataset rows on the device (replaces the per-sample CPU work of customDatasets/datasets.py:92-160).
 * ub_decode_masks: raw trimap bytes [n, pixels] (38 cat / 75 dog / 255 border) -> class ids {0,1,2} with the
 *   reference's per-image rule (`_deserialize_datapoint`, datasets.py:118-131).  In place is allowed.
 * ub_pack_dataset: the `.pt` cache / `__getitem__` form (fp32 [n,3,H,W] = k/255, int64 [n,H,W]; datasets.py:64-86,
 *   :115) -> the resident evaluation form (uint8 [n,H,W,3], uint8 [n,H,W]).  Either pair may be NULL. */
int ub_decode_masks(ub_ctx* ctx, const uint8_t* raw_masks, uint8_t* masks_out, int n, int64_t pixels_per_image,
                    void* stream);
int ub_pack_dataset(ub_ctx* ctx, const float* images_nchw, const int64_t* masks_i64, int n, int h, int w,
                    uint8_t* images_u8_nhwc, uint8_t* masks_u8, void* stream);

/* Metrics alone (models/losses.py:43-63,133-154 from materialised logits): fills cm / dice_sums as above. */
int ub_metrics(ub_ctx* ctx, const float* logits_nchw, const void* targets, int target_format, int n,
               int n_classes, int h, int w, int ref_batch, int64_t* cm, double* dice_sums, void* stream);

/* Test / profiling hook: one conv3x3+BN(folded)+ReLU layer (processing_blocks.py:43-45) by its
 * state_dict prefix, e.g. "enc2.block.0.conv.conv.3"; bf16 channel-blocked [n][C/8][h][w][8] in and out (the library's activation layout); pool = fuse MaxPool2d(2,2). */
int ub_debug_conv3x3(ub_ctx* ctx, const char* conv_key, const void* in_bf16_cg8, void* out_bf16_cg8,
                     int n, int h, int w, int pool, void* stream);
/* Test hook: ConvTranspose2d(k2,s2) (processing_blocks.py:102,106) by prefix, e.g. "dec2.up". */
int ub_debug_convt2x2(ub_ctx* ctx, const char* up_key, const void* in_bf16_cg8, void* out_bf16_cg8,
                      int n, int h, int w, void* stream);
/* The decoder's F.interpolate(mode="bilinear", align_corners=True) (models/processing_blocks.py:107) in isolation:
 * blocked bf16 [n][c/8][hi][wi][8] -> [n][c/8][ho][wo][8]. */
int ub_debug_resample(ub_ctx* ctx, const void* in_bf16_cg8, void* out_bf16_cg8, int n, int c, int hi, int wi,
                      int ho, int wo, void* stream);
/* Measurement hooks (bench.py's roofline leg): between begin and end every tensor-core GEMM launch is
 * bracketed by CUDA events on its own stream.  end() synchronises the device and returns the summed
 * GEMM time, the number of GEMM launches, and (optionally) one text line per layer:
 * "<state_dict key> <cin> <n_total> <ksize> <launches> <ms>". */
int ub_profile_begin(ub_ctx* ctx);
int ub_profile_end(ub_ctx* ctx, double* gemm_ms_total, int64_t* gemm_launches, char* report, int report_len);
/* Number of kernel launches issued by this ctx so far (bench.py's gpu_launches claim). */
int64_t ub_launch_count(const ub_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* UNET_B200_H_ */
