// ub_vit_host.inl — host side of the CLIP ViT image tower (included by ub_api.cu inside its anonymous namespace).
// Replaces models/processing_blocks.py:156-233 (CustomClipPreprocessor + CLIPModel.get_image_features); the state_dict keys
// are HuggingFace CLIPModel's ("vision_model.*", "visual_projection.weight").

struct VitPlanN {  // everything that depends on the number of images of a chunk
  int n = 0, tokens = 0, tp = 0, ptokens = 0, tp_p = 0;
  std::vector<ConvParams> gemm;  // [0] patch embedding, then per layer: qkv, out_proj, fc1, fc2
};

struct VitState {
  int D = 0, L = 0, I = 0, P = 0, G = 0, S = 0, ntok = 0, E = 0;
  int chunk = 0;  // images per internal pass (vit_pick_chunk)
  bool pair = false;  // UB_VIT_PAIR=1 (experiment): CTA-pair tiles for the token GEMMs
  float eps = 1e-5f;
  std::vector<int> lin;  // indices into ctx->layers, same order as VitPlanN::gemm
  float *cls = nullptr, *pos = nullptr, *pre_g = nullptr, *pre_b = nullptr, *post_g = nullptr, *post_b = nullptr,
        *w_proj = nullptr;
  std::vector<float*> ln1_g, ln1_b, ln2_g, ln2_b;
  __nv_bfloat16 *patches = nullptr, *patch_out = nullptr, *ln_out = nullptr, *qkv = nullptr, *attn = nullptr,
                *branch = nullptr, *mlp = nullptr;
  float* resid = nullptr;
  std::map<int, VitPlanN> plans;
  struct Taps { ub::ResizeTap *ry, *rx; int max_count; };
  std::map<std::pair<int, int>, Taps> taps;  // (H, W) -> device tables (rows, columns) and their largest tap count
};

// Launch with the programmatic-stream-serialization attribute (the kernels of ub_vit.cuh wait with griddepcontrol.wait
// before their first global access); UB_NO_PDL=1: plain stream order.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(const ub_ctx* ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
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
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

int upload_f32(ub_ctx* ctx, const std::vector<float>& v, float** out) {
  int rc;
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(out), v.size() * 4))) return rc;
  UB_CUDA(ctx, cudaMemcpy(*out, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
  return UB_OK;
}

int upload_key(ub_ctx* ctx, const std::string& key, size_t count, float** out) {
  const HostTensor* t = find(ctx, key);
  if (!t) return fail(ctx, UB_ERR_INVALID, "missing state_dict entry %s", key.c_str());
  if (t->data.size() != count) return fail(ctx, UB_ERR_INVALID, "%s has %zu elements, expected %zu", key.c_str(), t->data.size(), count);
  return upload_f32(ctx, t->data, out);
}

// nn.Linear weights [out][in] (or the patch conv's [D,3,P,P] = [D][3*P*P]) stacked row-wise into one GEMM B operand
// W2[n_total][K]; biases stacked likewise (none: zeros).
int pack_linear(ub_ctx* ctx, const std::string& key, const std::vector<std::string>& w_keys,
                const std::vector<std::string>& b_keys, int act, int* index) {
  GemmLayer L;
  L.key = key;
  L.ksize = 1;
  L.relu = 0;
  L.act = act;
  { const char* e = getenv("UB_VIT_PAIR"); L.no_pair = !(e && e[0] == '1'); }
  std::vector<float> w, b;
  for (const std::string& k : w_keys) {
    const HostTensor* t = find(ctx, k);
    if (!t || t->shape.size() < 2) return fail(ctx, UB_ERR_INVALID, "missing or malformed state_dict entry %s", k.c_str());
    const int out = static_cast<int>(t->shape[0]);
    const int in = static_cast<int>(t->data.size() / static_cast<size_t>(out));
    if (L.cin && in != L.cin) return fail(ctx, UB_ERR_INVALID, "%s: input width %d differs from %d", k.c_str(), in, L.cin);
    L.cin = in;
    L.cout += out;
    w.insert(w.end(), t->data.begin(), t->data.end());
  }
  for (const std::string& k : b_keys) {
    const HostTensor* t = find(ctx, k);
    if (!t) return fail(ctx, UB_ERR_INVALID, "missing state_dict entry %s", k.c_str());
    b.insert(b.end(), t->data.begin(), t->data.end());
  }
  if (b.empty()) b.assign(L.cout, 0.0f);
  if (static_cast<int>(b.size()) != L.cout) return fail(ctx, UB_ERR_INVALID, "%s: %zu biases for %d outputs", key.c_str(), b.size(), L.cout);
  L.n_total = L.cout;
  L.block_n = pick_block_n(L.n_total);
  { const char* e = getenv("UB_VIT_BN"); if (e && atoi(e) == 128) L.block_n = 128; }  // (experiment: 128-column tiles)
  L.bk = 64;
  if (L.cin % 64 != 0 || L.cout % 256 != 0)
    return fail(ctx, UB_ERR_UNSUPPORTED, "%s: [%d -> %d]: inputs must be a multiple of 64 and outputs of 256", key.c_str(), L.cin, L.cout);
  std::vector<__nv_bfloat16> packed(w.size());
  for (size_t i = 0; i < w.size(); ++i) packed[i] = __float2bfloat16(w[i]);
  int rc;
  if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&L.w), packed.size() * 2))) return rc;
  UB_CUDA(ctx, cudaMemcpy(L.w, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
  if ((rc = upload_f32(ctx, b, &L.bias))) return rc;
  L.h_bias = b;
  *index = static_cast<int>(ctx->layers.size());
  ctx->layer_by_key[L.key] = *index;
  ctx->layers.push_back(L);
  return UB_OK;
}

int alloc_zeroed(ub_ctx* ctx, void** p, size_t bytes) {
  int rc;
  if ((rc = dev_alloc(ctx, p, bytes))) return rc;
  UB_CUDA(ctx, cudaMemset(*p, 0, bytes));
  return UB_OK;
}

int vit_pick_chunk(const ub_ctx* ctx, const VitState* v);

int pack_vit(ub_ctx* ctx) {
  std::unique_ptr<VitState> v(new VitState());
  const std::string vm = "vision_model.";
  const HostTensor* pw = find(ctx, vm + "embeddings.patch_embedding.weight");
  const HostTensor* pe = find(ctx, vm + "embeddings.position_embedding.weight");
  const HostTensor* vp = find(ctx, "visual_projection.weight");
  if (!pw || !pe || !vp || pw->shape.size() != 4 || pe->shape.size() != 2 || vp->shape.size() != 2)
    return fail(ctx, UB_ERR_INVALID, "missing vision_model.embeddings.* / visual_projection.weight entries (HuggingFace CLIPModel keys)");
  v->D = static_cast<int>(pw->shape[0]);
  v->P = static_cast<int>(pw->shape[2]);
  v->ntok = static_cast<int>(pe->shape[0]);
  v->E = static_cast<int>(vp->shape[0]);
  v->G = 1;
  while ((v->G + 1) * (v->G + 1) + 1 <= v->ntok) v->G++;
  v->S = v->G * v->P;
  if (pw->shape[1] != 3 || pw->shape[3] != v->P || v->G * v->G + 1 != v->ntok || pe->shape[1] != v->D || vp->shape[1] != v->D)
    return fail(ctx, UB_ERR_UNSUPPORTED, "unexpected CLIP vision embedding shapes");
  if (v->D % 256 != 0 || v->D > ub::kVitMaxDim || v->ntok > ub::kVitMaxTokens || v->P % 8 != 0)
    return fail(ctx, UB_ERR_UNSUPPORTED, "CLIP vision tower: hidden size %d (multiple of 256, <= %d), %d tokens (<= %d), patch %d (multiple of 8)",
                v->D, ub::kVitMaxDim, v->ntok, ub::kVitMaxTokens, v->P);
  while (find(ctx, vm + "encoder.layers." + std::to_string(v->L) + ".self_attn.q_proj.weight")) v->L++;
  if (v->L == 0) return fail(ctx, UB_ERR_INVALID, "no vision_model.encoder.layers.* entries");
  int rc;
  const size_t D = static_cast<size_t>(v->D);
  if ((rc = upload_key(ctx, vm + "embeddings.class_embedding", D, &v->cls))) return rc;
  if ((rc = upload_key(ctx, vm + "embeddings.position_embedding.weight", D * v->ntok, &v->pos))) return rc;
  if ((rc = upload_key(ctx, vm + "pre_layrnorm.weight", D, &v->pre_g))) return rc;  // (sic: HuggingFace's spelling)
  if ((rc = upload_key(ctx, vm + "pre_layrnorm.bias", D, &v->pre_b))) return rc;
  if ((rc = upload_key(ctx, vm + "post_layernorm.weight", D, &v->post_g))) return rc;
  if ((rc = upload_key(ctx, vm + "post_layernorm.bias", D, &v->post_b))) return rc;
  if ((rc = upload_key(ctx, "visual_projection.weight", D * v->E, &v->w_proj))) return rc;
  int idx = -1;
  if ((rc = pack_linear(ctx, vm + "embeddings.patch_embedding", {vm + "embeddings.patch_embedding.weight"}, {}, ub::ACT_NONE, &idx))) return rc;
  v->lin.push_back(idx);
  v->ln1_g.resize(v->L); v->ln1_b.resize(v->L); v->ln2_g.resize(v->L); v->ln2_b.resize(v->L);
  for (int l = 0; l < v->L; ++l) {
    const std::string p = vm + "encoder.layers." + std::to_string(l) + ".";
    if ((rc = upload_key(ctx, p + "layer_norm1.weight", D, &v->ln1_g[l]))) return rc;
    if ((rc = upload_key(ctx, p + "layer_norm1.bias", D, &v->ln1_b[l]))) return rc;
    if ((rc = upload_key(ctx, p + "layer_norm2.weight", D, &v->ln2_g[l]))) return rc;
    if ((rc = upload_key(ctx, p + "layer_norm2.bias", D, &v->ln2_b[l]))) return rc;
    if ((rc = pack_linear(ctx, p + "self_attn.qkv",
                          {p + "self_attn.q_proj.weight", p + "self_attn.k_proj.weight", p + "self_attn.v_proj.weight"},
                          {p + "self_attn.q_proj.bias", p + "self_attn.k_proj.bias", p + "self_attn.v_proj.bias"}, ub::ACT_NONE, &idx)))
      return rc;
    v->lin.push_back(idx);
    if (ctx->layers[idx].cout != 3 * v->D || ctx->layers[idx].cin != v->D) return fail(ctx, UB_ERR_UNSUPPORTED, "%s: not [3D, D]", p.c_str());
    if ((rc = pack_linear(ctx, p + "self_attn.out_proj", {p + "self_attn.out_proj.weight"}, {p + "self_attn.out_proj.bias"}, ub::ACT_NONE, &idx))) return rc;
    v->lin.push_back(idx);
    if ((rc = pack_linear(ctx, p + "mlp.fc1", {p + "mlp.fc1.weight"}, {p + "mlp.fc1.bias"}, ub::ACT_QUICK_GELU, &idx))) return rc;
    v->lin.push_back(idx);
    if (l == 0) v->I = ctx->layers[idx].cout;
    if (ctx->layers[idx].cout != v->I) return fail(ctx, UB_ERR_UNSUPPORTED, "%s: MLP width changes between layers", p.c_str());
    if ((rc = pack_linear(ctx, p + "mlp.fc2", {p + "mlp.fc2.weight"}, {p + "mlp.fc2.bias"}, ub::ACT_NONE, &idx))) return rc;
    v->lin.push_back(idx);
  }
  // token buffers for max_mb images, rounded up to whole 128-token tiles (zeroed once: the rows after the last
  // token are GEMM operand rows whose results nobody reads)
  const size_t tp = (static_cast<size_t>(ctx->max_mb) * v->ntok + 255) / 256 * 256;
  const size_t tpp = (static_cast<size_t>(ctx->max_mb) * v->G * v->G + 255) / 256 * 256;
  if ((rc = alloc_zeroed(ctx, reinterpret_cast<void**>(&v->patches), tpp * 3 * v->P * v->P * 2))) return rc;
  if ((rc = alloc_zeroed(ctx, reinterpret_cast<void**>(&v->patch_out), tpp * D * 2))) return rc;
  if ((rc = alloc_zeroed(ctx, reinterpret_cast<void**>(&v->resid), tp * D * 4))) return rc;
  if ((rc = alloc_zeroed(ctx, reinterpret_cast<void**>(&v->ln_out), tp * D * 2))) return rc;
  if ((rc = alloc_zeroed(ctx, reinterpret_cast<void**>(&v->qkv), tp * 3 * D * 2))) return rc;
  if ((rc = alloc_zeroed(ctx, reinterpret_cast<void**>(&v->attn), tp * D * 2))) return rc;
  if ((rc = alloc_zeroed(ctx, reinterpret_cast<void**>(&v->branch), tp * D * 2))) return rc;
  if ((rc = alloc_zeroed(ctx, reinterpret_cast<void**>(&v->mlp), tp * static_cast<size_t>(v->I) * 2))) return rc;
  UB_CUDA(ctx, cudaDeviceSynchronize());
  { const char* e = getenv("UB_VIT_PAIR"); v->pair = e && e[0] == '1'; }
  v->chunk = vit_pick_chunk(ctx, v.get());
  { const char* e = getenv("UB_VIT_CHUNK"); if (e && atoi(e) >= 1 && atoi(e) <= ctx->max_mb) v->chunk = atoi(e); }  // (experiment)
  ctx->vit = v.release();
  ctx->weights_ready = true;
  return UB_OK;
}

// torchvision Resize on a tensor = F.interpolate(bilinear, antialias=True, align_corners=False): per-axis taps as ATen's
// _compute_indices_weights_aa builds them (triangle filter stretched by the scale when shrinking, normalised weights).
int make_resize_taps(ub_ctx* ctx, int in, int out, std::vector<ub::ResizeTap>* taps) {
  const float scale = static_cast<float>(in) / static_cast<float>(out);
  const float support = scale >= 1.0f ? scale : 1.0f;  // interp_size (2) * 0.5 (* scale)
  const float invscale = scale >= 1.0f ? 1.0f / scale : 1.0f;
  const int max_interp = static_cast<int>(ceilf(support)) * 2 + 1;
  taps->resize(out);
  for (int i = 0; i < out; ++i) {
    const float center = static_cast<float>(scale * (i + 0.5));
    int64_t xmin = static_cast<int64_t>(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int64_t hi = static_cast<int64_t>(center + support + 0.5);
    if (hi > in) hi = in;
    int64_t xsize = hi - xmin;
    if (xsize < 0) xsize = 0;
    if (xsize > max_interp) xsize = max_interp;
    if (xsize > ub::kClipMaxTaps)
      return fail(ctx, UB_ERR_UNSUPPORTED, "resize %d -> %d needs %d taps per axis (limit %d: inputs up to ~5x the CLIP size)", in, out,
                  static_cast<int>(xsize), ub::kClipMaxTaps);
    ub::ResizeTap& t = (*taps)[i];
    t.start = static_cast<int>(xmin);
    t.count = static_cast<int>(xsize);
    float total = 0.0f;
    for (int j = 0; j < t.count; ++j) {
      float a = static_cast<float>((j + xmin - center + 0.5) * invscale);
      a = a < 0.0f ? -a : a;
      t.w[j] = a < 1.0f ? 1.0f - a : 0.0f;
      total += t.w[j];
    }
    for (int j = 0; j < ub::kClipMaxTaps; ++j) t.w[j] = (j < t.count && total != 0.0f) ? t.w[j] / total : 0.0f;
  }
  return UB_OK;
}

// CLIP's normalisation constants (processing_blocks.py:204-205)
constexpr float kClipMean[3] = {0.48145466f, 0.4578275f, 0.40821073f};
constexpr float kClipStd[3] = {0.26862954f, 0.26130258f, 0.27577711f};

int vit_taps(ub_ctx* ctx, int H, int W, ub::ResizeTap** ry, ub::ResizeTap** rx, int* max_count = nullptr) {
  VitState* v = ctx->vit;
  auto it = v->taps.find({H, W});
  if (it == v->taps.end()) {
    if (v->taps.size() >= 32) {  // a caller sweeping input sizes: drop the cached tables once nothing can be reading them
      UB_CUDA(ctx, cudaDeviceSynchronize());
      for (auto& kv : v->taps) {
        for (void* p : {static_cast<void*>(kv.second.ry), static_cast<void*>(kv.second.rx)}) {
          cudaFree(p);
          ctx->owned.erase(std::find(ctx->owned.begin(), ctx->owned.end(), p));
        }
      }
      v->taps.clear();
    }
    std::vector<ub::ResizeTap> hy, hx;
    int rc;
    if ((rc = make_resize_taps(ctx, H, v->S, &hy))) return rc;
    if ((rc = make_resize_taps(ctx, W, v->S, &hx))) return rc;
    int mc = 1;
    for (const auto& t : hy) mc = t.count > mc ? t.count : mc;
    for (const auto& t : hx) mc = t.count > mc ? t.count : mc;
    ub::ResizeTap *dy = nullptr, *dx = nullptr;
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&dy), hy.size() * sizeof(ub::ResizeTap)))) return rc;
    if ((rc = dev_alloc(ctx, reinterpret_cast<void**>(&dx), hx.size() * sizeof(ub::ResizeTap)))) return rc;
    UB_CUDA(ctx, cudaMemcpy(dy, hy.data(), hy.size() * sizeof(ub::ResizeTap), cudaMemcpyHostToDevice));
    UB_CUDA(ctx, cudaMemcpy(dx, hx.data(), hx.size() * sizeof(ub::ResizeTap), cudaMemcpyHostToDevice));
    UB_CUDA(ctx, cudaDeviceSynchronize());
    it = v->taps.emplace(std::make_pair(H, W), VitState::Taps{dy, dx, mc}).first;
  }
  *ry = it->second.ry;
  *rx = it->second.rx;
  if (max_count) *max_count = it->second.max_count;
  return UB_OK;
}

template <int T>
int launch_patch(ub_ctx* ctx, const VitState* v, const float* x, int n, int H, int W, const ub::ResizeTap* ry, const ub::ResizeTap* rx,
                 int tp_p, cudaStream_t st) {
  const int smem = ub::clip_patch_smem_bytes(v->P, v->G);
  const void* fn = reinterpret_cast<const void*>(&ub::clip_patch_kernel<T, false>);
  if (smem > 48 * 1024 && !ctx->smem_opt_in.count(fn)) {
    UB_CUDA(ctx, cudaFuncSetAttribute(ub::clip_patch_kernel<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    ctx->smem_opt_in.insert(fn);
  }
  if (T == 3) {  // UB_PATCH_SEP=1 (experiment, the 256 -> 224 instantiation only): separable variant, rows resampled along x first
    const char* e = getenv("UB_PATCH_SEP");
    if (e && e[0] == '1') {
      const int nr_max = 48;  // 32 output rows x 256/224 + taps
      const int smem2 = smem + nr_max * v->G * v->P * 4;
      const void* fn2 = reinterpret_cast<const void*>(&ub::clip_patch_kernel<3, true>);
      if (!ctx->smem_opt_in.count(fn2)) {
        UB_CUDA(ctx, cudaFuncSetAttribute(ub::clip_patch_kernel<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
        ctx->smem_opt_in.insert(fn2);
      }
      UB_CUDA(ctx, launch_pdl(ctx, ub::clip_patch_kernel<3, true>, dim3(v->G, 3, n), dim3(256), smem2, st, x, n, H, W, ry, rx, v->P, v->G,
                              kClipMean[0], kClipMean[1], kClipMean[2], 1.0f / kClipStd[0], 1.0f / kClipStd[1], 1.0f / kClipStd[2],
                              v->patches, tp_p, nr_max));
      ctx->launches++;
      return UB_OK;
    }
  }
  UB_CUDA(ctx, launch_pdl(ctx, ub::clip_patch_kernel<T, false>, dim3(v->G, 3, n), dim3(256), smem, st, x, n, H, W, ry, rx, v->P, v->G,
                          kClipMean[0], kClipMean[1], kClipMean[2], 1.0f / kClipStd[0], 1.0f / kClipStd[1], 1.0f / kClipStd[2],
                          v->patches, tp_p, 0));
  ctx->launches++;
  return UB_OK;
}


int vit_plan(ub_ctx* ctx, int n, VitPlanN** out) {
  VitState* v = ctx->vit;
  auto it = v->plans.find(n);
  if (it == v->plans.end()) {
    VitPlanN pl;
    pl.n = n;
    pl.tokens = n * v->ntok;
    const int tm = v->pair ? 256 : 128;  // whole tiles (pairs: the CTA-pair kernel pads odd tile counts with a phantom image)
    pl.tp = (pl.tokens + tm - 1) / tm * tm;
    pl.ptokens = n * v->G * v->G;
    pl.tp_p = (pl.ptokens + tm - 1) / tm * tm;
    pl.gemm.resize(v->lin.size());
    int rc;
    // the token axis as an "image" of tp/8 rows x 8 columns, one image per launch
    auto gemm = [&](size_t i, const __nv_bfloat16* src, __nv_bfloat16* dst, int tp) {
      const GemmLayer& L = ctx->layers[v->lin[i]];
      return fill_gemm(ctx, &pl.gemm[i], L, src, L.cin / 8, 1, tp / 8, 8, ub::EPI_STORE, dst, L.cout / 8, 0);
    };
    if ((rc = gemm(0, v->patches, v->patch_out, pl.tp_p))) return rc;
    for (int l = 0; l < v->L; ++l) {
      if ((rc = gemm(1 + 4 * l + 0, v->ln_out, v->qkv, pl.tp))) return rc;
      if ((rc = gemm(1 + 4 * l + 1, v->attn, v->branch, pl.tp))) return rc;
      if ((rc = gemm(1 + 4 * l + 2, v->ln_out, v->mlp, pl.tp))) return rc;
      if ((rc = gemm(1 + 4 * l + 3, v->mlp, v->branch, pl.tp))) return rc;
    }
    it = v->plans.emplace(n, std::move(pl)).first;
  }
  *out = &it->second;
  return UB_OK;
}

template <int MODE>
int vit_launch_ln(ub_ctx* ctx, int ns, int blocks, cudaStream_t st, float* resid, const __nv_bfloat16* branch, const float* g,
                  const float* b, float eps, __nv_bfloat16* out, int tokens, int tp, const __nv_bfloat16* patches, int tp_p,
                  const float* cls, const float* pos, int ntok) {
  cudaError_t e = cudaSuccess;
  switch (ns) {
    case 1: e = launch_pdl(ctx, ub::vit_ln_kernel<1, MODE>, dim3(blocks), dim3(256), 0, st, resid, branch, g, b, eps, out, tokens, tp, patches, tp_p, cls, pos, ntok); break;
    case 2: e = launch_pdl(ctx, ub::vit_ln_kernel<2, MODE>, dim3(blocks), dim3(256), 0, st, resid, branch, g, b, eps, out, tokens, tp, patches, tp_p, cls, pos, ntok); break;
    case 3: e = launch_pdl(ctx, ub::vit_ln_kernel<3, MODE>, dim3(blocks), dim3(256), 0, st, resid, branch, g, b, eps, out, tokens, tp, patches, tp_p, cls, pos, ntok); break;
    case 4: e = launch_pdl(ctx, ub::vit_ln_kernel<4, MODE>, dim3(blocks), dim3(256), 0, st, resid, branch, g, b, eps, out, tokens, tp, patches, tp_p, cls, pos, ntok); break;
    default: return fail(ctx, UB_ERR_UNSUPPORTED, "hidden size %d", 256 * ns);
  }
  UB_CUDA(ctx, e);
  ctx->launches++;
  return UB_OK;
}

// Images per chunk.  The four GEMMs of a layer are small (ViT-B/32, 125 images: 49 token tiles x 3 / 9 / 12 column tiles), so
// how many whole waves of the SMs their tile counts make decides a third of their time: 128 images give 150 tiles for the two
// D -> D maps (two waves, the second one 2/148 full), 125 images give 147 (one wave).  Pick the chunk size <= max_mb that
// minimises the waves spent per image, FLOP-weighted over the four shapes.
int vit_pick_chunk(const ub_ctx* ctx, const VitState* v) {
  const int tile_m = v->pair ? 256 : 128;
  const int slots = (v->pair ? ctx->num_sms / 2 : ctx->num_sms) > 0 ? (v->pair ? ctx->num_sms / 2 : ctx->num_sms) : 1;
  const int nt[4] = {3 * v->D / 256, v->D / 256, v->I / 256, v->D / 256};                  // n-tiles: qkv, out_proj, fc1, fc2
  const double kw[4] = {double(v->D), double(v->D), double(v->D), double(v->I)};           // time of one tile ~ K
  int best = ctx->max_mb;
  double best_cost = 1e300;
  for (int n = ctx->max_mb; n >= 1 && n >= ctx->max_mb / 2; --n) {
    const int mt = (n * v->ntok + tile_m - 1) / tile_m;
    double cost = 0.0;
    for (int i = 0; i < 4; ++i) cost += kw[i] * ((mt * nt[i] + slots - 1) / slots);
    cost /= n;
    if (cost < best_cost * (1.0 - 1e-9)) { best_cost = cost; best = n; }
  }
  return best;
}

// One chunk of at most max_mb images: x fp32 [n,3,H,W] (values as the caller's model sees them, i.e. [0,1]) -> feats [n,E].
int vit_forward_chunk(ub_ctx* ctx, const float* x, int n, int H, int W, float* feats, cudaStream_t st) {
  VitState* v = ctx->vit;
  int rc;
  ub::ResizeTap *ry = nullptr, *rx = nullptr;
  int max_taps = ub::kClipMaxTaps;
  if ((rc = vit_taps(ctx, H, W, &ry, &rx, &max_taps))) return rc;
  VitPlanN* pl = nullptr;
  if ((rc = vit_plan(ctx, n, &pl))) return rc;
  const int ns = v->D / 256;
  if (max_taps <= 2) rc = launch_patch<2>(ctx, v, x, n, H, W, ry, rx, pl->tp_p, st);        // identity / enlarging
  else if (max_taps <= 3) rc = launch_patch<3>(ctx, v, x, n, H, W, ry, rx, pl->tp_p, st);   // 256 -> 224 and other mild shrinks
  else if (max_taps <= 4) rc = launch_patch<4>(ctx, v, x, n, H, W, ry, rx, pl->tp_p, st);
  else if (max_taps <= 6) rc = launch_patch<6>(ctx, v, x, n, H, W, ry, rx, pl->tp_p, st);
  else if (max_taps <= 8) rc = launch_patch<8>(ctx, v, x, n, H, W, ry, rx, pl->tp_p, st);
  else rc = launch_patch<ub::kClipMaxTaps>(ctx, v, x, n, H, W, ry, rx, pl->tp_p, st);
  if (rc) return rc;
  auto run_gemm = [&](size_t i) {
    const GemmLayer& L = ctx->layers[v->lin[i]];
    return launch_gemm(ctx, pl->gemm[i], L.block_n, L.bk, st, v->lin[i]);
  };
  if ((rc = run_gemm(0))) return rc;
  const int ln_blocks = (pl->tokens + 7) / 8;
  if ((rc = vit_launch_ln<ub::VIT_LN_EMBED>(ctx, ns, ln_blocks, st, v->resid, nullptr, v->pre_g, v->pre_b, v->eps, nullptr, pl->tokens,
                                            pl->tp, v->patch_out, pl->tp_p, v->cls, v->pos, v->ntok)))
    return rc;
  for (int l = 0; l < v->L; ++l) {
    // layer_norm1 (the previous layer's MLP output joins the residual stream here)
    if ((rc = vit_launch_ln<ub::VIT_LN_ADD>(ctx, ns, ln_blocks, st, v->resid, l ? v->branch : nullptr, v->ln1_g[l], v->ln1_b[l], v->eps,
                                            v->ln_out, pl->tokens, pl->tp, nullptr, 0, nullptr, nullptr, v->ntok)))
      return rc;
    if ((rc = run_gemm(1 + 4 * l + 0))) return rc;
    UB_CUDA(ctx, launch_pdl(ctx, ub::vit_attention_kernel, dim3(v->D / ub::kVitHeadDim, n), dim3(128), 0, st, v->qkv, v->attn, v->D,
                            v->ntok, pl->tp));
    ctx->launches++;
    if ((rc = run_gemm(1 + 4 * l + 1))) return rc;
    if ((rc = vit_launch_ln<ub::VIT_LN_ADD>(ctx, ns, ln_blocks, st, v->resid, v->branch, v->ln2_g[l], v->ln2_b[l], v->eps, v->ln_out,
                                            pl->tokens, pl->tp, nullptr, 0, nullptr, nullptr, v->ntok)))
      return rc;
    if ((rc = run_gemm(1 + 4 * l + 2))) return rc;
    if ((rc = run_gemm(1 + 4 * l + 3))) return rc;
  }
  UB_CUDA(ctx, launch_pdl(ctx, ub::vit_pool_proj_kernel, dim3(n, (v->E + 63) / 64), dim3(256), 0, st, v->resid, v->branch, v->D, v->ntok,
                          pl->tp, v->post_g, v->post_b, v->eps, v->w_proj, v->E, feats));
  ctx->launches++;
  return UB_OK;
}
