// ub_conv_tc.cuh — the tensor-core kernel of the path: conv3x3(pad 1)+foldedBN+ReLU as an implicit
// GEMM on tcgen05 with TMEM accumulators, fed by TMA from channel-blocked bf16 activations.  The same
// kernel (taps = 1) is the ConvTranspose2d(k2,s2) GEMM with a pixel-shuffle scatter epilogue, and with
// the HEAD epilogue it fuses the final 1x1 conv, softmax/argmax and the confusion-matrix histogram.
//
// Replaces: models/processing_blocks.py:43-48 (Conv2d+BatchNorm2d+ReLU), :73 (MaxPool2d, POOL
// epilogue), :102 (ConvTranspose2d), :108 (cat: operands live in one concat buffer), models/UNet.py:73
// (out 1x1) and models/losses.py:48-58 (softmax/argmax/counts).
//
// Activation layout in HBM: [B][C/8][H][W][8] bf16 ("channel groups" cg of 8 channels = 16 bytes).
//   * a concat along channels is a concat of cg planes: a consumer reads any contiguous cg range of a
//     buffer through one tensor map, two producers fill disjoint plane ranges of the same buffer;
//   * a warp of the epilogue (32 pixels = 4 tile rows x 8 columns) stores 4 x 128 contiguous bytes per
//     16-byte store instruction instead of 32 separate lines (NHWC): the store path stays off the
//     critical path.
//
// GEMM view:  D[M = 128 pixels (16 rows x 8 cols of one image)][N = BLOCK_N output channels]
//             K = taps * Cin, walked as (channel chunk c of BK) x (tap ky,kx) x (BK/16 MMAs).
// A operand:  ONE TMA box per channel chunk: {8*(8+2) elements, 16+2 rows, BK/8 cg} (halo included, image
//             border = TMA out-of-bounds zero fill = the conv's zero padding).  In shared memory it is the
//             no-swizzle K-major "interleaved" UMMA layout: pixel rows 16 bytes apart, 8-pixel groups
//             (one tile row) SBO = 10*16 bytes apart, the two K halves of an MMA LBO = one cg plane apart.
//             Because pixels are linear in memory, tap (ky,kx) is just the descriptor start address moved
//             by (ky*10 + kx) * 16 bytes: all nine taps read the same box — activations are fetched
//             1.4x (halo) instead of 9x, and written to shared memory once.
// B operand:  weights W2[N][K] (K contiguous, 128B/64B swizzle), one TMA box {BK, BLOCK_N} per (c, tap):
//             either streamed through a ring, or — when all taps of an n-tile fit beside the A ring (the
//             thin, high-resolution layers) — loaded ONCE per persistent CTA and kept resident (WS).
// KXS ("kx-stacked") variant for the thin 3x3 layers (Cout <= 64), where the 4 KB A read per MMA — not the
//             tensor pipe — is the limit (shared-memory bandwidth): the three kx taps become GEMM columns,
//             N = 3*Cout, D[p][kx*Cout+co] = sum_{ky,ci} X[y+ky-1][x'][ci] W[co][ci][ky][kx] over INPUT column
//             x', and the epilogue adds the three column blocks of the neighbouring pixels (two warp
//             shuffles per value): out[x] = D0[x-1] + D1[x] + D2[x+1].  One A read then feeds 3x the FLOPs.
//             Tiles are TW = 32 (or 16 for the pooled layer) pixels wide with one halo column on each
//             side (TW-2 outputs); the A box has no x halo, so 8-pixel groups stay 128 bytes apart.
// CG2 (CTA pair, tcgen05 cta_group::2) variant for the wide streaming layers (N >= 128): two CTAs of a cluster
//             compute two 128-pixel tiles of the same n-tile with ONE M = 256 MMA stream issued by the even
//             CTA.  Each CTA stages its own A box and HALF of every weight tile (N/2 rows); the tensor
//             cores of both SMs read both halves.  Weight traffic from L2 and the shared-memory bytes
//             written and read per MMA are halved — these layers were bound by exactly that.
// Warp roles: kEpiGroups groups of 4 epilogue warps FIRST (warps 0 .. 4*kEpiGroups-1), then the TMA producer, the MMA
// issuer and the TMEM allocator as the LAST warps of the CTA: a scheduler that has several eligible warps picks the one
// with the highest warp id, and the single MMA-issuing thread must never queue behind epilogue warps (or behind warps
// spinning on a barrier).  kAcc TMEM accumulators (up to 4): the MMA warp runs ahead of the epilogues by kAcc tiles.
#pragma once
#include "ub_ptx.cuh"
#include "ub_head.cuh"

namespace ub {

enum EpiMode : int { EPI_STORE = 0, EPI_POOL = 1, EPI_CONVT = 2, EPI_HEAD = 3,
                     EPI_CONVT_BLEND = 4 };  // ConvTranspose + the align_corners resample back to the input grid, in the epilogue

// F.interpolate(bilinear, align_corners=True) from 2n back to n samples (models/processing_blocks.py:106-107, the first
// decoder stage: ConvTranspose doubles the un-pooled bottleneck, the resize brings it back to the skip's size).  Output
// sample o reads source positions i0 = floor(o * (2n-1)/(n-1)) = 2o + floor(o/(n-1)) and i0 + 1: BOTH are children of
// ConvTranspose input sample o (its parities 0 and 1), so the two layers together are a per-pixel blend of the four
// parity products of ONE input pixel.  Same fp32 arithmetic as resample_bilinear_kernel (ub_aux.cuh); w0 / w1 = total
// weight on the even / odd child.  Returns false when a tap would leave input sample o (the host then keeps two kernels).
__host__ __device__ inline bool convt_resample_weights(int o, int n, float* w0, float* w1) {
  const int ni = 2 * n;
  const float s = n > 1 ? static_cast<float>(ni - 1) / static_cast<float>(n - 1) : 0.0f;
  const float f = s * static_cast<float>(o);
  const int i0 = static_cast<int>(f);
  const int i1 = i0 + (i0 < ni - 1 ? 1 : 0);
  const float l = f - static_cast<float>(i0), h = 1.0f - l;
  *w0 = ((i0 & 1) ? 0.0f : h) + ((i1 & 1) ? 0.0f : l);
  *w1 = ((i0 & 1) ? h : 0.0f) + ((i1 & 1) ? l : 0.0f);
  return (i0 >> 1) == o && (i1 >> 1) == o;
}
enum EpiAct : int { ACT_NONE = 0, ACT_QUICK_GELU = 1 };  // STORE epilogue of the 1x1 instantiations, after the bias (ReLU is `relu`)

__device__ __forceinline__ float2 quick_gelu2(float2 v) {
  v.x = __fdividef(v.x, 1.0f + __expf(-1.702f * v.x));
  v.y = __fdividef(v.y, 1.0f + __expf(-1.702f * v.y));
  return v;
}

constexpr int kTileM = 128;
constexpr int kMaxEpiGroups = 4;
constexpr int kMaxClasses = kHeadMaxClasses;
// tile geometry: TW = 8 is the standard 16-row x 8-column tile; TW = 16 / 32 are the kx-stacked variants
__host__ __device__ constexpr bool is_kxs(int tw) { return tw != 8; }
__host__ __device__ constexpr int tile_rows(int tw) { return kTileM / tw; }
__host__ __device__ constexpr int tile_xstep(int tw) { return is_kxs(tw) ? tw - 2 : tw; }  // output columns per tile
// epilogue groups of 4 warps: the thin layers (N <= 64) have short main loops and long per-pixel epilogues
// (latency-bound single warps), so they get 4 groups; the wide ones have 2 accumulators anyway
__host__ __device__ constexpr int epi_groups(int block_n, int tw, int ksize) {
  // never more groups than TMEM accumulators (a group may only be one barrier phase ahead): N = 256 has 2
  return (void)ksize, is_kxs(tw) ? (block_n <= 96 ? 4 : 2) : (block_n <= 32 ? 4 : 2);
}
__host__ __device__ constexpr int conv_threads(int block_n, int tw, int ksize) { return 128 + 128 * epi_groups(block_n, tw, ksize); }

struct ConvParams {
  CUtensorMap tmA;  // activations: dims {8*W, H, Cin/8, B}, box {8*PXW, ROWS_IN, BK/8, 1} (ConvCfg)
  CUtensorMap tmB;  // weights:     dims {Ktot, Ntot}, box {BK, BLOCK_N}
  int B, H, W;      // input extent (= output extent for conv; output is 2H x 2W for ConvT)
  int cin;          // channels of the A source = K per tap
  int ksize;        // 3 or 1
  int n_total;      // GEMM N (Cout, or 4*Cout for ConvT)
  int mode;         // EpiMode
  int relu;
  const float* bias;  // [n_total] (conv) / [cout] (ConvT)
  int n_bias;
  __nv_bfloat16* out;  // [B][out_cgs][Ho][Wo][8] bf16
  int out_cgs;         // channel groups per image of the output buffer (concat buffers are wider)
  int out_cg0;         // first channel group written
  int cout;            // ConvT: channels per 2x2 quadrant
  int cout_shift;      // log2(cout) (ConvT quadrant = n >> cout_shift)
  // HEAD epilogue: the 1x1 head and this layer's 32 biases travel by value in the kernel parameters, so the
  // epilogue reads them as constant-bank operands (no shared-memory traffic beside the MMA operand reads)
  float head_w[kMaxClasses * 32];  // [n_classes][32]
  float head_b[kMaxClasses];
  float bias32[32];
  int n_classes;
  const void* targets;  // uint8 or int64 [B,H,W], may be null (no counts)
  int target_fmt;       // 0 = int64, 1 = uint8
  int ref_batch;        // images per reference batch
  int img_base;         // index (within the whole call) of image 0 of this micro-batch
  unsigned long long* cm;  // [n_batches][C][C]
  double* dice;            // [n_batches][C][3] or null
  uint8_t* pred_out;       // [B,H,W] or null
  float* logits_out;       // [B,C,H,W] or null
  // pipeline shape chosen by the host (ub_api.cu: choose_pipeline)
  int dbg; // UB_DBG bit mask (experiments only): 1 = epilogue only drains, 2 = no MMAs issued, 8 = head always evaluates the softmax
  int pairq;  // CTA pairs: pair-granular tile order with the n-tile fastest (weight-stationary pairs, n_tiles > 1)
  int cg2; // 1 = CTA-pair kernel (host: tmB box holds N/2 rows, even grid, cluster launch)
  int tw;  // tile width chosen by the host (8 standard, 16/32 kx-stacked); the kernel has it as a template parameter
  int k_steps;  // MMAs (16 channels each) issued per tap and chunk; fewer than BK/16 when the tail of the chunk is zero padding
  int rs;  // 2 = fused first encoder stage (ub_conv_rs2.cuh); 1 = row-streaming kernel (ub_conv_rs.cuh): tmA is the 5-D map over (pixel, row in quarter, quarter, plane, image)
  int hq;  // row-streaming: rows per stream = H / 4
  // row-streaming: the A operand is filled by cp.async from the raw blocked tensor (ub_conv_rs.cuh explains why not TMA)
  const __nv_bfloat16* a_src;  // plane 0 of image 0 of the layer's input view
  int a_cgs;                   // channel-group planes per image of that buffer
  int a_planes;                // planes that exist in the view (cin_src / 8); planes beyond read as zero
  int sa;  // A stages (activation boxes)
  int sb;  // B stages when streaming weights; unused when the kernel is the WS instantiation
  int act; // EpiAct (1x1 STORE only)
  const float* bias1;  // fused first encoder stage (ub_conv_rs2.cuh, rs == 2): conv.0's bias; `bias` is conv.3's, tmA holds conv.0's weights
  float bias64[2][64]; // ... and both by value: the epilogues read them as constant-bank operands (their shared-memory reads
                       // competed with the tensor core's operand reads: 23 % of the LSU wavefronts of the first version)
};

constexpr int kMaxSA = 8;
constexpr int kBiasStage = 1024;  // biases staged in shared memory (ConvCfg::TAIL_BIAS)
constexpr int kMaxSB = 8;
constexpr int kMaxAcc = 4;
constexpr int kSmemLimit = 232448;  // 227 KB, the sm_100 per-block maximum

template <int BLOCK_N, int BK, int KSIZE, int TW, bool CG2 = false>
struct ConvCfg {
  static constexpr bool KXS = is_kxs(TW);
  static constexpr int TH = tile_rows(TW);
  static constexpr int PADY = KSIZE / 2;
  static constexpr int PADX = KXS ? 0 : PADY;            // kx-stacked tiles carry their x halo inside the tile
  static constexpr int PXW = TW + 2 * PADX;              // pixels per box row
  static constexpr int ROWS_IN = TH + 2 * PADY;          // box rows
  static constexpr int SBO_BYTES = KXS ? 128 : PXW * 16; // distance between 8-pixel groups of the A tile
  static constexpr int TAPS = KXS ? KSIZE : KSIZE * KSIZE;            // B boxes per channel chunk
  static constexpr int COUT_T = KXS ? BLOCK_N / 3 : BLOCK_N;          // output channels per tile
  static constexpr int A_PLANE_BYTES = ROWS_IN * PXW * 16;            // one channel group of the box = LBO
  static constexpr int A_BOX_BYTES = (BK / 8) * A_PLANE_BYTES;        // what one TMA load delivers
  static constexpr int A_STAGE_BYTES = ((A_BOX_BYTES + 1023) / 1024) * 1024;
  static constexpr int SWZ_B = 2 * BK;                                // weight rows: 128 B or 64 B swizzle
  static constexpr int B_ROWS = CG2 ? BLOCK_N / 2 : BLOCK_N;          // weight rows staged by one CTA
  static constexpr int B_STAGE_BYTES = B_ROWS * SWZ_B;
  static constexpr int ACC_STRIDE = BLOCK_N <= 32 ? 32 : BLOCK_N <= 64 ? 64 : BLOCK_N <= 128 ? 128 : 256;  // TMEM columns
  static constexpr int ACC = (512 / ACC_STRIDE) < kMaxAcc ? (512 / ACC_STRIDE) : kMaxAcc;
  static constexpr int TMEM_COLS = ACC * ACC_STRIDE;
  // fixed-size tail in front of the two operand regions
  static constexpr int TAIL_BIAS = 0;
  static constexpr int TAIL_HEAD = TAIL_BIAS + kBiasStage * 4;
  static constexpr int TAIL_BAR = TAIL_HEAD + (kMaxClasses * 32 + kMaxClasses) * 4;
  static constexpr int N_BARS = 2 * kMaxSA + 2 * kMaxSB + 2 * kMaxAcc + 1;
  static constexpr int TAIL_CM = TAIL_BAR + N_BARS * 8;
  static constexpr int TAIL_DICE = TAIL_CM + kMaxEpiGroups * kMaxClasses * kMaxClasses * 8;
  static constexpr int TAIL_TMEM = TAIL_DICE + kMaxEpiGroups * kMaxClasses * 2 * 8;
  static constexpr int TAIL_BYTES = ((TAIL_TMEM + 16 + 1023) / 1024) * 1024;  // keeps the rings 1024B aligned
  static constexpr int OPERAND_BUDGET = kSmemLimit - 1024 - TAIL_BYTES;       // 1024 = alignment slack
  static_assert(BK == 64 || BK == 32 || BK == 16, "channel chunk");
  static_assert(!CG2 || (TW == 8 && BLOCK_N >= 128), "CTA pairs are for the wide standard-geometry layers");
  static_assert(!KXS || (KSIZE == 3 && BLOCK_N % 48 == 0 && (TW == 16 || TW == 32)), "kx-stacked geometry");
  static_assert(TMEM_COLS >= 32 && (TMEM_COLS & (TMEM_COLS - 1)) == 0 && TMEM_COLS <= 512, "TMEM columns");
  static_assert(B_STAGE_BYTES % 1024 == 0, "swizzle atom alignment");
};

// Walks the tiles t0, t0 + stride, ... of one warp role without per-tile integer divisions: the tile index
// is kept as mixed-radix digits (n-tile, tile column, tile row, image) and advanced by the pre-decomposed stride.
struct TileIter {
  int d0, d1, d2, d3;      // digits, d0 fastest; radices r0, r1, r2 (d3 unbounded)
  int i0, i1, i2, i3;      // the stride, decomposed the same way
  int r0, r1, r2;
  __device__ __forceinline__ void init(int t0, int stride, int r0_, int r1_, int r2_) {
    r0 = r0_; r1 = r1_; r2 = r2_;
    d0 = t0 % r0; int m = t0 / r0;
    d1 = m % r1; m /= r1;
    d2 = m % r2; d3 = m / r2;
    i0 = stride % r0; m = stride / r0;
    i1 = m % r1; m /= r1;
    i2 = m % r2; i3 = m / r2;
  }
  __device__ __forceinline__ void next() {
    d0 += i0;
    int c = d0 >= r0 ? 1 : 0;
    d0 -= c ? r0 : 0;
    d1 += i1 + c;
    c = d1 >= r1 ? 1 : 0;
    d1 -= c ? r1 : 0;
    d2 += i2 + c;
    c = d2 >= r2 ? 1 : 0;
    d2 -= c ? r2 : 0;
    d3 += i3 + c;
  }
};

__device__ __forceinline__ void epi_bar_sync(int grp) {  // named barrier 1..: the 128 threads of one epilogue group
  asm volatile("bar.sync %0, 128;" ::"r"(grp + 1) : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  const __nv_bfloat162 r = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&r);
}

template <int BLOCK_N, int BK, int KSIZE, bool WS, int TW, bool CG2 = false>
__global__ void __launch_bounds__(conv_threads(BLOCK_N, TW, KSIZE), 1)
conv_tc_kernel(const __grid_constant__ ConvParams p) {
  using Cfg = ConvCfg<BLOCK_N, BK, KSIZE, TW, CG2>;
  // CG2 && WS: each CTA of the pair keeps ITS half of every weight tile resident (single n-tile layers only)
  static_assert(epi_groups(BLOCK_N, TW, KSIZE) <= Cfg::ACC, "an epilogue group per accumulator at most");
  constexpr bool KXS = Cfg::KXS;
  constexpr int TAPS = Cfg::TAPS;
  constexpr int ACC = Cfg::ACC;
  constexpr int kTileRows = Cfg::TH;
  constexpr int kXStep = tile_xstep(TW);
  constexpr int kEpiGroups = epi_groups(BLOCK_N, TW, KSIZE);
  constexpr int kConvThreads = conv_threads(BLOCK_N, TW, KSIZE);
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // pad to 1024 B with pointer arithmetic on the __shared__ array (keeps the address space known to
  // the compiler: LDS/STS/ATOMS instead of generic accesses)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int k_chunks = p.cin / BK;
  const int n_taps = k_chunks * TAPS;  // B tiles per output tile
  // [fixed-size tail: bias, head weights, barriers, counters][A ring][B ring or resident weights]
  float* s_bias = reinterpret_cast<float*>(smem + Cfg::TAIL_BIAS);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::TAIL_BAR);
  uint64_t* a_full = bars;
  uint64_t* a_empty = bars + kMaxSA;
  uint64_t* b_full = bars + 2 * kMaxSA;
  uint64_t* b_empty = b_full + kMaxSB;
  uint64_t* t_full = b_empty + kMaxSB;
  uint64_t* t_empty = t_full + kMaxAcc;
  uint64_t* w_full = t_empty + kMaxAcc;  // WS: all resident weights have landed
  unsigned long long* s_cm_all = reinterpret_cast<unsigned long long*>(smem + Cfg::TAIL_CM);
  double* s_dice_all = reinterpret_cast<double*>(smem + Cfg::TAIL_DICE);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + Cfg::TAIL_TMEM);
  uint8_t* a_ring = smem + Cfg::TAIL_BYTES;
  uint8_t* b_ring = a_ring + p.sa * Cfg::A_STAGE_BYTES;

  // warp index through a shuffle so the compiler treats the role branches as warp-uniform
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  constexpr int kEpiWarps = 4 * kEpiGroups;
  const int role = warp - kEpiWarps;  // 0 TMA producer, 1 MMA issuer, 2 TMEM allocator, 3 spare; negative: epilogue warp

  const int tiles_x = (p.W + kXStep - 1) / kXStep;
  const int tiles_y = (p.H + kTileRows - 1) / kTileRows;
  const int n_tiles = KXS ? 1 : p.n_total / BLOCK_N;
  // Tile order.  Standard: (n-tile, column, row, image), n-tile fastest.  CTA pairs: (column, row, image, n-tile)
  // with the image count padded so that tiles 2k and 2k+1 — the two CTAs of a pair — always share the n-tile.
  // CTA pairs with p.pairq (weight-stationary pairs, several n-tiles): the PAIR index runs over (n-tile, column
  // pair, row, image), n-tile fastest, and CTA r of the pair takes column 2*pair_column + r; with the pair count a
  // multiple of n_tiles every pair keeps one n-tile (its halves of the weights stay resident) while the n-tiles of
  // one pixel tile are worked on by neighbouring pairs at the same time (the activations hit L2).
  const uint32_t cta_rank = CG2 ? cluster_ctarank() : 0u;
  const bool pairq = CG2 && p.pairq;
  const int tiles_xh = (tiles_x + 1) >> 1;
  const int b_pad = (CG2 && !pairq) ? p.B + ((tiles_x * tiles_y * p.B) & 1) : p.B;
  const int total_tiles = pairq ? n_tiles * tiles_xh * tiles_y * p.B : b_pad * tiles_x * tiles_y * n_tiles;
  const int t_first = pairq ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int t_step = pairq ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  auto tile_iter = [&](int t0, int stride) {
    TileIter ti;
    if (pairq) ti.init(t0, stride, n_tiles, tiles_xh, tiles_y);
    else if constexpr (CG2) ti.init(t0, stride, tiles_x, tiles_y, b_pad);
    else ti.init(t0, stride, n_tiles, tiles_x, tiles_y);
    return ti;
  };
  // tile coordinates from the iterator's digits: n-tile, tile column, tile row, image
  auto tile_nt = [&](const TileIter& ti) { return (CG2 && !pairq) ? ti.d3 : ti.d0; };
  auto tile_tx = [&](const TileIter& ti) { return pairq ? 2 * ti.d1 + static_cast<int>(cta_rank) : (CG2 ? ti.d0 : ti.d1); };
  auto tile_ty = [&](const TileIter& ti) { return (CG2 && !pairq) ? ti.d1 : ti.d2; };
  auto tile_b = [&](const TileIter& ti) { return (CG2 && !pairq) ? ti.d2 : ti.d3; };

  // ---------------- one-time setup (nothing here depends on the previous kernel's output)
  // (the staging area holds kBiasStage floats; wider 1x1 layers — the ViT's qkv and fc1 — read their bias from global memory)
  for (int i = threadIdx.x; i < (p.n_bias < kBiasStage ? p.n_bias : kBiasStage); i += kConvThreads) s_bias[i] = p.bias[i];
  if (p.mode == EPI_HEAD) {
    if (threadIdx.x < kEpiGroups * kMaxClasses * kMaxClasses) s_cm_all[threadIdx.x] = 0ull;
    if (threadIdx.x < kEpiGroups * kMaxClasses * 2) s_dice_all[threadIdx.x] = 0.0;
  }
  if (role == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
  }
  if (role == 1 && lane == 0) {
    for (int i = 0; i < kMaxSA; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < kMaxSB; ++i) { mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1); }
    for (int i = 0; i < kMaxAcc; ++i) { mbar_init(&t_full[i], 1); mbar_init(&t_empty[i], CG2 ? 256 : 128); }
    mbar_init(w_full, 1);
    fence_barrier_init();
  }
  if (role == 2) {
    if constexpr (CG2) { tmem_alloc_pair(s_tmem, Cfg::TMEM_COLS); tmem_relinquish_pair(); }
    else { tmem_alloc(s_tmem, Cfg::TMEM_COLS); tmem_relinquish(); }
  }
  tc_fence_before();
  if constexpr (CG2) cluster_sync_all();  // the peer's barriers are initialised before anything is signalled on them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  // Programmatic dependent launch, armed: once every CTA of this grid has got here (or exited) the next kernel of the
  // stream may be scheduled, so its CTAs take over an SM the moment one of ours leaves it and run their prologue
  // (barrier init, TMEM allocation, resident-weight loads) under our tail.  Safe because every consumer of another
  // kernel's output in this file sits behind griddepcontrol.wait (warp 0), which waits for the WHOLE prior grid and its
  // memory flush; epilogue stores only follow MMAs that follow those loads.  (UB_DBG bit 16 disarms it for A/B timing.)
  if (!(p.dbg & 16)) pdl_launch_dependents();

  if (role == 0) {
    // ================= TMA producer (whole warp runs the loop; one elected lane issues) =================
    const bool leader = elect_one();
    if constexpr (WS) {
      // weight-stationary: with gridDim a multiple of n_tiles every tile of this CTA has the same n-tile.
      // Weights are constants of the layer: their loads may precede the wait on the previous kernel.
      // every tile of this CTA (pair) has the same n-tile: the host makes the CTA (pair) count a multiple of n_tiles
      const int nt = CG2 ? static_cast<int>(blockIdx.x >> 1) % n_tiles : static_cast<int>(blockIdx.x) % n_tiles;
      if (leader) {
        if constexpr (CG2) {
          // both halves are counted on the even CTA's barrier (its MMA warp is the only reader)
          if (cta_rank == 0) mbar_arrive_expect_tx(w_full, static_cast<uint32_t>(2 * n_taps * Cfg::B_STAGE_BYTES));
        } else {
          mbar_arrive_expect_tx(w_full, static_cast<uint32_t>(n_taps * Cfg::B_STAGE_BYTES));
        }
#pragma unroll 1
        for (int j = 0; j < n_taps; ++j) {  // consumption order: c outer, tap inner
          const int c = j / TAPS, tap = j % TAPS;
          if constexpr (CG2)
            tma_load_2d_pair(b_ring + j * Cfg::B_STAGE_BYTES, &p.tmB, w_full, tap * p.cin + c * BK,
                             nt * BLOCK_N + static_cast<int>(cta_rank) * Cfg::B_ROWS);
          else if constexpr (KXS) tma_load_2d(b_ring + j * Cfg::B_STAGE_BYTES, &p.tmB, w_full, c * BK, tap * BLOCK_N);
          else tma_load_2d(b_ring + j * Cfg::B_STAGE_BYTES, &p.tmB, w_full, tap * p.cin + c * BK, nt * BLOCK_N);
        }
      }
      __syncwarp();
    }
    pdl_wait_prior_grid();  // activations of the previous layer are complete and visible from here on
    int sa = 0, sb = 0;
    uint32_t pa = 0, pb = 0;
    TileIter ti = tile_iter(t_first, t_step);
#pragma unroll 1
    for (int t = t_first; t < total_tiles; t += t_step, ti.next()) {
      const int nt = tile_nt(ti);
      const int x0 = tile_tx(ti) * kXStep - (KXS ? 1 : 0);  // input column of tile column 0 (may lie past the image: zero fill)
      const int y0 = tile_ty(ti) * kTileRows;
      const int b = tile_b(ti);  // may be the padding image (== B) of a CTA pair: TMA zero-fills
#pragma unroll 1
      for (int c = 0; c < k_chunks; ++c) {
        mbar_wait(&a_empty[sa], pa ^ 1u);
        if (leader) {
          if constexpr (CG2) {
            // both CTAs' boxes are counted on the leader's barrier; only the leader announces the bytes
            if (cta_rank == 0) mbar_arrive_expect_tx(&a_full[sa], static_cast<uint32_t>(2 * Cfg::A_BOX_BYTES));
            tma_load_4d_pair(a_ring + sa * Cfg::A_STAGE_BYTES, &p.tmA, &a_full[sa], (x0 - Cfg::PADX) * 8, y0 - Cfg::PADY,
                             c * (BK / 8), b);
          } else {
            mbar_arrive_expect_tx(&a_full[sa], static_cast<uint32_t>(Cfg::A_BOX_BYTES));
            tma_load_4d(a_ring + sa * Cfg::A_STAGE_BYTES, &p.tmA, &a_full[sa], (x0 - Cfg::PADX) * 8, y0 - Cfg::PADY,
                        c * (BK / 8), b);
          }
        }
        __syncwarp();
        if (++sa == p.sa) { sa = 0; pa ^= 1u; }
        if constexpr (!WS) {
#pragma unroll
          for (int tap = 0; tap < TAPS; ++tap) {
            mbar_wait(&b_empty[sb], pb ^ 1u);
            if (leader) {
              if constexpr (CG2) {
                if (cta_rank == 0) mbar_arrive_expect_tx(&b_full[sb], 2 * Cfg::B_STAGE_BYTES);
                tma_load_2d_pair(b_ring + sb * Cfg::B_STAGE_BYTES, &p.tmB, &b_full[sb], tap * p.cin + c * BK,
                                 nt * BLOCK_N + static_cast<int>(cta_rank) * Cfg::B_ROWS);
              } else {
                mbar_arrive_expect_tx(&b_full[sb], Cfg::B_STAGE_BYTES);
                if constexpr (KXS) tma_load_2d(b_ring + sb * Cfg::B_STAGE_BYTES, &p.tmB, &b_full[sb], c * BK, tap * BLOCK_N);
                else tma_load_2d(b_ring + sb * Cfg::B_STAGE_BYTES, &p.tmB, &b_full[sb], tap * p.cin + c * BK, nt * BLOCK_N);
              }
            }
            __syncwarp();
            if (++sb == p.sb) { sb = 0; pb ^= 1u; }
          }
        }
      }
    }
  } else if (role == 1 && cta_rank == 0) {
    // ================= MMA issuer (whole warp runs the loop; ONE fixed lane issues and commits) =================
    // (CTA pairs: the even CTA issues M = 256 MMAs for both; the odd CTA's warp 1 has nothing to do)
    const bool leader = elect_one();
    constexpr uint32_t idesc = umma_idesc_bf16(CG2 ? 2 * kTileM : kTileM, BLOCK_N);
    auto mma = [&](auto acc_tag, uint32_t d, uint32_t al, uint32_t ah, uint32_t bl, uint32_t bh) {
      constexpr bool kAcc = decltype(acc_tag)::value;
      if constexpr (CG2) umma_bf16_lo_pair<kAcc>(d, al, ah, bl, bh, idesc);
      else umma_bf16_lo<kAcc>(d, al, ah, bl, bh, idesc);
    };
    auto commit = [&](uint64_t* bar) {
      if constexpr (CG2) umma_commit_pair(bar);
      else umma_commit(bar);
    };
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    // descriptors advance by plain 32-bit adds on the low word ((address >> 4) field); high words are constant
    const uint64_t desc_a = umma_smem_desc_interleaved(smem_u32(a_ring), Cfg::A_PLANE_BYTES, Cfg::SBO_BYTES);
    const uint64_t desc_b = umma_smem_desc<Cfg::SWZ_B>(smem_u32(b_ring));
    const uint32_t a_hi = static_cast<uint32_t>(desc_a >> 32), b_hi = static_cast<uint32_t>(desc_b >> 32);
    const uint32_t a_lo_ring = static_cast<uint32_t>(desc_a), b_lo_ring = static_cast<uint32_t>(desc_b);
    constexpr uint32_t A_STAGE_U = Cfg::A_STAGE_BYTES >> 4;
    constexpr uint32_t A_KSTEP_U = (2 * Cfg::A_PLANE_BYTES) >> 4;  // one MMA = 16 channels = 2 cg planes
    constexpr uint32_t B_STAGE_U = Cfg::B_STAGE_BYTES >> 4;
    if constexpr (WS) {
      mbar_wait(w_full, 0);
      tc_fence_after();
    }
    int sa = 0, sb = 0;
    uint32_t pa = 0, pb = 0;
    int it = 0;
    const bool no_mma = (p.dbg & 2) != 0;  // ablation hook, hoisted out of the issue loop
    const int k_steps = p.k_steps > 0 ? p.k_steps : BK / 16;
#pragma unroll 1
    for (int t = t_first; t < total_tiles; t += t_step, ++it) {
      const int acc = it % ACC;
      const uint32_t use = static_cast<uint32_t>(it / ACC);
      mbar_wait(&t_empty[acc], (use & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_u + static_cast<uint32_t>(acc * Cfg::ACC_STRIDE);
#pragma unroll 1
      for (int c = 0; c < k_chunks; ++c) {
        mbar_wait(&a_full[sa], pa);
        tc_fence_after();
        const uint32_t a_lo = a_lo_ring + static_cast<uint32_t>(sa) * A_STAGE_U;
        if constexpr (WS) {
          // resident weights: nothing to wait for between the taps, so the elected lane issues the whole chunk
          // (TAPS * BK/16 MMAs), frees the A stage and — after the last chunk — publishes the accumulator in ONE
          // divergent region.  The thin layers have 6-18 MMAs per tile and were bound by this warp's own
          // instruction stream (ncu: ~110 instructions and ~850 cycles per tile in the head layer).
          if (leader) {
            if (!no_mma) {
#pragma unroll
              for (int tap = 0; tap < TAPS; ++tap) {
                const uint32_t a_tap = a_lo + static_cast<uint32_t>(KXS ? tap * Cfg::PXW : (tap / KSIZE) * Cfg::PXW + (tap % KSIZE));
                const uint32_t b_lo = b_lo_ring + static_cast<uint32_t>(c * TAPS + tap) * B_STAGE_U;
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                  if (k >= k_steps) continue;  // zero-padded tail of the chunk (host: fill_gemm)
                  const uint32_t al = a_tap + k * A_KSTEP_U, bl = b_lo + 2 * k;
                  if (tap == 0 && k == 0) {
                    if (c == 0) mma(std::false_type{}, d_tmem, al, a_hi, bl, b_hi);
                    else mma(std::true_type{}, d_tmem, al, a_hi, bl, b_hi);
                  } else {
                    mma(std::true_type{}, d_tmem, al, a_hi, bl, b_hi);
                  }
                }
              }
            }
            commit(&a_empty[sa]);
            if (c == k_chunks - 1) commit(&t_full[acc]);
          }
          __syncwarp();
        } else {
#pragma unroll
          for (int tap = 0; tap < TAPS; ++tap) {
            // kx-stacked: tap = ky (one box row down); else tap = ky*3+kx (row down, pixel right)
            const uint32_t a_tap = a_lo + static_cast<uint32_t>(KXS ? tap * Cfg::PXW : (tap / KSIZE) * Cfg::PXW + (tap % KSIZE));
            mbar_wait(&b_full[sb], pb);
            tc_fence_after();
            const uint32_t b_lo = b_lo_ring + static_cast<uint32_t>(sb) * B_STAGE_U;
            if (leader) {
              if (!no_mma) {
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                  if (k >= k_steps) continue;  // zero-padded tail of the chunk (host: fill_gemm)
                  const uint32_t al = a_tap + k * A_KSTEP_U, bl = b_lo + 2 * k;
                  if (tap == 0 && k == 0) {
                    if (c == 0) mma(std::false_type{}, d_tmem, al, a_hi, bl, b_hi);
                    else mma(std::true_type{}, d_tmem, al, a_hi, bl, b_hi);
                  } else {
                    mma(std::true_type{}, d_tmem, al, a_hi, bl, b_hi);
                  }
                }
              }
              commit(&b_empty[sb]);
              if (tap == TAPS - 1) commit(&a_empty[sa]);
            }
            __syncwarp();
            if (++sb == p.sb) { sb = 0; pb ^= 1u; }
          }
        }
        if (++sa == p.sa) { sa = 0; pa ^= 1u; }
      }
      if constexpr (!WS) {
        if (leader) commit(&t_full[acc]);
        __syncwarp();
      }
    }
  } else if (role < 0) {
    // ================= epilogue (kEpiGroups x 128 threads; warp w owns TMEM lanes 32*(w%4)..) =================
#include "ub_conv_epilogue.inl"
  }

  // ---------------- teardown
  __syncwarp();
  tc_fence_before();
  if constexpr (CG2) cluster_sync_all();  // the peer may still signal our barriers / read our operands until here
  else __syncthreads();
  if (role == 2) {
    tc_fence_after();
    if constexpr (CG2) tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace ub
