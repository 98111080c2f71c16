// ub_conv_rs.cuh — "row-streaming" conv3x3(pad 1)+foldedBN+ReLU for the thin layers (Cout = 32 or 64) at the high
// resolutions: the three ky taps are GEMM columns (N = 3*Cout) whose sums are formed INSIDE TMEM by accumulating
// consecutive image rows into a sliding window of column blocks, so the epilogue reads each output row once
// (Cout columns), adds nothing and shuffles nothing.
//
// Replaces for those layers: models/processing_blocks.py:43-48 (Conv2d+BatchNorm2d+ReLU), :73 (MaxPool2d, POOL epilogue),
// models/UNet.py:73 + models/losses.py:48-58 (HEAD epilogue: 1x1 head, argmax, confusion counts).
//
// Why.  Micro-benchmarks on the B200 (profiles/r02_call4_ubench_mma_cadence.txt): one tcgen05.mma M=128 K=16 costs
// max(N/2, 32 + N/4, 48) cycles — the shared-memory read of A (4 KB at 128 B/cycle) plus B, never less than 48 — so a
// 9-tap conv with N = 32 or 64 runs at a third / two thirds of the tensor rate, and stacking three taps along N
// (N = 96 / 192) restores it.  The kx-stacked kernel (ub_conv_tc.cuh) does that with the kx taps and pays in the
// epilogue: it reads 3*Cout TMEM columns per pixel and adds the neighbours' blocks with 2 warp shuffles per value; ncu
// showed those layers epilogue-bound (dec4.conv.3: 32 % tensor-pipe, enc1.conv.3: 78 %).  Here the ky taps are stacked:
//
//   tile tau = ONE image row y' of a 30-pixel-wide band (x 4 row-quarters of the image side by side in the 128 lanes);
//   MMA(A = row y' [shifted by kx-1 pixels], B = [W(ky=2); W(ky=1); W(ky=0)][kx])  ->  column blocks of rows y'-1, y', y'+1
//   TMEM holds R column blocks of Cout columns; row r lives in block r % R; tile tau accumulates into blocks
//   (tau-1, tau, tau+1) % R, overwrites (accumulate = 0) the block of row tau+1 with its first MMA, and after its last
//   MMA row tau-1 is complete: one tcgen05.commit hands that block to the epilogue, which reads Cout columns.
//
// The kx taps are three MMAs whose A descriptor starts one pixel earlier / later (pixels are 16 bytes apart in the
// box), as in the 9-tap kernel, but each now carries N = 3*Cout.  Per 128 output... per 120 useful lanes a row costs the
// same MMA time as the kx-stacked tile, the input is read 1.03x (one halo row per 64) instead of 1.5x (4+2 rows per
// 4), and the epilogue shrinks to: 1 tcgen05.ld per 32 channels, bias, ReLU, (pool | head), store.
//
// Geometry.  A task = (image b, band j): output columns [30j, 30j+30), all H rows.  The 128 accumulator lanes are 4
// streams x 32 pixels: stream q walks rows [q*H/4, (q+1)*H/4) — lanes 0 and 31 of a stream are halo columns (their
// results are discarded), and each stream reads one extra row above and below its quarter (tiles t = 0 and H/4+1), so
// a task is H/4 + 2 tiles.  A stage holds [plane][stream][32 px][8 ch] — the no-swizzle K-major UMMA layout with
// LBO = 2048 B (planes) and SBO = 128 B (8-pixel groups, uniform across the streams).
//
// Operand fill.  The first version used ONE 5-D TMA box per (tile, chunk) — {32 px * 8 ch, 1 row, 4 quarters, BK/8 planes}
// — and measured no faster than the kx-stacked kernel: both sat at ~33-48 cycles per 512-byte box row (enc1.conv.3:
// 1,600 cycles per row tile against 1,170 of MMA time; the head layer 930 against 380), i.e. the TMA unit's per-row
// cost on these short, 16-byte-misaligned rows bounds the thin layers, not the tensor pipe or the epilogue
// (profiles/r02_call5_*).  The A operand is therefore filled by two producer warps with cp.async (LDGSTS.128: one warp
// instruction = the 32 pixels x 16 bytes of one (plane, stream) row, 512 contiguous bytes of the blocked tensor),
// completion signalled on the stage's mbarrier (cp.async.mbarrier.arrive.noinc), zero fill (src-size 0) for x = -1,
// x >= W, the row above stream 0 / below stream 3 and channel planes the view does not have.  The MMA thread orders
// those generic-proxy writes before its async-proxy reads with fence.proxy.async after the barrier wait.  The
// resident weights are still TMA boxes (loaded once per CTA).
//
// Global counters.  Each CTA numbers its tiles tau = 0, 1, ... across its tasks; rows sigma = tau - 1 become complete in
// the same order; the two rows per task that are not outputs (the halo rows rho = 0 and rho = H/4+1) flow through the
// same protocol and are masked, so the barrier arithmetic has no special cases.  Epilogue group g takes the row PAIRS
// p = (sigma+1)/2 with p % 4 == g: the two rows of a 2x2 pooling window are consecutive rows of the same thread.
#pragma once
#include "ub_conv_tc.cuh"

namespace ub {

constexpr int kRsOut = 30;      // output columns per band (32 lanes per stream, one halo lane each side)
constexpr int kRsMaxGroups = 4;
constexpr int kRsMaxBlocks = 16;
constexpr int kRsMaxSA = 24;    // A stages: a row tile is small (4-16 KB), so the ring is deep — the bytes in flight, not the
                                // stage count, are what hides the memory latency
// epilogue groups of 4 warps: the 32-channel layers (the head among them) have short MMA phases per row and the longest
// per-pixel epilogue -> 4 groups; the 64-channel rows take three times as long on the tensor pipe -> 2 groups (and twice
// the registers per thread for the held pooling row)
__host__ __device__ constexpr int rs_groups(int cout) { return cout == 64 ? 2 : 4; }
__host__ __device__ constexpr int rs_threads(int cout) { return 128 + 128 * rs_groups(cout); }

template <int COUT, int BK>
struct RsCfg {
  static_assert(COUT == 32 || COUT == 64, "row-streaming kernel: Cout 32 or 64");
  static_assert(BK == 64 || BK == 32 || BK == 16, "channel chunk");
  static constexpr int N3 = 3 * COUT;
  static constexpr int R = (512 / COUT) > kRsMaxBlocks ? kRsMaxBlocks : (512 / COUT);  // row blocks in TMEM
  static constexpr int TMEM_COLS = 512;
  static constexpr int A_PLANE_BYTES = 4 * 32 * 16;               // [stream][pixel][8 ch] = LBO
  static constexpr int A_BOX_BYTES = (BK / 8) * A_PLANE_BYTES;    // one TMA box
  static constexpr int A_STAGE_BYTES = A_BOX_BYTES;               // 4 / 8 / 16 KB: keeps the ring 1024 B aligned
  static constexpr int SWZ_B = 2 * BK;
  static constexpr int B_STAGE_BYTES = N3 * SWZ_B;                // one (chunk, kx): rows (2-ky)*Cout + co
  static_assert(B_STAGE_BYTES % 1024 == 0, "swizzle atom alignment");
  static constexpr int G = rs_groups(COUT);
  static_assert((R / 2) % G == 0, "a block must always be drained by the same epilogue group (row pairs, period R)");
  // fixed-size tail in front of the operand regions
  static constexpr int TAIL_BIAS = 0;                              // [COUT] floats
  static constexpr int TAIL_BAR = 256;                             // barriers
  static constexpr int N_BARS = 2 * kRsMaxSA + 2 * kRsMaxBlocks + 1;
  static constexpr int TAIL_CM = TAIL_BAR + N_BARS * 8;
  static constexpr int TAIL_DICE = TAIL_CM + kRsMaxGroups * kMaxClasses * kMaxClasses * 8;
  static constexpr int TAIL_TMEM = TAIL_DICE + kRsMaxGroups * kMaxClasses * 2 * 8;
  static constexpr int TAIL_BYTES = ((TAIL_TMEM + 16 + 1023) / 1024) * 1024;
  static constexpr int OPERAND_BUDGET = kSmemLimit - 1024 - TAIL_BYTES;
};

// 16-byte asynchronous copy global -> shared, bypassing L1; `ok` false: nothing is read and 16 zero bytes are written
__device__ __forceinline__ void cp_async_16_zfill(uint32_t smem_dst, const void* gsrc, bool ok) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_dst), "l"(gsrc), "r"(ok ? 16 : 0) : "memory");
}
// the executing thread's arrival on `bar` is deferred until all of its earlier cp.async copies have completed; .noinc:
// the arrival counts against the barrier's initial count (64 = every producer thread arrives once per stage)
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// D[tmem] (+)= A * B with a run-time accumulate flag (the window's fresh block)
__device__ __forceinline__ void umma_bf16_lo_rt(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                                uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\tsetp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}

// The MMAs of one (row tile, channel chunk), straight-line: the single issuing thread's own instruction stream bounds
// the 32-channel layers (ncu source view of the first version: ~100 uniform-datapath instructions per row tile at
// ~9 cycles each = the 900-cycle tile period, with 7 MMAs = 390 cycles of tensor work), so everything that can be is a
// compile-time constant: CASE = where the three-block window sits (0: contiguous, 1: row tau+1 wraps to block 0, 2: rows
// tau and tau+1 wrap, 3: contiguous and every block untouched so far), FIRST = this chunk carries the tile's first contribution to the fresh block (accumulate = 0).
template <int COUT, int BK, int CASE, bool FIRST>
__device__ __forceinline__ void rs_issue_chunk(uint32_t d0, uint32_t d_wrap, uint32_t a_stage, uint32_t a_hi, uint32_t bl_c,
                                               uint32_t b_hi, bool skip_last_k) {
  using Cfg = RsCfg<COUT, BK>;
  constexpr uint32_t idesc1 = umma_idesc_bf16(kTileM, COUT);
  constexpr uint32_t idesc2 = umma_idesc_bf16(kTileM, 2 * COUT);
  constexpr uint32_t idesc3 = umma_idesc_bf16(kTileM, 3 * COUT);
  constexpr uint32_t A_KSTEP_U = (2 * Cfg::A_PLANE_BYTES) >> 4;   // one MMA = 16 channels = 2 planes
  constexpr uint32_t B_STAGE_U = Cfg::B_STAGE_BYTES >> 4;
  constexpr uint32_t B_BLOCK_U = (COUT * Cfg::SWZ_B) >> 4;        // Cout weight rows = one ky block (whole swizzle atoms)
#pragma unroll
  for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
    for (int k = 0; k < BK / 16; ++k) {
      if (k == BK / 16 - 1 && BK / 16 > 1 && skip_last_k) continue;   // zero-padded tail of the chunk (k_steps = BK/16 - 1)
      // tap kx reads pixel (lane + kx - 1) of the stream: the descriptor starts one 16-byte pixel earlier / later
      const uint32_t al = a_stage + static_cast<uint32_t>(k) * A_KSTEP_U + static_cast<uint32_t>(kx) - 1u;
      const uint32_t bl = bl_c + static_cast<uint32_t>(kx) * B_STAGE_U + 2u * static_cast<uint32_t>(k);
      const bool first = FIRST && kx == 0 && k == 0;
      if constexpr (CASE == 0) {
        if (!first) {
          umma_bf16_lo<true>(d0, al, a_hi, bl, b_hi, idesc3);
        } else {
          umma_bf16_lo<true>(d0, al, a_hi, bl, b_hi, idesc2);
          umma_bf16_lo<false>(d0 + 2 * COUT, al, a_hi, bl + 2 * B_BLOCK_U, b_hi, idesc1);
        }
      } else if constexpr (CASE == 1) {   // rows tau-1, tau in the last two blocks, row tau+1 in block 0
        umma_bf16_lo<true>(d0, al, a_hi, bl, b_hi, idesc2);
        if (!first) umma_bf16_lo<true>(d_wrap, al, a_hi, bl + 2 * B_BLOCK_U, b_hi, idesc1);
        else umma_bf16_lo<false>(d_wrap, al, a_hi, bl + 2 * B_BLOCK_U, b_hi, idesc1);
      } else if constexpr (CASE == 3) {   // linear window (ub_conv_rs2.cuh, overflow blocks): all three blocks are first touches
        if (!first) umma_bf16_lo<true>(d0, al, a_hi, bl, b_hi, idesc3);
        else umma_bf16_lo<false>(d0, al, a_hi, bl, b_hi, idesc3);
      } else {                            // row tau-1 in the last block, rows tau, tau+1 in blocks 0, 1
        umma_bf16_lo<true>(d0, al, a_hi, bl, b_hi, idesc1);
        if (!first) {
          umma_bf16_lo<true>(d_wrap, al, a_hi, bl + B_BLOCK_U, b_hi, idesc2);
        } else {
          umma_bf16_lo<true>(d_wrap, al, a_hi, bl + B_BLOCK_U, b_hi, idesc1);
          umma_bf16_lo<false>(d_wrap + COUT, al, a_hi, bl + 2 * B_BLOCK_U, b_hi, idesc1);
        }
      }
    }
  }
}

template <int COUT, int BK>
__global__ void __launch_bounds__(rs_threads(COUT), 1)
conv_rs_kernel(const __grid_constant__ ConvParams p) {
  using Cfg = RsCfg<COUT, BK>;
  constexpr int R = Cfg::R;
  constexpr int G = Cfg::G;
  constexpr int kRsThreads = rs_threads(COUT);
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  float* s_bias = reinterpret_cast<float*>(smem + Cfg::TAIL_BIAS);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::TAIL_BAR);
  uint64_t* a_full = bars;
  uint64_t* a_empty = bars + kRsMaxSA;
  uint64_t* blk_full = bars + 2 * kRsMaxSA;        // MMA -> epilogue: row sigma (block sigma % R) is complete
  uint64_t* blk_empty = blk_full + kRsMaxBlocks;   // epilogue -> MMA: the block's columns are in registers
  uint64_t* w_full = blk_empty + kRsMaxBlocks;
  unsigned long long* s_cm_all = reinterpret_cast<unsigned long long*>(smem + Cfg::TAIL_CM);
  double* s_dice_all = reinterpret_cast<double*>(smem + Cfg::TAIL_DICE);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + Cfg::TAIL_TMEM);
  uint8_t* a_ring = smem + Cfg::TAIL_BYTES;
  uint8_t* b_res = a_ring + p.sa * Cfg::A_STAGE_BYTES;   // resident weights: [chunk][kx] stages

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  // epilogue warps first, the producers / MMA issuer / TMEM allocator last: the scheduler favours the highest warp id
  const int role = warp - 4 * G;  // 0 producer (+ weights), 1 MMA issuer, 2 TMEM allocator, 3 producer; negative: epilogue
  const int k_chunks = p.cin / BK;
  const int hq = p.hq;                       // rows per stream (H / 4)
  const int tpt = hq + 2;                    // tiles per task
  const int bands = (p.W + kRsOut - 1) / kRsOut;
  const int n_tasks = p.B * bands;
  const int my_tasks = (n_tasks - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int my_tiles = my_tasks * tpt;

  // ---------------- one-time setup (nothing here depends on the previous kernel's output)
  for (int i = threadIdx.x; i < COUT; i += kRsThreads) s_bias[i] = p.bias[i];
  if (threadIdx.x < G * kMaxClasses * kMaxClasses) s_cm_all[threadIdx.x] = 0ull;
  if (threadIdx.x < G * kMaxClasses * 2) s_dice_all[threadIdx.x] = 0.0;
  if (role == 0 && lane == 0) tma_prefetch_desc(&p.tmB);
  if (role == 1 && lane == 0) {
    for (int i = 0; i < kRsMaxSA; ++i) { mbar_init(&a_full[i], 64); mbar_init(&a_empty[i], 1); }  // 64 = the two producer warps
    for (int i = 0; i < kRsMaxBlocks; ++i) { mbar_init(&blk_full[i], 1); mbar_init(&blk_empty[i], 128); }
    mbar_init(w_full, 1);
    fence_barrier_init();
  }
  if (role == 2) { tmem_alloc(s_tmem, Cfg::TMEM_COLS); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  if (!(p.dbg & 16)) pdl_launch_dependents();  // see conv_tc_kernel

  if (role == 0 || role == 3) {
    // ================= producers: warp 0 (also loads the weights) and warp 3 fill the A stages with cp.async =================
    if (role == 0) {
      const bool leader = elect_one();
      const int n_w = 3 * k_chunks;
      if (leader) {  // weights are constants of the layer: loaded before the wait on the previous kernel
        mbar_arrive_expect_tx(w_full, static_cast<uint32_t>(n_w * Cfg::B_STAGE_BYTES));
#pragma unroll 1
        for (int j = 0; j < n_w; ++j) {
          const int c = j / 3, kx = j % 3;
          tma_load_2d(b_res + j * Cfg::B_STAGE_BYTES, &p.tmB, w_full, c * BK, kx * Cfg::N3);
        }
      }
      __syncwarp();
    }
    pdl_wait_prior_grid();  // activations of the previous layer are complete and visible from here on
    const int half = role == 0 ? 0 : 1;      // which rows of a stage this warp copies
    constexpr int kRows = 4 * (BK / 8);      // (plane, stream) rows of one stage, 512 bytes each
    const size_t row_stride = static_cast<size_t>(p.W) * 8;            // elements per image row of one plane
    const size_t plane_stride = row_stride * p.H;
    const size_t image_stride = plane_stride * p.a_cgs;
    const size_t chunk_stride = plane_stride * (BK / 8);
    // this warp's rows: element offset of (plane pl, first row of quarter q) — the per-tile address is one 64-bit add
    size_t off[kRows / 2];
#pragma unroll
    for (int r = 0; r < kRows / 2; ++r) {
      const int row = 2 * r + half, pl = row >> 2, q = row & 3;
      off[r] = plane_stride * pl + row_stride * hq * q;
    }
    int sa = 0;
    uint32_t pa = 0;
#pragma unroll 1
    for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
      const int b = task / bands, j = task - b * bands;
      const int x = j * kRsOut - 1 + lane;   // this lane's pixel: lane 0 / 31 are the band's halo columns
      const bool x_ok = x >= 0 && x < p.W;
      // row -1 of the image (stream 0 at t = 0) is never dereferenced: those copies have src-size 0
      const __nv_bfloat16* img = p.a_src + image_stride * b + static_cast<size_t>(x_ok ? x : 0) * 8 - row_stride;
#pragma unroll 1
      for (int t = 0; t < tpt; ++t, img += row_stride) {
        // stream q reads image row q*hq - 1 + t: outside the image only for (q = 0, t = 0) and (q = 3, t = last)
        const bool top = t == 0, bottom = t == tpt - 1;
        const __nv_bfloat16* src_c = img;
#pragma unroll 1
        for (int c = 0; c < k_chunks; ++c, src_c += chunk_stride) {
          mbar_wait(&a_empty[sa], pa ^ 1u);
          const uint32_t dst0 = smem_u32(a_ring + sa * Cfg::A_STAGE_BYTES) + static_cast<uint32_t>(lane) * 16u +
                                static_cast<uint32_t>(half) * 512u;
          const int planes_left = p.a_planes - c * (BK / 8);   // planes of this chunk that exist in the view
#pragma unroll
          for (int r = 0; r < kRows / 2; ++r) {
            const int row = 2 * r + half, pl = row >> 2, q = row & 3;   // (pl and q are compile-time up to `half`)
            const bool ok = x_ok && pl < planes_left && !(top && q == 0) && !(bottom && q == 3);
            cp_async_16_zfill(dst0 + static_cast<uint32_t>(r) * 1024u, src_c + off[r], ok);
          }
          cp_async_mbar_arrive_noinc(&a_full[sa]);   // arrives once this thread's copies above have landed
          if (++sa == p.sa) { sa = 0; pa ^= 1u; }
        }
      }
    }
  } else if (role == 1) {
    // ================= MMA issuer (one elected lane issues and commits) =================
    const bool leader = elect_one();
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint64_t desc_a = umma_smem_desc_interleaved(smem_u32(a_ring), Cfg::A_PLANE_BYTES, 128);
    const uint64_t desc_b = umma_smem_desc<Cfg::SWZ_B>(smem_u32(b_res));
    const uint32_t a_hi = static_cast<uint32_t>(desc_a >> 32), b_hi = static_cast<uint32_t>(desc_b >> 32);
    const uint32_t a_lo_ring = static_cast<uint32_t>(desc_a), b_lo_res = static_cast<uint32_t>(desc_b);
    constexpr uint32_t A_STAGE_U = Cfg::A_STAGE_BYTES >> 4;
    constexpr uint32_t B_CHUNK_U = 3 * (Cfg::B_STAGE_BYTES >> 4);
    const bool skip_last_k = p.k_steps > 0 && p.k_steps < BK / 16;
    mbar_wait(w_full, 0);
    tc_fence_after();
    int sa = 0;
    uint32_t pa = 0;
    uint32_t a_stage = a_lo_ring;  // descriptor low word of stage `sa`
    int blk0 = R - 1;        // block of row tau - 1
    int fresh = 1;           // block of row tau + 1
    uint32_t fresh_par = 1;  // parity to wait for on the fresh block's `empty` barrier: (uses so far & 1) ^ 1
    const uint32_t d_wrap = tmem_u;  // block 0
    auto next_stage = [&]() {
      a_stage += A_STAGE_U;
      if (++sa == p.sa) { sa = 0; pa ^= 1u; a_stage = a_lo_ring; }
    };
#pragma unroll 1
    for (int tau = 0; tau < my_tiles; ++tau) {
      // the block about to be overwritten held row tau + 1 - R: its columns must be in the epilogue's registers
      mbar_wait(&blk_empty[fresh], fresh_par);
      const uint32_t d0 = tmem_u + static_cast<uint32_t>(blk0 * COUT);
      // chunk 0 carries the tile's first contribution to the fresh block
      mbar_wait(&a_full[sa], pa);
      fence_proxy_async();  // the stage was written through the generic proxy (cp.async); the MMA reads it through the async proxy
      tc_fence_after();
      if (leader) {
        if (blk0 <= R - 3) rs_issue_chunk<COUT, BK, 0, true>(d0, d_wrap, a_stage, a_hi, b_lo_res, b_hi, skip_last_k);
        else if (blk0 == R - 2) rs_issue_chunk<COUT, BK, 1, true>(d0, d_wrap, a_stage, a_hi, b_lo_res, b_hi, skip_last_k);
        else rs_issue_chunk<COUT, BK, 2, true>(d0, d_wrap, a_stage, a_hi, b_lo_res, b_hi, skip_last_k);
        umma_commit(&a_empty[sa]);
        if (k_chunks == 1 && tau >= 1) umma_commit(&blk_full[blk0]);  // row tau - 1 is complete
      }
      __syncwarp();
      next_stage();
#pragma unroll 1
      for (int c = 1; c < k_chunks; ++c) {
        mbar_wait(&a_full[sa], pa);
        fence_proxy_async();
        tc_fence_after();
        if (leader) {
          const uint32_t bl_c = b_lo_res + static_cast<uint32_t>(c) * B_CHUNK_U;
          if (blk0 <= R - 3) rs_issue_chunk<COUT, BK, 0, false>(d0, d_wrap, a_stage, a_hi, bl_c, b_hi, skip_last_k);
          else if (blk0 == R - 2) rs_issue_chunk<COUT, BK, 1, false>(d0, d_wrap, a_stage, a_hi, bl_c, b_hi, skip_last_k);
          else rs_issue_chunk<COUT, BK, 2, false>(d0, d_wrap, a_stage, a_hi, bl_c, b_hi, skip_last_k);
          umma_commit(&a_empty[sa]);
          if (c == k_chunks - 1 && tau >= 1) umma_commit(&blk_full[blk0]);
        }
        __syncwarp();
        next_stage();
      }
      blk0 = blk0 + 1 == R ? 0 : blk0 + 1;
      if (++fresh == R) { fresh = 0; fresh_par ^= 1u; }  // rows u*R .. u*R + R - 1 are the u-th occupants of their blocks
    }
  } else if (role < 0) {
    // ================= epilogue: G groups x 128 threads; thread (wq, lane) owns TMEM lane 32*wq + lane =================
    const int wq = warp & 3;                 // stream = row quarter
    const int grp = warp >> 2;
    const int et = threadIdx.x & 127;
    unsigned long long* s_cm = s_cm_all + grp * (kMaxClasses * kMaxClasses);
    double* s_dice = s_dice_all + grp * (2 * kMaxClasses);
    HeadState hs;
    hs.init();
    uint32_t held[COUT / 2];                 // POOL: the first row of a pooling pair, packed bf16x2 (bias added)
#pragma unroll
    for (int i = 0; i < COUT / 2; ++i) held[i] = 0u;
    const int n_rows = my_tiles - 1;         // rows sigma = 0 .. my_tiles - 2 are committed (by tiles 1 .. my_tiles - 1)
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(wq * 32) << 16);
    const int oh = p.mode == EPI_POOL ? p.H >> 1 : p.H, ow = p.mode == EPI_POOL ? p.W >> 1 : p.W;
    const size_t plane = static_cast<size_t>(oh) * ow;
#pragma unroll 1
    for (int pr = grp; 2 * pr - 1 < n_rows; pr += G) {
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        const int sigma = 2 * pr - 1 + half;
        if (sigma < 0 || sigma >= n_rows) continue;
        const int ti = sigma / tpt, rho = sigma - ti * tpt;       // task (of this CTA) and row within the task
        const int task = static_cast<int>(blockIdx.x) + ti * static_cast<int>(gridDim.x);
        const int b = task / bands, j = task - b * bands;
        const int y = wq * hq + rho - 1, x = j * kRsOut + lane - 1;
        const bool valid_row = rho >= 1 && rho <= hq;
        const bool inb = valid_row && lane >= 1 && lane <= kRsOut && x < p.W && y < p.H;
        const int blk = sigma % R;
        const uint32_t use = static_cast<uint32_t>(sigma / R);
        int tgt_pre = -1;
        if (COUT == 32 && p.mode == EPI_HEAD && p.targets && valid_row) {
          head_begin_tile(p, hs, (p.img_base + b) / p.ref_batch, grp, et, lane, s_cm, s_dice);
          if (inb) tgt_pre = head_load_target(p, b, y, x);
        }
        mbar_wait(&blk_full[blk], use & 1u);
        tc_fence_after();
        const uint32_t t_addr = lane_addr + static_cast<uint32_t>(blk * COUT);
        auto release_blk = [&]() {
          tc_fence_before();
          mbar_arrive(&blk_empty[blk]);
        };
        if (!valid_row) {  // a halo row of the task: nothing to emit, the block goes straight back
          release_blk();
          continue;
        }
        if (COUT == 32 && p.mode == EPI_HEAD) {
          if constexpr (COUT == 32) {
            uint32_t r[32];
            tmem_ld_32x32(t_addr, r);
            tmem_ld_wait();
            release_blk();
            if (p.n_classes == 3) head_pixel<3>(p, hs, r, inb, b, y, x, tgt_pre);
            else head_pixel<0>(p, hs, r, inb, b, y, x, tgt_pre);
          }
          continue;
        }
        // ---- STORE / POOL: bias -> bf16 pack -> (vertical max with the held row) -> (horizontal max) -> ReLU -> store
#pragma unroll
        for (int ch = 0; ch < COUT; ch += 32) {
          uint32_t r[32];
          tmem_ld_32x32(t_addr + static_cast<uint32_t>(ch), r);
          tmem_ld_wait();
          if (ch + 32 >= COUT) release_blk();
          uint32_t pk[16];
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 bz = *reinterpret_cast<const float4*>(s_bias + ch + 4 * j4);
            const float2 a = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 0]), __uint_as_float(r[4 * j4 + 1])),
                                        make_float2(bz.x, bz.y));
            const float2 c2 = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 2]), __uint_as_float(r[4 * j4 + 3])),
                                         make_float2(bz.z, bz.w));
            pk[2 * j4 + 0] = pack_bf16(a.x, a.y);
            pk[2 * j4 + 1] = pack_bf16(c2.x, c2.y);
          }
          if (p.mode != EPI_POOL) {
            // bf16 rounding is monotone and keeps the sign: ReLU on the packed pairs (max with +0) gives the same bits
            if (p.relu) {
#pragma unroll
              for (int i = 0; i < 16; ++i) pk[i] = bf16x2_max(pk[i], 0u);
            }
            if (inb) {
              uint4* dst = reinterpret_cast<uint4*>(p.out) +
                           ((static_cast<size_t>(b) * p.out_cgs + p.out_cg0 + (ch >> 3)) * plane + static_cast<size_t>(y) * ow + x);
#pragma unroll
              for (int g4 = 0; g4 < 4; ++g4)
                dst[static_cast<size_t>(g4) * plane] = make_uint4(pk[4 * g4], pk[4 * g4 + 1], pk[4 * g4 + 2], pk[4 * g4 + 3]);
            }
          } else if (half == 0) {
            // MaxPool2d(2,2): y is even here (rho odd, quarter height even): keep the row, its partner is this thread's
            // next row.  (max commutes with the monotone bf16 rounding, so pooling the packed values equals pooling in fp32)
#pragma unroll
            for (int i = 0; i < 16; ++i) held[ch / 2 + i] = pk[i];
          } else {
            // second row of the window: vertical max in registers, then the horizontal pair (x even = odd lane, x + 1)
            // as a reduce-scatter: each lane keeps half of the channel groups and stores them
            const bool odd = lane & 1;
            const int partner = odd ? ((lane + 1) & 31) : ((lane + 31) & 31);
            uint32_t m[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const uint32_t v_keep = bf16x2_max(odd ? pk[i] : pk[8 + i], odd ? held[ch / 2 + i] : held[ch / 2 + 8 + i]);
              const uint32_t v_send = bf16x2_max(odd ? pk[8 + i] : pk[i], odd ? held[ch / 2 + 8 + i] : held[ch / 2 + i]);
              m[i] = bf16x2_max(v_keep, __shfl_sync(0xffffffffu, v_send, partner));
              if (p.relu) m[i] = bf16x2_max(m[i], 0u);
            }
            if (inb) {  // both lanes of a pair are inside together (W even, lanes 1..30 pair up as (1,2) ... (29,30))
              uint4* dst = reinterpret_cast<uint4*>(p.out) +
                           ((static_cast<size_t>(b) * p.out_cgs + p.out_cg0 + (ch >> 3) + (odd ? 0 : 2)) * plane +
                            static_cast<size_t>(y >> 1) * ow + (x >> 1));
              dst[0] = make_uint4(m[0], m[1], m[2], m[3]);
              dst[plane] = make_uint4(m[4], m[5], m[6], m[7]);
            }
          }
        }
      }
    }
    if (COUT == 32 && p.mode == EPI_HEAD && p.targets && hs.cur_batch >= 0)
      head_flush(p, hs, hs.cur_batch, grp, et, lane, s_cm, s_dice);
  }

  // ---------------- teardown
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (role == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace ub
