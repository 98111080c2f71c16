// ub_conv_rs2.cuh — a full-resolution conv pair as ONE kernel, the tensor between the two convolutions living only in shared
// memory.  Two instantiations, both the default:
//   the first encoder stage  enc1.conv.0 (+ folded stem) -> enc1.conv.3 -> MaxPool2d(2,2)   (UB_NO_FUSE_ENC1=1: two kernels)
//   the last decoder stage   dec4.conv.0 (+ folded stem) -> dec4.conv.3 -> 1x1 head, argmax, counts   (UB_NO_FUSE_DEC4=1)
// Replaces models/processing_blocks.py:43-48 twice per stage, :73, models/UNet.py:73 and models/losses.py:48-58.  The text below
// describes the encoder instantiation; the decoder one differs in widths (32 channels, 8 KB rows) and in its stage-2 epilogue.
//
// Why.  enc1.conv.0 writes 8.4 MB per image that enc1.conv.3 reads straight back: 2.4 GB of the 12.7 GB a micro-batch of
// 148 images moves through HBM, and conv.0 alone is bound by HBM WRITE bandwidth (1.18 GB in 359 us against a
// 3.9 TB/s write ceiling) while the tensor pipe idles.  Keeping that tensor in L2 by sub-batching saved the energy but
// not the time (profiles/r02_call41_*): the saving needs one kernel.
//
// How.  Two chained row-streaming convolutions (ub_conv_rs.cuh) in one persistent CTA.  A task = (image, band of 28
// output columns); the 128 accumulator lanes are 4 streams (row quarters) x 32 pixels: lane l of a band is column
// 28*j - 2 + l.  Stage 1 is exact on lanes 1..30 (it reads lanes 0..31 of the packed image, its kx taps reach one pixel to
// either side), stage 2 on lanes 2..29.  Per step t of a task (t = 0 .. H/4 + 3, input row q*H/4 - 2 + t of stream q):
//
//   producers   cp.async the packed-image row (2 channel planes x 4 streams x 32 px x 16 B = 4 KB) into the A1 ring
//   MMA         stage 1: 3 MMAs (kx) N = 192 (ky-stacked), K = 16 into the S1 window (4 TMEM blocks of 64 columns);
//               row t-1 of stage 1 is then complete -> commit
//   S1 epilogue TMEM -> bias, ReLU, bf16; zero outside the image (conv.3's zero padding) -> the A2 ring in shared memory
//               in the UMMA operand layout [8 planes][4 streams][32 px][16 B]  (generic-proxy stores, fence.proxy.async,
//               mbarrier arrive)
//   MMA         stage 2 on that row: 3 x 4 MMAs N = 192, K = 64 into the S2 window (TMEM columns 256..511); the row before
//               it is complete -> commit
//   S2 epilogue bias, 2x2 max with the held row / the neighbour lane, ReLU, store — as in conv_rs_kernel.
//
// The stage-1 rows and the stage-2 rows run through the same sliding-window protocol as conv_rs_kernel (global row
// counters across a CTA's tasks; the rows at a task's edge that take contributions from the neighbouring task are not
// outputs and are masked); each stage has its own issuing thread, coupled only through the shared-memory row ring.
#pragma once
#include <cstdio>
#include "ub_conv_rs.cuh"
#include "ub_head.cuh"

namespace ub {

constexpr int kRs2Out = 28;        // output columns per band
constexpr int kRs2MaxSA1 = 12;     // A1 stages
constexpr int kRs2MaxSA2 = 8;      // A2 stages
constexpr int kRs2MaxWin = 8;      // TMEM blocks per stage window (barrier arrays)

// <C1, C2, BK1, HEAD>: stage 1 = conv3x3 BK1 (-> K per tap) -> C1 channels, stage 2 = conv3x3 C1 -> C2 channels with
//   HEAD = false: MaxPool2d(2,2) + store            (the first encoder stage: <64, 64, 16, false>)
//   HEAD = true : 1x1 head, argmax, confusion counts (the last decoder stage, dec4.conv.0 -> conv.3 -> out: <32, 32, 64, true>)
// SPARE: two overflow blocks behind each stage's ring of W TMEM blocks ("linear window").  Input row tau adds to the output rows
//   tau-1, tau, tau+1 = ring blocks h, h+1, h+2 (h = (tau-1) mod W) with ONE N = 3C MMA per k-step.  Without overflow blocks
//   the two positions per lap where that span crosses the end of the ring split every MMA in two (N = 2C + C, 48 cycles each
//   instead of 56 for N = 96: 29 % of the last decoder stage's tensor-pipe time).  With them the span simply runs on into
//   blocks W, W+1: a row whose home is block 0 (1) collects its first two (one) contributions in overflow block W (W+1) and
//   the rest at home, and its epilogue adds the two.  Barriers are unchanged: an overflow block is drained together with its
//   home block, by the same thread, before that thread arrives on the home block's `empty` barrier.
template <int C1_, int C2_, int BK1_, bool HEAD_, bool SPARE_ = false>
struct Rs2Cfg {
  static constexpr int C1 = C1_, C2 = C2_, BK1 = BK1_, BK2 = C1_;
  static constexpr bool HEAD = HEAD_, SPARE = SPARE_;
  using S1 = RsCfg<C1, BK1>;
  using S2 = RsCfg<C2, BK2>;
  static constexpr int W1 = 256 / C1 - (SPARE ? 2 : 0);  // TMEM ring blocks of the stage-1 window (columns [0, 256), overflow blocks included)
  static constexpr int W2 = HEAD ? 6 : 256 / C2;       // ... of the stage-2 window (from column 256); HEAD: 6 blocks for 3 groups
  static constexpr int G1 = C1 == 64 ? 2 : 1;          // stage-1 epilogue groups (rows alternate)
  static constexpr int G2 = HEAD ? 3 : 2;              // stage-2 epilogue groups (HEAD: rows alternate; pool: row pairs alternate)
  static_assert(W1 % G1 == 0 && W1 <= kRs2MaxWin && W2 <= kRs2MaxWin, "a block is always drained by the same group");
  static_assert(HEAD ? (W2 % G2 == 0) : ((W2 / 2) % G2 == 0), "a block is always drained by the same group");
  static_assert((W1 + (SPARE ? 2 : 0)) * C1 == 256 && 256 + (W2 + (SPARE ? 2 : 0)) * C2 <= 512, "TMEM columns");
  static_assert(!SPARE || (W1 >= 4 && W2 >= 4), "three open rows and one being drained");
  static_assert(!SPARE || HEAD, "the pooling epilogue does not read overflow blocks (C = 64 leaves no TMEM columns for them)");
  static constexpr int THREADS = 128 + 128 * (G1 + G2);   // 4 role warps + the epilogue warps
  static constexpr int A1_STAGE = S1::A_STAGE_BYTES;
  static constexpr int A2_STAGE = S2::A_STAGE_BYTES;
  static constexpr int B1_BYTES = 3 * S1::B_STAGE_BYTES;
  static constexpr int B2_BYTES = 3 * S2::B_STAGE_BYTES;
  static constexpr int TAIL_BAR = 0;
  static constexpr int N_BARS = 2 * kRs2MaxSA1 + 2 * kRs2MaxSA2 + 4 * kRs2MaxWin + 1;
  static constexpr int TAIL_CM = TAIL_BAR + N_BARS * 8;
  static constexpr int TAIL_DICE = TAIL_CM + 4 * kMaxClasses * kMaxClasses * 8;
  static constexpr int TAIL_TMEM = TAIL_DICE + 4 * kMaxClasses * 2 * 8;
  static constexpr int TAIL_BYTES = ((TAIL_TMEM + 16 + 1023) / 1024) * 1024;
  static constexpr int OPERAND_BUDGET = kSmemLimit - 1024 - TAIL_BYTES;
};

template <int C1_, int C2_, int BK1_, bool HEAD_, bool SPARE_ = false>
__global__ void __launch_bounds__((Rs2Cfg<C1_, C2_, BK1_, HEAD_, SPARE_>::THREADS), 1)
conv_rs2_kernel(const __grid_constant__ ConvParams p) {
  using Cfg = Rs2Cfg<C1_, C2_, BK1_, HEAD_, SPARE_>;
  constexpr int C1 = Cfg::C1, C2 = Cfg::C2, W1 = Cfg::W1, W2 = Cfg::W2, G1 = Cfg::G1, G2 = Cfg::G2;
  constexpr bool HEAD = Cfg::HEAD, SPARE = Cfg::SPARE;
  constexpr int kThreads = Cfg::THREADS;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::TAIL_BAR);
  uint64_t* a1_full = bars;
  uint64_t* a1_empty = a1_full + kRs2MaxSA1;
  uint64_t* a2_full = a1_empty + kRs2MaxSA1;    // S1 epilogue -> MMA: the stage-1 row is in shared memory
  uint64_t* a2_empty = a2_full + kRs2MaxSA2;    // MMA -> S1 epilogue (tcgen05.commit)
  uint64_t* s1_full = a2_empty + kRs2MaxSA2;    // MMA -> S1 epilogue: stage-1 row complete in TMEM
  uint64_t* s1_empty = s1_full + kRs2MaxWin;
  uint64_t* s2_full = s1_empty + kRs2MaxWin;
  uint64_t* s2_empty = s2_full + kRs2MaxWin;
  uint64_t* w_full = s2_empty + kRs2MaxWin;
  unsigned long long* s_cm_all = reinterpret_cast<unsigned long long*>(smem + Cfg::TAIL_CM);
  double* s_dice_all = reinterpret_cast<double*>(smem + Cfg::TAIL_DICE);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + Cfg::TAIL_TMEM);
  uint8_t* a1_ring = smem + Cfg::TAIL_BYTES;
  uint8_t* a2_ring = a1_ring + p.sa * Cfg::A1_STAGE;
  uint8_t* b1_res = a2_ring + p.sb * Cfg::A2_STAGE;    // (p.sb = A2 stages here)
  uint8_t* b2_res = b1_res + Cfg::B1_BYTES;

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  // warps [0, 4*G2): S2 epilogue groups, then the S1 epilogue groups, then: producer (+ weights), stage-1 MMA, TMEM allocator +
  // stage-2 MMA, producer
  constexpr int kS2Warps = 4 * G2, kEpiWarps = 4 * (G1 + G2);
  const int role = warp - kEpiWarps;
  const int hq = p.hq;
  const int tpt = hq + 4;                     // steps per task
  const int bands = (p.W + kRs2Out - 1) / kRs2Out;
  const int n_tasks = p.B * bands;
  const int my_tasks = (n_tasks - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int my_tiles = my_tasks * tpt;        // stage-1 tiles tau = 0 .. my_tiles-1; stage-1 rows sigma1 = 0 .. my_tiles-2 (= stage-2
                                              // tiles); stage-2 rows sigma2 = 0 .. my_tiles-3

  if (role == 0 && lane == 0) { tma_prefetch_desc(&p.tmA); tma_prefetch_desc(&p.tmB); }
  if (threadIdx.x < 4 * kMaxClasses * kMaxClasses) s_cm_all[threadIdx.x] = 0ull;
  if (threadIdx.x < 4 * kMaxClasses * 2) s_dice_all[threadIdx.x] = 0.0;
  if (role == 1 && lane == 0) {
    for (int i = 0; i < kRs2MaxSA1; ++i) { mbar_init(&a1_full[i], 64); mbar_init(&a1_empty[i], 1); }
    for (int i = 0; i < kRs2MaxSA2; ++i) { mbar_init(&a2_full[i], 128); mbar_init(&a2_empty[i], 1); }
    for (int i = 0; i < kRs2MaxWin; ++i) {
      mbar_init(&s1_full[i], 1); mbar_init(&s1_empty[i], 128);
      mbar_init(&s2_full[i], 1); mbar_init(&s2_empty[i], 128);
    }
    mbar_init(w_full, 1);
    fence_barrier_init();
  }
  if (role == 2) { tmem_alloc(s_tmem, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  if (!(p.dbg & 16)) pdl_launch_dependents();

  if (role == 0 || role == 3) {
    // ================= producers: the packed image rows (two planes of 8 channels) =================
    if (role == 0) {
      if (elect_one()) {  // both layers' weights: constants, loaded before the wait on the previous kernel
        mbar_arrive_expect_tx(w_full, static_cast<uint32_t>(Cfg::B1_BYTES + Cfg::B2_BYTES));
#pragma unroll 1
        for (int kx = 0; kx < 3; ++kx) {
          tma_load_2d(b1_res + kx * Cfg::S1::B_STAGE_BYTES, &p.tmA, w_full, 0, kx * Cfg::S1::N3);   // tmA: stage-1 weights
          tma_load_2d(b2_res + kx * Cfg::S2::B_STAGE_BYTES, &p.tmB, w_full, 0, kx * Cfg::S2::N3);
        }
      }
      __syncwarp();
    }
    pdl_wait_prior_grid();
    const int half = role == 0 ? 0 : 1;
    constexpr int kRows = 4 * (Cfg::BK1 / 8);   // (plane, stream) rows of 512 bytes per stage
    const size_t row_stride = static_cast<size_t>(p.W) * 8;
    const size_t plane_stride = row_stride * p.H;
    const size_t image_stride = plane_stride * p.a_cgs;
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
      const int x = j * kRs2Out - 2 + lane;
      const bool x_ok = x >= 0 && x < p.W;
      // rows -2, -1 (stream 0, t = 0, 1) and H, H+1 (stream 3, the last two steps) are never dereferenced (src-size 0);
      // channel planes the view does not have (dec4.conv.0: 6 of 8) are zero-filled (their MMAs are not issued)
      const __nv_bfloat16* img = p.a_src + image_stride * b + static_cast<size_t>(x_ok ? x : 0) * 8;
#pragma unroll 1
      for (int t = 0; t < tpt; ++t) {
        mbar_wait(&a1_empty[sa], pa ^ 1u);
        const uint32_t dst0 = smem_u32(a1_ring + sa * Cfg::A1_STAGE) + static_cast<uint32_t>(lane) * 16u + static_cast<uint32_t>(half) * 512u;
#pragma unroll
        for (int r = 0; r < kRows / 2; ++r) {
          const int row = 2 * r + half, pl = row >> 2, q = row & 3;
          const int y = q * hq - 2 + t;
          const bool ok = x_ok && pl < p.a_planes && y >= 0 && y < p.H;
          const __nv_bfloat16* src = img + off[r] + (ok ? static_cast<ptrdiff_t>(t - 2) * static_cast<ptrdiff_t>(row_stride) : 0);
          cp_async_16_zfill(dst0 + static_cast<uint32_t>(r) * 1024u, ok ? src : img, ok);
        }
        cp_async_mbar_arrive_noinc(&a1_full[sa]);
        if (++sa == p.sa) { sa = 0; pa ^= 1u; }
      }
    }
  } else if (role == 1 || role == 2) {
    // ================= MMA issuers: warp 13 issues stage 1, warp 14 (the TMEM allocator) stage 2 =================
    // One thread issuing both stages was the limit of the first version: ~200 instructions per step at ~9 cycles each in
    // one warp = the 2,000-cycle step it measured, against 1,440 cycles of MMA work (15 MMAs, N = 192).  The stages share
    // nothing but mbarriers (tcgen05.commit tracks the issuing thread's own MMAs), so they get a thread each.
    const bool leader = elect_one();
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    mbar_wait(w_full, 0);
    tc_fence_after();
    if (role == 1) {
      const uint64_t d_a1 = umma_smem_desc_interleaved(smem_u32(a1_ring), Cfg::S1::A_PLANE_BYTES, 128);
      const uint64_t d_b1 = umma_smem_desc<Cfg::S1::SWZ_B>(smem_u32(b1_res));
      const uint32_t a1_hi = static_cast<uint32_t>(d_a1 >> 32), b1_hi = static_cast<uint32_t>(d_b1 >> 32);
      const uint32_t a1_lo0 = static_cast<uint32_t>(d_a1), b1_lo = static_cast<uint32_t>(d_b1);
      const uint32_t win1 = tmem_u;
      const bool skip1 = p.k_steps > 0 && p.k_steps < Cfg::BK1 / 16;   // the tail of the chunk is zero padding (dec4.conv.0: 48 of 64)
      int sa1 = 0, blk1 = W1 - 1, fresh1 = 1;   // ring position; block of row tau-1; fresh block = row tau+1 ...
      uint32_t pa1 = 0, fpar1 = 1;              // ... and the parity of its drain barrier
#pragma unroll 1
      for (int tau = 0; tau < my_tiles; ++tau) {
        mbar_wait(&s1_empty[fresh1], fpar1);
        mbar_wait(&a1_full[sa1], pa1);
        fence_proxy_async();
        tc_fence_after();
        if (leader) {
          const uint32_t d0 = win1 + static_cast<uint32_t>(blk1 * C1);
          const uint32_t a_st = a1_lo0 + static_cast<uint32_t>(sa1) * (Cfg::A1_STAGE >> 4);
          if constexpr (SPARE) {   // blocks blk1 .. blk1+2 are physically contiguous (W1, W1+1: overflow); at blk1 = 0 all are first touches
            if (blk1 == 0) rs_issue_chunk<C1, Cfg::BK1, 3, true>(d0, win1, a_st, a1_hi, b1_lo, b1_hi, skip1);
            else rs_issue_chunk<C1, Cfg::BK1, 0, true>(d0, win1, a_st, a1_hi, b1_lo, b1_hi, skip1);
          } else if (blk1 <= W1 - 3) rs_issue_chunk<C1, Cfg::BK1, 0, true>(d0, win1, a_st, a1_hi, b1_lo, b1_hi, skip1);
          else if (blk1 == W1 - 2) rs_issue_chunk<C1, Cfg::BK1, 1, true>(d0, win1, a_st, a1_hi, b1_lo, b1_hi, skip1);
          else rs_issue_chunk<C1, Cfg::BK1, 2, true>(d0, win1, a_st, a1_hi, b1_lo, b1_hi, skip1);
          umma_commit(&a1_empty[sa1]);
          if (tau >= 1) umma_commit(&s1_full[blk1]);   // stage-1 row tau-1 is complete
        }
        __syncwarp();
        if (++sa1 == p.sa) { sa1 = 0; pa1 ^= 1u; }
        blk1 = blk1 + 1 == W1 ? 0 : blk1 + 1;
        if (++fresh1 == W1) { fresh1 = 0; fpar1 ^= 1u; }
      }
    } else {
      const uint64_t d_a2 = umma_smem_desc_interleaved(smem_u32(a2_ring), Cfg::S2::A_PLANE_BYTES, 128);
      const uint64_t d_b2 = umma_smem_desc<Cfg::S2::SWZ_B>(smem_u32(b2_res));
      const uint32_t a2_hi = static_cast<uint32_t>(d_a2 >> 32), b2_hi = static_cast<uint32_t>(d_b2 >> 32);
      const uint32_t a2_lo0 = static_cast<uint32_t>(d_a2), b2_lo = static_cast<uint32_t>(d_b2);
      const uint32_t win2 = tmem_u + 256u;
      int sa2 = 0, blk2 = W2 - 1, fresh2 = 1;
      uint32_t pa2 = 0, fpar2 = 1;
#pragma unroll 1
#ifdef UB_RS2_PROFILE   // (compile with -DUB_RS2_PROFILE and run with UB_DBG=512: where this thread's cycles go, printed by CTA 0;
                        // compiled out by default — the extra instructions sit on the issuing thread's critical path)
      long long t_w1 = 0, t_w2 = 0, t_f = 0, t_i = 0;
      const bool prof = (p.dbg & 512) != 0;
#else
      constexpr bool prof = false;
      long long t_w1 = 0, t_w2 = 0, t_f = 0, t_i = 0;
#endif
      for (int tau2 = 0; tau2 < my_tiles - 1; ++tau2) {   // stage-2 tile = stage-1 row
        const long long c0 = prof ? clock64() : 0;
        mbar_wait(&s2_empty[fresh2], fpar2);
        const long long c1 = prof ? clock64() : 0;
        mbar_wait(&a2_full[sa2], pa2);
        const long long c2 = prof ? clock64() : 0;
        // (no proxy fence here: the row's writers execute fence.proxy.async before they arrive on a2_full; with one the step
        // measured 0.6 % slower — every instruction of this thread sits between two MMA bursts)
        tc_fence_after();
        const long long c3 = prof ? clock64() : 0;
        if (leader) {
          const uint32_t d0 = win2 + static_cast<uint32_t>(blk2 * C2);
          const uint32_t a_st = a2_lo0 + static_cast<uint32_t>(sa2) * (Cfg::A2_STAGE >> 4);
          if (p.dbg & 64) {}  // (experiment: stage 2 issues no MMAs — timing only)
          else if constexpr (SPARE) {
            if (blk2 == 0) rs_issue_chunk<C2, Cfg::BK2, 3, true>(d0, win2, a_st, a2_hi, b2_lo, b2_hi, false);
            else rs_issue_chunk<C2, Cfg::BK2, 0, true>(d0, win2, a_st, a2_hi, b2_lo, b2_hi, false);
          } else if (blk2 <= W2 - 3) rs_issue_chunk<C2, Cfg::BK2, 0, true>(d0, win2, a_st, a2_hi, b2_lo, b2_hi, false);
          else if (blk2 == W2 - 2) rs_issue_chunk<C2, Cfg::BK2, 1, true>(d0, win2, a_st, a2_hi, b2_lo, b2_hi, false);
          else rs_issue_chunk<C2, Cfg::BK2, 2, true>(d0, win2, a_st, a2_hi, b2_lo, b2_hi, false);
          umma_commit(&a2_empty[sa2]);
          if (tau2 >= 1) umma_commit(&s2_full[blk2]);   // stage-2 row tau2-1 is complete
        }
        __syncwarp();
        if (prof) { const long long c4 = clock64(); t_w1 += c1 - c0; t_w2 += c2 - c1; t_f += c3 - c2; t_i += c4 - c3; }
        if (++sa2 == p.sb) { sa2 = 0; pa2 ^= 1u; }
        blk2 = blk2 + 1 == W2 ? 0 : blk2 + 1;
        if (++fresh2 == W2) { fresh2 = 0; fpar2 ^= 1u; }
      }
      if (prof && blockIdx.x == 0 && lane == 0)
        printf("[rs2 C1=%d stage-2 issuer] steps %d: wait window %lld, wait row %lld, fences %lld, issue+commit %lld cycles per step\n", C1,
               my_tiles - 1, t_w1 / (my_tiles - 1), t_w2 / (my_tiles - 1), t_f / (my_tiles - 1), t_i / (my_tiles - 1));
    }
  } else if (role < 0 && warp >= kS2Warps) {
    // ================= S1 epilogue (one group): stage-1 row sigma1 -> the A2 ring =================
    const int wq = warp & 3;
    const int g1 = (warp - kS2Warps) >> 2;   // this group takes the rows sigma % G1 == g1 (W1 is a multiple of G1)
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(wq * 32) << 16);
    const int n_rows = my_tiles - 1;
#pragma unroll 1
    for (int sigma = g1; sigma < n_rows; sigma += G1) {
      const int sa2 = sigma % p.sb;                                   // ring slot and how often it has been used before
      const uint32_t pa2 = static_cast<uint32_t>(sigma / p.sb) & 1u;
      const int ti = sigma / tpt, rho = sigma - ti * tpt;
      const int task = static_cast<int>(blockIdx.x) + ti * static_cast<int>(gridDim.x);
      const int j = task % bands;
      const int y = wq * hq - 2 + rho, x = j * kRs2Out - 2 + lane;
      // exact only on lanes 1..30; outside the image the NEXT conv sees its zero padding
      const bool keep = lane >= 1 && lane <= 30 && x >= 0 && x < p.W && y >= 0 && y < p.H;
      const int blk = sigma % W1;
      const uint32_t use = static_cast<uint32_t>(sigma / W1);
      mbar_wait(&s1_full[blk], use & 1u);
      tc_fence_after();
      const uint32_t t_addr = lane_addr + static_cast<uint32_t>(blk * C1);
      uint32_t pk[C1 / 2];
      if (p.dbg & 32) {  // (experiment: no stage-1 epilogue work — timing only, results are wrong)
        tc_fence_before(); mbar_arrive(&s1_empty[blk]);
#pragma unroll
        for (int i = 0; i < C1 / 2; ++i) pk[i] = 0u;
      } else
#pragma unroll
      for (int ch = 0; ch < C1; ch += 32) {
        uint32_t r[32];
        tmem_ld_32x32(t_addr + static_cast<uint32_t>(ch), r);
        if constexpr (SPARE) {
          if (blk < 2) {   // the row's first contributions sit in its overflow block
            uint32_t o[32];
            tmem_ld_32x32(lane_addr + static_cast<uint32_t>((W1 + blk) * C1 + ch), o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const float2 a = __fadd2_rn(make_float2(__uint_as_float(r[i]), __uint_as_float(r[i + 1])),
                                          make_float2(__uint_as_float(o[i]), __uint_as_float(o[i + 1])));
              r[i] = __float_as_uint(a.x); r[i + 1] = __float_as_uint(a.y);
            }
          }
        }
        tmem_ld_wait();
        if (ch + 32 >= C1) { tc_fence_before(); mbar_arrive(&s1_empty[blk]); }
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 bz = make_float4(p.bias64[0][ch + 4 * j4], p.bias64[0][ch + 4 * j4 + 1], p.bias64[0][ch + 4 * j4 + 2], p.bias64[0][ch + 4 * j4 + 3]);
          const float2 a = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 0]), __uint_as_float(r[4 * j4 + 1])), make_float2(bz.x, bz.y));
          const float2 c2 = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 2]), __uint_as_float(r[4 * j4 + 3])), make_float2(bz.z, bz.w));
          pk[ch / 2 + 2 * j4 + 0] = keep ? bf16x2_max(pack_bf16(a.x, a.y), 0u) : 0u;    // ReLU on the packed pairs (monotone rounding)
          pk[ch / 2 + 2 * j4 + 1] = keep ? bf16x2_max(pack_bf16(c2.x, c2.y), 0u) : 0u;
        }
      }
      mbar_wait(&a2_empty[sa2], pa2 ^ 1u);
      uint4* dst = reinterpret_cast<uint4*>(a2_ring + sa2 * Cfg::A2_STAGE + wq * 512 + lane * 16);
#pragma unroll
      for (int g = 0; g < C1 / 8; ++g) dst[g * (Cfg::S2::A_PLANE_BYTES / 16)] = make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
      fence_proxy_async();          // generic-proxy stores -> visible to the tensor core's async-proxy reads
      mbar_arrive(&a2_full[sa2]);
    }
  } else if (role < 0) {
    const int wq = warp & 3;
    const int grp = warp >> 2;
    const uint32_t lane_addr2 = tmem_base + (static_cast<uint32_t>(wq * 32) << 16) + 256u;
    const int n_rows = my_tiles - 2;
    if constexpr (!HEAD) {
    // ================= S2 epilogue: bias, 2x2 max-pool, ReLU, store (row pairs alternate between the groups) =================
    uint32_t held[C2 / 2];
#pragma unroll
    for (int i = 0; i < C2 / 2; ++i) held[i] = 0u;
    const int oh = p.H >> 1, ow = p.W >> 1;
    const size_t plane = static_cast<size_t>(oh) * ow;
#pragma unroll 1
    for (int pr = grp; 2 * pr < n_rows; pr += G2) {
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        const int sigma = 2 * pr + half;
        if (sigma >= n_rows) continue;
        const int ti = sigma / tpt, rho = sigma - ti * tpt;
        const int task = static_cast<int>(blockIdx.x) + ti * static_cast<int>(gridDim.x);
        const int b = task / bands, j = task - b * bands;
        const int y = wq * hq - 2 + rho, x = j * kRs2Out - 2 + lane;
        const bool valid_row = rho >= 2 && rho <= hq + 1;
        const bool inb = valid_row && lane >= 2 && lane <= 29 && x < p.W;
        const int blk = sigma % W2;
        const uint32_t use = static_cast<uint32_t>(sigma / W2);
        mbar_wait(&s2_full[blk], use & 1u);
        tc_fence_after();
        const uint32_t t_addr = lane_addr2 + static_cast<uint32_t>(blk * C2);
        if (!valid_row) {
          tc_fence_before();
          mbar_arrive(&s2_empty[blk]);
          continue;
        }
#pragma unroll
        for (int ch = 0; ch < C2; ch += 32) {
          uint32_t r[32];
          tmem_ld_32x32(t_addr + static_cast<uint32_t>(ch), r);
          tmem_ld_wait();
          if (ch + 32 >= C2) { tc_fence_before(); mbar_arrive(&s2_empty[blk]); }
          uint32_t pk[16];
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 bz = make_float4(p.bias64[1][ch + 4 * j4], p.bias64[1][ch + 4 * j4 + 1], p.bias64[1][ch + 4 * j4 + 2], p.bias64[1][ch + 4 * j4 + 3]);
            const float2 a = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 0]), __uint_as_float(r[4 * j4 + 1])), make_float2(bz.x, bz.y));
            const float2 c2 = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 2]), __uint_as_float(r[4 * j4 + 3])), make_float2(bz.z, bz.w));
            pk[2 * j4 + 0] = pack_bf16(a.x, a.y);
            pk[2 * j4 + 1] = pack_bf16(c2.x, c2.y);
          }
          if (half == 0) {   // y even (rho even, quarter height even): hold the row for its partner
#pragma unroll
            for (int i = 0; i < 16; ++i) held[ch / 2 + i] = pk[i];
          } else {
            // x even = even lane here (x = 28j - 2 + lane): the pair is (even lane, lane + 1); reduce-scatter as in conv_rs_kernel
            const bool first = !(lane & 1);
            const int partner = first ? ((lane + 1) & 31) : ((lane + 31) & 31);
            uint32_t m[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const uint32_t v_keep = bf16x2_max(first ? pk[i] : pk[8 + i], first ? held[ch / 2 + i] : held[ch / 2 + 8 + i]);
              const uint32_t v_send = bf16x2_max(first ? pk[8 + i] : pk[i], first ? held[ch / 2 + 8 + i] : held[ch / 2 + i]);
              m[i] = bf16x2_max(v_keep, __shfl_sync(0xffffffffu, v_send, partner));
              if (p.relu) m[i] = bf16x2_max(m[i], 0u);
            }
            if (inb) {
              uint4* dst = reinterpret_cast<uint4*>(p.out) +
                           ((static_cast<size_t>(b) * p.out_cgs + p.out_cg0 + (ch >> 3) + (first ? 0 : 2)) * plane +
                            static_cast<size_t>(y >> 1) * ow + (x >> 1));
              dst[0] = make_uint4(m[0], m[1], m[2], m[3]);
              dst[plane] = make_uint4(m[4], m[5], m[6], m[7]);
            }
          }
        }
      }
    }
    } else {
      // ================= S2 epilogue, HEAD: bias, ReLU, 1x1 head, argmax, confusion counts (rows alternate between the groups) ====
      const int et = threadIdx.x & 127;
      unsigned long long* s_cm = s_cm_all + grp * (kMaxClasses * kMaxClasses);
      double* s_dice = s_dice_all + grp * (2 * kMaxClasses);
      HeadState hs;
      hs.init();
#pragma unroll 1
      for (int sigma = grp; sigma < n_rows; sigma += G2) {
        const int ti = sigma / tpt, rho = sigma - ti * tpt;
        const int task = static_cast<int>(blockIdx.x) + ti * static_cast<int>(gridDim.x);
        const int b = task / bands, j = task - b * bands;
        const int y = wq * hq - 2 + rho, x = j * kRs2Out - 2 + lane;
        const bool valid_row = rho >= 2 && rho <= hq + 1;
        const bool inb = valid_row && lane >= 2 && lane <= 29 && x < p.W;
        const int blk = sigma % W2;
        const uint32_t use = static_cast<uint32_t>(sigma / W2);
        int tgt_pre = -1;
        if (p.targets && valid_row) {
          head_begin_tile(p, hs, (p.img_base + b) / p.ref_batch, grp, et, lane, s_cm, s_dice);
          if (inb) tgt_pre = head_load_target(p, b, y, x);
        }
        mbar_wait(&s2_full[blk], use & 1u);
        tc_fence_after();
        if (!valid_row) {
          tc_fence_before();
          mbar_arrive(&s2_empty[blk]);
          continue;
        }
        if (p.dbg & 128) {  // (experiment: no head arithmetic — timing only)
          tc_fence_before();
          mbar_arrive(&s2_empty[blk]);
          continue;
        }
        if constexpr (C2 == 32) {
          uint32_t r[32];
          tmem_ld_32x32(lane_addr2 + static_cast<uint32_t>(blk * C2), r);
          if constexpr (SPARE) {
            if (blk < 2) {   // the row's first contributions sit in its overflow block
              uint32_t o[32];
              tmem_ld_32x32(lane_addr2 + static_cast<uint32_t>((W2 + blk) * C2), o);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                const float2 a = __fadd2_rn(make_float2(__uint_as_float(r[i]), __uint_as_float(r[i + 1])),
                                            make_float2(__uint_as_float(o[i]), __uint_as_float(o[i + 1])));
                r[i] = __float_as_uint(a.x); r[i + 1] = __float_as_uint(a.y);
              }
            }
          }
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(&s2_empty[blk]);
          if (p.n_classes == 3) head_pixel<3>(p, hs, r, inb, b, y, x, tgt_pre);
          else head_pixel<0>(p, hs, r, inb, b, y, x, tgt_pre);
        }
      }
      if (p.targets && hs.cur_batch >= 0) head_flush(p, hs, hs.cur_batch, grp, et, lane, s_cm, s_dice);
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (role == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace ub
