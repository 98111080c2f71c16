// Epilogue body of conv_tc_kernel (included inside the `role < 0` branch: warps 0 .. 4*kEpiGroups-1).
// Thread (wq, lane) owns accumulator row m = 32*wq + lane = pixel (ty = m / TW, tx = m % TW) of the
// (128/TW) x TW tile and reads its fp32 columns from TMEM 32 at a time (double-buffered: the next
// tcgen05.ld is in flight while the current 32 channels are converted and stored).  In the kx-stacked
// variants tx is an INPUT column: columns 0 and TW-1 are halo, and output channel co of column tx is
// D[tx-1][co] + D[tx][Cout+co] + D[tx+1][2*Cout+co].
{
  constexpr bool kHead = (Cfg::COUT_T == 32);  // the fused head exists only in the 32-channel instantiations
  const int wq = warp & 3;
  const int grp = warp >> 2;                   // tiles it with it % kEpiGroups == grp
  const int ty = (32 * wq + lane) / TW;
  const int tx = (32 * wq + lane) % TW;
  const int et = threadIdx.x & 127;            // 0..127 within the group
  unsigned long long* s_cm = s_cm_all + grp * (kMaxClasses * kMaxClasses);
  double* s_dice = s_dice_all + grp * (2 * kMaxClasses);
  // HEAD: every thread counts its own pixels in registers (ub_head.cuh); they are reduced (warp shuffle -> shared
  // memory -> global atomics) only when the reference batch changes, every kHeadFlushTiles tiles, and once at the end.
  HeadState hs;
  hs.init();

  int it = grp;
  TileIter ti = tile_iter(t_first + grp * t_step, kEpiGroups * t_step);
  // first image of the NEXT reference batch, relative to this micro-batch (HEAD): b only grows
  int batch_now = (p.img_base + tile_b(ti)) / p.ref_batch;
  int batch_end = (batch_now + 1) * p.ref_batch - p.img_base;
#pragma unroll 1
  for (int t = t_first + grp * t_step; t < total_tiles; t += kEpiGroups * t_step, it += kEpiGroups, ti.next()) {
    const int acc = it % ACC;
    const uint32_t use = static_cast<uint32_t>(it / ACC);
    const int nt = tile_nt(ti);
    const int x = tile_tx(ti) * kXStep - (KXS ? 1 : 0) + tx;
    const int y = tile_ty(ti) * kTileRows + ty;
    const int b = tile_b(ti);
    const bool inb = (y < p.H) && (x < p.W) && (!KXS || (tx >= 1 && tx <= TW - 2)) && (!CG2 || b < p.B);
    const int n0 = nt * BLOCK_N;
    int tgt_pre = -1;

    if (kHead && p.mode == EPI_HEAD && p.targets) {
      // counts accumulate in registers across tiles and go to HBM only when the reference batch
      // changes (and once at the end): <= C*C global atomics per batch per CTA instead of per tile
      while (b >= batch_end) { ++batch_now; batch_end += p.ref_batch; }
      head_begin_tile(p, hs, batch_now, grp, et, lane, s_cm, s_dice);
      // the target is fetched before waiting for the accumulator so its latency hides behind the MMAs
      if (inb) tgt_pre = head_load_target(p, b, y, x);
    }

    mbar_wait(&t_full[acc], use & 1u);
    tc_fence_after();
    const uint32_t t_addr =
        tmem_base + (static_cast<uint32_t>(wq * 32) << 16) + static_cast<uint32_t>(acc * Cfg::ACC_STRIDE);
    auto release_acc = [&]() {  // every column of this accumulator is in registers: hand it back to the MMA warp
      tc_fence_before();
      if constexpr (CG2) mbar_arrive_leader(&t_empty[acc]);  // the pair's MMA warp lives in the even CTA
      else mbar_arrive(&t_empty[acc]);
    };

    // kx-stacked: 32 channels [ch, ch+32) of this thread's output column = D1[tx] + D0[tx-1] + D2[tx+1]; two loads
    // in flight, the third reuses the first one's registers (the 4-group variants have 96 registers per thread)
    auto kxs_sum = [&](uint32_t (&sum)[32], int ch, bool last) {
      constexpr int CO = Cfg::COUT_T;
      uint32_t d[32];
      tmem_ld_32x32(t_addr + static_cast<uint32_t>(CO + ch), sum);
      tmem_ld_32x32(t_addr + static_cast<uint32_t>(ch), d);
      tmem_ld_wait();
      // the adds are packed fp32x2 (add.rn.f32x2: two IEEE adds per issue slot — these epilogues are issue-bound)
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        const float2 nb = make_float2(__shfl_up_sync(0xffffffffu, __uint_as_float(d[j]), 1, TW),
                                      __shfl_up_sync(0xffffffffu, __uint_as_float(d[j + 1]), 1, TW));
        const float2 a = __fadd2_rn(make_float2(__uint_as_float(sum[j]), __uint_as_float(sum[j + 1])), nb);
        sum[j] = __float_as_uint(a.x); sum[j + 1] = __float_as_uint(a.y);
      }
      tmem_ld_32x32(t_addr + static_cast<uint32_t>(2 * CO + ch), d);
      tmem_ld_wait();
      if (last) release_acc();
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        const float2 nb = make_float2(__shfl_down_sync(0xffffffffu, __uint_as_float(d[j]), 1, TW),
                                      __shfl_down_sync(0xffffffffu, __uint_as_float(d[j + 1]), 1, TW));
        const float2 a = __fadd2_rn(make_float2(__uint_as_float(sum[j]), __uint_as_float(sum[j + 1])), nb);
        sum[j] = __float_as_uint(a.x); sum[j + 1] = __float_as_uint(a.y);
      }
    };

    if (p.dbg & 1) {
      uint32_t r[32];
      tmem_ld_32x32(t_addr, r);
      tmem_ld_wait();
      release_acc();
      if (r[0] == 0x12345678u && p.out) p.out[0] = __float2bfloat16(1.0f);
    } else if (kHead && p.mode == EPI_HEAD) {
      // ---- dec4.conv.3 epilogue: bias+ReLU (fp32) -> 1x1 head -> softmax -> argmax -> counts
      uint32_t r[32];
      if constexpr (KXS) {
        kxs_sum(r, 0, true);
      } else {
        tmem_ld_32x32(t_addr, r);
        tmem_ld_wait();
        release_acc();
      }
      // the per-pixel head, specialised for the usual 3 classes (NC = 0: class count known at run time only)
      if (p.n_classes == 3) head_pixel<3>(p, hs, r, inb, b, y, x, tgt_pre);
      else head_pixel<0>(p, hs, r, inb, b, y, x, tgt_pre);
    } else {
      // ---- STORE / POOL / CONVT: bias (+ReLU) -> bf16 -> channel-blocked global store, 32 channels at a time
      // geometry of this thread's output pixel (fixed per tile)
      int oy = y, ox = x, oh = p.H, ow = p.W;
      bool writer = inb;
      if (p.mode == EPI_POOL) {
        // MaxPool2d(2,2), H and W even.  Standard tile: partners are lane^1 (x) and lane^8 (y).  Kx-stacked TW = 16
        // tile: outputs start at an even x in tile column 1, so pairs are (odd tx, tx+1) and rows are lane^16.
        // (every lane of a window writes its share of the pooled pixel, see store_chunk)
        oy = y >> 1; ox = x >> 1; oh = p.H >> 1; ow = p.W >> 1;
      } else if (p.mode == EPI_CONVT) {
        oh = 2 * p.H; ow = 2 * p.W;
      }
      const size_t plane = static_cast<size_t>(oh) * ow;  // pixels per channel-group plane
      auto store_chunk = [&](const uint32_t (&r)[32], int cb) {
        const int n = n0 + cb;
        int bias_off = n, cg = p.out_cg0 + (n >> 3), py = oy, px = ox;
        if (p.mode == EPI_CONVT) {
          const int q = n >> p.cout_shift;
          bias_off = n - (q << p.cout_shift);
          py = 2 * y + (q >> 1);
          px = 2 * x + (q & 1);
          cg = p.out_cg0 + (bias_off >> 3);
        }
        // bias: packed fp32x2 adds.  bf16 rounding is monotone and keeps the sign, so ReLU and max-pool commute with
        // it: pack first, then ReLU (max with +0) and pool on the packed pairs — half the instructions, same result.
        uint32_t pk[16];
        const float* bias_src = s_bias + bias_off;
        if constexpr (KSIZE == 1) {
          if (p.n_bias > kBiasStage) bias_src = p.bias + bias_off;  // wider than the staging area: straight from global (L1-resident)
        }
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 bz = *reinterpret_cast<const float4*>(bias_src + 4 * j4);
          float2 a = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 0]), __uint_as_float(r[4 * j4 + 1])),
                                make_float2(bz.x, bz.y));
          float2 c = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 2]), __uint_as_float(r[4 * j4 + 3])),
                                make_float2(bz.z, bz.w));
          if constexpr (KSIZE == 1) {
            // QuickGELU x * sigmoid(1.702 x) (CLIP's MLP activation; only the 1x1 instantiations carry the branch)
            if (p.act == ACT_QUICK_GELU) { a = quick_gelu2(a); c = quick_gelu2(c); }
          }
          pk[2 * j4 + 0] = pack_bf16(a.x, a.y);
          pk[2 * j4 + 1] = pack_bf16(c.x, c.y);
        }
        if (p.relu && p.mode != EPI_POOL) {  // (pooled layers: ReLU commutes with max, applied to the pooled quarter below)
#pragma unroll
          for (int j = 0; j < 16; ++j) pk[j] = bf16x2_max(pk[j], 0u);
        }
        if (p.mode == EPI_POOL) {
          // 2x2 max as a reduce-scatter over the four lanes of a pooling window: stage 1 trades HALF of the packed
          // registers with the x-partner, stage 2 a quarter with the y-partner; every lane ends with one channel
          // group (8 channels, 16 B) of the pooled pixel and stores it.  12 shuffles per 32 channels instead of 32.
          const bool sx = KXS ? !(tx & 1) : (lane & 1);    // keeps groups 2,3 in stage 1
          const bool sy = KXS ? (lane >= 16) : (lane & 8);  // keeps the odd group in stage 2
          const int partner = KXS ? ((tx & 1) ? lane + 1 : lane - 1) : (lane ^ 1);
          uint32_t m[8], q[4];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t send = sx ? pk[j] : pk[8 + j], keep = sx ? pk[8 + j] : pk[j];
            m[j] = bf16x2_max(keep, __shfl_sync(0xffffffffu, send, partner));
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t send = sy ? m[j] : m[4 + j], keep = sy ? m[4 + j] : m[j];
            q[j] = bf16x2_max(keep, __shfl_xor_sync(0xffffffffu, send, KXS ? 16 : 8));
            if (p.relu) q[j] = bf16x2_max(q[j], 0u);
          }
          if (writer) {
            const int g = (sx ? 2 : 0) + (sy ? 1 : 0);
            uint4* dst = reinterpret_cast<uint4*>(p.out) +
                         ((static_cast<size_t>(b) * p.out_cgs + cg + g) * plane + static_cast<size_t>(py) * ow + px);
            *dst = make_uint4(q[0], q[1], q[2], q[3]);
          }
          return;
        }
        if (writer) {
          uint4* dst = reinterpret_cast<uint4*>(p.out) +
                       ((static_cast<size_t>(b) * p.out_cgs + cg) * plane + static_cast<size_t>(py) * ow + px);
#pragma unroll
          for (int g = 0; g < 4; ++g)
            dst[static_cast<size_t>(g) * plane] = make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
        }
      };
      bool blended = false;
      if constexpr (KSIZE == 1 && !KXS) {
        if (p.mode == EPI_CONVT_BLEND) {
          // ---- ConvTranspose2d(k2,s2) + F.interpolate(align_corners) back to the input grid (dec1), as one epilogue: the
          // weights are packed with the four parities of a channel in adjacent columns (n = 4*co + 2a + b), so 32 columns
          // are 8 channels x 4 parity products of THIS pixel; their blend (convt_resample_weights) + bias is the resized
          // pixel, rounded to bf16 once.  One 16-byte store per chunk; the 2H x 2W tensor never exists.
          blended = true;
          float wy0, wy1, wx0, wx1;
          convt_resample_weights(y, p.H, &wy0, &wy1);
          convt_resample_weights(x, p.W, &wx0, &wx1);
          const float c00 = wy0 * wx0, c01 = wy0 * wx1, c10 = wy1 * wx0, c11 = wy1 * wx1;
          uint4* dst0 = reinterpret_cast<uint4*>(p.out) +
                        ((static_cast<size_t>(b) * p.out_cgs + p.out_cg0) * plane + static_cast<size_t>(y) * ow + x);
          auto blend_chunk = [&](const uint32_t (&r)[32], int cb) {
            const int c0 = (n0 + cb) >> 2;  // first of the chunk's 8 output channels
            const float* bias_src = (p.n_bias > kBiasStage ? p.bias : s_bias) + c0;
            const float4 b0 = *reinterpret_cast<const float4*>(bias_src), b1 = *reinterpret_cast<const float4*>(bias_src + 4);
            const float bz[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k)
              o[k] = fmaf(c00, __uint_as_float(r[4 * k + 0]),
                          fmaf(c01, __uint_as_float(r[4 * k + 1]),
                               fmaf(c10, __uint_as_float(r[4 * k + 2]), c11 * __uint_as_float(r[4 * k + 3])))) + bz[k];
            if (inb)
              dst0[static_cast<size_t>(c0 >> 3) * plane] =
                  make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
          };
          uint32_t ra[32], rb[32];
          tmem_ld_32x32(t_addr, ra);
#pragma unroll 1
          for (int cb = 0; cb < BLOCK_N; cb += 64) {
            tmem_ld_wait();
            if (cb + 32 < BLOCK_N) tmem_ld_32x32(t_addr + static_cast<uint32_t>(cb + 32), rb);
            else release_acc();
            blend_chunk(ra, cb);
            if (cb + 32 < BLOCK_N) {
              tmem_ld_wait();
              if (cb + 64 < BLOCK_N) tmem_ld_32x32(t_addr + static_cast<uint32_t>(cb + 64), ra);
              else release_acc();
              blend_chunk(rb, cb + 32);
            }
          }
        }
      }
      if (blended) {
      } else if constexpr (KXS) {
        // phase 1 (needs TMEM): read the three kx blocks of every 32-channel chunk and add the neighbours'
        // shares; the accumulator is released as soon as the last block is in registers.  phase 2: bias,
        // ReLU, pool, store from registers while the MMA warp already refills the accumulator.
        constexpr int NCH = Cfg::COUT_T / 32;
        uint32_t sum[NCH][32];
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) kxs_sum(sum[ch], 32 * ch, ch == NCH - 1);
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) store_chunk(sum[ch], 32 * ch);
      } else if (p.mode == EPI_CONVT && BLOCK_N >= 2 * p.cout) {
        // ConvTranspose scatter with the two x-quadrants of a pixel paired: columns [q*Cout + cb, +32) (b = 0) and
        // [(q+1)*Cout + cb, +32) (b = 1) are the same channels of output pixels (2y+a, 2x) and (2y+a, 2x+1), which
        // are adjacent in the blocked layout -> one 32-byte store per channel group, full sectors.
        const int cout = p.cout;
        const int n_chunks = BLOCK_N / 64;  // pairs of 32-column chunks in this n-tile
        uint32_t ra[32], rb[32];
#pragma unroll 1
        for (int i = 0; i < n_chunks; ++i) {
          const int per_q = cout >> 5;                      // 32-channel chunks per quadrant
          const int off = (i / per_q) * 2 * cout, cb = (i % per_q) * 32;
          tmem_ld_32x32(t_addr + static_cast<uint32_t>(off + cb), ra);
          tmem_ld_32x32(t_addr + static_cast<uint32_t>(off + cout + cb), rb);
          tmem_ld_wait();
          if (i == n_chunks - 1) release_acc();
          const int q = (n0 + off) >> p.cout_shift;          // even quadrant: a = q >> 1, b = 0
          uint32_t pa[16], pb[16];
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 bz = *reinterpret_cast<const float4*>(s_bias + cb + 4 * j4);
            const float2 b01 = make_float2(bz.x, bz.y), b23 = make_float2(bz.z, bz.w);
            auto add_pack = [](uint32_t r0, uint32_t r1, float2 bb) {
              const float2 t = __fadd2_rn(make_float2(__uint_as_float(r0), __uint_as_float(r1)), bb);
              return pack_bf16(t.x, t.y);
            };
            pa[2 * j4 + 0] = add_pack(ra[4 * j4 + 0], ra[4 * j4 + 1], b01);
            pa[2 * j4 + 1] = add_pack(ra[4 * j4 + 2], ra[4 * j4 + 3], b23);
            pb[2 * j4 + 0] = add_pack(rb[4 * j4 + 0], rb[4 * j4 + 1], b01);
            pb[2 * j4 + 1] = add_pack(rb[4 * j4 + 2], rb[4 * j4 + 3], b23);
          }
          if (inb) {
            uint4* dst = reinterpret_cast<uint4*>(p.out) +
                         ((static_cast<size_t>(b) * p.out_cgs + p.out_cg0 + (cb >> 3)) * plane +
                          static_cast<size_t>(2 * y + (q >> 1)) * ow + 2 * x);
#pragma unroll
            for (int g = 0; g < 4; ++g)
              st_global_256(dst + static_cast<size_t>(g) * plane, make_uint4(pa[4 * g], pa[4 * g + 1], pa[4 * g + 2], pa[4 * g + 3]),
                            make_uint4(pb[4 * g], pb[4 * g + 1], pb[4 * g + 2], pb[4 * g + 3]));
          }
        }
      } else if constexpr (kEpiGroups == 4) {
        // four groups share the register file (96 registers per thread): one chunk in flight
        uint32_t ra[32];
#pragma unroll 1
        for (int cb = 0; cb < BLOCK_N; cb += 32) {
          tmem_ld_32x32(t_addr + static_cast<uint32_t>(cb), ra);
          tmem_ld_wait();
          if (cb + 32 >= BLOCK_N) release_acc();
          store_chunk(ra, cb);
        }
      } else {
        uint32_t ra[32], rb[32];
        tmem_ld_32x32(t_addr, ra);
#pragma unroll 1
        for (int cb = 0; cb < BLOCK_N; cb += 64) {
          tmem_ld_wait();
          if (cb + 32 < BLOCK_N) tmem_ld_32x32(t_addr + static_cast<uint32_t>(cb + 32), rb);
          else release_acc();
          store_chunk(ra, cb);
          if (cb + 32 < BLOCK_N) {
            tmem_ld_wait();
            if (cb + 64 < BLOCK_N) tmem_ld_32x32(t_addr + static_cast<uint32_t>(cb + 64), ra);
            else release_acc();
            store_chunk(rb, cb + 32);
          }
        }
      }
    }
  }
  if (kHead && p.mode == EPI_HEAD && p.targets && hs.cur_batch >= 0) head_flush(p, hs, hs.cur_batch, grp, et, lane, s_cm, s_dice);
}
