// Epilogue body of conv_tc_kernel (included inside the `warp >= 4` branch).
// Thread (wq, lane) owns accumulator row m = 32*wq + lane = pixel (ty = m / TW, tx = m % TW) of the
// (128/TW) x TW tile and reads its fp32 columns from TMEM 32 at a time (double-buffered: the next
// tcgen05.ld is in flight while the current 32 channels are converted and stored).  In the kx-stacked
// variants tx is an INPUT column: columns 0 and TW-1 are halo, and output channel co of column tx is
// D[tx-1][co] + D[tx][Cout+co] + D[tx+1][2*Cout+co].
{
  constexpr bool kHead = (Cfg::COUT_T == 32);  // the fused head exists only in the 32-channel instantiations
  const int wq = warp & 3;
  const int grp = (warp - 4) >> 2;             // tiles it with it % kEpiGroups == grp
  const int ty = (32 * wq + lane) / TW;
  const int tx = (32 * wq + lane) % TW;
  const int et = (threadIdx.x - 128) & 127;    // 0..127 within the group
  unsigned long long* s_cm = s_cm_all + grp * (kMaxClasses * kMaxClasses);
  double* s_dice = s_dice_all + grp * (2 * kMaxClasses);
  int cur_batch = -1;
  // HEAD: every thread counts its own pixels in registers (cell = target*C + prediction, and the
  // soft-Dice partial sums); they are reduced (warp shuffle -> shared memory -> global atomics) only
  // when the reference batch changes, every kFlushTiles tiles, and once at the end.
  constexpr int kCells = kMaxClasses * kMaxClasses;
  constexpr int kFlushTiles = 250;  // < 255: the packed 8-bit count fields cannot overflow between flushes
  // counts of target class t live in pk[t]: kMaxClasses 8-bit fields indexed by the prediction (4 registers
  // instead of 16, one shift and kMaxClasses predicated adds per pixel)
  static_assert(kMaxClasses * 8 <= 32, "packed count fields");
  uint32_t pk[kMaxClasses];
  float dsum[2 * kMaxClasses];
#pragma unroll
  for (int i = 0; i < kMaxClasses; ++i) pk[i] = 0u;
#pragma unroll
  for (int i = 0; i < 2 * kMaxClasses; ++i) dsum[i] = 0.0f;
  int tiles_since_flush = 0;
  auto flush_counts = [&](int batch) {
    const int C = p.n_classes;
#pragma unroll
    for (int i = 0; i < kCells; ++i) {
      uint32_t v = (pk[i / kMaxClasses] >> (8 * (i % kMaxClasses))) & 255u;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && v) atomicAdd(&s_cm[i], static_cast<unsigned long long>(v));
    }
#pragma unroll
    for (int i = 0; i < kMaxClasses; ++i) pk[i] = 0u;
    if (p.dice) {
#pragma unroll
      for (int i = 0; i < 2 * kMaxClasses; ++i) {
        float v = dsum[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) atomicAdd(&s_dice[i], static_cast<double>(v));
        dsum[i] = 0.0f;
      }
    }
    epi_bar_sync(grp);  // all four warps of the group have added their share
    if (et < C * C) {
      const int tc = et / C, pc = et % C;
      const unsigned long long v = atomicExch(&s_cm[tc * kMaxClasses + pc], 0ull);
      if (v) atomicAdd(&p.cm[static_cast<size_t>(batch) * C * C + et], v);
    } else if (p.dice && et >= 32 && et < 32 + 2 * C) {
      const int i = et - 32;
      const double v = s_dice[i];
      s_dice[i] = 0.0;
      atomicAdd(&p.dice[(static_cast<size_t>(batch) * C + (i >> 1)) * 3 + (i & 1)], v);
    }
    epi_bar_sync(grp);  // shared counters are zero again before the next batch is added
    tiles_since_flush = 0;
  };

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
      const int batch = batch_now;
      if (batch != cur_batch || tiles_since_flush >= kFlushTiles) {
        if (cur_batch >= 0) flush_counts(cur_batch);
        cur_batch = batch;
      }
      ++tiles_since_flush;
      // the target is fetched before waiting for the accumulator so its latency hides behind the MMAs
      if (inb) {
        const size_t tp = (static_cast<size_t>(b) * p.H + y) * p.W + x;
        tgt_pre = p.target_fmt == 1 ? static_cast<int>(reinterpret_cast<const uint8_t*>(p.targets)[tp])
                                    : static_cast<int>(reinterpret_cast<const long long*>(p.targets)[tp]);
      }
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
      auto head_pixel = [&](auto nc_tag) {
        constexpr int NC = decltype(nc_tag)::value;
        const int C = NC ? NC : p.n_classes;
      float z[kMaxClasses];
#pragma unroll
        for (int c = 0; c < kMaxClasses; ++c) z[c] = (c < C) ? p.head_b[c] : -INFINITY;
        // packed fp32x2 arithmetic (two IEEE operations per issue slot): even and odd input channels accumulate
        // separately and are added at the end
        float2 z2[kMaxClasses];
#pragma unroll
        for (int c = 0; c < kMaxClasses; ++c) z2[c] = make_float2(z[c], 0.0f);
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float2 v = __fadd2_rn(make_float2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])),
                                make_float2(p.bias32[j], p.bias32[j + 1]));
          if (p.relu) { v.x = fmaxf(v.x, 0.0f); v.y = fmaxf(v.y, 0.0f); }
#pragma unroll
          for (int c = 0; c < kMaxClasses; ++c)
            if (c < C) z2[c] = __ffma2_rn(make_float2(p.head_w[c * 32 + j], p.head_w[c * 32 + j + 1]), v, z2[c]);
        }
#pragma unroll
        for (int c = 0; c < kMaxClasses; ++c)
          if (c < C) z[c] = z2[c].x + z2[c].y;
        // argmax(softmax(z)) with the first index winning ties (models/losses.py:48,50).  When the largest logit
        // leads the runner-up by more than 1e-4 the softmax cannot change the order: e_best = expf(0) = 1 exactly and
        // every other e_c <= expf(-1e-4) < 1 - 9e-5, far more than the rounding of the division by the common sum, so
        // pr_best > pr_c strictly and the prediction is the (unique) argmax of z.  Only near-ties, non-finite logits
        // and the soft-Dice sums (which need the probabilities) evaluate the softmax itself.
        float zmax = z[0];
#pragma unroll
        for (int c = 1; c < kMaxClasses; ++c) zmax = fmaxf(zmax, z[c]);
        int pred = 0;
        float second = -INFINITY;
        if constexpr (NC == 3) {
          const float m12 = fmaxf(z[1], z[2]), m02 = fmaxf(z[0], z[2]), m01 = fmaxf(z[0], z[1]);
          pred = (z[0] == zmax) ? 0 : ((z[1] == zmax) ? 1 : 2);
          second = (pred == 0) ? m12 : ((pred == 1) ? m02 : m01);
        } else {
#pragma unroll
          for (int c = kMaxClasses - 1; c >= 0; --c) if (c < C && z[c] == zmax) pred = c;
#pragma unroll
          for (int c = 0; c < kMaxClasses; ++c) if (c < C && c != pred) second = fmaxf(second, z[c]);
        }
        const float gap = zmax - second;
        const bool decisive = (gap > 1e-4f) && (gap <= 3.0e38f) && !(p.dbg & 8);  // false for NaN / infinite logits
        float pr[kMaxClasses];
        if (!decisive || p.dice) {
          float psum = 0.0f;
#pragma unroll
          for (int c = 0; c < kMaxClasses; ++c) { pr[c] = (c < C) ? expf(z[c] - zmax) : 0.0f; psum += pr[c]; }
          int pred_s = 0;
          float best = -1.0f;
#pragma unroll
          for (int c = 0; c < kMaxClasses; ++c) {
            pr[c] = pr[c] / psum;
            if (c < C && pr[c] > best) { best = pr[c]; pred_s = c; }
          }
          pred = pred_s;
        }
        const size_t pix = (static_cast<size_t>(b) * p.H + y) * p.W + x;
        if (inb) {
          if (p.pred_out) p.pred_out[pix] = static_cast<uint8_t>(pred);
          if (p.logits_out) {
#pragma unroll
            for (int c = 0; c < kMaxClasses; ++c)
              if (c < C) p.logits_out[((static_cast<size_t>(b) * C + c) * p.H + y) * p.W + x] = z[c];
          }
        }
        if (p.targets) {
          const int tgt = (tgt_pre >= 0 && tgt_pre < C) ? tgt_pre : -1;
          const uint32_t inc = 1u << (8 * pred);
#pragma unroll
          for (int tc = 0; tc < kMaxClasses; ++tc)
            if (tc < C) pk[tc] += (tgt == tc) ? inc : 0u;
          if (p.dice) {
            // smp DiceLoss(from_logits=True) applies softmax again to the softmax output (losses.py:100)
            float q[kMaxClasses], qmax = pr[0], qsum = 0.0f;
#pragma unroll
            for (int c = 1; c < kMaxClasses; ++c) if (c < C) qmax = fmaxf(qmax, pr[c]);
#pragma unroll
            for (int c = 0; c < kMaxClasses; ++c) { q[c] = (c < C) ? expf(pr[c] - qmax) : 0.0f; qsum += q[c]; }
#pragma unroll
            for (int c = 0; c < kMaxClasses; ++c) {
              const float pc = (inb && c < C) ? q[c] / qsum : 0.0f;
              dsum[2 * c + 0] += (tgt == c) ? pc : 0.0f;
              dsum[2 * c + 1] += pc;
            }
          }
        }
      };
      if (p.n_classes == 3) head_pixel(std::integral_constant<int, 3>{});
      else head_pixel(std::integral_constant<int, 0>{});
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
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 bz = *reinterpret_cast<const float4*>(s_bias + bias_off + 4 * j4);
          const float2 a = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 0]), __uint_as_float(r[4 * j4 + 1])),
                                      make_float2(bz.x, bz.y));
          const float2 c = __fadd2_rn(make_float2(__uint_as_float(r[4 * j4 + 2]), __uint_as_float(r[4 * j4 + 3])),
                                      make_float2(bz.z, bz.w));
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
      if constexpr (KXS) {
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
  if (kHead && p.mode == EPI_HEAD && p.targets && cur_batch >= 0) flush_counts(cur_batch);
}
