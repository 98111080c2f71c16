// Epilogue body of conv_tc_kernel (included inside the `warp >= 4` branch).
// Thread (wq, lane) owns accumulator row m = 32*wq + lane = pixel (ty = 2*wq + lane/16, tx = lane%16)
// of the tile and reads its BLOCK_N fp32 columns from TMEM 32 at a time.
{
  const int wq = warp & 3;
  const int grp = (warp - 4) >> 2;             // 0: even tiles of this CTA / accumulator 0, 1: odd / accumulator 1
  const int ty = 2 * wq + (lane >> 4);
  const int tx = lane & 15;
  const int et = (threadIdx.x - 128) & 127;    // 0..127 within the group
  unsigned long long* s_cm = s_cm_all + grp * (kMaxClasses * kMaxClasses);
  double* s_dice = s_dice_all + grp * (2 * kMaxClasses);
  int it = 0;
  int cur_batch = -1;
  // HEAD: every thread counts its own pixels in registers (cell = target*C + prediction, and the
  // soft-Dice partial sums); they are reduced (warp shuffle -> shared memory -> global atomics) only
  // when the reference batch changes, every kFlushTiles tiles, and once at the end.
  constexpr int kCells = kMaxClasses * kMaxClasses;
  constexpr int kFlushTiles = 4096;
  uint32_t cnt[kCells];
  float dsum[2 * kMaxClasses];
#pragma unroll
  for (int i = 0; i < kCells; ++i) cnt[i] = 0u;
#pragma unroll
  for (int i = 0; i < 2 * kMaxClasses; ++i) dsum[i] = 0.0f;
  int tiles_since_flush = 0;
  auto flush_counts = [&](int batch) {
    const int C = p.n_classes;
#pragma unroll
    for (int i = 0; i < kCells; ++i) {
      uint32_t v = cnt[i];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && v) atomicAdd(&s_cm[i], static_cast<unsigned long long>(v));
      cnt[i] = 0u;
    }
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
  for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
    if ((it & 1) != grp) continue;
    const int acc = grp;
    const uint32_t use = static_cast<uint32_t>(it >> 1);
    const int nt = t % n_tiles;
    int mm = t / n_tiles;
    const int x = (mm % tiles_x) * kTileCols + tx;
    mm /= tiles_x;
    const int y = (mm % tiles_y) * kTileRows + ty;
    const int b = mm / tiles_y;
    const bool inb = (y < p.H) && (x < p.W);
    const int n0 = nt * BLOCK_N;
    int tgt_pre = -1;

    if (p.mode == EPI_HEAD && p.targets) {
      // counts accumulate in shared memory across tiles and go to HBM only when the reference batch
      // changes (and once at the end): <= C*C global atomics per batch per CTA instead of per tile
      const int batch = (p.img_base + b) / p.ref_batch;
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
        tmem_base + (static_cast<uint32_t>(wq * 32) << 16) + static_cast<uint32_t>(acc * BLOCK_N);

    if (p.mode == EPI_HEAD) {
      // ---- dec4.conv.3 epilogue: bias+ReLU (fp32) -> 1x1 head -> softmax -> argmax -> counts
      uint32_t r[32];
      tmem_ld_32x32(t_addr, r);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&t_empty[acc]);
      const int C = p.n_classes;
      float z[kMaxClasses];
#pragma unroll
      for (int c = 0; c < kMaxClasses; ++c) z[c] = (c < C) ? s_head[kMaxClasses * 32 + c] : -INFINITY;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float v = __uint_as_float(r[j]) + s_bias[j];
        v = p.relu ? fmaxf(v, 0.0f) : v;
#pragma unroll
        for (int c = 0; c < kMaxClasses; ++c)
          if (c < C) z[c] = fmaf(s_head[c * 32 + j], v, z[c]);
      }
      // softmax (models/losses.py:48), then first-index argmax (:50)
      float zmax = z[0];
#pragma unroll
      for (int c = 1; c < kMaxClasses; ++c) zmax = fmaxf(zmax, z[c]);
      float pr[kMaxClasses], psum = 0.0f;
#pragma unroll
      for (int c = 0; c < kMaxClasses; ++c) { pr[c] = (c < C) ? expf(z[c] - zmax) : 0.0f; psum += pr[c]; }
      int pred = 0;
      float best = -1.0f;
#pragma unroll
      for (int c = 0; c < kMaxClasses; ++c) {
        pr[c] = pr[c] / psum;
        if (c < C && pr[c] > best) { best = pr[c]; pred = c; }
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
#pragma unroll
        for (int tc = 0; tc < kMaxClasses; ++tc)
#pragma unroll
          for (int pc = 0; pc < kMaxClasses; ++pc) cnt[tc * kMaxClasses + pc] += (tgt == tc && pred == pc) ? 1u : 0u;
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
    } else {
      // ---- STORE / POOL / CONVT: bias (+ReLU) -> bf16 -> NHWC global store, 32 channels at a time
#pragma unroll 1
      for (int cb = 0; cb < BLOCK_N; cb += 32) {
        uint32_t r[32];
        tmem_ld_32x32(t_addr + static_cast<uint32_t>(cb), r);
        tmem_ld_wait();
        if (cb + 32 >= BLOCK_N) {  // last read of this accumulator: hand it back to the MMA warp
          tc_fence_before();
          mbar_arrive(&t_empty[acc]);
        }
        const int n = n0 + cb;
        int bias_off = n;
        int oy = y, ox = x, oh = p.H, ow = p.W, och = p.out_choff + n;
        if (p.mode == EPI_CONVT) {
          const int q = n / p.cout;
          bias_off = n - q * p.cout;
          oy = 2 * y + (q >> 1);
          ox = 2 * x + (q & 1);
          oh = 2 * p.H;
          ow = 2 * p.W;
          och = p.out_choff + bias_off;
        }
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          v[j] = __uint_as_float(r[j]) + s_bias[bias_off + j];
          if (p.relu) v[j] = fmaxf(v[j], 0.0f);
        }
        bool writer = inb;
        if (p.mode == EPI_POOL) {
          // MaxPool2d(2,2): partners are lane^1 (x) and lane^16 (y) of the same warp
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            v[j] = fmaxf(v[j], __shfl_xor_sync(0xffffffffu, v[j], 1));
            v[j] = fmaxf(v[j], __shfl_xor_sync(0xffffffffu, v[j], 16));
          }
          writer = inb && ((lane & 17) == 0);
          oy = y >> 1;
          ox = x >> 1;
          oh = p.H >> 1;
          ow = p.W >> 1;
        }
        if (writer) {
          __nv_bfloat16* dst =
              p.out + ((static_cast<size_t>(b) * oh + oy) * ow + ox) * p.out_cstride + och;
          uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 o;
            o.x = pack_bf16(v[8 * g + 0], v[8 * g + 1]);
            o.y = pack_bf16(v[8 * g + 2], v[8 * g + 3]);
            o.z = pack_bf16(v[8 * g + 4], v[8 * g + 5]);
            o.w = pack_bf16(v[8 * g + 6], v[8 * g + 7]);
            d4[g] = o;
          }
        }
      }
    }
  }
  if (p.mode == EPI_HEAD && p.targets && cur_batch >= 0) flush_counts(cur_batch);
}
