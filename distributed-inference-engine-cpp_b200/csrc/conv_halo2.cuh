// 3x3 / stride 1 / pad 1 with one halo load per tile AND two M tiles per weight tile ("dual-M halo").
// (included by conv_umma.cu inside namespace rn::{anonymous}, after conv3x3_halo_kernel whose constants it shares)
//
// When a layer's weight slab does not fit in shared memory (128 -> 128 channels: 288 KB), conv3x3_halo_kernel streams
// all 18 weight tiles (16 KB each) through a ring again for EVERY 112-pixel tile: 288 KB of weights against 40 KB of
// activations per tile.  Measured with the MMAs switched off (B200RN_DBG_SKIP=1) those layers still take 0.062 of
// their 0.076 ms: they run at the pace of the L2 -> shared-memory feed (~50 B/clk/SM), not of the tensor core.
// Here an item is TWO M tiles (two halo patches per ring slot); every weight tile that lands is used by eight MMAs --
// four into each tile's accumulator -- so the weight bytes per output tile halve.  Four accumulators (2 item parities
// x 2 tiles x N columns) fill tensor memory at N = 128; epilogue group h drains tile h of every item (own staging
// buffer, named barrier and store thread), so the epilogue of item i overlaps the MMAs of item i + 1.
struct Halo2Params {
  int cblks, n_tiles, m_tiles, total_items, tiles_w, tiles_h;
  int halo_stages, b_stages, relu;
  const float* bias;
};

template <typename T, int BLOCK_N>
__global__ void __launch_bounds__(kTwoGroupThreads, 1)
conv3x3_halo2_kernel(const __grid_constant__ CUtensorMap tmHalo, const __grid_constant__ CUtensorMap tmW,
                     const __grid_constant__ CUtensorMap tmOutRow, const Halo2Params p) {
  constexpr int kChunks = BLOCK_N / 64;
  constexpr uint32_t B_BYTES = BLOCK_N * 128;
  constexpr uint32_t OUT_BYTES = kChunks * kChunkBytes;
  constexpr uint32_t SLOT_BYTES = 2 * kHaloStageBytes;             // two patches per ring slot

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int HS = p.halo_stages, SB = p.b_stages;
  uint8_t* halo = smem;
  uint8_t* bbuf = halo + HS * SLOT_BYTES;                          // ring of SB weight tiles
  uint8_t* outbuf = bbuf + SB * B_BYTES;                           // 2 x [kChunks x 16 KB], one per epilogue group
  float* s_bias_all = reinterpret_cast<float*>(outbuf + 2 * OUT_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_bias_all + 2 * BLOCK_N);
  uint64_t* halo_full = bars;                 // [HS]
  uint64_t* halo_empty = halo_full + HS;      // [HS]
  uint64_t* b_full = halo_empty + HS;         // [SB]
  uint64_t* b_empty = b_full + SB;            // [SB]
  uint64_t* tmem_full = b_empty + SB;         // [2]  per item parity
  uint64_t* tmem_empty = tmem_full + 2;       // [2]  both groups' threads
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int my_items = (int(blockIdx.x) < p.total_items) ? (p.total_items - 1 - int(blockIdx.x)) / int(gridDim.x) + 1 : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmHalo);
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmOutRow);
    for (int i = 0; i < HS; ++i) { mbar_init(smem_u32(&halo_full[i]), 1); mbar_init(smem_u32(&halo_empty[i]), 1); }
    for (int i = 0; i < SB; ++i) { mbar_init(smem_u32(&b_full[i]), 1); mbar_init(smem_u32(&b_empty[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&tmem_full[i]), 1); mbar_init(smem_u32(&tmem_empty[i]), 2 * kEpiThreads); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 4 * BLOCK_N);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  // item -> (N tile, pair of M tiles); M tile h of the pair -> patch origin.  An M tile index past the last one decodes
  // to an image beyond the batch: its TMA loads read zeros and its TMA stores are clipped.
  auto item_coord = [&](int i, int h, int& x0, int& y0, int& n, int& col0) {
    const int item = int(blockIdx.x) + i * int(gridDim.x);
    col0 = (item % p.n_tiles) * BLOCK_N;
    int mt = 2 * (item / p.n_tiles) + h;
    const int tw = mt % p.tiles_w; mt /= p.tiles_w;
    const int th = mt % p.tiles_h;
    n = mt / p.tiles_h;
    x0 = tw * kHaloBW; y0 = th * kHaloBH;
  };

  if (warp == 0) {
    // ===== TMA producer: halo patch pairs run ahead of the weight tiles =====
    if (elect_one() && my_items > 0) {
      const int Q = my_items * p.cblks;         // units: (item, channel block)
      auto issue_halo = [&](int hq) {
        const int i = hq / p.cblks, cb = hq % p.cblks;
        const int slot = hq % HS;
        const uint32_t fb = smem_u32(&halo_full[slot]);
        mbar_expect_tx(fb, 2 * kHaloLoadBytes);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          int x0, y0, n, col0;
          item_coord(i, h, x0, y0, n, col0);
          tma_load_4d(smem_u32(halo + slot * SLOT_BYTES + h * kHaloStageBytes), &tmHalo, fb, cb * 64, x0 - 1, y0 - 1, n);
        }
      };
      int hq = 0, bstage = 0;
      uint32_t bphase = 0;
      for (int q = 0; q < Q; ++q) {
        // keep up to HS patch pairs in flight without ever blocking on a slot that is still being read
        while (hq < Q && hq < q + HS) {
          const int slot = hq % HS;
          const uint32_t par = (uint32_t(hq / HS) & 1) ^ 1;
          if (hq == q) mbar_wait(smem_u32(&halo_empty[slot]), par);               // this one is needed now
          else if (hq >= HS && !mbar_try_wait(smem_u32(&halo_empty[slot]), par)) break;
          issue_halo(hq);
          ++hq;
        }
        const int i = q / p.cblks, cb = q % p.cblks;
        int x0, y0, n, col0;
        item_coord(i, 0, x0, y0, n, col0);
        for (int tap = 0; tap < 9; ++tap) {
          mbar_wait(smem_u32(&b_empty[bstage]), bphase ^ 1);
          const uint32_t fb = smem_u32(&b_full[bstage]);
          mbar_expect_tx(fb, B_BYTES);
          tma_load_2d(smem_u32(bbuf + bstage * B_BYTES), &tmW, fb, (tap * p.cblks + cb) * 64, col0);
          if (++bstage == SB) { bstage = 0; bphase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one elected thread; every weight tile feeds both M tiles of the item =====
    constexpr uint32_t idesc = make_idesc<T, BLOCK_N>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (elect_one()) {
      const uint32_t halo_lo = smem_desc_lo(smem_u32(halo));
      const uint32_t bbuf_lo = smem_desc_lo(smem_u32(bbuf));
      // the next unit's barriers are probed (non-blocking) before this unit's last two taps are issued (see conv3x3_halo_kernel)
      const int units = my_items * p.cblks;
      int bstage = 0, slot = 0, i = 0, cb = 0;
      uint32_t bphase = 0, hphase = 0;
      bool ready = false;
      for (int q = 0; q < units; ++q) {
        const int buf = i & 1;
        if (!ready) {
          if (cb == 0) mbar_wait(smem_u32(&tmem_empty[buf]), ((uint32_t(i) >> 1) & 1) ^ 1);
          mbar_wait(smem_u32(&halo_full[slot]), hphase);
        }
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + uint32_t(buf * 2 * BLOCK_N);
        const uint32_t a_lo = halo_lo + uint32_t(slot) * (SLOT_BYTES >> 4);
        const bool last_cb = cb == p.cblks - 1;
        int nslot = slot + 1;
        uint32_t nphase = hphase;
        if (nslot == HS) { nslot = 0; nphase ^= 1; }
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          if (tap == 7)                         // look ahead, non-blocking
            ready = q + 1 < units && mbar_test(smem_u32(&halo_full[nslot]), nphase) &&
                    (!last_cb || mbar_test(smem_u32(&tmem_empty[(i + 1) & 1]), ((uint32_t(i + 1) >> 1) & 1) ^ 1));
          mbar_wait(smem_u32(&b_full[bstage]), bphase);
          tc_fence_after();
          const uint32_t b_lo = bbuf_lo + uint32_t(bstage) * (B_BYTES >> 4);
          // row-shifted view of a patch: tap (ky,kx) starts ky*16 + kx rows (of 128 bytes) further on
          const uint32_t shift = uint32_t((tap / 3) * kHaloPW + (tap % 3)) * 8;
#pragma unroll
          for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_f16_lo(tmem_d + uint32_t(h * BLOCK_N), a_lo + uint32_t(h) * (kHaloStageBytes >> 4) + shift + 2 * k, b_lo + 2 * k, desc_hi,
                          idesc, (tap | k) != 0 ? 1u : uint32_t(cb != 0));
          umma_commit(smem_u32(&b_empty[bstage]));
          if (++bstage == SB) { bstage = 0; bphase ^= 1; }
        }
        umma_commit(smem_u32(&halo_empty[slot]));
        if (last_cb) umma_commit(smem_u32(&tmem_full[buf]));
        slot = nslot; hphase = nphase;
        if (last_cb) { cb = 0; ++i; } else ++cb;
      }
    }
  } else {
    // ===== epilogue: group h (8 warps) takes M tile h of every item: bias + ReLU, the 14 valid pixels of each row packed
    // densely, one TMA store per 64-channel chunk =====
    const int grp = (warp - 2) >> 3;
    const int q4 = warp & 3;
    const int g = ((warp - 2) & 7) >> 2;
    float* s_bias = s_bias_all + grp * BLOCK_N;
    const int row = q4 * 32 + lane;                       // TMEM lane = y*16 + x
    const bool valid = (row & 15) < kHaloBW;              // x = 14, 15 are the junk columns
    const int drow = (row >> 4) * kHaloBW + (row & 15);   // dense staging row y*14 + x
    const int sw = drow & 7;
    const int et = threadIdx.x - 64 - grp * kEpiThreads;
    const bool store_thread = (((warp - 2) & 7) == 0 && lane == 0);
    const int bar_id = 1 + grp;
    uint8_t* out_tile = outbuf + grp * OUT_BYTES;
    int cur_col0 = -1;
    for (int i = 0; i < my_items; ++i) {
      int x0, y0, n, col0;
      item_coord(i, grp, x0, y0, n, col0);
      const int buf = i & 1;
      if (col0 != cur_col0) {
        cur_col0 = col0;
        for (int k = et; k < BLOCK_N; k += kEpiThreads) s_bias[k] = __ldg(p.bias + col0 + k);
      }
      if (store_thread && i > 0) tma_store_wait_read();   // this group's previous tile has left its staging buffer
      asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
      mbar_wait(smem_u32(&tmem_full[buf]), (uint32_t(i) >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(tmem_base + (uint32_t(q4 * 32) << 16) + uint32_t((buf * 2 + grp) * BLOCK_N + c * 64 + g * 32), acc);
        tmem_ld_wait();
        if (c == kChunks - 1) {
          tc_fence_before();
          mbar_arrive(smem_u32(&tmem_empty[buf]));
        }
        uint8_t* stage = out_tile + c * kChunkBytes;
        uint4* out_row = reinterpret_cast<uint4*>(stage + drow * 128);
        if (valid) {
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const int j = g * 4 + jj;
            const int col = c * 64 + j * 8;
            const float4 b0 = *reinterpret_cast<const float4*>(s_bias + col);
            const float4 b1 = *reinterpret_cast<const float4*>(s_bias + col + 4);
            float v[8];
            v[0] = __uint_as_float(acc[jj * 8 + 0]) + b0.x; v[1] = __uint_as_float(acc[jj * 8 + 1]) + b0.y;
            v[2] = __uint_as_float(acc[jj * 8 + 2]) + b0.z; v[3] = __uint_as_float(acc[jj * 8 + 3]) + b0.w;
            v[4] = __uint_as_float(acc[jj * 8 + 4]) + b1.x; v[5] = __uint_as_float(acc[jj * 8 + 5]) + b1.y;
            v[6] = __uint_as_float(acc[jj * 8 + 6]) + b1.z; v[7] = __uint_as_float(acc[jj * 8 + 7]) + b1.w;
            if (p.relu) {
#pragma unroll
              for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
            }
            out_row[j ^ sw] = make_uint4(pack2<T>(v[0], v[1]), pack2<T>(v[2], v[3]), pack2<T>(v[4], v[5]), pack2<T>(v[6], v[7]));
          }
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
        if (store_thread) {
          tma_store_4d(&tmOutRow, smem_u32(stage), col0 + c * 64, x0, y0, n);   // box {64 ch, 14, 8, 1}
          tma_store_commit();
        }
      }
    }
    if (store_thread) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 4 * BLOCK_N);
}
