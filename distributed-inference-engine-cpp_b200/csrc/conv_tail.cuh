// 1x1 conv + residual add, and the NEXT bottleneck block's first 1x1 conv, in one kernel.
// (included by conv_umma.cu inside namespace rn::{anonymous}, after the PTX wrappers)
//
// In a pre-activation ResNet the block output T' = conv3(U2) + T is written twice today: raw (the next
// block's shortcut) and as relu(bn(T')) (the next block's conv1 operand), and that second copy is read
// straight back by a 1x1 conv.  On the 56x56 stage these two layers are pure HBM traffic (T' is 4x wider
// than anything else in the block).  Here the activated copy never leaves the SM:
//
//   per 128-pixel tile, per 64-channel chunk c of the wide tensor:
//     acc3[:, c] = U2[128 x Cin] * W3[c]^T                        (tcgen05, TMEM cols 64c..64c+63)
//     v = acc3 + bias3 (+ReLU) + T      -> raw chunk, TMA-stored from the residual's own staging slot
//     a = relu?(s*v + t) as 16-bit      -> shared memory, swizzle-128B = the K-major A operand of K block c
//     acc1[128 x N1] += a * W1[:, c]^T  (tcgen05, TMEM cols 256.., double-buffered)
//   u = relu?(acc1 + bias1)             -> TMA-stored: the next block's conv1 output
//
// HBM per tile: reads Cin*256 B + 64 KB (T), writes 64 KB (T') + N1*256 B; the unfused pair moves 128 KB more.
//
// Nothing in the epilogue is a CTA-wide barrier.  Every buffer has its own mbarrier pair and the warps drift:
//   residual / raw ring slot : r_full (TMA landed)            -> epilogue adds in place -> staged (8 warp arrivals)
//                              -> store warp: TMA store, and once the store has read the slot -> r_empty -> TMA refills
//   activated chunk buffer   : staged -> store warp forwards act_ready -> MMA warp runs the tail GEMM's K block,
//                              tcgen05.commit -> act_free -> epilogue may overwrite
//   acc3 slice c             : after the tail MMAs of chunk c the MMA warp issues the NEXT tile's conv3 slice c
//                              (act_ready(c) implies every warp has drained that slice), commit -> acc3_full[c]
//   acc1[tile & 1]           : drained by the epilogue one chunk into the next tile, so the MMA latency is hidden
// The thread that issues stores is in a warp of its own: measured on the first version of this kernel, the
// issue + "previous store has been read" wait cost the epilogue ~500 of 2270 cycles per chunk when one of the
// epilogue threads did it between two bar.syncs.
constexpr int kTailN3 = 256;
constexpr int kTailEpiWarps = 16;                    // two groups of 8: group 0 takes chunks 0 and 2, group 1 chunks 1 and 3
constexpr int kTailGroupWarps = 8;
constexpr int kTailThreads = (3 + kTailEpiWarps) * 32;   // TMA warp, MMA warp, store warp, epilogue warps

struct TailParams {
  int cblks, bw, bh, bn, tiles_w, tiles_h, total_tiles;
  int a_stages, r_slots, act_bufs;
  int relu, act_relu, tail_relu;
  int has_res;                 // 0: no residual tensor (two-input conv: the shortcut is part of the GEMM); the ring is then staging only
  int k1_blocks, stride_b;     // two-input conv: K blocks [0, k1_blocks) of a tile come from tmA, the rest from tmA2 at stride_b
  int pf_dist;                 // residual L2 prefetch distance in tiles (0 = off)
  int pf_in_b;                 // second-input L2 prefetch distance in tiles (0 = off)
  uint32_t rows;
  const float* bias;
  const float* act_scale;
  const float* act_shift;
  const float* tail_bias;
  unsigned long long* trace;   // timeline experiment (B200RN_DBG_TRACE): CTA 0, first epilogue warp, clock64 per chunk
};

struct TailTile { int w0, h0, n0; };
__device__ __forceinline__ TailTile tail_decode(const TailParams& p, int tile) {
  TailTile t;
  const int tw = tile % p.tiles_w; tile /= p.tiles_w;
  const int th = tile % p.tiles_h;
  t.w0 = tw * p.bw; t.h0 = th * p.bh; t.n0 = (tile / p.tiles_h) * p.bn;
  return t;
}

template <typename T, int N1>
__global__ void __launch_bounds__(kTailThreads, 1)
conv_tail_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
                 const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmRes,
                 const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmOut1, const TailParams p) {
  constexpr int kChunks = kTailN3 / 64;                 // 64-channel chunks of the wide tensor = K blocks of the tail GEMM
  constexpr int kChunks1 = N1 / 64;
  constexpr uint32_t W3_BLOCK = kTailN3 * 128;          // one K block of W3 (256 rows x 128 B)
  constexpr uint32_t W3_SLICE = 64 * 128;               // the 64 output channels of one chunk within it
  constexpr uint32_t W1_BLOCK = N1 * 128;               // one K block of W1
  constexpr uint32_t A_BLOCK = kTileM * 128;
  constexpr uint32_t kAcc1Col = kTailN3;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int SA = p.a_stages, SR = p.r_slots, AB = p.act_bufs;
  uint8_t* w3 = smem;
  uint8_t* w1 = w3 + p.cblks * W3_BLOCK;
  uint8_t* a_ring = w1 + kChunks * W1_BLOCK;
  uint8_t* r_ring = a_ring + SA * p.cblks * A_BLOCK;
  uint8_t* actb = r_ring + SR * kChunkBytes;
  uint8_t* out1 = actb + AB * kChunkBytes;
  float* s_bias = reinterpret_cast<float*>(out1 + kChunks1 * kChunkBytes);
  float* s_scale = s_bias + kTailN3;
  float* s_shift = s_scale + kTailN3;
  float* s_bias1 = s_shift + kTailN3;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_bias1 + N1);
  uint64_t* w_bar = bars;                   // [1]   weights landed
  uint64_t* a_full = w_bar + 1;             // [SA]  TMA -> MMA
  uint64_t* a_empty = a_full + SA;          // [SA]  MMA -> TMA
  uint64_t* r_full = a_empty + SA;          // [SR]  TMA -> epilogue (residual chunk landed)
  uint64_t* r_empty = r_full + SR;          // [SR]  store warp -> TMA (the chunk's store has read the slot)
  uint64_t* staged = r_empty + SR;          // [SR]  epilogue -> store warp (raw + activated chunk written; 8 arrivals)
  uint64_t* act_ready = staged + SR;        // [AB]  store warp -> MMA
  uint64_t* act_free = act_ready + AB;      // [AB]  MMA -> epilogue (tail MMAs on the chunk done)
  uint64_t* acc3_full = act_free + AB;      // [4]   MMA -> epilogue, per 64-column slice
  uint64_t* acc1_full = acc3_full + kChunks;  // [2] MMA -> epilogue
  uint64_t* out1_staged = acc1_full + 2;    // [1]   epilogue -> store warp (8 arrivals)
  uint64_t* out1_free = out1_staged + 1;    // [1]   store warp -> epilogue
  uint64_t* acc1_empty = out1_free + 1;     // [2]   epilogue -> MMA (8 arrivals: the group that drained it)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc1_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmA2); prefetch_tmap(&tmW); prefetch_tmap(&tmW1);
    prefetch_tmap(&tmRes); prefetch_tmap(&tmOut); prefetch_tmap(&tmOut1);
    mbar_init(smem_u32(w_bar), 1);
    for (int i = 0; i < SA; ++i) { mbar_init(smem_u32(&a_full[i]), 1); mbar_init(smem_u32(&a_empty[i]), 1); }
    for (int i = 0; i < SR; ++i) {
      mbar_init(smem_u32(&r_full[i]), 1); mbar_init(smem_u32(&r_empty[i]), 1); mbar_init(smem_u32(&staged[i]), kTailGroupWarps);
    }
    for (int i = 0; i < AB; ++i) { mbar_init(smem_u32(&act_ready[i]), 1); mbar_init(smem_u32(&act_free[i]), 1); }
    for (int i = 0; i < kChunks; ++i) mbar_init(smem_u32(&acc3_full[i]), 1);
    mbar_init(smem_u32(&acc1_full[0]), 1); mbar_init(smem_u32(&acc1_full[1]), 1);
    mbar_init(smem_u32(out1_staged), kTailGroupWarps);
    // no residual tensor: the ring is only the raw output's staging, nothing has to land in a slot.  Its "ready"
    // barrier is then completed by the store warp itself when a store has read the slot (and once here, for the
    // first use), so the operand producer is not held back by the ring at all.
    if (!p.has_res)
      for (int i = 0; i < SR; ++i) mbar_arrive(smem_u32(&r_full[i]));
    mbar_init(smem_u32(&acc1_empty[0]), kTailGroupWarps); mbar_init(smem_u32(&acc1_empty[1]), kTailGroupWarps);
    mbar_init(smem_u32(out1_free), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 512);
  for (int k = threadIdx.x; k < kTailN3; k += blockDim.x) {
    s_bias[k] = __ldg(p.bias + k);
    s_scale[k] = __ldg(p.act_scale + k);
    s_shift[k] = __ldg(p.act_shift + k);
    if (k < N1) s_bias1[k] = __ldg(p.tail_bias + k);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();
  const int my_tiles = (p.total_tiles - int(blockIdx.x) + int(gridDim.x) - 1) / int(gridDim.x);   // >= 1: grid <= tiles

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      const uint32_t wb = smem_u32(w_bar);
      mbar_expect_tx(wb, uint32_t(p.cblks) * W3_BLOCK + kChunks * W1_BLOCK);
      for (int kb = 0; kb < p.cblks; ++kb) tma_load_2d(smem_u32(w3 + kb * W3_BLOCK), &tmW, wb, kb * 64, 0);
      for (int c = 0; c < kChunks; ++c) tma_load_2d(smem_u32(w1 + c * W1_BLOCK), &tmW1, wb, c * 64, 0);
      int stage = 0;
      uint32_t phase = 0, cc = 0;
      // experiment knob (B200RN_TAIL_PF, default off): L2 prefetch of a later tile's residual; measured slower
      auto prefetch_tile = [&](int t) {
        if (t >= p.total_tiles || p.pf_dist == 0 || !p.has_res) return;
        const TailTile pt = tail_decode(p, t);
        for (int c = 0; c < kChunks; ++c) tma_prefetch_4d(&tmRes, c * 64, pt.w0, pt.h0, pt.n0);
      };
      for (int j = 0; j < p.pf_dist; ++j) prefetch_tile(int(blockIdx.x) + j * int(gridDim.x));
      // Two-input conv: the second input was written several layers ago and comes from HBM, and with one A stage its
      // load cannot be issued until the previous tile's conv3 MMAs are issued -- the MMA thread then sat ~5000 cycles
      // per tile on that load (timeline).  Its 16 KB per tile are pulled into L2 two tiles ahead.
      auto prefetch_in_b = [&](int t) {
        if (p.k1_blocks >= p.cblks || t >= p.total_tiles || p.pf_in_b == 0) return;
        const TailTile pt = tail_decode(p, t);
        for (int kb = p.k1_blocks; kb < p.cblks; ++kb)
          tma_prefetch_4d(&tmA2, (kb - p.k1_blocks) * 64, pt.w0 * p.stride_b, pt.h0 * p.stride_b, pt.n0);
      };
      for (int j = 0; j < p.pf_in_b; ++j) prefetch_in_b(int(blockIdx.x) + j * int(gridDim.x));
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const TailTile tc = tail_decode(p, tile);
        prefetch_tile(tile + p.pf_dist * int(gridDim.x));
        prefetch_in_b(tile + p.pf_in_b * int(gridDim.x));
        mbar_wait(smem_u32(&a_empty[stage]), phase ^ 1);
        const uint32_t fb = smem_u32(&a_full[stage]);
        mbar_expect_tx(fb, uint32_t(p.cblks) * p.rows * 128u);
        for (int kb = 0; kb < p.cblks; ++kb) {
          const uint32_t dst = smem_u32(a_ring + (stage * p.cblks + kb) * A_BLOCK);
          if (kb < p.k1_blocks) tma_load_4d(dst, &tmA, fb, kb * 64, tc.w0, tc.h0, tc.n0);
          else tma_load_4d(dst, &tmA2, fb, (kb - p.k1_blocks) * 64, tc.w0 * p.stride_b, tc.h0 * p.stride_b, tc.n0);
        }
        if (++stage == SA) { stage = 0; phase ^= 1; }
        for (int c = 0; p.has_res && c < kChunks; ++c, ++cc) {
          const uint32_t slot = cc % uint32_t(SR);
          mbar_wait(smem_u32(&r_empty[slot]), ((cc / uint32_t(SR)) & 1) ^ 1);
          const uint32_t bar = smem_u32(&r_full[slot]);
          mbar_expect_tx(bar, p.rows * 128u);
          tma_load_4d(smem_u32(r_ring + slot * kChunkBytes), &tmRes, bar, c * 64, tc.w0, tc.h0, tc.n0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    // tile i: for each chunk c, as the epilogue hands over the activated chunk: the tail GEMM's K block c,
    // then conv3's slice c of tile i+1 (the epilogue is done with those 64 accumulator columns).
    constexpr uint32_t idesc3 = make_idesc<T, 64>();
    constexpr uint32_t idesc1 = make_idesc<T, N1>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (elect_one()) {
      const uint32_t w3_lo = smem_desc_lo(smem_u32(w3)), w1_lo = smem_desc_lo(smem_u32(w1));
      const uint32_t ring_lo = smem_desc_lo(smem_u32(a_ring)), act_lo = smem_desc_lo(smem_u32(actb));
      auto conv3_slice = [&](uint32_t a_lo, int c) {
        for (int kb = 0; kb < p.cblks; ++kb) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_lo(tmem_base + uint32_t(c * 64), a_lo + uint32_t(kb) * (A_BLOCK >> 4) + 2 * k,
                        w3_lo + uint32_t(kb) * (W3_BLOCK >> 4) + uint32_t(c) * (W3_SLICE >> 4) + 2 * k, desc_hi, idesc3,
                        (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&acc3_full[c]));
      };
      mbar_wait(smem_u32(w_bar), 0);
      int stage = 0;
      uint32_t phase = 0, cc = 0;
      mbar_wait(smem_u32(&a_full[0]), 0);
      tc_fence_after();
      for (int c = 0; c < kChunks; ++c) conv3_slice(ring_lo, c);
      umma_commit(smem_u32(&a_empty[0]));
      if (++stage == SA) { stage = 0; phase ^= 1; }
      for (int i = 0; i < my_tiles; ++i) {
        const bool has_next = i + 1 < my_tiles;
        const uint32_t acc1 = tmem_base + kAcc1Col + uint32_t((i & 1) * N1);
        uint32_t a_next = 0;
        for (int c = 0; c < kChunks; ++c, ++cc) {
          const uint32_t ab = cc % uint32_t(AB);
          mbar_wait(smem_u32(&act_ready[ab]), (cc / uint32_t(AB)) & 1);
          if (c == 0 && i >= 2) mbar_wait(smem_u32(&acc1_empty[i & 1]), uint32_t((i >> 1) - 1) & 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_lo(acc1, act_lo + ab * (kChunkBytes >> 4) + 2 * k, w1_lo + uint32_t(c) * (W1_BLOCK >> 4) + 2 * k,
                        desc_hi, idesc1, (c | k) != 0 ? 1u : 0u);
          umma_commit(smem_u32(&act_free[ab]));
          if (c == kChunks - 1) umma_commit(smem_u32(&acc1_full[i & 1]));
          if (has_next) {
            if (c == 0) {
              mbar_wait(smem_u32(&a_full[stage]), phase);
              tc_fence_after();
              a_next = ring_lo + uint32_t(stage * p.cblks) * (A_BLOCK >> 4);
            }
            conv3_slice(a_next, c);
            if (c == kChunks - 1) {
              umma_commit(smem_u32(&a_empty[stage]));
              if (++stage == SA) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 2) {
    // ===== store warp (one thread): forwards "chunk staged" to the MMA warp, issues the TMA stores, and hands
    // each staging buffer back once its store has read it =====
    if (elect_one()) {
      uint32_t cc = 0;
      // A buffer goes back as soon as its own store has read it (this thread has nothing better to do than wait):
      // the residual ring is short, and every cycle a slot sits here is a cycle its next load is not in flight.
      auto store_out1 = [&](const TailTile& tc, int j) {
        mbar_wait(smem_u32(out1_staged), uint32_t(j) & 1);
#pragma unroll
        for (int c1 = 0; c1 < kChunks1; ++c1) tma_store_4d(&tmOut1, smem_u32(out1 + c1 * kChunkBytes), c1 * 64, tc.w0, tc.h0, tc.n0);
        tma_store_commit();
        tma_store_wait_read();
        mbar_arrive(smem_u32(out1_free));
      };
      TailTile prev = {0, 0, 0};
      int i = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++i) {
        const TailTile tc = tail_decode(p, tile);
        for (int c = 0; c < kChunks; ++c, ++cc) {
          const uint32_t slot = cc % uint32_t(SR), ab = cc % uint32_t(AB);
          mbar_wait(smem_u32(&staged[slot]), (cc / uint32_t(SR)) & 1);
          mbar_arrive(smem_u32(&act_ready[ab]));
          tma_store_4d(&tmOut, smem_u32(r_ring + slot * kChunkBytes), c * 64, tc.w0, tc.h0, tc.n0);
          tma_store_commit();
          tma_store_wait_read();
          mbar_arrive(smem_u32(p.has_res ? &r_empty[slot] : &r_full[slot]));
          if (c == 1 && i > 0) store_out1(prev, i - 1);
        }
        prev = tc;
      }
      store_out1(prev, i - 1);
      tma_store_wait_read();
    }
  } else {
    // ===== epilogue: two groups of 8 warps, each group owns every other chunk; thread = one pixel (TMEM lane) x 32
    // of the chunk's 64 channels.  A warp's chunk is a serial latency chain (barrier polls, TMEM load, smem loads,
    // ~250 dependent ALU ops, smem stores, fence) of ~2400 cycles -- measured -- so two chunks are kept in flight
    // per SM; that, not arithmetic throughput, is what the second group buys. =====
    const int ew = warp - 3;
    const int grp = ew >> 3;                // chunk parity this warp works on
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    const int g = (ew & 7) >> 2;            // which half of every 64-channel chunk
    const int row = q * 32 + lane;
    const int sw = row & 7;
    const uint32_t lane_addr = tmem_base + (uint32_t(q * 32) << 16);
    const bool tracer = p.trace && blockIdx.x == 0 && lane == 0 && (ew & 7) == 0;
    unsigned long long* trace = p.trace + grp * 32 * 8;          // [group][chunk < 32][8 stamps]
    const uint32_t s_bias_a = smem_u32(s_bias), s_scale_a = smem_u32(s_scale), s_shift_a = smem_u32(s_shift);
    const uint32_t row_off = uint32_t(row) * 128u;

    // the tail conv's output tile j: acc1[j & 1] -> +bias, ReLU -> staging -> (store warp) TMA store
    auto finish_tail = [&](int j) {
      mbar_wait(smem_u32(&acc1_full[j & 1]), uint32_t(j >> 1) & 1);
      tc_fence_after();
      if (j > 0) mbar_wait(smem_u32(out1_free), uint32_t(j - 1) & 1);      // the previous tile's store has read the staging
#pragma unroll 1
      for (int c1 = 0; c1 < kChunks1; ++c1) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(lane_addr + kAcc1Col + uint32_t((j & 1) * N1 + c1 * 64 + g * 32), acc);
        tmem_ld_wait();
        if (c1 == kChunks1 - 1) {              // acc1[j & 1] is drained: the MMA warp may start tile j + 2 in it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&acc1_empty[j & 1]));
        }
        const uint32_t out_row = smem_u32(out1 + c1 * kChunkBytes) + row_off;
        const uint32_t b_a = smem_u32(s_bias1) + uint32_t(c1 * 64 + g * 32) * 4;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const float4 b0 = lds_f4(b_a + jj * 32), b1 = lds_f4(b_a + jj * 32 + 16);
          float v[8];
          v[0] = __uint_as_float(acc[jj * 8 + 0]) + b0.x; v[1] = __uint_as_float(acc[jj * 8 + 1]) + b0.y;
          v[2] = __uint_as_float(acc[jj * 8 + 2]) + b0.z; v[3] = __uint_as_float(acc[jj * 8 + 3]) + b0.w;
          v[4] = __uint_as_float(acc[jj * 8 + 4]) + b1.x; v[5] = __uint_as_float(acc[jj * 8 + 5]) + b1.y;
          v[6] = __uint_as_float(acc[jj * 8 + 6]) + b1.z; v[7] = __uint_as_float(acc[jj * 8 + 7]) + b1.w;
          if (p.tail_relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
          }
          sts_u4(out_row + uint32_t(((g * 4 + jj) ^ sw) << 4),
                 make_uint4(pack2<T>(v[0], v[1]), pack2<T>(v[2], v[3]), pack2<T>(v[4], v[5]), pack2<T>(v[6], v[7])));
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(out1_staged));
    };

    for (int i = 0; i < my_tiles; ++i) {
#pragma unroll 1
      for (int c = grp; c < kChunks; c += 2) {
        const uint32_t cc = uint32_t(i) * kChunks + uint32_t(c);
        const uint32_t slot = cc % uint32_t(SR), ab = cc % uint32_t(AB);
        const bool tr = tracer && cc < 32;
        if (tr) trace[cc * 8 + 0] = clock64();
        {
          // the three conditions are usually all true already: poll them back to back (one round trip, not three)
          const uint32_t b0 = smem_u32(&acc3_full[c]), p0 = uint32_t(i) & 1;
          const uint32_t b1 = smem_u32(&r_full[slot]), p1 = (cc / uint32_t(SR)) & 1;
          const uint32_t b2 = smem_u32(&act_free[ab]), p2 = ((cc / uint32_t(AB)) - 1) & 1;
          const bool need2 = cc >= uint32_t(AB);
          const bool ok0 = mbar_try_wait(b0, p0), ok1 = mbar_try_wait(b1, p1), ok2 = need2 ? mbar_try_wait(b2, p2) : true;
          if (!(ok0 && ok1 && ok2)) {
            mbar_wait(b0, p0);
            mbar_wait(b1, p1);
            if (need2) mbar_wait(b2, p2);
          }
        }
        tc_fence_after();
        if (tr) trace[cc * 8 + 1] = clock64();
        uint32_t acc[32];
        tmem_ld_32x32b_x32(lane_addr + uint32_t(c * 64 + g * 32), acc);
        const uint32_t raw_row = smem_u32(r_ring + slot * kChunkBytes) + row_off;
        const uint32_t act_row = smem_u32(actb + ab * kChunkBytes) + row_off;
        const uint32_t pcol = uint32_t(c * 64 + g * 32) * 4;       // byte offset of this thread's first channel in the parameter rows
        uint4 res[4];
        if (p.has_res) {
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) res[jj] = lds_u4(raw_row + uint32_t(((g * 4 + jj) ^ sw) << 4));
        }
        tmem_ld_wait();
        if (tr) trace[cc * 8 + 4] = clock64();
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const uint32_t slot16 = uint32_t(((g * 4 + jj) ^ sw) << 4);   // SWIZZLE_128B: 16-byte chunk index XOR (row mod 8)
          const float4 b0 = lds_f4(s_bias_a + pcol + jj * 32), b1 = lds_f4(s_bias_a + pcol + jj * 32 + 16);
          float v[8];
          v[0] = __uint_as_float(acc[jj * 8 + 0]) + b0.x; v[1] = __uint_as_float(acc[jj * 8 + 1]) + b0.y;
          v[2] = __uint_as_float(acc[jj * 8 + 2]) + b0.z; v[3] = __uint_as_float(acc[jj * 8 + 3]) + b0.w;
          v[4] = __uint_as_float(acc[jj * 8 + 4]) + b1.x; v[5] = __uint_as_float(acc[jj * 8 + 5]) + b1.y;
          v[6] = __uint_as_float(acc[jj * 8 + 6]) + b1.z; v[7] = __uint_as_float(acc[jj * 8 + 7]) + b1.w;
          if (p.relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
          }
          if (p.has_res) {
            const float2 r0 = unpack2<T>(res[jj].x), r1 = unpack2<T>(res[jj].y), r2 = unpack2<T>(res[jj].z), r3 = unpack2<T>(res[jj].w);
            v[0] += r0.x; v[1] += r0.y; v[2] += r1.x; v[3] += r1.y;
            v[4] += r2.x; v[5] += r2.y; v[6] += r3.x; v[7] += r3.y;
          }
          sts_u4(raw_row + slot16, make_uint4(pack2<T>(v[0], v[1]), pack2<T>(v[2], v[3]), pack2<T>(v[4], v[5]), pack2<T>(v[6], v[7])));
          const float4 s0 = lds_f4(s_scale_a + pcol + jj * 32), s1 = lds_f4(s_scale_a + pcol + jj * 32 + 16);
          const float4 t0 = lds_f4(s_shift_a + pcol + jj * 32), t1 = lds_f4(s_shift_a + pcol + jj * 32 + 16);
          float a[8];
          a[0] = fmaf(s0.x, v[0], t0.x); a[1] = fmaf(s0.y, v[1], t0.y); a[2] = fmaf(s0.z, v[2], t0.z); a[3] = fmaf(s0.w, v[3], t0.w);
          a[4] = fmaf(s1.x, v[4], t1.x); a[5] = fmaf(s1.y, v[5], t1.y); a[6] = fmaf(s1.z, v[6], t1.z); a[7] = fmaf(s1.w, v[7], t1.w);
          if (p.act_relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = fmaxf(a[k], 0.f);
          }
          sts_u4(act_row + slot16, make_uint4(pack2<T>(a[0], a[1]), pack2<T>(a[2], a[3]), pack2<T>(a[4], a[5]), pack2<T>(a[6], a[7])));
        }
        if (tr) trace[cc * 8 + 5] = clock64();
        tc_fence_before();                   // this thread's acc3 reads are ordered before the signal below
        fence_proxy_async_smem();            // generic-proxy writes -> visible to the TMA store and to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&staged[slot]));
        if (tr) trace[cc * 8 + 6] = clock64();
        if (c == grp && i > 0 && ((i - 1) & 1) == grp) finish_tail(i - 1);
        if (tr) trace[cc * 8 + 7] = clock64();
      }
    }
    if (((my_tiles - 1) & 1) == grp) finish_tail(my_tiles - 1);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}
