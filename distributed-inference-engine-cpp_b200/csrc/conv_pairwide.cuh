// Two-SM convolution with the wide kernel's epilogue: the 1x1 convs that END a bottleneck block or join it with its
// projection shortcut (residual add, raw and/or BN-ReLU'd copy written; optionally a second input concatenated along
// K) when their K is large and their weights stream.  (included by conv_umma.cu inside namespace rn::{anonymous})
//
// Why: conv_wide_kernel gives 112 KB of shared memory to its residual / activated-copy rings, which leaves TWO 48 KB
// operand stages; with K >= 512 its main loop then waits for every K block (timeline, B200RN_DBG_TRACE: ~2000 cycles
// per block, 15000 per tile against 4100 cycles of MMAs on 512 -> 2048 @7x7), and a third stage costs the rings more
// than it gains.  As a CTA pair (conv_pair.cuh) each SM stages its own A tile and HALF of the weight tile -- 32 KB a
// stage -- so three or four stages fit beside the same rings, and each byte of B feeds 256 output rows.
//
// Main loop = conv_pair_kernel's (TMA producer in both CTAs signalling the leader's "full" barrier, one MMA thread in
// the leader, tcgen05.mma.cta_group::2, multicast commits); epilogue = conv_wide_kernel's, run independently by each
// CTA on its own 128 rows (store warp, residual-ring warp, sixteen epilogue warps in two groups on alternating
// chunks).  A cluster keeps one N tile for its lifetime, so bias / scale / shift sit in shared memory.
// The TMA loads here read this GPU's memory only and are not multicast.
constexpr int kPairWideThreads = kWideThreads;

template <typename T, int N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairWideThreads, 1)
conv_pairwide_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
                     const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmRes,
                     const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmAct, const WideParams p) {
  constexpr int kChunks = N / 64;
  constexpr uint32_t A_BYTES = kTileM * 128;
  constexpr uint32_t BH_BYTES = (N / 2) * 128;              // this CTA's half of the weight tile
  // a stage holds an A tile and, unless the pair's weight slab is stationary, this CTA's half of the matching B tile
  const uint32_t STAGE_BYTES = p.b_stationary ? A_BYTES : A_BYTES + BH_BYTES;

  extern __shared__ uint8_t smem_raw[];
  // identical carve-up in both CTAs (the MMA and the multicast commits address the peer by the same offsets)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = p.stages, SR = p.r_slots, SC = p.c_slots;
  const bool use_r = p.has_res || p.has_out;
  const int SS = use_r ? SR : SC;
  uint8_t* bstat = smem;                          // [num_k_iters x half B tile] when b_stationary
  uint8_t* pipe = bstat + (p.b_stationary ? p.num_k_iters * BH_BYTES : 0);
  uint8_t* r_ring = pipe + S * STAGE_BYTES;
  uint8_t* c_ring = r_ring + (use_r ? SR : 0) * kChunkBytes;
  float* s_bias = reinterpret_cast<float*>(c_ring + (p.has_act ? SC : 0) * kChunkBytes);
  float* s_scale = s_bias + N;
  float* s_shift = s_scale + (p.has_act ? N : 0);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_shift + (p.has_act ? N : 0));
  uint64_t* full_bar = bars;                 // [S]   (used in the leader) both CTAs' operand TMA -> MMA
  uint64_t* empty_bar = full_bar + S;        // [S]   MMA -> operand TMA (multicast commit)
  uint64_t* tmem_full = empty_bar + S;       // [2]   MMA -> epilogue (multicast commit)
  uint64_t* tmem_empty = tmem_full + 2;      // [2]   (used in the leader) one arrival per epilogue warp of both CTAs
  uint64_t* r_full = tmem_empty + 2;         // [SR]  residual TMA -> epilogue
  uint64_t* r_empty = r_full + 8;            // [SR]  store warp -> residual TMA
  uint64_t* c_free = r_empty + 8;            // [SC]  store warp -> epilogue
  uint64_t* staged = c_free + 8;             // [SS]  epilogue -> store warp (8 warp arrivals: the owning group)
  uint64_t* bstat_bar = staged + 8;          // [1]   (used in the leader) both CTAs' stationary weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bstat_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = int(blockIdx.x) >> 1, n_clusters = int(gridDim.x) >> 1;   // n_clusters is a multiple of n_tiles
  const int n_tile = cluster_id % p.n_tiles;
  const int col0 = n_tile * N;
  const int pair_tiles = (p.m_tiles + 1) >> 1;
  const int pm_first = cluster_id / p.n_tiles, pm_step = n_clusters / p.n_tiles;
  const int my_tiles = pm_first < pair_tiles ? (pair_tiles - pm_first + pm_step - 1) / pm_step : 0;
  // this CTA's M tile of pair tile `pm`; an index past the last tile lands beyond the batch (TMA loads read zeros
  // there and TMA stores are clipped)
  auto my_tile = [&](int i) { return wide_decode(p, 2 * (pm_first + i * pm_step) + int(rank)); };

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmW);
    if (p.k1_iters < p.num_k_iters) prefetch_tmap(&tmA2);
    if (p.has_res) prefetch_tmap(&tmRes);
    if (p.has_out) prefetch_tmap(&tmOut);
    if (p.has_act) prefetch_tmap(&tmAct);
    for (int i = 0; i < S; ++i) { mbar_init(smem_u32(&full_bar[i]), 1); mbar_init(smem_u32(&empty_bar[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&tmem_full[i]), 1); mbar_init(smem_u32(&tmem_empty[i]), 2 * kWideEpiWarps); }
    for (int i = 0; i < 8; ++i) {
      mbar_init(smem_u32(&r_full[i]), 1); mbar_init(smem_u32(&r_empty[i]), 1);
      mbar_init(smem_u32(&c_free[i]), 1); mbar_init(smem_u32(&staged[i]), kWideGroupWarps);
    }
    mbar_init(smem_u32(bstat_bar), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair(smem_u32(tmem_slot), 2 * N);
  for (int k = threadIdx.x; k < N; k += blockDim.x) {
    s_bias[k] = __ldg(p.bias + col0 + k);
    if (p.has_act) {
      s_scale[k] = __ldg(p.act_scale + col0 + k);
      s_shift[k] = __ldg(p.act_shift + col0 + k);
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // the peer's barriers exist before anything is signalled across the pair
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  if (warp == 0) {
    // ===== operand TMA producer (both CTAs) =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const int bcol = col0 + int(rank) * (N / 2);
      if (p.b_stationary && my_tiles > 0) {      // the pair's N tile never changes: each CTA fetches its half of the slab once
        const uint32_t bb_local = smem_u32(bstat_bar);
        const uint32_t bb = mapa_cluster(bb_local, 0);
        if (rank == 0) mbar_expect_tx(bb_local, 2u * uint32_t(p.num_k_iters) * BH_BYTES);
        for (int it = 0; it < p.num_k_iters; ++it) tma_load_2d_pair(smem_u32(bstat + it * BH_BYTES), &tmW, bb, it * 64, bcol);
      }
      for (int i = 0; i < my_tiles; ++i) {
        const WideTile tc = my_tile(i);
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          const uint32_t fb_local = smem_u32(&full_bar[stage]);
          const uint32_t fb = mapa_cluster(fb_local, 0);
          if (rank == 0) mbar_expect_tx(fb_local, p.b_stationary ? 2 * p.a_bytes : 2 * (p.a_bytes + BH_BYTES));
          const uint32_t dst = smem_u32(pipe + stage * STAGE_BYTES);
          if (it < p.k1_iters) tma_load_4d_pair(dst, &tmA, fb, it * 64, tc.w0 * p.stride, tc.h0 * p.stride, tc.n0);
          else tma_load_4d_pair(dst, &tmA2, fb, (it - p.k1_iters) * 64, tc.w0 * p.stride_b, tc.h0 * p.stride_b, tc.n0);
          if (!p.b_stationary) tma_load_2d_pair(dst + A_BYTES, &tmW, fb, it * 64, bcol);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
      // tail: every commit aimed at this CTA's "empty" barriers has landed before the CTA may exit
      const int total = my_tiles * p.num_k_iters;
      for (int k = 0; k < (total < S ? total : S); ++k) {
        mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one thread of the leader CTA =====
    constexpr uint32_t idesc = make_idesc_pair<T, N>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (rank == 0 && elect_one()) {
      const uint32_t pipe_lo = smem_desc_lo(smem_u32(pipe));
      const uint32_t bstat_lo = smem_desc_lo(smem_u32(bstat));
      int stage = 0;
      uint32_t phase = 0;
      if (p.b_stationary && my_tiles > 0) mbar_wait(smem_u32(bstat_bar), 0);
      for (int i = 0; i < my_tiles; ++i) {
        const int buf = i & 1;
        mbar_wait(smem_u32(&tmem_empty[buf]), ((uint32_t(i) >> 1) & 1) ^ 1);   // both CTAs' epilogues drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + uint32_t(buf * N);
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t a_lo = pipe_lo + uint32_t(stage) * (STAGE_BYTES >> 4);
          const uint32_t b_lo = p.b_stationary ? bstat_lo + uint32_t(it) * (BH_BYTES >> 4) : a_lo + (A_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_f16_lo_pair(tmem_d, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (it | k) != 0 ? 1u : 0u);
          umma_commit_pair(smem_u32(&empty_bar[stage]));
          if (it == p.num_k_iters - 1) umma_commit_pair(smem_u32(&tmem_full[buf]));
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 2) {
    // ===== store warp (one thread): TMA-stores each staged chunk and hands its slots back once the store has read them =====
    if (elect_one()) {
      uint32_t cc = 0;
      for (int i = 0; i < my_tiles; ++i) {
        const WideTile tc = my_tile(i);
        for (int c = 0; c < kChunks; ++c, ++cc) {
          const uint32_t slot = use_r ? cc % uint32_t(SR) : 0u, cslot = p.has_act ? cc % uint32_t(SC) : 0u;
          mbar_wait(smem_u32(&staged[cc % uint32_t(SS)]), (cc / uint32_t(SS)) & 1);
          if (p.has_out) tma_store_4d(&tmOut, smem_u32(r_ring + slot * kChunkBytes), col0 + c * 64, tc.w0, tc.h0, tc.n0);
          if (p.has_act) tma_store_4d(&tmAct, smem_u32(c_ring + cslot * kChunkBytes), col0 + c * 64, tc.w0, tc.h0, tc.n0);
          tma_store_commit();
          tma_store_wait_read();
          if (use_r) mbar_arrive(smem_u32(&r_empty[slot]));
          if (p.has_act) mbar_arrive(smem_u32(&c_free[cslot]));
        }
      }
    }
  } else if (warp == 3) {
    // ===== residual TMA producer (one thread): keeps the raw / residual ring full, a whole tile ahead of the epilogue =====
    if (use_r && elect_one()) {
      uint32_t cc = 0;
      // Experiment knob (B200RN_WIDE_PF, default off): pull the residual of a later tile into L2 ahead of the ring's
      // own load.  Measured at batch 256 it makes every residual layer 3-20 % slower, so it stays off.
      auto prefetch_tile = [&](int j) {
        if (!p.has_res || j >= my_tiles || p.pf_dist == 0) return;
        const WideTile t = my_tile(j);
        for (int c = 0; c < kChunks; ++c) tma_prefetch_4d(&tmRes, col0 + c * 64, t.w0, t.h0, t.n0);
      };
      for (int j = 0; j < p.pf_dist; ++j) prefetch_tile(j);
      for (int i = 0; i < my_tiles; ++i) {
        const WideTile tc = my_tile(i);
        prefetch_tile(i + p.pf_dist);
        for (int c = 0; c < kChunks; ++c, ++cc) {
          const uint32_t slot = cc % uint32_t(SR);
          mbar_wait(smem_u32(&r_empty[slot]), ((cc / uint32_t(SR)) & 1) ^ 1);
          const uint32_t bar = smem_u32(&r_full[slot]);
          if (p.has_res) {
            mbar_expect_tx(bar, p.rows * 128u);
            tma_load_4d(smem_u32(r_ring + slot * kChunkBytes), &tmRes, bar, col0 + c * 64, tc.w0, tc.h0, tc.n0);
          } else {
            mbar_arrive(bar);                  // no residual: the slot is only the raw output's staging buffer
          }
        }
      }
    }
  } else {
    // ===== epilogue: two groups of 8 warps on alternating chunks; thread = one pixel (TMEM lane) x 32 channels =====
    const int ew = warp - 4;
    const int grp = ew >> 3;
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    const int g = (ew & 7) >> 2;            // which half of every 64-channel chunk
    const int row = q * 32 + lane;
    const int sw = row & 7;
    const uint32_t lane_addr = tmem_base + (uint32_t(q * 32) << 16);
    const uint32_t s_bias_a = smem_u32(s_bias), s_scale_a = smem_u32(s_scale), s_shift_a = smem_u32(s_shift);
    const uint32_t row_off = uint32_t(row) * 128u;
    const uint32_t tmem_empty_leader = mapa_cluster(smem_u32(&tmem_empty[0]), 0);

    for (int i = 0; i < my_tiles; ++i) {
      const int buf = i & 1;
      mbar_wait(smem_u32(&tmem_full[buf]), (uint32_t(i) >> 1) & 1);
#pragma unroll 1
      for (int c = grp; c < kChunks; c += 2) {
        const uint32_t cc = uint32_t(i) * kChunks + uint32_t(c);
        const uint32_t slot = use_r ? cc % uint32_t(SR) : 0u, cslot = p.has_act ? cc % uint32_t(SC) : 0u;
        {
          // usually both are already true: poll back to back (one round trip instead of two)
          const uint32_t b1 = smem_u32(&r_full[slot]), p1 = (cc / uint32_t(use_r ? SR : 1)) & 1;
          const uint32_t b2 = smem_u32(&c_free[cslot]), p2 = ((cc / uint32_t(p.has_act ? SC : 1)) - 1) & 1;
          const bool need2 = p.has_act && cc >= uint32_t(SC);
          const bool ok1 = use_r ? mbar_try_wait(b1, p1) : true, ok2 = need2 ? mbar_try_wait(b2, p2) : true;
          if (!(ok1 && ok2)) {
            if (use_r) mbar_wait(b1, p1);
            if (need2) mbar_wait(b2, p2);
          }
        }
        tc_fence_after();
        uint32_t acc[32];
        tmem_ld_32x32b_x32(lane_addr + uint32_t(buf * N + c * 64 + g * 32), acc);
        const uint32_t raw_row = smem_u32(r_ring + slot * kChunkBytes) + row_off;
        const uint32_t act_row = smem_u32(c_ring + cslot * kChunkBytes) + row_off;
        const uint32_t pcol = uint32_t(c * 64 + g * 32) * 4;       // byte offset of this thread's first channel in the parameter rows
        uint4 res[4];
        if (p.has_res) {
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) res[jj] = lds_u4(raw_row + uint32_t(((g * 4 + jj) ^ sw) << 4));
        }
        tmem_ld_wait();
        if (c + 2 >= kChunks) {             // this warp's last chunk of the tile: its part of the accumulator is drained
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(tmem_empty_leader + uint32_t(buf) * 8u);
        }
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
          if (p.has_out)
            sts_u4(raw_row + slot16, make_uint4(pack2<T>(v[0], v[1]), pack2<T>(v[2], v[3]), pack2<T>(v[4], v[5]), pack2<T>(v[6], v[7])));
          if (p.has_act) {
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
        }
        fence_proxy_async_smem();            // generic-proxy writes -> visible to the TMA stores
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&staged[cc % uint32_t(SS)]));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // neither CTA leaves (or frees tensor memory) while the other may still signal it
  if (warp == 1) tmem_dealloc_pair(tmem_base, 2 * N);
}
