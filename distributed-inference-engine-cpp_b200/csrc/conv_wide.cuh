// Convolution with a wide output (256-channel N tile) whose cost is the epilogue, not the MMAs: the 1x1 convs
// that end a bottleneck block (Cin -> 4*Cin channels + residual, raw and/or BN-ReLU'd copy written) and the
// projection shortcuts.  (included by conv_umma.cu inside namespace rn::{anonymous})
//
// Same implicit GEMM as conv_umma_kernel (TMA operand ring, one thread issuing tcgen05.mma into a double-buffered
// TMEM accumulator), but the epilogue is organised like conv_tail_kernel's, because measured on these layers the
// generic epilogue -- eight warps walking the four 64-channel chunks of a tile in lock step, bar.sync around every
// TMA store, whole-tile staging buffers -- is a ~2400-cycle serial latency chain per chunk and the layer runs at
// that pace, not at HBM's:
//   * sixteen epilogue warps in two groups; group 0 takes chunks 0 and 2 of every tile, group 1 chunks 1 and 3,
//     so two chains are in flight per SM;
//   * no CTA-wide barrier: every staging buffer has its own mbarrier pair;
//   * staging is a ring of 16 KB chunk slots instead of whole tiles (the residual chunk lands in a slot, the sum is
//     written in place and TMA-stored from there; the activated copy has its own, shorter ring), which leaves room
//     for the 256-wide N tile the generic kernel had to give up on these layers;
//   * TMA stores are issued by a warp of their own, and the residual ring is fed by another, so neither the
//     operand producer nor the epilogue ever waits for a store to be read.
// A CTA keeps one N tile for its lifetime (grid is a multiple of the N tile count), so bias / scale / shift sit in
// shared memory from the start.
constexpr int kWideN = 256;
constexpr int kWideEpiWarps = 16;
constexpr int kWideGroupWarps = 8;
constexpr int kWideThreads = (4 + kWideEpiWarps) * 32;   // operand TMA, MMA, store, residual TMA, epilogue warps

struct WideParams {
  int num_k_iters, cblks, kw, stride, pad;
  int k1_iters, stride_b;      // two-input conv: K blocks [0, k1_iters) come from tmA, the rest from tmA2 at stride_b
  int bw, bh, bn, tiles_w, tiles_h, m_tiles, n_tiles;
  int stages, r_slots, c_slots, b_stationary;
  int has_res, has_out, has_act, relu, act_relu;
  int pf_dist;                 // residual L2 prefetch distance in tiles (0 = off)
  uint32_t rows, a_bytes;
  const float* bias;
  const float* act_scale;
  const float* act_shift;
  unsigned long long* trace;   // timeline experiment (B200RN_DBG_TRACE): CTA 0; [0..31] MMA thread per tile, [32..] epilogue groups per chunk
  int dbg_skip;                // timing experiments (B200RN_WIDE_DBG; conv_wide_kernel only): 1 no MMAs, 2 no epilogue arithmetic / staging
                               // writes, 4 no TMA stores, 8 no residual loads, 16 no operand loads, 32 no TMEM read
};

struct WideTile { int w0, h0, n0; };
__device__ __forceinline__ WideTile wide_decode(const WideParams& p, int mt) {
  WideTile t;
  const int tw = mt % p.tiles_w; mt /= p.tiles_w;
  const int th = mt % p.tiles_h;
  t.w0 = tw * p.bw; t.h0 = th * p.bh; t.n0 = (mt / p.tiles_h) * p.bn;
  return t;
}

template <typename T, int N>
__global__ void __launch_bounds__(kWideThreads, 1)
conv_wide_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
                 const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmRes,
                 const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmAct, const WideParams p) {
  constexpr int kChunks = N / 64;
  constexpr uint32_t A_BYTES = kTileM * 128;
  constexpr uint32_t B_BYTES = N * 128;
  // a stage holds an A tile and, unless the CTA's weight slab is stationary, the matching B tile
  const uint32_t STAGE_BYTES = p.b_stationary ? A_BYTES : A_BYTES + B_BYTES;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = p.stages, SR = p.r_slots, SC = p.c_slots;
  const bool use_r = p.has_res || p.has_out;      // the raw / residual ring exists
  const int SS = use_r ? SR : SC;                 // "chunk staged" barriers
  uint8_t* bstat = smem;                          // [num_k_iters x B tile] when b_stationary
  uint8_t* pipe = bstat + (p.b_stationary ? p.num_k_iters * B_BYTES : 0);
  uint8_t* r_ring = pipe + S * STAGE_BYTES;
  uint8_t* c_ring = r_ring + (use_r ? SR : 0) * kChunkBytes;
  float* s_bias = reinterpret_cast<float*>(c_ring + (p.has_act ? SC : 0) * kChunkBytes);
  float* s_scale = s_bias + N;                             // scale / shift rows exist only with an activated output
  float* s_shift = s_scale + (p.has_act ? N : 0);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_shift + (p.has_act ? N : 0));
  uint64_t* full_bar = bars;                 // [S]   operand TMA -> MMA
  uint64_t* empty_bar = full_bar + S;        // [S]   MMA -> operand TMA
  uint64_t* tmem_full = empty_bar + S;       // [2]   MMA -> epilogue
  uint64_t* tmem_empty = tmem_full + 2;      // [2]   epilogue -> MMA (16 warp arrivals)
  uint64_t* r_full = tmem_empty + 2;         // [SR]  residual TMA -> epilogue (chunk landed / slot handed over)
  uint64_t* r_empty = r_full + 8;            // [SR]  store warp -> residual TMA
  uint64_t* c_free = r_empty + 8;            // [SC]  store warp -> epilogue (activated-copy slot reusable)
  uint64_t* staged = c_free + 8;             // [SS]  epilogue -> store warp (8 warp arrivals: the owning group)
  uint64_t* bstat_bar = staged + 8;          // [1]   stationary weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bstat_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = int(blockIdx.x) % p.n_tiles;
  const int col0 = n_tile * N;
  const int m_first = int(blockIdx.x) / p.n_tiles, m_step = int(gridDim.x) / p.n_tiles;
  const int my_tiles = m_first < p.m_tiles ? (p.m_tiles - m_first + m_step - 1) / m_step : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmW);
    if (p.k1_iters < p.num_k_iters) prefetch_tmap(&tmA2);
    if (p.has_res) prefetch_tmap(&tmRes);
    if (p.has_out) prefetch_tmap(&tmOut);
    if (p.has_act) prefetch_tmap(&tmAct);
    for (int i = 0; i < S; ++i) { mbar_init(smem_u32(&full_bar[i]), 1); mbar_init(smem_u32(&empty_bar[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&tmem_full[i]), 1); mbar_init(smem_u32(&tmem_empty[i]), kWideEpiWarps); }
    for (int i = 0; i < 8; ++i) {
      mbar_init(smem_u32(&r_full[i]), 1); mbar_init(smem_u32(&r_empty[i]), 1);
      mbar_init(smem_u32(&c_free[i]), 1); mbar_init(smem_u32(&staged[i]), kWideGroupWarps);
    }
    mbar_init(smem_u32(bstat_bar), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 2 * N);
  for (int k = threadIdx.x; k < N; k += blockDim.x) {
    s_bias[k] = __ldg(p.bias + col0 + k);
    if (p.has_act) {
      s_scale[k] = __ldg(p.act_scale + col0 + k);
      s_shift[k] = __ldg(p.act_shift + col0 + k);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  if (warp == 0) {
    // ===== operand TMA producer =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      if (p.b_stationary && my_tiles > 0) {      // the CTA's N tile never changes: fetch its weights once
        const uint32_t bb = smem_u32(bstat_bar);
        mbar_expect_tx(bb, uint32_t(p.num_k_iters) * B_BYTES);
        for (int it = 0; it < p.num_k_iters; ++it) tma_load_2d(smem_u32(bstat + it * B_BYTES), &tmW, bb, it * 64, col0);
      }
      for (int i = 0; i < my_tiles; ++i) {
        const WideTile tc = wide_decode(p, m_first + i * m_step);
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          if (p.dbg_skip & 16) { mbar_arrive(fb); if (++stage == S) { stage = 0; phase ^= 1; } continue; }
          mbar_expect_tx(fb, p.b_stationary ? p.a_bytes : p.a_bytes + B_BYTES);
          const uint32_t a_dst = smem_u32(pipe + stage * STAGE_BYTES);
          if (it < p.k1_iters) {
            const int tap = it / p.cblks, cb = it - tap * p.cblks;
            const int ky = tap / p.kw, kx = tap - ky * p.kw;
            tma_load_4d(a_dst, &tmA, fb, cb * 64, tc.w0 * p.stride - p.pad + kx, tc.h0 * p.stride - p.pad + ky, tc.n0);
          } else {                         // second input of a two-input 1x1 conv
            tma_load_4d(a_dst, &tmA2, fb, (it - p.k1_iters) * 64, tc.w0 * p.stride_b, tc.h0 * p.stride_b, tc.n0);
          }
          if (!p.b_stationary) tma_load_2d(a_dst + A_BYTES, &tmW, fb, it * 64, col0);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    constexpr uint32_t idesc = make_idesc<T, N>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t pipe_lo = smem_desc_lo(smem_u32(pipe));
      const uint32_t bstat_lo = smem_desc_lo(smem_u32(bstat));
      if (p.b_stationary && my_tiles > 0) mbar_wait(smem_u32(bstat_bar), 0);
      for (int i = 0; i < my_tiles; ++i) {
        const int buf = i & 1;
        const bool tr = p.trace && blockIdx.x == 0 && i < 32;
        if (tr) p.trace[i * 8 + 0] = clock64();
        mbar_wait(smem_u32(&tmem_empty[buf]), ((uint32_t(i) >> 1) & 1) ^ 1);   // epilogue drained this accumulator
        tc_fence_after();
        if (tr) p.trace[i * 8 + 1] = clock64();
        const uint32_t tmem_d = tmem_base + uint32_t(buf * N);
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          if (tr && it < 5) p.trace[i * 8 + 2 + it] = clock64();
          const uint32_t a_lo = pipe_lo + uint32_t(stage) * (STAGE_BYTES >> 4);
          const uint32_t b_lo = p.b_stationary ? bstat_lo + uint32_t(it) * (B_BYTES >> 4) : a_lo + (A_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (!(p.dbg_skip & 1)) umma_f16_lo(tmem_d, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (it | k) != 0 ? 1u : 0u);
          umma_commit(smem_u32(&empty_bar[stage]));
          if (it + 1 == p.num_k_iters) umma_commit(smem_u32(&tmem_full[buf]));
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 2) {
    // ===== store warp (one thread): TMA-stores each staged chunk and hands its slots back once the store has read them =====
    if (elect_one()) {
      uint32_t cc = 0;
      for (int i = 0; i < my_tiles; ++i) {
        const WideTile tc = wide_decode(p, m_first + i * m_step);
        for (int c = 0; c < kChunks; ++c, ++cc) {
          const uint32_t slot = use_r ? cc % uint32_t(SR) : 0u, cslot = p.has_act ? cc % uint32_t(SC) : 0u;
          mbar_wait(smem_u32(&staged[cc % uint32_t(SS)]), (cc / uint32_t(SS)) & 1);
          if (p.has_out && !(p.dbg_skip & 4)) tma_store_4d(&tmOut, smem_u32(r_ring + slot * kChunkBytes), col0 + c * 64, tc.w0, tc.h0, tc.n0);
          if (p.has_act && !(p.dbg_skip & 4)) tma_store_4d(&tmAct, smem_u32(c_ring + cslot * kChunkBytes), col0 + c * 64, tc.w0, tc.h0, tc.n0);
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
        const WideTile t = wide_decode(p, m_first + j * m_step);
        for (int c = 0; c < kChunks; ++c) tma_prefetch_4d(&tmRes, col0 + c * 64, t.w0, t.h0, t.n0);
      };
      for (int j = 0; j < p.pf_dist; ++j) prefetch_tile(j);
      for (int i = 0; i < my_tiles; ++i) {
        const WideTile tc = wide_decode(p, m_first + i * m_step);
        prefetch_tile(i + p.pf_dist);
        for (int c = 0; c < kChunks; ++c, ++cc) {
          const uint32_t slot = cc % uint32_t(SR);
          mbar_wait(smem_u32(&r_empty[slot]), ((cc / uint32_t(SR)) & 1) ^ 1);
          const uint32_t bar = smem_u32(&r_full[slot]);
          if (p.has_res && !(p.dbg_skip & 8)) {
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

    const bool tracer = p.trace && blockIdx.x == 0 && lane == 0 && (ew & 7) == 0;
    for (int i = 0; i < my_tiles; ++i) {
      const int buf = i & 1;
      if (tracer && i < 16) p.trace[(32 + i * 4 + grp) * 8 + 7] = clock64();
      mbar_wait(smem_u32(&tmem_full[buf]), (uint32_t(i) >> 1) & 1);
#pragma unroll 1
      for (int c = grp; c < kChunks; c += 2) {
        const uint32_t cc = uint32_t(i) * kChunks + uint32_t(c);
        unsigned long long* trc = p.trace + (32 + cc) * 8;
        const bool tr = tracer && cc < 64;
        if (tr) trc[0] = clock64();
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
        if (tr) trc[1] = clock64();
        uint32_t acc[32];
        if (!(p.dbg_skip & 32)) tmem_ld_32x32b_x32(lane_addr + uint32_t(buf * N + c * 64 + g * 32), acc);
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
          if (lane == 0) mbar_arrive(smem_u32(&tmem_empty[buf]));
        }
#pragma unroll
        for (int jj = 0; jj < 4 && !(p.dbg_skip & 2); ++jj) {
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
        if (tr) trc[2] = clock64();
        fence_proxy_async_smem();            // generic-proxy writes -> visible to the TMA stores
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&staged[cc % uint32_t(SS)]));
        if (tr) trc[3] = clock64();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 2 * N);
}
