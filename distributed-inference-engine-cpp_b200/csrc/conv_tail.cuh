// 1x1 conv + residual add, and the NEXT bottleneck block's first 1x1 conv, in one kernel.
// (included by conv_umma.cu inside namespace rn::{anonymous}, after the PTX wrappers)
//
// In a pre-activation ResNet the block output T' = conv3(U2) + T is written twice today: raw (the next
// block's shortcut) and as relu(bn(T')) (the next block's conv1 operand), and that second copy is read
// straight back by a 1x1 conv.  On the 56x56 stage these two layers are pure HBM traffic (T' is 4x wider
// than anything else in the block).  Here the activated copy never leaves the SM:
//
//   per 128-pixel tile:  acc3[128 x 256] = U2[128 x Cin] * W3^T                 (tcgen05, TMEM cols 0..255)
//     per 64-channel chunk c:  v = acc3 + bias3 (+ReLU) + T     -> raw chunk, TMA-stored (in place in the
//                                                                  residual's staging slot)
//                              a = relu?(s*v + t) as 16-bit     -> shared memory, swizzle-128B, which is exactly
//                                                                  the K-major A operand layout of K block c
//                              acc1[128 x N1] += a * W1[:, c]^T  (tcgen05, TMEM cols 256..256+N1)
//     u = relu?(acc1 + bias1)                                   -> TMA-stored: the next block's conv1 output
//
// HBM per tile: reads Cin*256 B + 64 KB (T), writes 64 KB (T') + N1*256 B; the unfused pair moves 128 KB more.
// W3 and W1 stay in shared memory for the CTA's lifetime.  The residual / raw staging is a ring of 16 KB
// chunk slots (load -> add in place -> store -> refill), the activated chunks alternate between act_bufs
// buffers that the MMA warp hands back with tcgen05.commit.  Neither accumulator is double-buffered: the
// epilogue's own order (drain acc1 of tile i-1, then acc3 of tile i chunk by chunk) already guarantees each
// is empty before the MMA warp -- which is gated on the "chunk ready" barriers -- overwrites it.
constexpr int kTailN3 = 256;

struct TailParams {
  int cblks, bw, bh, bn, tiles_w, tiles_h, total_tiles;
  int a_stages, r_slots, act_bufs;
  int relu, act_relu, tail_relu;
  uint32_t rows;
  const float* bias;
  const float* act_scale;
  const float* act_shift;
  const float* tail_bias;
  unsigned long long* trace;   // timeline experiment (B200RN_DBG_TRACE): CTA 0's store thread stamps clock64 per chunk
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
__global__ void __launch_bounds__(kBaseThreads, 1)
conv_tail_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmRes,
                 const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmOut1, const TailParams p) {
  constexpr int kChunks = kTailN3 / 64;                 // 64-channel chunks of the wide tensor = K blocks of the tail GEMM
  constexpr int kChunks1 = N1 / 64;
  constexpr uint32_t W3_BLOCK = kTailN3 * 128;          // one K block of W3 (256 rows x 128 B)
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
  uint64_t* r_empty = r_full + SR;          // [SR]  epilogue -> TMA (the chunk's store has read the slot)
  uint64_t* act_ready = r_empty + SR;       // [AB]  epilogue -> MMA (activated chunk written)
  uint64_t* act_free = act_ready + AB;      // [AB]  MMA -> epilogue (tail MMAs on the chunk done)
  uint64_t* acc3_full = act_free + AB;      // [1]
  uint64_t* acc1_full = acc3_full + 1;      // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc1_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmW); prefetch_tmap(&tmW1);
    prefetch_tmap(&tmRes); prefetch_tmap(&tmOut); prefetch_tmap(&tmOut1);
    mbar_init(smem_u32(w_bar), 1);
    for (int i = 0; i < SA; ++i) { mbar_init(smem_u32(&a_full[i]), 1); mbar_init(smem_u32(&a_empty[i]), 1); }
    for (int i = 0; i < SR; ++i) { mbar_init(smem_u32(&r_full[i]), 1); mbar_init(smem_u32(&r_empty[i]), 1); }
    for (int i = 0; i < AB; ++i) { mbar_init(smem_u32(&act_ready[i]), 1); mbar_init(smem_u32(&act_free[i]), 1); }
    mbar_init(smem_u32(acc3_full), 1);
    mbar_init(smem_u32(acc1_full), 1);
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

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one() && int(blockIdx.x) < p.total_tiles) {
      const uint32_t wb = smem_u32(w_bar);
      mbar_expect_tx(wb, uint32_t(p.cblks) * W3_BLOCK + kChunks * W1_BLOCK);
      for (int kb = 0; kb < p.cblks; ++kb) tma_load_2d(smem_u32(w3 + kb * W3_BLOCK), &tmW, wb, kb * 64, 0);
      for (int c = 0; c < kChunks; ++c) tma_load_2d(smem_u32(w1 + c * W1_BLOCK), &tmW1, wb, c * 64, 0);
      int stage = 0;
      uint32_t phase = 0, cc = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const TailTile tc = tail_decode(p, tile);
        mbar_wait(smem_u32(&a_empty[stage]), phase ^ 1);
        const uint32_t fb = smem_u32(&a_full[stage]);
        mbar_expect_tx(fb, uint32_t(p.cblks) * p.rows * 128u);
        for (int kb = 0; kb < p.cblks; ++kb)
          tma_load_4d(smem_u32(a_ring + (stage * p.cblks + kb) * A_BLOCK), &tmA, fb, kb * 64, tc.w0, tc.h0, tc.n0);
        if (++stage == SA) { stage = 0; phase ^= 1; }
        for (int c = 0; c < kChunks; ++c, ++cc) {
          const uint32_t slot = cc % uint32_t(SR);
          mbar_wait(smem_u32(&r_empty[slot]), ((cc / uint32_t(SR)) & 1) ^ 1);
          const uint32_t bar = smem_u32(&r_full[slot]);
          mbar_expect_tx(bar, p.rows * 128u);
          tma_load_4d(smem_u32(r_ring + slot * kChunkBytes), &tmRes, bar, c * 64, tc.w0, tc.h0, tc.n0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread): conv3 of tile i, then the tail GEMM chunk by chunk as the epilogue feeds it =====
    constexpr uint32_t idesc3 = make_idesc<T, kTailN3>();
    constexpr uint32_t idesc1 = make_idesc<T, N1>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (elect_one() && int(blockIdx.x) < p.total_tiles) {
      const uint32_t w3_lo = smem_desc_lo(smem_u32(w3)), w1_lo = smem_desc_lo(smem_u32(w1));
      const uint32_t ring_lo = smem_desc_lo(smem_u32(a_ring)), act_lo = smem_desc_lo(smem_u32(actb));
      mbar_wait(smem_u32(w_bar), 0);
      int stage = 0;
      uint32_t phase = 0, cc = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        // acc3 is free: for tile > first, "chunk 3 ready" of the previous tile was waited on below, and the
        // epilogue threads had all read acc3 before signalling it
        mbar_wait(smem_u32(&a_full[stage]), phase);
        tc_fence_after();
        const uint32_t a_lo = ring_lo + uint32_t(stage * p.cblks) * (A_BLOCK >> 4);
        for (int kb = 0; kb < p.cblks; ++kb) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_lo(tmem_base, a_lo + uint32_t(kb) * (A_BLOCK >> 4) + 2 * k, w3_lo + uint32_t(kb) * (W3_BLOCK >> 4) + 2 * k,
                        desc_hi, idesc3, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&a_empty[stage]));
        umma_commit(smem_u32(acc3_full));
        if (++stage == SA) { stage = 0; phase ^= 1; }
        for (int c = 0; c < kChunks; ++c, ++cc) {
          const uint32_t ab = cc % uint32_t(AB);
          mbar_wait(smem_u32(&act_ready[ab]), (cc / uint32_t(AB)) & 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_lo(tmem_base + kAcc1Col, act_lo + ab * (kChunkBytes >> 4) + 2 * k, w1_lo + uint32_t(c) * (W1_BLOCK >> 4) + 2 * k,
                        desc_hi, idesc1, (c | k) != 0 ? 1u : 0u);
          umma_commit(smem_u32(&act_free[ab]));
          if (c == kChunks - 1) umma_commit(smem_u32(acc1_full));
        }
      }
    }
  } else {
    // ===== epilogue: 8 warps; thread = one pixel (TMEM lane) x 32 of a chunk's 64 channels =====
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    const int g = (warp - 2) >> 2;          // which half of every 64-channel chunk
    const int row = q * 32 + lane;
    const int sw = row & 7;
    const bool store_thread = (warp == 2 && lane == 0);
    const uint32_t lane_addr = tmem_base + (uint32_t(q * 32) << 16);
    int prev_slot = -1;                     // store thread: ring slot whose store is the newest bulk group
    uint32_t cc = 0;
    int i = 0;
    TailTile prev = {0, 0, 0};

    // the tail conv's output tile: acc1 -> +bias, ReLU -> staging -> TMA store
    auto finish_tail = [&](const TailTile& tc, int j) {
      mbar_wait(smem_u32(acc1_full), uint32_t(j) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c1 = 0; c1 < kChunks1; ++c1) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(lane_addr + kAcc1Col + uint32_t(c1 * 64 + g * 32), acc);
        tmem_ld_wait();
        uint4* out_row = reinterpret_cast<uint4*>(out1 + c1 * kChunkBytes + row * 128);
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int j8 = g * 4 + jj;
          const float4 b0 = *reinterpret_cast<const float4*>(s_bias1 + c1 * 64 + j8 * 8);
          const float4 b1 = *reinterpret_cast<const float4*>(s_bias1 + c1 * 64 + j8 * 8 + 4);
          float v[8];
          v[0] = __uint_as_float(acc[jj * 8 + 0]) + b0.x; v[1] = __uint_as_float(acc[jj * 8 + 1]) + b0.y;
          v[2] = __uint_as_float(acc[jj * 8 + 2]) + b0.z; v[3] = __uint_as_float(acc[jj * 8 + 3]) + b0.w;
          v[4] = __uint_as_float(acc[jj * 8 + 4]) + b1.x; v[5] = __uint_as_float(acc[jj * 8 + 5]) + b1.y;
          v[6] = __uint_as_float(acc[jj * 8 + 6]) + b1.z; v[7] = __uint_as_float(acc[jj * 8 + 7]) + b1.w;
          if (p.tail_relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
          }
          out_row[j8 ^ sw] = make_uint4(pack2<T>(v[0], v[1]), pack2<T>(v[2], v[3]), pack2<T>(v[4], v[5]), pack2<T>(v[6], v[7]));
        }
      }
      tc_fence_before();                     // acc1 is read: ordered before the "chunk ready" that lets the MMA warp overwrite it
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (store_thread) {
#pragma unroll
        for (int c1 = 0; c1 < kChunks1; ++c1) tma_store_4d(&tmOut1, smem_u32(out1 + c1 * kChunkBytes), c1 * 64, tc.w0, tc.h0, tc.n0);
        tma_store_commit();
        tma_store_wait_read_pending<1>();    // every group but this one has been read
        if (prev_slot >= 0) mbar_arrive(smem_u32(&r_empty[prev_slot]));
        prev_slot = -1;
      }
    };

    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++i) {
      const TailTile tc = tail_decode(p, tile);
      if (i > 0) finish_tail(prev, i - 1);
      prev = tc;
      mbar_wait(smem_u32(acc3_full), uint32_t(i) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c, ++cc) {
        const uint32_t slot = cc % uint32_t(SR), ab = cc % uint32_t(AB);
        const bool tr = p.trace && blockIdx.x == 0 && store_thread && cc < 64;
        if (tr) p.trace[cc * 8 + 0] = clock64();
        uint32_t acc[32];
        tmem_ld_32x32b_x32(lane_addr + uint32_t(c * 64 + g * 32), acc);
        mbar_wait(smem_u32(&r_full[slot]), (cc / uint32_t(SR)) & 1);
        if (cc >= uint32_t(AB)) mbar_wait(smem_u32(&act_free[ab]), ((cc / uint32_t(AB)) - 1) & 1);
        tmem_ld_wait();
        if (tr) p.trace[cc * 8 + 1] = clock64();
        uint4* raw_row = reinterpret_cast<uint4*>(r_ring + slot * kChunkBytes + row * 128);
        uint4* act_row = reinterpret_cast<uint4*>(actb + ab * kChunkBytes + row * 128);
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int j8 = g * 4 + jj;
          const int col = c * 64 + j8 * 8;
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
          const int slot16 = j8 ^ sw;          // SWIZZLE_128B: 16-byte chunk index XOR (row mod 8)
          const uint4 r = raw_row[slot16];
          const float2 r0 = unpack2<T>(r.x), r1 = unpack2<T>(r.y), r2 = unpack2<T>(r.z), r3 = unpack2<T>(r.w);
          v[0] += r0.x; v[1] += r0.y; v[2] += r1.x; v[3] += r1.y;
          v[4] += r2.x; v[5] += r2.y; v[6] += r3.x; v[7] += r3.y;
          raw_row[slot16] = make_uint4(pack2<T>(v[0], v[1]), pack2<T>(v[2], v[3]), pack2<T>(v[4], v[5]), pack2<T>(v[6], v[7]));
          const float4 s0 = *reinterpret_cast<const float4*>(s_scale + col);
          const float4 s1 = *reinterpret_cast<const float4*>(s_scale + col + 4);
          const float4 t0 = *reinterpret_cast<const float4*>(s_shift + col);
          const float4 t1 = *reinterpret_cast<const float4*>(s_shift + col + 4);
          float a[8];
          a[0] = fmaf(s0.x, v[0], t0.x); a[1] = fmaf(s0.y, v[1], t0.y); a[2] = fmaf(s0.z, v[2], t0.z); a[3] = fmaf(s0.w, v[3], t0.w);
          a[4] = fmaf(s1.x, v[4], t1.x); a[5] = fmaf(s1.y, v[5], t1.y); a[6] = fmaf(s1.z, v[6], t1.z); a[7] = fmaf(s1.w, v[7], t1.w);
          if (p.act_relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = fmaxf(a[k], 0.f);
          }
          act_row[slot16] = make_uint4(pack2<T>(a[0], a[1]), pack2<T>(a[2], a[3]), pack2<T>(a[4], a[5]), pack2<T>(a[6], a[7]));
        }
        if (tr) p.trace[cc * 8 + 2] = clock64();
        tc_fence_before();                   // this thread's acc3 / acc1 reads are ordered before the signal below
        fence_proxy_async_smem();            // generic-proxy writes -> visible to the TMA store and to the tensor core
        if (tr) p.trace[cc * 8 + 3] = clock64();
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (tr) p.trace[cc * 8 + 4] = clock64();
        if (store_thread) {
          mbar_arrive(smem_u32(&act_ready[ab]));
          tma_store_4d(&tmOut, smem_u32(r_ring + slot * kChunkBytes), c * 64, tc.w0, tc.h0, tc.n0);
          tma_store_commit();
          tma_store_wait_read_pending<1>();
          if (prev_slot >= 0) mbar_arrive(smem_u32(&r_empty[prev_slot]));
          prev_slot = int(slot);
        }
        if (tr) p.trace[cc * 8 + 5] = clock64();
      }
    }
    if (i > 0) finish_tail(prev, i - 1);
    if (store_thread) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}
