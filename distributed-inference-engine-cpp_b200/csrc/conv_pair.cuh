// Two-SM convolution: one 256 x N output tile per CTA PAIR (thread-block cluster of 2), tcgen05.mma.cta_group::2.
// (included by conv_umma.cu inside namespace rn::{anonymous})
//
// Why: with one CTA per tile the tensor core re-reads both operands from shared memory for every MMA, and
// (128 + N) x 32 B per M128 x N x K16 instruction is 96 B/clk at N = 256 on top of the 96 B/clk the TMA unit writes --
// streamed operands cap a single SM at ~67 % of the MMA rate.  In a CTA pair each SM stages its own
// 128-row A tile and only HALF of the weight tile (N/2 rows); the M = 256 instruction reads A locally and B from both
// SMs, so per SM the shared-memory traffic drops to 64 B/clk read + 64 B/clk written at N = 256.  The accumulator is
// 128 lanes x N columns in EACH SM's tensor memory (CTA r holds rows 128 r ... of the pair tile), double-buffered, so a
// tile's epilogue overlaps the next tile's MMAs -- which the dual-M kernel could not do (both accumulators in use).
//
// Roles per CTA (576 threads): warp 0 = TMA producer (both CTAs; every load signals the LEADER's "full" barrier),
// warp 1 = TMEM allocation in both CTAs + the single MMA-issuing thread in the leader (rank 0), warps 2..17 = epilogue
// in two groups of eight that take alternating 64-channel chunks (bias, optional ReLU, one output, TMA store).
// Barriers:  full[s]       leader only; 1 arrival (leader's expect_tx of both CTAs' bytes) + both CTAs' TMA bytes
//            empty[s]      both CTAs; tcgen05.commit multicast from the leader's MMA thread
//            tmem_full[b]  both CTAs; tcgen05.commit multicast
//            tmem_empty[b] leader only; one arrival per epilogue warp of BOTH CTAs (remote mbarrier.arrive)
// The TMA loads here read this GPU's memory only and are not multicast.
constexpr int kPairThreads = 64 + 2 * kEpiThreads;

struct PairParams {
  int num_k_iters, cblks, kw, stride, pad;
  int bw, bh, bn, tiles_w, tiles_h, m_tiles, n_tiles, total_items, num_stages, relu;
  uint32_t a_bytes;
  const float* bias;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_cluster(uint32_t cta_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA loads of a CTA pair: the data lands in the executing CTA's shared memory, the bytes are counted on the barrier
// at `bar_cluster` (a shared::cluster address: the leader's)
__device__ __forceinline__ void tma_load_4d_pair(uint32_t dst, const CUtensorMap* m, uint32_t bar_cluster, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* m, uint32_t bar_cluster, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t slot_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// completion of all MMAs issued so far by this thread -> one arrival on the barrier at this offset in both CTAs
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
  asm volatile(
      "{\n\t"
      ".reg .b16 mask;\n\t"
      "mov.b16 mask, 3;\n\t"
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], mask;\n\t"
      "}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_f16_lo_pair(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p;\n\t"
      "}" ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
// instruction descriptor of the pair MMA: M = 256 across the two CTAs
template <typename T, int N>
__device__ __forceinline__ constexpr uint32_t make_idesc_pair() {
  return (1u << 4) | (Fmt<T>::id << 7) | (Fmt<T>::id << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(256 >> 4) << 24);
}

template <typename T, int BLOCK_N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
conv_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmOut, const PairParams p) {
  constexpr int kChunks = BLOCK_N / 64;
  constexpr uint32_t A_BYTES = kTileM * 128;
  constexpr uint32_t BH_BYTES = (BLOCK_N / 2) * 128;        // this CTA's half of the weight tile
  constexpr uint32_t STAGE_BYTES = A_BYTES + BH_BYTES;

  extern __shared__ uint8_t smem_raw[];
  // identical carve-up in both CTAs (the MMA and the multicast commits address the peer by the same offsets)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = p.num_stages;
  uint8_t* pipe = smem;
  uint8_t* outbuf = pipe + S * STAGE_BYTES;        // two 16 KB chunk buffers, one per epilogue group
  uint64_t* bars = reinterpret_cast<uint64_t*>(outbuf + 2 * kChunkBytes);
  uint64_t* full_bar = bars;                  // [S] (used in the leader)
  uint64_t* empty_bar = bars + S;             // [S]
  uint64_t* tmem_full = empty_bar + S;        // [2]
  uint64_t* tmem_empty = tmem_full + 2;       // [2] (used in the leader)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = int(blockIdx.x) >> 1, n_clusters = int(gridDim.x) >> 1;
  const int my_items = cluster_id < p.total_items ? (p.total_items - 1 - cluster_id) / n_clusters + 1 : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmOut);
    for (int i = 0; i < S; ++i) { mbar_init(smem_u32(&full_bar[i]), 1); mbar_init(smem_u32(&empty_bar[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&tmem_full[i]), 1); mbar_init(smem_u32(&tmem_empty[i]), 2 * 2 * kEpiWarps); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair(smem_u32(tmem_slot), 2 * BLOCK_N);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // the peer's barriers exist before anything is signalled across the pair
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  // this CTA's M tile of pair `pm`; an index past the last tile lands beyond the batch (TMA loads read zeros there
  // and TMA stores are clipped)
  auto m_coord = [&](int pm, int& w0, int& h0, int& n0) {
    int mt = 2 * pm + int(rank);
    const int tw = mt % p.tiles_w; mt /= p.tiles_w;
    const int th = mt % p.tiles_h;
    w0 = tw * p.bw; h0 = th * p.bh; n0 = (mt / p.tiles_h) * p.bn;
  };

  if (warp == 0) {
    // ===== TMA producer (both CTAs) =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_items; ++i) {
        const int item = cluster_id + i * n_clusters;
        const int col0 = (item % p.n_tiles) * BLOCK_N + int(rank) * (BLOCK_N / 2);
        int w0, h0, n0;
        m_coord(item / p.n_tiles, w0, h0, n0);
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          const uint32_t fb_local = smem_u32(&full_bar[stage]);
          const uint32_t fb = mapa_cluster(fb_local, 0);
          if (rank == 0) mbar_expect_tx(fb_local, 2 * (p.a_bytes + BH_BYTES));
          const uint32_t dst = smem_u32(pipe + stage * STAGE_BYTES);
          const int tap = it / p.cblks, cb = it - tap * p.cblks;
          const int ky = tap / p.kw, kx = tap - ky * p.kw;
          tma_load_4d_pair(dst, &tmA, fb, cb * 64, w0 * p.stride - p.pad + kx, h0 * p.stride - p.pad + ky, n0);
          tma_load_2d_pair(dst + A_BYTES, &tmW, fb, it * 64, col0);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
      // tail: every commit aimed at this CTA's "empty" barriers has landed before the CTA may exit
      const int total = my_items * p.num_k_iters;
      for (int k = 0; k < (total < S ? total : S); ++k) {
        mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one thread of the leader CTA =====
    constexpr uint32_t idesc = make_idesc_pair<T, BLOCK_N>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (rank == 0 && elect_one()) {
      const uint32_t pipe_lo = smem_desc_lo(smem_u32(pipe));
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_items; ++i) {
        const int buf = i & 1;
        mbar_wait(smem_u32(&tmem_empty[buf]), ((uint32_t(i) >> 1) & 1) ^ 1);   // both CTAs' epilogues drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + uint32_t(buf * BLOCK_N);
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t a_lo = pipe_lo + uint32_t(stage) * (STAGE_BYTES >> 4);
          const uint32_t b_lo = a_lo + (A_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_f16_lo_pair(tmem_d, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (it | k) != 0 ? 1u : 0u);
          umma_commit_pair(smem_u32(&empty_bar[stage]));
          if (it == p.num_k_iters - 1) umma_commit_pair(smem_u32(&tmem_full[buf]));
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===== epilogue: two groups of 8 warps on alternating chunks; thread = one pixel (TMEM lane) x 32 channels =====
    const int grp = (warp - 2) >> 3;
    const int q4 = warp & 3;
    const int g = ((warp - 2) & 7) >> 2;
    const int row = q4 * 32 + lane;
    const int sw = row & 7;
    const bool store_thread = (((warp - 2) & 7) == 0 && lane == 0);
    const int bar_id = 1 + grp;
    uint8_t* stage_out = outbuf + grp * kChunkBytes;
    const uint32_t out_row = smem_u32(stage_out) + uint32_t(row) * 128u;
    const uint32_t lane_addr = tmem_base + (uint32_t(q4 * 32) << 16);
    const uint32_t tmem_empty_leader = mapa_cluster(smem_u32(&tmem_empty[0]), 0);
    int cc = 0;                                  // this group's running chunk counter
    for (int i = 0; i < my_items; ++i) {
      const int item = cluster_id + i * n_clusters;
      const int col0 = (item % p.n_tiles) * BLOCK_N;
      int w0, h0, n0;
      m_coord(item / p.n_tiles, w0, h0, n0);
      const int buf = i & 1;
      mbar_wait(smem_u32(&tmem_full[buf]), (uint32_t(i) >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = grp; c < kChunks; c += 2, ++cc) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(lane_addr + uint32_t(buf * BLOCK_N + c * 64 + g * 32), acc);
        const float4* bsrc = reinterpret_cast<const float4*>(p.bias + col0 + c * 64 + g * 32);
        float4 bv[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) bv[k] = __ldg(bsrc + k);
        tmem_ld_wait();
        if (c + 2 >= kChunks) {             // this warp's last chunk of the tile: its part of the accumulator is drained
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(tmem_empty_leader + uint32_t(buf) * 8u);
        }
        if (store_thread && cc > 0) tma_store_wait_read();   // the group's previous store has read the staging buffer
        asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const float4 b0 = bv[2 * jj], b1 = bv[2 * jj + 1];
          float v[8];
          v[0] = __uint_as_float(acc[jj * 8 + 0]) + b0.x; v[1] = __uint_as_float(acc[jj * 8 + 1]) + b0.y;
          v[2] = __uint_as_float(acc[jj * 8 + 2]) + b0.z; v[3] = __uint_as_float(acc[jj * 8 + 3]) + b0.w;
          v[4] = __uint_as_float(acc[jj * 8 + 4]) + b1.x; v[5] = __uint_as_float(acc[jj * 8 + 5]) + b1.y;
          v[6] = __uint_as_float(acc[jj * 8 + 6]) + b1.z; v[7] = __uint_as_float(acc[jj * 8 + 7]) + b1.w;
          if (p.relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
          }
          sts_u4(out_row + uint32_t(((g * 4 + jj) ^ sw) << 4),
                 make_uint4(pack2<T>(v[0], v[1]), pack2<T>(v[2], v[3]), pack2<T>(v[4], v[5]), pack2<T>(v[6], v[7])));
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
        if (store_thread) {
          tma_store_4d(&tmOut, smem_u32(stage_out), col0 + c * 64, w0, h0, n0);
          tma_store_commit();
        }
      }
    }
    if (store_thread) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();            // neither CTA leaves (or frees tensor memory) while the other may still signal it
  if (warp == 1) tmem_dealloc_pair(tmem_base, 2 * BLOCK_N);
}
