// Exact-mode convolution on the tensor cores: fp32 activations and weights, 3 x TF32 split products accumulated in
// fp32 tensor memory.  (included by conv_umma.cu inside namespace rn::{anonymous})
//
// tcgen05 has no fp32-input MMA; kind::tf32 keeps 10 mantissa bits of each operand.  Each fp32 value is therefore
// split exactly into hi = the value with its low 13 mantissa bits cleared and lo = value - hi (again cut to TF32), and
//     a * w  ~=  a_hi * w_hi  +  a_hi * w_lo  +  a_lo * w_hi          (the dropped a_lo * w_lo is ~2^-22 relative)
// The weights are split once on the host ([Cout][K] hi and lo arrays).  The activation tile arrives by TMA as raw fp32
// and four transform warps write its lo part into a second buffer of the same swizzled layout before the MMA thread
// sees the stage.  The raw tile itself serves as A_hi: the instruction ignores the low 13 mantissa bits of its operands,
// which is exactly the cut that defines hi (measured: bit-for-bit the same op results as with hi written in place, and
// a third less shared-memory traffic in the split; B200RN_TF32_RAW_HI=0 writes hi as the first version did).  Per K step of 8 channels three M128 x N x K8 instructions are issued:
//     MAIN += A_hi * W_hi^T          LO += A_hi * W_lo^T          LO += A_lo * W_hi^T
// Two-level accumulation.  The tensor core's fp32 accumulate truncates instead of rounding, a one-sided error of about
// half an ulp per instruction that grows linearly with the length of the chain: measured 1.5e-5 relative at K = 2304
// (288 K steps) with everything accumulated in tensor memory, 1.4e-3 max-abs on the logits -- over the 1e-3 bar.  So
// MAIN only ever accumulates a chunk of kT32ChunkBlocks K blocks (16 K steps); it alternates between two tensor-memory
// buffers and the epilogue warps add every finished chunk into fp32 REGISTERS (N/2 per thread) with properly rounded
// adds while the next chunk is being multiplied.  LO is 2^-11 of MAIN, its truncation error is negligible, and it stays
// in tensor memory for the whole K (double-buffered across tiles).  v = registers + LO + bias.
// Fused hi products (Tf32Params::fused, layers with >= 16 K blocks).  A_hi * W_hi^T and A_hi * W_lo^T read the same A
// tile; stacking W_hi over W_lo (they are adjacent in a stage) makes them ONE instruction of width 2N whose result columns
// are [MAIN | LO] side by side, so A_hi is read from shared memory once instead of twice per K step (20 KB instead of
// 24 KB at N = 128, 14 instead of 18 KB at N = 64: these layers are bound by shared-memory bandwidth).  For that the LO
// accumulator is tied to the MAIN buffer it sits beside -- tensor memory holds [MAIN0 | LO0 | MAIN1 | LO1], a chunk of
// parity m accumulates its lo terms into LO_m and the epilogue adds LO0 + LO1 -- which gives up the double-buffering of
// LO across tiles (the next tile's first MMA waits until the epilogue has read both; a few hundred cycles per tile, hence
// only for long K).  The one instruction has one accumulate flag, and MAIN restarts at every chunk while LO_m carries on,
// so the first K step of a chunk is issued as the three separate instructions.
// Epilogue (two groups of 8 warps on alternating 32-channel chunks): bias, ReLU, residual, raw and/or BN-ReLU'd output
// staged in swizzled shared memory (one buffer per output and group) and written by TMA stores, 32 channels (128 bytes)
// per chunk.  The residual chunk arrives by TMA in a landing buffer per group, requested one chunk ahead by the group's
// store thread (res_tma); the first version read it straight from global memory, 64 contiguous bytes per thread, which
// is 32 separate row pieces per warp instruction and ran the residual layers 10-50 % slower (kept as the fallback).
// The 7x7/2 stem (3 channels stored as 4) reads a zero-framed fp32 input through a 5-D overlapping view: K block ky is the
// 8 pixels x 4 channels = 32 contiguous floats at frame row 2y + ky, pixel 2x (the 8th pixel meets zero weights), so the
// stem is 7 K blocks of the same pipeline.
// K is ordered (ky, kx, cin) in blocks of 32 channels; the operand box is {32 ch, bw, bh, bn} fp32 with the conv's
// stride as element stride and TMA zero fill as padding, exactly like the 16-bit kernels.
constexpr int kT32XfWarps = 4;
constexpr int kT32ChunkBlocks = 4;          // K blocks (of 32 channels) per MAIN chunk
constexpr int kT32EpiWarps = 16;
constexpr int kT32Threads = 64 + kT32XfWarps * 32 + kT32EpiWarps * 32;   // TMA, MMA, 4 transform warps, 2 x 8 epilogue warps

struct Tf32Params {
  int num_k_iters, cblks, kw, stride, pad;
  int stem;                       // 7x7/2 stem on a zero-framed 4-channel input: K block ky = 8 pixels x 4 channels of frame row 2y + ky
  int bw, bh, bn, tiles_w, tiles_h, n_tiles, total_tiles, num_stages;
  int has_res, has_out, has_act, relu, act_relu;
  int cout;                       // row pitch (floats) of the residual tensor
  int Ho, Wo, B;
  uint32_t rows, a_bytes;
  const float* bias;
  const float* act_scale;
  const float* act_shift;
  const float* residual;          // fp32 NHWC [B, Ho, Wo, cout] or null
  int fused;                      // one MMA of width 2N per K step for both A_hi products ([MAIN_m | LO_m] side by side in tensor memory):
                                  // see "Fused hi products" above; needs raw_hi-style operands only in that W_lo follows W_hi in the stage
  int raw_hi;                     // the transform warps leave the raw fp32 tile in place as A_hi (the MMA truncates it itself)
  int res_tma;                    // residual chunks arrive by TMA in a 16 KB buffer per epilogue group (one chunk ahead) instead of
                                  // per-thread global loads of 64-byte row pieces
};

// kind::tf32 instruction descriptor: D fp32 (bit 4), A and B formats = 2 (TF32), both K-major, N >> 3, M >> 4
template <int N>
__device__ __forceinline__ constexpr uint32_t make_idesc_tf32() {
  return (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(kTileM >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32_lo(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %4, p;\n\t"
      "}" ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

template <int BLOCK_N, bool RES_TMA>
__global__ void __launch_bounds__(kT32Threads, 1)
conv_tf32x3_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmWhi,
                   const __grid_constant__ CUtensorMap tmWlo, const __grid_constant__ CUtensorMap tmOut,
                   const __grid_constant__ CUtensorMap tmAct, const __grid_constant__ CUtensorMap tmRes, const Tf32Params p) {
  constexpr int kChunks = BLOCK_N / 32;               // 32 fp32 channels = one 128-byte staging row
  constexpr int kGroupChunks = kChunks / 2;           // chunks per epilogue group and tile
  constexpr uint32_t A_BYTES = kTileM * 128;          // 128 rows x 32 fp32
  constexpr uint32_t B_BYTES = BLOCK_N * 128;         // one of the two weight tiles
  constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;   // A_hi | A_lo | W_hi | W_lo
  // tensor memory: MAIN chunk buffers 0/1 at columns [0, N) and [N, 2N); LO accumulators 0/1 at [2N, 3N) and [3N, 4N)
  constexpr uint32_t TMEM_COLS = 4 * BLOCK_N;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = p.num_stages;
  uint8_t* pipe = smem;
  uint8_t* outbuf = pipe + S * STAGE_BYTES;                       // raw-output staging: one 16 KB chunk per epilogue group
  uint8_t* actbuf = outbuf + (p.has_out ? 2 * kChunkBytes : 0);   // activated-output staging: one chunk per group
  uint8_t* resbuf = actbuf + (p.has_act ? 2 * kChunkBytes : 0);   // residual landing buffers: one chunk per group (res_tma)
  uint64_t* bars = reinterpret_cast<uint64_t*>(resbuf + (RES_TMA ? 2 * kChunkBytes : 0));
  uint64_t* full_bar = bars;                  // [S]  TMA -> transform warps
  uint64_t* xf_bar = full_bar + S;            // [S]  transform warps -> MMA (4 warp arrivals)
  uint64_t* empty_bar = xf_bar + S;           // [S]  MMA -> TMA
  uint64_t* main_full = empty_bar + S;        // [2]  MMA -> epilogue: a MAIN chunk is complete
  uint64_t* main_empty = main_full + 2;       // [2]  epilogue -> MMA (16 warp arrivals): chunk added into registers
  uint64_t* lo_full = main_empty + 2;         // [2]  MMA -> epilogue: a tile's LO accumulator is complete
  uint64_t* lo_empty = lo_full + 2;           // [2]  epilogue -> MMA (16 warp arrivals)
  uint64_t* res_full = lo_empty + 2;          // [2]  residual TMA -> epilogue group
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_full + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int my_tiles = (int(blockIdx.x) < p.total_tiles) ? (p.total_tiles - 1 - int(blockIdx.x)) / int(gridDim.x) + 1 : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmWhi); prefetch_tmap(&tmWlo);
    if (p.has_out) prefetch_tmap(&tmOut);
    if (p.has_act) prefetch_tmap(&tmAct);
    if (RES_TMA) prefetch_tmap(&tmRes);
    for (int i = 0; i < S; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1);
      mbar_init(smem_u32(&xf_bar[i]), kT32XfWarps);
      mbar_init(smem_u32(&empty_bar[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&main_full[i]), 1); mbar_init(smem_u32(&main_empty[i]), kT32EpiWarps);
      mbar_init(smem_u32(&lo_full[i]), 1); mbar_init(smem_u32(&lo_empty[i]), kT32EpiWarps);
      mbar_init(smem_u32(&res_full[i]), 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  auto tile_coord = [&](int i, int& w0, int& h0, int& n0, int& col0) {
    const int tile = int(blockIdx.x) + i * int(gridDim.x);
    col0 = (tile % p.n_tiles) * BLOCK_N;
    int mt = tile / p.n_tiles;
    const int tw = mt % p.tiles_w; mt /= p.tiles_w;
    const int th = mt % p.tiles_h;
    w0 = tw * p.bw; h0 = th * p.bh; n0 = (mt / p.tiles_h) * p.bn;
  };

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_tiles; ++i) {
        int w0, h0, n0, col0;
        tile_coord(i, w0, h0, n0, col0);
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_expect_tx(fb, p.a_bytes + 2 * B_BYTES);
          const uint32_t dst = smem_u32(pipe + stage * STAGE_BYTES);
          const int tap = it / p.cblks, cb = it - tap * p.cblks;
          const int ky = tap / p.kw, kx = tap - ky * p.kw;
          if (p.stem) tma_load_5d(dst, &tmA, fb, 0, w0, it & 1, h0 + (it >> 1), n0);   // (32 elems, out col, row parity, row pair, image)
          else tma_load_4d(dst, &tmA, fb, cb * 32, w0 * p.stride - p.pad + kx, h0 * p.stride - p.pad + ky, n0);
          tma_load_2d(dst + 2 * A_BYTES, &tmWhi, fb, it * 32, col0);
          tma_load_2d(dst + 2 * A_BYTES + B_BYTES, &tmWlo, fb, it * 32, col0);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    constexpr uint32_t idesc = make_idesc_tf32<BLOCK_N>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (elect_one()) {
      const uint32_t pipe_lo = smem_desc_lo(smem_u32(pipe));
      int stage = 0;
      uint32_t phase = 0;
      uint32_t chunk = 0;                   // running MAIN chunk counter over all of this CTA's tiles
      constexpr uint32_t idesc2 = make_idesc_tf32<(2 * BLOCK_N <= 256 ? 2 * BLOCK_N : BLOCK_N)>();
      for (int i = 0; i < my_tiles; ++i) {
        // fused: one LO pair for all tiles (barrier 0, a phase per tile); otherwise LO double-buffered across tiles
        const int lbuf = p.fused ? 0 : (i & 1);
        const uint32_t lo_parity = p.fused ? (uint32_t(i) & 1) : ((uint32_t(i) >> 1) & 1);
        mbar_wait(smem_u32(&lo_empty[lbuf]), lo_parity ^ 1);
        tc_fence_after();
        uint32_t tile_chunk = 0;            // MAIN chunk index within this tile
        for (int it = 0; it < p.num_k_iters; ++it) {
          const int in_chunk = it % kT32ChunkBlocks;
          const uint32_t mbuf = chunk & 1;
          if (in_chunk == 0) {
            mbar_wait(smem_u32(&main_empty[mbuf]), ((chunk >> 1) & 1) ^ 1);   // the epilogue has taken this buffer's last chunk
            tc_fence_after();
          }
          // tensor memory: [MAIN0 | MAIN1 | LO0 | LO1], or fused [MAIN0 | LO0 | MAIN1 | LO1]
          const uint32_t tmem_main = tmem_base + (p.fused ? 2 * mbuf : mbuf) * BLOCK_N;
          const uint32_t tmem_lo = p.fused ? tmem_main + BLOCK_N : tmem_base + uint32_t(2 + lbuf) * BLOCK_N;
          mbar_wait(smem_u32(&xf_bar[stage]), phase);
          tc_fence_after();
          const uint32_t ahi = pipe_lo + uint32_t(stage) * (STAGE_BYTES >> 4);
          const uint32_t alo = ahi + (A_BYTES >> 4);
          const uint32_t bhi = ahi + 2 * (A_BYTES >> 4);
          const uint32_t blo = bhi + (B_BYTES >> 4);
          // lo terms: the first instruction into an LO buffer within a tile overwrites it
          const bool lo_first = p.fused ? (tile_chunk < 2 && in_chunk == 0) : (it == 0);
#pragma unroll
          for (int k = 0; k < 4; ++k) {       // K = 8 fp32 = 32 bytes per instruction
            if (p.fused && (in_chunk | k) != 0) {
              umma_tf32_lo(tmem_main, ahi + 2 * k, bhi + 2 * k, desc_hi, idesc2, 1u);      // [MAIN | LO] += A_hi * [W_hi ; W_lo]^T
            } else {
              umma_tf32_lo(tmem_main, ahi + 2 * k, bhi + 2 * k, desc_hi, idesc, (in_chunk | k) != 0 ? 1u : 0u);
              umma_tf32_lo(tmem_lo, ahi + 2 * k, blo + 2 * k, desc_hi, idesc, (lo_first && k == 0) ? 0u : 1u);
            }
            umma_tf32_lo(tmem_lo, alo + 2 * k, bhi + 2 * k, desc_hi, idesc, 1u);
          }
          umma_commit(smem_u32(&empty_bar[stage]));
          const bool last = it == p.num_k_iters - 1;
          if (in_chunk == kT32ChunkBlocks - 1 || last) { umma_commit(smem_u32(&main_full[mbuf])); ++chunk; ++tile_chunk; }
          if (last) umma_commit(smem_u32(&lo_full[lbuf]));
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp < 2 + kT32XfWarps) {
    // ===== operand split: A -> A_hi (in place) and A_lo, 16 bytes per thread per step, layout-preserving =====
    const int t = threadIdx.x - 64;                 // 0..127
    int stage = 0;
    uint32_t phase = 0;
    const int total = my_tiles * p.num_k_iters;
    for (int q = 0; q < total; ++q) {
      mbar_wait(smem_u32(&full_bar[stage]), phase);
      const uint32_t a_hi = smem_u32(pipe + stage * STAGE_BYTES);
#pragma unroll
      for (int j = 0; j < int(A_BYTES / 16 / 128); ++j) {
        const uint32_t off = uint32_t(j * 128 + t) * 16u;
        const float4 a = lds_f4(a_hi + off);
        float4 h, l;
        h.x = tf32_hi(a.x); h.y = tf32_hi(a.y); h.z = tf32_hi(a.z); h.w = tf32_hi(a.w);
        l.x = a.x - h.x; l.y = a.y - h.y; l.z = a.z - h.z; l.w = a.w - h.w;      // exact in fp32; the MMA cuts it to TF32 itself
        // kind::tf32 ignores the low 13 mantissa bits of its operands, so the raw tile already IS A_hi to the MMA: only A_lo
        // is written (raw_hi; a third less shared-memory traffic for the split)
        if (!p.raw_hi) sts_u4(a_hi + off, make_uint4(__float_as_uint(h.x), __float_as_uint(h.y), __float_as_uint(h.z), __float_as_uint(h.w)));
        sts_u4(a_hi + A_BYTES + off, make_uint4(__float_as_uint(l.x), __float_as_uint(l.y), __float_as_uint(l.z), __float_as_uint(l.w)));
      }
      fence_proxy_async_smem();            // generic-proxy writes -> visible to the tensor core's reads
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&xf_bar[stage]));
      if (++stage == S) { stage = 0; phase ^= 1; }
    }
  } else {
    // ===== epilogue: two groups of 8 warps; group grp owns the 32-channel chunks grp, grp + 2, ... of every tile (its own
    // staging buffers, named barrier and store thread), so two chunk chains are in flight per SM -- one chain at a time
    // (TMEM load, residual, math, smem, proxy fence, barrier, TMA store, "store has read" wait) ran the residual layers at a
    // third of the HBM rate.  thread = one output pixel (TMEM lane) x 16 of a chunk's 32 channels =====
    const int ew = warp - 2 - kT32XfWarps;  // 0..15
    const int grp = ew >> 3;
    const int q4 = warp & 3;                // TMEM lane quarter this warp may access
    const int g = (ew & 7) >> 2;            // which half of every 32-channel chunk (0/1)
    const int row = q4 * 32 + lane;
    const int sw = row & 7;
    const bool store_thread = ((ew & 7) == 0 && lane == 0);
    const int bar_id = 1 + grp;
    const uint32_t lane_addr = tmem_base + (uint32_t(q4 * 32) << 16);
    const uint32_t out_stage = smem_u32(outbuf) + uint32_t(grp) * kChunkBytes;
    const uint32_t act_stage = smem_u32(actbuf) + uint32_t(grp) * kChunkBytes;
    const uint32_t out_row = out_stage + uint32_t(row) * 128u, act_row = act_stage + uint32_t(row) * 128u;
    int cc = 0;                                 // this group's running chunk counter
    uint32_t mchunk = 0;                        // running MAIN chunk counter (same sequence as the MMA thread's)
    const uint32_t res_stage = smem_u32(resbuf) + uint32_t(grp) * kChunkBytes, res_row = res_stage + uint32_t(row) * 128u;
    const uint32_t res_bar = smem_u32(&res_full[grp]);
    // res_tma: the group's store thread requests chunk j of tile t into the group's landing buffer (rows past the
    // batch are zero-filled by TMA); one request is outstanding per group, issued as soon as the previous chunk's
    // residual has been read by all of the group's threads
    auto request_residual = [&](int t, int j) {
      if (t >= my_tiles) return;
      int rw0, rh0, rn0, rcol0;
      tile_coord(t, rw0, rh0, rn0, rcol0);
      mbar_expect_tx(res_bar, p.rows * 128u);
      tma_load_4d(res_stage, &tmRes, res_bar, rcol0 + (grp + 2 * j) * 32, rw0, rh0, rn0);
    };
    if (RES_TMA && store_thread) request_residual(0, 0);
    for (int i = 0; i < my_tiles; ++i) {
      int w0, h0, n0, col0;
      tile_coord(i, w0, h0, n0, col0);
      const int buf = i & 1;
      // this thread's output pixel (for the residual read): row -> (image, y, x) inside the patch
      const int px = row % p.bw, py = (row / p.bw) % p.bh, pn = row / (p.bw * p.bh);
      const int ox = w0 + px, oy = h0 + py, on = n0 + pn;
      const bool in_range = row < int(p.rows) && ox < p.Wo && oy < p.Ho && on < p.B;
      const float* res_px = p.residual + ((size_t(on) * p.Ho + oy) * p.Wo + ox) * size_t(p.cout) + col0 + g * 16;
      // residual of the group's first chunk: requested now, needed only after the whole K loop
      float4 rv[4];
      if (p.has_res && !RES_TMA) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          rv[k] = in_range ? __ldg(reinterpret_cast<const float4*>(res_px + grp * 32) + k) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      // ---- level 2 of the accumulation: every finished MAIN chunk is added into registers ----
      float sum[kGroupChunks * 16];
#pragma unroll
      for (int k = 0; k < kGroupChunks * 16; ++k) sum[k] = 0.f;
      const int n_chunks = (p.num_k_iters + kT32ChunkBlocks - 1) / kT32ChunkBlocks;
      for (int kc = 0; kc < n_chunks; ++kc, ++mchunk) {
        const uint32_t mbuf = mchunk & 1;
        mbar_wait(smem_u32(&main_full[mbuf]), (mchunk >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int j = 0; j < kGroupChunks; ++j) {
          uint32_t d[16];
          tmem_ld_32x32b_x16(lane_addr + (p.fused ? 2 * mbuf : mbuf) * BLOCK_N + uint32_t((grp + 2 * j) * 32 + g * 16), d);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 16; ++k) sum[j * 16 + k] += __uint_as_float(d[k]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&main_empty[mbuf]));
      }
      // lo terms: one accumulator per tile parity, or (fused) LO0 + LO1 of the single pair, a barrier phase per tile
      const int lbar = p.fused ? 0 : buf;
      mbar_wait(smem_u32(&lo_full[lbar]), p.fused ? (uint32_t(i) & 1) : ((uint32_t(i) >> 1) & 1));
      tc_fence_after();
#pragma unroll
      for (int j = 0; j < kGroupChunks; ++j, ++cc) {
        const int c = grp + 2 * j;
        const int ch0 = c * 32 + g * 16;            // first of this thread's 16 channels within the N tile
        uint32_t d1[16];
        if (p.fused) {
          uint32_t d2[16];
          tmem_ld_32x32b_x16(lane_addr + uint32_t(BLOCK_N + ch0), d1);
          tmem_ld_32x32b_x16(lane_addr + uint32_t(3 * BLOCK_N + ch0), d2);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 16; ++k) d1[k] = __float_as_uint(__uint_as_float(d1[k]) + __uint_as_float(d2[k]));
        } else {
          tmem_ld_32x32b_x16(lane_addr + uint32_t(2 + buf) * BLOCK_N + uint32_t(ch0), d1);
        }
        float4 bv[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) bv[k] = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + ch0) + k);
        tmem_ld_wait();
        if (j == kGroupChunks - 1) {        // this warp's part of the LO accumulator is read: hand it back to the MMA thread
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&lo_empty[lbar]));
        }
        if (store_thread && cc > 0) tma_store_wait_read();   // the group's previous stores have read its staging buffers
        asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
        if (RES_TMA) {
          mbar_wait(res_bar, uint32_t(cc) & 1);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint4 r = lds_u4(res_row + uint32_t(((g * 4 + k) ^ sw) << 4));
            rv[k] = make_float4(__uint_as_float(r.x), __uint_as_float(r.y), __uint_as_float(r.z), __uint_as_float(r.w));
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float v[4];
          v[0] = sum[j * 16 + 4 * k + 0] + __uint_as_float(d1[4 * k + 0]) + bv[k].x;
          v[1] = sum[j * 16 + 4 * k + 1] + __uint_as_float(d1[4 * k + 1]) + bv[k].y;
          v[2] = sum[j * 16 + 4 * k + 2] + __uint_as_float(d1[4 * k + 2]) + bv[k].z;
          v[3] = sum[j * 16 + 4 * k + 3] + __uint_as_float(d1[4 * k + 3]) + bv[k].w;
          if (p.relu) { v[0] = fmaxf(v[0], 0.f); v[1] = fmaxf(v[1], 0.f); v[2] = fmaxf(v[2], 0.f); v[3] = fmaxf(v[3], 0.f); }
          if (p.has_res) { v[0] += rv[k].x; v[1] += rv[k].y; v[2] += rv[k].z; v[3] += rv[k].w; }
          const uint32_t slot16 = uint32_t(((g * 4 + k) ^ sw) << 4);      // SWIZZLE_128B: 16-byte slot XOR (row mod 8)
          if (p.has_out)
            sts_u4(out_row + slot16, make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3])));
          if (p.has_act) {
            const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.act_scale + col0 + ch0) + k);
            const float4 t4 = __ldg(reinterpret_cast<const float4*>(p.act_shift + col0 + ch0) + k);
            float a[4] = {fmaf(s4.x, v[0], t4.x), fmaf(s4.y, v[1], t4.y), fmaf(s4.z, v[2], t4.z), fmaf(s4.w, v[3], t4.w)};
            if (p.act_relu) { a[0] = fmaxf(a[0], 0.f); a[1] = fmaxf(a[1], 0.f); a[2] = fmaxf(a[2], 0.f); a[3] = fmaxf(a[3], 0.f); }
            sts_u4(act_row + slot16, make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]), __float_as_uint(a[3])));
          }
        }
        if (p.has_res && !RES_TMA && j + 1 < kGroupChunks) {       // the group's next chunk's residual (64 channels further on)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            rv[k] = in_range ? __ldg(reinterpret_cast<const float4*>(res_px + (c + 2) * 32) + k) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
        if (store_thread) {
          // every thread of the group has read this chunk's residual: the landing buffer takes the next one
          if (RES_TMA) { if (j + 1 < kGroupChunks) request_residual(i, j + 1); else request_residual(i + 1, 0); }
          if (p.has_out) tma_store_4d(&tmOut, out_stage, col0 + c * 32, w0, h0, n0);
          if (p.has_act) tma_store_4d(&tmAct, act_stage, col0 + c * 32, w0, h0, n0);
          tma_store_commit();
        }
      }
    }
    if (store_thread) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}
