// Implicit-GEMM convolution on the 5th-generation tensor cores (sm_100a): tcgen05.mma with the
// accumulator in TMEM, operands staged in shared memory by TMA, fused epilogue
// (bias, ReLU, residual add, optional second "BN+ReLU'd" output) written back with TMA stores.
//
//   D[M = pixels][N = Cout] = A[M][K] * W[N][K]^T,   K ordered (ky, kx, cin)
//
// * No im2col buffer exists anywhere.  The activation tensor is NHWC; one K-block of A for tap
//   (ky,kx) and channel block cb is a 4-D TMA box {64 ch, bw, bh, bn} of the input at the shifted
//   (and, for stride 2, element-strided) position.  Out-of-bounds coordinates are zero-filled by
//   the TMA unit, which *is* the conv's zero padding.  The M tile is therefore a spatial patch
//   bw x bh of bn images (bw*bh*bn <= 128), chosen per layer and batch to fill the 128-row MMA.
// * The 7x7/2 stem (3 input channels, stored as 4) reads a zero-framed [B, H+6, Wp, 4] input: stem7x7_kernel fetches one
//   37-row patch per 8x16-output tile and feeds the MMA through un-swizzled overlapping descriptors (the sliding
//   window is the descriptor's 16-byte LBO); the generic kernel's 5-D overlapping-stride view (one 64-byte K block
//   per filter row, SWIZZLE_64B) remains as B200RN_STEM_V2=0 and for the allocation-free entry point.
// * Every kernel here is persistent (one CTA -- or one CTA of a 2-CTA cluster -- per SM walking a static tile list)
//   and warp-specialised: warp 0 = TMA producer, warp 1 = TMEM allocation + the single tcgen05.mma-issuing thread,
//   then 8 or 16 epilogue warps (TMEM -> registers -> swizzled shared memory -> TMA store), plus kernel-specific
//   roles (operand-transform warps, store warp, residual producer).  Accumulators are double-buffered in TMEM so
//   a tile's epilogue overlaps the next tile's MMAs.  Kernels in this file and its includes:
//     conv_umma_kernel     general implicit GEMM, compile-time epilogue variants, stationary small weight slabs
//     conv3x3_halo_kernel  3x3/s1 on 64/128 channels, one halo patch per tile, row-shifted A descriptors
//     stem7x7_kernel       the stem (above)
//     conv_dualm_kernel    two M tiles per weight tile (large K, N = 128)
//     conv_pair_kernel     conv_pair.cuh: one 256 x 256 tile per 2-CTA cluster, tcgen05.mma.cta_group::2
//     conv_wide_kernel     conv_wide.cuh: block-ending 1x1 convs, chunk rings, store + residual warps, two-input form
//     conv_tail_kernel     conv_tail.cuh: block-ending 1x1 conv + the next block's 1x1 conv in one launch
//     conv_tf32x3_kernel   conv_tf32.cuh: exact mode, fp32 storage, 3 x TF32 split products, two-level accumulation
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "kernels.h"

namespace rn {
namespace {

constexpr int kEpiWarps = 8;                       // two warps per TMEM lane quarter, 32 columns of a chunk each
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kBaseThreads = 64 + kEpiThreads;     // TMA warp, MMA warp, epilogue warps
constexpr int kThreads = kBaseThreads + 128;       // + 4 operand-transform warps (layers with a prologue only)
constexpr int kTwoGroupThreads = 64 + 2 * kEpiThreads;   // kernels whose epilogue runs as two 8-warp groups on alternating tiles
constexpr int kTileM = 128;
constexpr int kChunkBytes = kTileM * 128;  // one 64-channel output chunk of a tile (16 KB)

struct UmmaParams {
  int num_k_iters, cblks, kw, stride, pad, stem;
  int bw, bh, bn, tiles_w, tiles_h;
  int n_tiles, num_stages, total_tiles, raw_bufs, act_bufs;
  int relaxed_store_wait;   // staging is double-buffered: the previous tile's TMA stores may still be reading
  int has_res, has_out, has_act, relu, act_relu;
  int has_pre, pre_relu, cin;
  int dbg_ashift;   // descriptor experiment (B200RN_DBG_ASHIFT, tools/dbg_ashift.py), 0 in production
  unsigned long long* trace;   // timeline experiment (B200RN_DBG_TRACE): CTA 0 stamps clock64 per tile and actor
  int kgroup;         // K blocks per pipeline stage / barrier (fewer waits and commits for the MMA thread)
  int b_stationary;   // the CTA's whole weight slab (all K blocks of its N tile) is loaded once and stays in smem
  uint32_t a_bytes, b_bytes, rows;
  const float* bias;
  const float* act_scale;
  const float* act_shift;
  const float* pre_scale;
  const float* pre_shift;
  float* out_f32;           // fp32 row output instead of 16-bit tiles (1x1 spatial outputs: row = image)
  int f32_ld, f32_cols, f32_rows;
};

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Set when a barrier wait gives up (a mis-sized TMA transaction would otherwise hang the GPU);
// the host checks it after the stream is synchronised and reports an error instead of results.
__device__ unsigned int g_umma_timeout = 0;

__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, P1;\n\t"
      "}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
// non-blocking probe (mbarrier.test_wait never suspends the thread): used to look ahead at a barrier that is normally
// complete already without stalling on it when it is not
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, P1;\n\t"
      "}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 1000000000LL) {  // ~0.5 s at 2 GHz: never reached by a healthy launch
      atomicExch(&g_umma_timeout, 1u);
      return;
    }
  }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// Programmatic dependent launch: every tcgen05 kernel is launched with programmaticStreamSerializationAllowed, lets
// its successor be scheduled at once (launch_dependents) and only then waits for its predecessor's memory
// (griddepcontrol.wait).  Barrier init, TMEM allocation, descriptor prefetch and the parameter loads of kernel N+1
// thus run while kernel N's last CTAs are still finishing.  Everything that reads or WRITES an activation buffer
// comes after the wait: the arena reuses a dead tensor's memory for the next layer's output.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait_prior_grid() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2, int c3, int c4) {
  asm volatile("cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
}
// L2 prefetch of a box (no shared-memory destination, no barrier): hides HBM latency for data whose landing buffer
// is not free yet
__device__ __forceinline__ void tma_prefetch_4d(const CUtensorMap* m, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global [%0, {%1, %2, %3, %4}];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void tma_store_wait_read_pending() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kPending) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// Same instruction, descriptors given as their low words (start address >> 4 | LBO); the high word (SBO,
// version, swizzle mode) is the same for A and B.  The single issuing thread is latency-bound on its own
// instruction stream, so descriptor arithmetic is kept to one 32-bit add per operand.
__device__ __forceinline__ void umma_f16_lo(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t"
      "}" ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
template <int KB>
__device__ __forceinline__ constexpr uint32_t smem_desc_hi() {
  return ((8u * KB) >> 4) | (1u << 14) | ((KB == 128 ? 2u : 4u) << 29);
}
__device__ __forceinline__ uint32_t smem_desc_lo(uint32_t smem_addr) { return ((smem_addr & 0x3FFFF) >> 4) | (1u << 16); }

__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}
// shared-memory accesses by 32-bit shared-space address (LDS / STS instead of generic loads through a 64-bit pointer)
__device__ __forceinline__ uint4 lds_u4(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_u4(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t"
      ".reg .b32 rx;\n\t"
      ".reg .pred px;\n\t"
      "elect.sync rx|px, %1;\n\t"
      "@px mov.s32 %0, 1;\n\t"
      "}" : "+r"(pred) : "r"(0xffffffffu));
  return pred != 0;
}

// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
//  [0,14) start>>4 | [16,30) LBO>>4 (unused for swizzled K-major) | [32,46) SBO>>4 = 8 rows of the
//  swizzle atom | [46,48) version = 1 | [61,64) layout: 2 = SWIZZLE_128B, 4 = SWIZZLE_64B
template <int KB>
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  constexpr uint64_t layout = (KB == 128) ? 2 : 4;
  constexpr uint64_t sbo = (8 * KB) >> 4;
  return uint64_t((smem_addr & 0x3FFFF) >> 4) | (uint64_t(1) << 16) | (sbo << 32) | (uint64_t(1) << 46) | (layout << 61);
}

template <typename T> struct Fmt;
template <> struct Fmt<__nv_bfloat16> { static constexpr uint32_t id = 1; };
template <> struct Fmt<__half> { static constexpr uint32_t id = 0; };

// cute::UMMA::InstrDescriptor: c_format F32 (bit 4) | a_format [7,10) | b_format [10,13) |
// a/b K-major (bits 15,16 = 0) | N>>3 at [17,23) | M>>4 at [24,29)
template <typename T, int N>
__device__ __forceinline__ constexpr uint32_t make_idesc() {
  return (1u << 4) | (Fmt<T>::id << 7) | (Fmt<T>::id << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(kTileM >> 4) << 24);
}

template <typename T> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
template <typename T> __device__ __forceinline__ float2 unpack2(uint32_t u);
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t u) {
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
}
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t u) {
  return __half22float2(*reinterpret_cast<__half2*>(&u));
}

// ---- the kernel -------------------------------------------------------------------------------
// Persistent: one CTA per SM walks the tile list (tile = blockIdx.x + i*gridDim.x, N-tile fastest so
// CTAs that share an A tile run side by side).  The accumulator is double-buffered in TMEM
// (2 x BLOCK_N columns), so the epilogue of tile i overlaps the MMAs of tile i+1; the residual /
// raw-output staging buffer is double-buffered in shared memory for the same reason.
//
// Dynamic shared memory (1024-byte aligned):
//   [pipeline: num_stages x (A tile | B tile)]
//   [rawbuf:   raw_bufs (1|2) x BLOCK_N/64 x 16 KB]   residual lands here by TMA; the raw output is staged in place
//   [actbuf:   BLOCK_N/64 x 16 KB]       only when has_act
//   [params:   3 x BLOCK_N floats][barriers][tmem slot]
struct TileCoord { int w0, h0, n0, col0, n_tile; };

__device__ __forceinline__ TileCoord decode_tile(const UmmaParams& p, int tile, int block_n) {
  TileCoord t;
  t.n_tile = tile % p.n_tiles;
  int mt = tile / p.n_tiles;
  const int tw = mt % p.tiles_w; mt /= p.tiles_w;
  const int th = mt % p.tiles_h;
  const int tn = mt / p.tiles_h;
  t.w0 = tw * p.bw; t.h0 = th * p.bh; t.n0 = tn * p.bn;
  t.col0 = t.n_tile * block_n;
  return t;
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// EPI >= 0: epilogue flags fixed at compile time (bit 0 ReLU, 1 residual, 2 raw output, 3 activated
// output, 4 ReLU on the activated output) so that unused math is not even issued -- the epilogue warps
// are issue-bound on the memory-bound layers.  EPI < 0: flags read from the parameters (any combination).
enum : int { EPI_RELU = 1, EPI_RES = 2, EPI_OUT = 4, EPI_ACT = 8, EPI_ACT_RELU = 16 };

template <typename T, int BLOCK_N, int KB, int EPI>
__global__ void __launch_bounds__(kThreads, 1)
conv_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmOut,
                 const __grid_constant__ CUtensorMap tmAct, const UmmaParams p) {
  constexpr int kChunks = BLOCK_N / 64;
  constexpr int kMmaPerBlock = KB / 32;        // UMMA_K = 16 elements = 32 bytes
  constexpr uint32_t A_BYTES = kTileM * KB;
  constexpr uint32_t B_BYTES = BLOCK_N * KB;
  // a pipeline stage holds kgroup K blocks: each an A tile and, unless the weights are stationary, its B tile
  const uint32_t SLOT_BYTES = p.b_stationary ? A_BYTES : A_BYTES + B_BYTES;
  const uint32_t STAGE_BYTES = uint32_t(p.kgroup) * SLOT_BYTES;
  constexpr uint32_t RAW_BYTES = kChunks * kChunkBytes;   // one tile's raw / residual staging

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = p.num_stages;
  uint8_t* pipe = smem;
  uint8_t* bstat = pipe + S * STAGE_BYTES;      // [num_k_iters x B tile] when b_stationary
  uint8_t* rawbuf = bstat + (p.b_stationary ? p.num_k_iters * B_BYTES : 0);
  uint8_t* actbuf = rawbuf + ((p.has_out || p.has_res) ? p.raw_bufs * RAW_BYTES : 0);
  uint8_t* tail = actbuf + (p.has_act ? p.act_bufs * RAW_BYTES : 0);
  float* s_bias = reinterpret_cast<float*>(tail);
  float* s_scale = s_bias + BLOCK_N;
  float* s_shift = s_scale + BLOCK_N;
  float* s_pre_scale = s_shift + BLOCK_N;      // [cin] each, only when has_pre
  float* s_pre_shift = s_pre_scale + (p.has_pre ? p.cin : 0);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_pre_shift + (p.has_pre ? p.cin : 0));
  uint64_t* full_bar = bars;                  // [S]  TMA -> MMA (or -> transform warps when has_pre)
  uint64_t* empty_bar = bars + S;             // [S]  MMA -> TMA
  uint64_t* tmem_full = bars + 2 * S;         // [2]  MMA -> epilogue
  uint64_t* tmem_empty = tmem_full + 2;       // [2]  epilogue -> MMA (128 arrivals)
  uint64_t* res_full = tmem_empty + 2;        // [2]  TMA -> epilogue
  uint64_t* res_empty = res_full + 2;         // [2]  epilogue -> TMA
  uint64_t* xf_bar = res_empty + 2;           // [S]  transform warps -> MMA (128 arrivals)
  uint64_t* bstat_bar = xf_bar + S;           // [1]  stationary weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bstat_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool f_relu = EPI >= 0 ? bool(EPI & EPI_RELU) : bool(p.relu);
  const bool f_res = EPI >= 0 ? bool(EPI & EPI_RES) : bool(p.has_res);
  const bool f_out = EPI >= 0 ? bool(EPI & EPI_OUT) : bool(p.has_out);
  const bool f_act = EPI >= 0 ? bool(EPI & EPI_ACT) : bool(p.has_act);
  const bool f_act_relu = EPI >= 0 ? bool(EPI & EPI_ACT_RELU) : bool(p.act_relu);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW);
    if (p.has_res) prefetch_tmap(&tmRes);
    if (p.has_out) prefetch_tmap(&tmOut);
    if (p.has_act) prefetch_tmap(&tmAct);
    for (int i = 0; i < S; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1);
      mbar_init(smem_u32(&empty_bar[i]), 1);
      mbar_init(smem_u32(&xf_bar[i]), 128);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&tmem_full[i]), 1);
      mbar_init(smem_u32(&tmem_empty[i]), kEpiThreads);
      mbar_init(smem_u32(&res_full[i]), 1);
      mbar_init(smem_u32(&res_empty[i]), 1);
    }
    mbar_init(smem_u32(bstat_bar), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 2 * BLOCK_N);
  if (p.has_pre)
    for (int k = threadIdx.x; k < p.cin; k += blockDim.x) {
      s_pre_scale[k] = __ldg(p.pre_scale + k);
      s_pre_shift[k] = __ldg(p.pre_shift + k);
    }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int i = 0;
      if (p.b_stationary && int(blockIdx.x) < p.total_tiles) {
        // gridDim.x is a multiple of n_tiles, so this CTA only ever sees one N tile: fetch its weights once
        const int col0 = (int(blockIdx.x) % p.n_tiles) * BLOCK_N;
        const uint32_t bb = smem_u32(bstat_bar);
        mbar_expect_tx(bb, uint32_t(p.num_k_iters) * p.b_bytes);
        for (int it = 0; it < p.num_k_iters; ++it)
          tma_load_2d(smem_u32(bstat + it * B_BYTES), &tmW, bb, it * (KB / 2), col0);
      }
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++i) {
        const TileCoord tc = decode_tile(p, tile, BLOCK_N);
        for (int it0 = 0; it0 < p.num_k_iters; it0 += p.kgroup) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          if (p.trace && blockIdx.x == 0 && i < 64 && it0 == 0) p.trace[i * 8 + 0] = clock64();
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_expect_tx(fb, uint32_t(p.kgroup) * (p.b_stationary ? p.a_bytes : p.a_bytes + p.b_bytes));
          for (int j = 0; j < p.kgroup; ++j) {
            const int it = it0 + j;
            const uint32_t a_dst = smem_u32(pipe + stage * STAGE_BYTES + j * SLOT_BYTES);
            if (p.stem) {
              tma_load_5d(a_dst, &tmA, fb, 0, tc.w0, it & 1, tc.h0 + (it >> 1), tc.n0);
            } else {
              const int tap = it / p.cblks, cb = it - tap * p.cblks;
              const int ky = tap / p.kw, kx = tap - ky * p.kw;
              tma_load_4d(a_dst, &tmA, fb, cb * 64, tc.w0 * p.stride - p.pad + kx, tc.h0 * p.stride - p.pad + ky, tc.n0);
            }
            if (!p.b_stationary) tma_load_2d(a_dst + A_BYTES, &tmW, fb, it * (KB / 2), tc.col0);
          }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        if (p.has_res) {
          // after this tile's operands, so a late epilogue never delays the main loop
          const int rb = (p.raw_bufs == 2) ? (i & 1) : 0;
          const uint32_t rpar = (p.raw_bufs == 2) ? ((i >> 1) & 1) : (i & 1);
          mbar_wait(smem_u32(&res_empty[rb]), rpar ^ 1);
          const uint32_t bar = smem_u32(&res_full[rb]);
          mbar_expect_tx(bar, p.rows * 128u * kChunks);
#pragma unroll
          for (int c = 0; c < kChunks; ++c)
            tma_load_4d(smem_u32(rawbuf + rb * RAW_BYTES + c * kChunkBytes), &tmRes, bar, tc.col0 + c * 64, tc.w0, tc.h0, tc.n0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one elected thread runs the whole loop (waits, tcgen05.mma, commits) =====
    constexpr uint32_t idesc = make_idesc<T, BLOCK_N>();
    constexpr uint32_t desc_hi = smem_desc_hi<KB>();
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int i = 0;
      const uint32_t pipe_lo = smem_desc_lo(smem_u32(pipe));
      const uint32_t bstat_lo = smem_desc_lo(smem_u32(bstat));
      const uint32_t dbg_off = uint32_t(p.dbg_ashift) * 8;   // descriptor experiment only (tools/dbg_ashift.py)
      if (p.b_stationary && int(blockIdx.x) < p.total_tiles) mbar_wait(smem_u32(bstat_bar), 0);
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++i) {
        const int buf = i & 1;
        const bool tr = p.trace && blockIdx.x == 0 && i < 64;
        if (tr) p.trace[i * 8 + 1] = clock64();
        mbar_wait(smem_u32(&tmem_empty[buf]), ((i >> 1) & 1) ^ 1);   // epilogue drained this accumulator
        tc_fence_after();
        if (tr) p.trace[i * 8 + 2] = clock64();
        const uint32_t tmem_d = tmem_base + uint32_t(buf * BLOCK_N);
        for (int it0 = 0; it0 < p.num_k_iters; it0 += p.kgroup) {
          mbar_wait(smem_u32(p.has_pre ? &xf_bar[stage] : &full_bar[stage]), phase);
          tc_fence_after();
          if (tr && it0 == 0) p.trace[i * 8 + 3] = clock64();
          const uint32_t st_lo = pipe_lo + uint32_t(stage) * (STAGE_BYTES >> 4) + dbg_off;
          for (int j = 0; j < p.kgroup; ++j) {
            const int it = it0 + j;
            const uint32_t a_lo = st_lo + uint32_t(j) * (SLOT_BYTES >> 4);
            const uint32_t b_lo = p.b_stationary ? bstat_lo + uint32_t(it) * (B_BYTES >> 4) : a_lo + (A_BYTES >> 4);
#pragma unroll
            for (int k = 0; k < kMmaPerBlock; ++k)
              umma_f16_lo(tmem_d, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (it | k) != 0 ? 1u : 0u);
          }
          umma_commit(smem_u32(&empty_bar[stage]));                                             // frees the smem slot
          if (it0 + p.kgroup >= p.num_k_iters) umma_commit(smem_u32(&tmem_full[buf]));           // accumulator complete
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        if (tr) p.trace[i * 8 + 4] = clock64();
      }
    }
  } else if (warp >= 2 + kEpiWarps) {
    // ===== operand transform (has_pre only): A' = relu?(pre_scale[c] * A + pre_shift[c]) in place =====
    // (these four warps are only launched for layers with a prologue)
    // thread = one 16-byte channel chunk (8 channels) of every 16th row, so its 8 scale/shift pairs stay in
    // registers for the whole K block; a warp touches 4 complete 128-byte rows per access (conflict-free).
    // Chunk j of row r sits at 16-byte slot (j ^ (r & 7)) of the row (SWIZZLE_128B).
    const int t = threadIdx.x - kBaseThreads;
    const int j = t & 7;
    const int r0 = t >> 3;                   // rows r0, r0+16, ..., r0+112
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      for (int it = 0; it < p.num_k_iters; ++it) {
        const int cb = it % p.cblks;
        const float4 s0 = *reinterpret_cast<const float4*>(s_pre_scale + cb * 64 + j * 8);
        const float4 s1 = *reinterpret_cast<const float4*>(s_pre_scale + cb * 64 + j * 8 + 4);
        const float4 t0 = *reinterpret_cast<const float4*>(s_pre_shift + cb * 64 + j * 8);
        const float4 t1 = *reinterpret_cast<const float4*>(s_pre_shift + cb * 64 + j * 8 + 4);
        mbar_wait(smem_u32(&full_bar[stage]), phase);
        uint8_t* a_tile = pipe + stage * STAGE_BYTES;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = r0 + i * 16;
          uint4* ptr = reinterpret_cast<uint4*>(a_tile + r * 128 + ((j ^ (r & 7)) << 4));
          const uint4 q = *ptr;
          float2 x0 = unpack2<T>(q.x), x1 = unpack2<T>(q.y), x2 = unpack2<T>(q.z), x3 = unpack2<T>(q.w);
          x0.x = fmaf(s0.x, x0.x, t0.x); x0.y = fmaf(s0.y, x0.y, t0.y);
          x1.x = fmaf(s0.z, x1.x, t0.z); x1.y = fmaf(s0.w, x1.y, t0.w);
          x2.x = fmaf(s1.x, x2.x, t1.x); x2.y = fmaf(s1.y, x2.y, t1.y);
          x3.x = fmaf(s1.z, x3.x, t1.z); x3.y = fmaf(s1.w, x3.y, t1.w);
          if (p.pre_relu) {
            x0.x = fmaxf(x0.x, 0.f); x0.y = fmaxf(x0.y, 0.f); x1.x = fmaxf(x1.x, 0.f); x1.y = fmaxf(x1.y, 0.f);
            x2.x = fmaxf(x2.x, 0.f); x2.y = fmaxf(x2.y, 0.f); x3.x = fmaxf(x3.x, 0.f); x3.y = fmaxf(x3.y, 0.f);
          }
          *ptr = make_uint4(pack2<T>(x0.x, x0.y), pack2<T>(x1.x, x1.y), pack2<T>(x2.x, x2.y), pack2<T>(x3.x, x3.y));
        }
        fence_proxy_async_smem();            // generic-proxy writes -> visible to the tensor core's reads
        mbar_arrive(smem_u32(&xf_bar[stage]));
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue: 8 warps; thread = one output pixel (TMEM lane) x 32 of a chunk's 64 channels =====
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    const int g = (warp - 2) >> 2;          // which half of every 64-channel chunk (16-byte slots 4g..4g+3)
    const int row = q * 32 + lane;
    const int sw = row & 7;
    const int et = threadIdx.x - 64;        // 0..255
    const bool store_thread = (warp == 2 && lane == 0);
    int cur_ntile = -1;
    int i = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++i) {
      const TileCoord tc = decode_tile(p, tile, BLOCK_N);
      const int buf = i & 1;
      const uint32_t par = (i >> 1) & 1;
      if (tc.n_tile != cur_ntile) {         // per-channel parameters of this N tile
        cur_ntile = tc.n_tile;
        for (int k = et; k < BLOCK_N; k += kEpiThreads) {
          s_bias[k] = __ldg(p.bias + tc.col0 + k);
          if (f_act) {
            s_scale[k] = __ldg(p.act_scale + tc.col0 + k);
            s_shift[k] = __ldg(p.act_shift + tc.col0 + k);
          }
        }
      }
      if (store_thread && i > 0) {
        // staging buffers about to be rewritten must have been read by their TMA stores: all of them, or --
        // with double-buffered staging -- all but the previous tile's kChunks store groups
        if (p.relaxed_store_wait) tma_store_wait_read_pending<kChunks>();
        else tma_store_wait_read();
        if (f_res) mbar_arrive(smem_u32(&res_empty[(p.raw_bufs == 2) ? ((i - 1) & 1) : 0]));
      }
      const bool etr = p.trace && blockIdx.x == 0 && store_thread && i < 64;
      if (etr) p.trace[i * 8 + 5] = clock64();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      mbar_wait(smem_u32(&tmem_full[buf]), par);
      tc_fence_after();
      if (etr) p.trace[i * 8 + 6] = clock64();
      const int rb = (p.raw_bufs == 2) ? (i & 1) : 0;
      if (f_res) mbar_wait(smem_u32(&res_full[rb]), (p.raw_bufs == 2) ? par : uint32_t(i & 1));
      uint8_t* raw_tile = rawbuf + rb * RAW_BYTES;
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(tmem_base + (uint32_t(q * 32) << 16) + uint32_t(buf * BLOCK_N + c * 64 + g * 32), acc);
        tmem_ld_wait();
        if (c == kChunks - 1) {             // accumulator fully read: hand it back to the MMA warp
          tc_fence_before();
          mbar_arrive(smem_u32(&tmem_empty[buf]));
        }
        uint8_t* raw_stage = raw_tile + c * kChunkBytes;
        uint8_t* act_stage = actbuf + ((p.act_bufs == 2) ? (i & 1) * RAW_BYTES : 0) + c * kChunkBytes;
        uint4* raw_row = reinterpret_cast<uint4*>(raw_stage + row * 128);
        uint4* act_row = reinterpret_cast<uint4*>(act_stage + row * 128);
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {      // 8 channels (16 bytes) at a time
          const int j = g * 4 + jj;
          const int col = c * 64 + j * 8;
          const float4 b0 = *reinterpret_cast<const float4*>(s_bias + col);
          const float4 b1 = *reinterpret_cast<const float4*>(s_bias + col + 4);
          float v[8];
          v[0] = __uint_as_float(acc[jj * 8 + 0]) + b0.x; v[1] = __uint_as_float(acc[jj * 8 + 1]) + b0.y;
          v[2] = __uint_as_float(acc[jj * 8 + 2]) + b0.z; v[3] = __uint_as_float(acc[jj * 8 + 3]) + b0.w;
          v[4] = __uint_as_float(acc[jj * 8 + 4]) + b1.x; v[5] = __uint_as_float(acc[jj * 8 + 5]) + b1.y;
          v[6] = __uint_as_float(acc[jj * 8 + 6]) + b1.z; v[7] = __uint_as_float(acc[jj * 8 + 7]) + b1.w;
          if (f_relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
          }
          const int slot = j ^ sw;            // SWIZZLE_128B: 16-byte chunk index XOR (row mod 8)
          if (f_res) {
            const uint4 r = raw_row[slot];
            const float2 r0 = unpack2<T>(r.x), r1 = unpack2<T>(r.y), r2 = unpack2<T>(r.z), r3 = unpack2<T>(r.w);
            v[0] += r0.x; v[1] += r0.y; v[2] += r1.x; v[3] += r1.y;
            v[4] += r2.x; v[5] += r2.y; v[6] += r3.x; v[7] += r3.y;
          }
          if (p.out_f32) {
            const int img = tc.n0 + row, cg = tc.col0 + col;
            if (row < int(p.rows) && img < p.f32_rows) {
              float* dst = p.out_f32 + size_t(img) * p.f32_ld + cg;
              if (cg + 8 <= p.f32_cols && (p.f32_ld & 3) == 0) {
                *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
                *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
              } else {
#pragma unroll
                for (int k = 0; k < 8; ++k)
                  if (cg + k < p.f32_cols) dst[k] = v[k];
              }
            }
          }
          if (f_out)
            raw_row[slot] = make_uint4(pack2<T>(v[0], v[1]), pack2<T>(v[2], v[3]), pack2<T>(v[4], v[5]), pack2<T>(v[6], v[7]));
          if (f_act) {
            const float4 s0 = *reinterpret_cast<const float4*>(s_scale + col);
            const float4 s1 = *reinterpret_cast<const float4*>(s_scale + col + 4);
            const float4 t0 = *reinterpret_cast<const float4*>(s_shift + col);
            const float4 t1 = *reinterpret_cast<const float4*>(s_shift + col + 4);
            float a[8];
            a[0] = fmaf(s0.x, v[0], t0.x); a[1] = fmaf(s0.y, v[1], t0.y); a[2] = fmaf(s0.z, v[2], t0.z); a[3] = fmaf(s0.w, v[3], t0.w);
            a[4] = fmaf(s1.x, v[4], t1.x); a[5] = fmaf(s1.y, v[5], t1.y); a[6] = fmaf(s1.z, v[6], t1.z); a[7] = fmaf(s1.w, v[7], t1.w);
            if (f_act_relu) {
#pragma unroll
              for (int k = 0; k < 8; ++k) a[k] = fmaxf(a[k], 0.f);
            }
            act_row[slot] = make_uint4(pack2<T>(a[0], a[1]), pack2<T>(a[2], a[3]), pack2<T>(a[4], a[5]), pack2<T>(a[6], a[7]));
          }
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (store_thread) {
          if (f_out) tma_store_4d(&tmOut, smem_u32(raw_stage), tc.col0 + c * 64, tc.w0, tc.h0, tc.n0);
          if (f_act) tma_store_4d(&tmAct, smem_u32(act_stage), tc.col0 + c * 64, tc.w0, tc.h0, tc.n0);
          tma_store_commit();
        }
      }
      if (etr) p.trace[i * 8 + 7] = clock64();
    }
    if (store_thread) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 2 * BLOCK_N);
}

// ---- 3x3 / stride 1 / pad 1 with ONE halo load per tile ------------------------------------------
// For the 64- and 128-channel 3x3 layers the generic kernel is bound by the shared-memory fill rate: it
// fetches the A tile nine times (once per filter tap).  Here a tile is 14 x 8 output pixels of one image,
// laid out as 8 rows of 16 (two junk columns per row), so M row r = y*16 + x.  The (14+2) x (8+2) input
// patch of one 64-channel block is fetched ONCE (TMA zero-fills outside the image = the conv's padding) as
// 160 rows of 128 bytes, and the A operand of tap (ky,kx) is simply that buffer read from row ky*16+kx on:
// row r of the tile then reads patch row (y+ky)*16 + (x+kx).  This works because the tensor core applies the
// 128-byte swizzle to absolute shared-memory addresses, so a descriptor may start at any 128-byte row
// (measured: tools/dbg_ashift.py).  Outputs in the junk columns are computed and never stored.
// Weights stay in shared memory for the CTA's lifetime when the slab fits, else stream through their own ring.
constexpr int kHaloBW = 14, kHaloBH = 8, kHaloPW = 16, kHaloPH = 10;
constexpr uint32_t kHaloLoadBytes = kHaloPW * kHaloPH * 128;     // 160 rows
constexpr uint32_t kHaloStageBytes = 21 * 1024;                   // + junk rows read by the junk columns

struct HaloParams {
  int cblks, n_tiles, total_tiles, tiles_w, tiles_h;
  int halo_stages, b_stages, b_stationary, relu;
  unsigned long long* trace;   // timeline experiment (B200RN_DBG_TRACE): CTA 0 stamps clock64 per tile and actor
  int dbg_skip;   // bottleneck experiment (B200RN_DBG_SKIP): 1 = issue no MMA, 2 = no epilogue math/stores, 4 = one MMA per tap, 8 = taps ignore their kx row shift (atom-aligned descriptors; wrong results)
  const float* bias;
};

template <typename T, int BLOCK_N>
__global__ void __launch_bounds__(kTwoGroupThreads, 1)
conv3x3_halo_kernel(const __grid_constant__ CUtensorMap tmHalo, const __grid_constant__ CUtensorMap tmW,
                    const __grid_constant__ CUtensorMap tmOutRow, const HaloParams p) {
  constexpr int kChunks = BLOCK_N / 64;
  constexpr uint32_t B_BYTES = BLOCK_N * 128;
  constexpr uint32_t OUT_BYTES = kChunks * kChunkBytes;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int HS = p.halo_stages, SB = p.b_stages;
  const int kblocks = 9 * p.cblks;
  uint8_t* halo = smem;
  uint8_t* bbuf = halo + HS * kHaloStageBytes;                     // stationary slab or ring of SB tiles
  uint8_t* outbuf = bbuf + (p.b_stationary ? kblocks : SB) * B_BYTES;  // 2 x [kChunks x 16 KB]
  float* s_bias_all = reinterpret_cast<float*>(outbuf + 2 * OUT_BYTES);     // one copy per epilogue group
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_bias_all + 2 * BLOCK_N);
  uint64_t* halo_full = bars;                 // [HS]
  uint64_t* halo_empty = halo_full + HS;      // [HS]
  uint64_t* b_full = halo_empty + HS;         // [SB] (b_full[0] doubles as "slab landed" when stationary)
  uint64_t* b_empty = b_full + SB;            // [SB]
  uint64_t* tmem_full = b_empty + SB;         // [2]
  uint64_t* tmem_empty = tmem_full + 2;       // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int my_tiles = (int(blockIdx.x) < p.total_tiles) ? (p.total_tiles - 1 - int(blockIdx.x)) / int(gridDim.x) + 1 : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmHalo);
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmOutRow);
    for (int i = 0; i < HS; ++i) { mbar_init(smem_u32(&halo_full[i]), 1); mbar_init(smem_u32(&halo_empty[i]), 1); }
    for (int i = 0; i < SB; ++i) { mbar_init(smem_u32(&b_full[i]), 1); mbar_init(smem_u32(&b_empty[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&tmem_full[i]), 1); mbar_init(smem_u32(&tmem_empty[i]), kEpiThreads); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 2 * BLOCK_N);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  auto tile_coord = [&](int tile, int& x0, int& y0, int& n, int& col0) {
    const int nt = tile % p.n_tiles;
    int mt = tile / p.n_tiles;
    const int tw = mt % p.tiles_w; mt /= p.tiles_w;
    const int th = mt % p.tiles_h;
    n = mt / p.tiles_h;
    x0 = tw * kHaloBW; y0 = th * kHaloBH; col0 = nt * BLOCK_N;
  };

  if (warp == 0) {
    // ===== TMA producer: halo patches run ahead of the weight tiles =====
    if (elect_one() && my_tiles > 0) {
      const int Q = my_tiles * p.cblks;         // items: (tile, channel block)
      if (p.b_stationary) {
        const int col0 = (int(blockIdx.x) % p.n_tiles) * BLOCK_N;
        mbar_expect_tx(smem_u32(&b_full[0]), uint32_t(kblocks) * B_BYTES);
        for (int kb = 0; kb < kblocks; ++kb)
          tma_load_2d(smem_u32(bbuf + kb * B_BYTES), &tmW, smem_u32(&b_full[0]), kb * 64, col0);
      }
      auto issue_halo = [&](int hq) {
        const int tile = int(blockIdx.x) + (hq / p.cblks) * int(gridDim.x);
        const int cb = hq % p.cblks;
        int x0, y0, n, col0;
        tile_coord(tile, x0, y0, n, col0);
        const int slot = hq % HS;
        const uint32_t fb = smem_u32(&halo_full[slot]);
        mbar_expect_tx(fb, kHaloLoadBytes);
        tma_load_4d(smem_u32(halo + slot * kHaloStageBytes), &tmHalo, fb, cb * 64, x0 - 1, y0 - 1, n);
        if (p.trace && blockIdx.x == 0 && hq < 64) p.trace[hq * 8 + 0] = clock64();
      };
      int hq = 0, bstage = 0;
      uint32_t bphase = 0;
      for (int q = 0; q < Q; ++q) {
        // keep up to HS patches in flight without ever blocking on a slot that is still being read
        while (hq < Q && hq < q + HS) {
          const int slot = hq % HS;
          const uint32_t par = (uint32_t(hq / HS) & 1) ^ 1;
          if (hq == q) mbar_wait(smem_u32(&halo_empty[slot]), par);               // this one is needed now
          else if (hq >= HS && !mbar_try_wait(smem_u32(&halo_empty[slot]), par)) break;
          issue_halo(hq);
          ++hq;
        }
        if (!p.b_stationary) {
          const int tile = int(blockIdx.x) + (q / p.cblks) * int(gridDim.x);
          const int cb = q % p.cblks;
          int x0, y0, n, col0;
          tile_coord(tile, x0, y0, n, col0);
          for (int tap = 0; tap < 9; ++tap) {
            mbar_wait(smem_u32(&b_empty[bstage]), bphase ^ 1);
            const uint32_t fb = smem_u32(&b_full[bstage]);
            mbar_expect_tx(fb, B_BYTES);
            tma_load_2d(smem_u32(bbuf + bstage * B_BYTES), &tmW, fb, (tap * p.cblks + cb) * 64, col0);
            if (++bstage == SB) { bstage = 0; bphase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one elected thread =====
    constexpr uint32_t idesc = make_idesc<T, BLOCK_N>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (elect_one()) {
      const uint32_t halo_lo = smem_desc_lo(smem_u32(halo));
      const uint32_t bbuf_lo = smem_desc_lo(smem_u32(bbuf));
      if (p.b_stationary && my_tiles > 0) mbar_wait(smem_u32(&b_full[0]), 0);
      // The issuing thread must not leave the tensor core without queued work between tiles: a tile here is only 36
      // short MMAs (~1600 cycles), and the timeline (B200RN_DBG_TRACE) showed ~900 cycles per tile spent in this thread
      // on barrier polls (~90 cycles each even when already complete), commits and ring arithmetic with nothing issued.
      // So the NEXT unit's barriers (its patch; for a new tile also its accumulator) are probed before the last two
      // taps of the current unit are issued, while earlier taps are still queued -- without blocking: a blocking
      // look-ahead cost the dual-M form 12 % when the next patch had not landed yet -- and the ring position is kept
      // incrementally (no division by the runtime stage count).
      const int units = my_tiles * p.cblks;
      int bstage = 0, slot = 0, i = 0, cb = 0;
      uint32_t bphase = 0, hphase = 0;
      bool ready = false;                        // this unit's barriers were already seen complete by the look-ahead
      for (int q = 0; q < units; ++q) {
        const int buf = i & 1;
        const bool tr = p.trace && blockIdx.x == 0 && q < 64;
        if (!ready) {
          if (cb == 0) mbar_wait(smem_u32(&tmem_empty[buf]), ((uint32_t(i) >> 1) & 1) ^ 1);
          mbar_wait(smem_u32(&halo_full[slot]), hphase);
        }
        tc_fence_after();
        if (tr) p.trace[q * 8 + 2] = clock64();
        const uint32_t tmem_d = tmem_base + uint32_t(buf * BLOCK_N);
        const uint32_t a_lo = halo_lo + uint32_t(slot) * (kHaloStageBytes >> 4);
        const bool last_cb = cb == p.cblks - 1;
        int nslot = slot + 1;
        uint32_t nphase = hphase;
        if (nslot == HS) { nslot = 0; nphase ^= 1; }
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          if (tap == 7) {                       // look ahead (non-blocking: if not complete yet, the top of the next unit waits)
            if (tr) p.trace[q * 8 + 1] = clock64();
            ready = q + 1 < units && mbar_test(smem_u32(&halo_full[nslot]), nphase) &&
                    (!last_cb || mbar_test(smem_u32(&tmem_empty[(i + 1) & 1]), ((uint32_t(i + 1) >> 1) & 1) ^ 1));
            if (tr) p.trace[q * 8 + 7] = clock64();
          }
          uint32_t b_lo;
          if (p.b_stationary) {
            b_lo = bbuf_lo + uint32_t(tap * p.cblks + cb) * (B_BYTES >> 4);
          } else {
            mbar_wait(smem_u32(&b_full[bstage]), bphase);
            tc_fence_after();
            b_lo = bbuf_lo + uint32_t(bstage) * (B_BYTES >> 4);
          }
          // row-shifted view of the patch: tap (ky,kx) starts ky*16 + kx rows (of 128 bytes) further on
          const uint32_t a_tap = a_lo + uint32_t((tap / 3) * kHaloPW + ((p.dbg_skip & 8) ? 0 : (tap % 3))) * 8;   // dbg 8: no kx shift (timing only)
          if (!(p.dbg_skip & 1)) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (k == 0 || !(p.dbg_skip & 4))
                umma_f16_lo(tmem_d, a_tap + 2 * k, b_lo + 2 * k, desc_hi, idesc, (tap | k) != 0 ? 1u : uint32_t(cb != 0));
          }
          if (!p.b_stationary) {
            umma_commit(smem_u32(&b_empty[bstage]));
            if (++bstage == SB) { bstage = 0; bphase ^= 1; }
          }
        }
        if (tr) p.trace[q * 8 + 3] = clock64();
        umma_commit(smem_u32(&halo_empty[slot]));
        if (tr) p.trace[q * 8 + 4] = clock64();
        if (last_cb) umma_commit(smem_u32(&tmem_full[buf]));
        if (tr) p.trace[q * 8 + 5] = clock64();
        slot = nslot; hphase = nphase;
        if (last_cb) { cb = 0; ++i; } else ++cb;
      }
    }
  } else {
    // ===== epilogue: bias + ReLU; the 14 valid pixels of each row are packed densely, one TMA store per chunk =====
    // Two groups of 8 warps; group k takes the tiles with (i & 1) == k, i.e. it owns TMEM accumulator k and staging
    // buffer k.  One tile's epilogue is a ~2300-cycle latency chain (TMEM load, ~150 dependent ALU ops per thread,
    // smem stores, proxy fence, barrier, store issue) that the MMA thread used to wait ~700 cycles per tile for
    // (timeline, B200RN_DBG_TRACE); with two chains in flight it no longer does.
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
    int cur_ntile = -1;
    for (int i = grp; i < my_tiles; i += 2) {
      const int tile = int(blockIdx.x) + i * int(gridDim.x);
      int x0, y0, n, col0;
      tile_coord(tile, x0, y0, n, col0);
      const int buf = grp;
      if (tile % p.n_tiles != cur_ntile) {
        cur_ntile = tile % p.n_tiles;
        for (int k = et; k < BLOCK_N; k += kEpiThreads) s_bias[k] = __ldg(p.bias + col0 + k);
      }
      const bool tr = false;
      if (tr) p.trace[i * 8 + 4] = clock64();
      if (store_thread && i > grp) tma_store_wait_read();   // this group's previous tile has left its staging buffer
      asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
      if (tr) p.trace[i * 8 + 5] = clock64();
      mbar_wait(smem_u32(&tmem_full[buf]), (i >> 1) & 1);
      tc_fence_after();
      if (tr) p.trace[i * 8 + 6] = clock64();
      uint8_t* out_tile = outbuf + buf * OUT_BYTES;
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c) {
        uint32_t acc[32];
        tmem_ld_32x32b_x32(tmem_base + (uint32_t(q4 * 32) << 16) + uint32_t(buf * BLOCK_N + c * 64 + g * 32), acc);
        tmem_ld_wait();
        if (c == kChunks - 1) {
          tc_fence_before();
          mbar_arrive(smem_u32(&tmem_empty[buf]));
        }
        uint8_t* stage = out_tile + c * kChunkBytes;
        uint4* out_row = reinterpret_cast<uint4*>(stage + drow * 128);
        if (valid && !(p.dbg_skip & 2)) {
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
        if (store_thread && !(p.dbg_skip & 2)) {
          tma_store_4d(&tmOutRow, smem_u32(stage), col0 + c * 64, x0, y0, n);   // box {64 ch, 14, 8, 1}
          tma_store_commit();
        }
      }
      if (tr) p.trace[i * 8 + 7] = clock64();
    }
    if (store_thread) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 2 * BLOCK_N);
}

// packed 16-bit max (exact; NaN operands are ignored like fmaxf does) and the lowest finite value, per type
template <typename T> __device__ __forceinline__ uint32_t max2_u32(uint32_t a, uint32_t b);
template <> __device__ __forceinline__ uint32_t max2_u32<__half>(uint32_t a, uint32_t b) {
  __half2 r = __hmax2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
template <> __device__ __forceinline__ uint32_t max2_u32<__nv_bfloat16>(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
template <typename T> __device__ __forceinline__ uint32_t lowest2_u32();
template <> __device__ __forceinline__ uint32_t lowest2_u32<__half>() { return 0xFBFFFBFFu; }
template <> __device__ __forceinline__ uint32_t lowest2_u32<__nv_bfloat16>() { return 0xFF7FFF7Fu; }

// ---- 7x7 / stride 2 / pad 3 stem, one input patch per tile ------------------------------------------
// The generic kernel fetches the stem's A operand as 7 boxes of 128 rows x 64 bytes per tile; the TMA unit
// needs ~4 cycles per such row, which made the stem the slowest layer of the network (measured timeline:
// 3850 cycles per tile, all of it waiting for those 896 rows).  Here a tile is 8 x 16 output pixels
// (M row r = y*8 + x) and the zero-framed fp16 input patch it needs -- 37 rows of 22 pixels x 4 channels =
// 176 bytes -- is fetched ONCE (37 TMA rows).  Both operands use the un-swizzled K-major layout, whose
// "core matrix" is 8 rows x 16 bytes at a 16-byte row pitch: for the A operand of filter row ky the 8 rows
// of a core matrix are 8 consecutive output pixels (input pixels 2x apart = 16 bytes apart), the next 16-byte
// K chunk lies 16 bytes further (LBO = 16: the chunks of neighbouring pixels overlap, which is exactly the
// sliding window), and the next output row is two patch rows down (SBO = 352).  No im2col copy exists even
// in shared memory.  K per filter row = 8 pixels x 4 channels = 32 (two K=16 MMAs); the 8th pixel and the 4th
// channel carry zero weights.  Weights (28 KB per 64 output channels) are loaded once per CTA by a bulk copy.
constexpr int kStemTW = 8, kStemTH = 16;                       // output tile
constexpr int kStemPatchRows = 2 * kStemTH + 5;                // 37
constexpr int kStemPatchPx = 2 * kStemTW + 6;                  // 22
constexpr uint32_t kStemPatchPitch = kStemPatchPx * 8;         // 176 bytes
constexpr uint32_t kStemPatchBytes = kStemPatchRows * kStemPatchPitch;   // 6512
constexpr uint32_t kStemStageBytes = 6656;                     // 128-byte multiple
constexpr uint32_t kStemWBytes = 7 * 4096;                     // one 64-channel N tile of packed weights

struct Stem2Params {
  int tiles_w, tiles_h, n_tiles, total_tiles, stages, relu;
  int groups;                // epilogue warp groups (8 warps each) working on consecutive tiles: 2, or 3 with the fused pool
  int dbg_skip;              // timing experiments only (B200RN_STEM_DBG): 1 no MMAs, 2 no pool phase / store, 4 no TMEM read + staging,
                             // 8 no patch loads, 16 no look-ahead
  unsigned long long* trace; // timeline experiment (B200RN_DBG_TRACE): CTA 0 stamps clock64 per tile and actor
  const float* bias;
  const void* w_packed;
  // Fused 3x3 / stride-2 / pad-1 max-pool (pool != 0): a tile is still 8 x 16 conv outputs, but placed at conv
  // (6*tw - 1, 14*th - 1) so that it holds every tap of a 3 x 7 block of pooled pixels; the conv tile only ever
  // exists in shared memory and the pooled values are written straight to global memory.
  int pool, conv_w, conv_h, pool_w, pool_h, cout, act_relu;
  void* pool_out;            // pooled raw output or null
  void* pool_act;            // relu?(act_scale * pooled + act_shift) or null
  const float* act_scale;
  const float* act_shift;
};

// un-swizzled K-major descriptor: start | LBO (K-adjacent core matrices) | SBO (8-row groups) | version 1
__device__ __forceinline__ uint64_t make_desc_noswz(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return uint64_t((addr & 0x3FFFF) >> 4) | (uint64_t(lbo_bytes >> 4) << 16) | (uint64_t(sbo_bytes >> 4) << 32) | (uint64_t(1) << 46);
}
__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(64 + 3 * kEpiThreads, 1)
stem7x7_kernel(const __grid_constant__ CUtensorMap tmPatch, const __grid_constant__ CUtensorMap tmOut, const Stem2Params p) {
  constexpr int BLOCK_N = 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = p.stages;
  uint8_t* wbuf = smem;                                   // 28 KB, stationary
  const int G = p.groups;                                 // epilogue groups = accumulators = staging buffers (2 or 3)
  uint8_t* outbuf = wbuf + kStemWBytes;                   // G x 16 KB staging (SWIZZLE_128B rows)
  uint8_t* patch = outbuf + G * kChunkBytes;              // S x kStemStageBytes
  float* s_bias = reinterpret_cast<float*>(patch + S * kStemStageBytes);
  float* s_scale = s_bias + BLOCK_N;    // pool mode: the activated output's scale / shift
  float* s_shift = s_scale + BLOCK_N;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_shift + BLOCK_N);
  uint64_t* full_bar = bars;            // [S]
  uint64_t* empty_bar = bars + S;       // [S]
  uint64_t* w_bar = empty_bar + S;      // [1]
  uint64_t* tmem_full = w_bar + 1;      // [G <= 4]
  uint64_t* tmem_empty = tmem_full + 4; // [G <= 4]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int my_tiles = (int(blockIdx.x) < p.total_tiles) ? (p.total_tiles - 1 - int(blockIdx.x)) / int(gridDim.x) + 1 : 0;
  const int n_tile = int(blockIdx.x) % p.n_tiles;         // gridDim.x is a multiple of n_tiles

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmPatch);
    prefetch_tmap(&tmOut);
    for (int i = 0; i < S; ++i) { mbar_init(smem_u32(&full_bar[i]), 1); mbar_init(smem_u32(&empty_bar[i]), 1); }
    mbar_init(smem_u32(w_bar), 1);
    for (int i = 0; i < G; ++i) { mbar_init(smem_u32(&tmem_full[i]), 1); mbar_init(smem_u32(&tmem_empty[i]), kEpiThreads); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), G > 2 ? 4 * BLOCK_N : 2 * BLOCK_N);
  if (warp >= 2)
    for (int k = threadIdx.x - 64; k < BLOCK_N; k += G * kEpiThreads) {
      s_bias[k] = __ldg(p.bias + n_tile * BLOCK_N + k);
      s_scale[k] = (p.pool && p.pool_act) ? __ldg(p.act_scale + n_tile * BLOCK_N + k) : 1.f;
      s_shift[k] = (p.pool && p.pool_act) ? __ldg(p.act_shift + n_tile * BLOCK_N + k) : 0.f;
    }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  // Tile i of this CTA is M tile blockIdx.x / n_tiles + i * (gridDim.x / n_tiles), decomposed as (tw, th, image).  The
  // decomposition is kept incrementally (five runtime divisions per tile cost the producer thread ~800 cycles, more than
  // the tile's MMAs: timeline, B200RN_DBG_TRACE): `TileWalk` holds the coordinates of one tile and the per-step carry.
  struct TileWalk {
    int tw, th, n, sw, sh, sn, tiles_w, tiles_h;
    __device__ __forceinline__ void init(int first, int step, int tiles_w_, int tiles_h_) {
      tiles_w = tiles_w_; tiles_h = tiles_h_;
      tw = first % tiles_w; first /= tiles_w; th = first % tiles_h; n = first / tiles_h;
      sw = step % tiles_w; step /= tiles_w; sh = step % tiles_h; sn = step / tiles_h;
    }
    __device__ __forceinline__ void advance() {
      tw += sw;
      int c = 0;
      if (tw >= tiles_w) { tw -= tiles_w; c = 1; }
      th += sh + c;
      c = 0;
      if (th >= tiles_h) { th -= tiles_h; c = 1; }
      n += sn + c;
    }
  };
  const int walk_first = int(blockIdx.x) / p.n_tiles, walk_step = int(gridDim.x) / p.n_tiles;
  auto tile_origin = [&](const TileWalk& t, int& x0, int& y0) {
    if (p.pool) { x0 = t.tw * 6 - 1; y0 = t.th * 14 - 1; }
    else { x0 = t.tw * kStemTW; y0 = t.th * kStemTH; }
  };

  if (warp == 0) {
    if (elect_one() && my_tiles > 0) {
      mbar_expect_tx(smem_u32(w_bar), kStemWBytes);
      bulk_load_1d(smem_u32(wbuf), static_cast<const uint8_t*>(p.w_packed) + size_t(n_tile) * kStemWBytes, kStemWBytes, smem_u32(w_bar));
      int stage = 0;
      uint32_t phase = 0;
      TileWalk t;
      t.init(walk_first, walk_step, p.tiles_w, p.tiles_h);
      for (int i = 0; i < my_tiles; ++i, t.advance()) {
        int x0, y0;
        tile_origin(t, x0, y0);
        int n = t.n;
        if (p.dbg_skip & 64) {                  // timing experiment: the five divisions the walk replaced
          int mt = (int(blockIdx.x) + i * int(gridDim.x)) / p.n_tiles;
          const int tw = mt % p.tiles_w; mt /= p.tiles_w;
          const int th = mt % p.tiles_h;
          n = mt / p.tiles_h;
          x0 = p.pool ? tw * 6 - 1 : tw * kStemTW; y0 = p.pool ? th * 14 - 1 : th * kStemTH;
        }
        mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
        if (p.trace && blockIdx.x == 0 && i < 64) p.trace[i * 8 + 0] = clock64();
        const uint32_t fb = smem_u32(&full_bar[stage]);
        if (p.dbg_skip & 8) { mbar_arrive(fb); if (++stage == S) { stage = 0; phase ^= 1; } continue; }
        mbar_expect_tx(fb, kStemPatchBytes);
        // frame coordinates: pixel (2*x0, 2*y0) of the zero-framed input = tap (0,0) of output (x0, y0)
        tma_load_3d(smem_u32(patch + stage * kStemStageBytes), &tmPatch, fb, 2 * x0 * 4, 2 * y0, n);
        if (++stage == S) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc<T, BLOCK_N>();
    if (elect_one() && my_tiles > 0) {
      mbar_wait(smem_u32(w_bar), 0);
      const uint32_t w_addr = smem_u32(wbuf);
      // A tile is 14 short MMAs: the next tile's barriers (patch landed, accumulator drained) are probed before the last
      // four are issued, while the earlier ones are still queued, so the tensor core is not left idle over the polls
      // (~90 cycles each even when complete) and commits between tiles; ring and accumulator indices are kept incrementally.
      int stage = 0, buf = 0;
      uint32_t phase = 0, eround = 0;           // eround = i / G
      bool ready = false;
      for (int i = 0; i < my_tiles; ++i) {
        const bool tr = p.trace && blockIdx.x == 0 && i < 64;
        if (tr) p.trace[i * 8 + 1] = clock64();
        if (!ready) {
          mbar_wait(smem_u32(&tmem_empty[buf]), (eround & 1) ^ 1);
          mbar_wait(smem_u32(&full_bar[stage]), phase);
        }
        tc_fence_after();
        if (tr) p.trace[i * 8 + 2] = clock64();
        const uint32_t tmem_d = tmem_base + uint32_t(buf * BLOCK_N);
        const uint32_t a_addr = smem_u32(patch + stage * kStemStageBytes);
        int nstage = stage + 1, nbuf = buf + 1;
        uint32_t nphase = phase, nround = eround;
        if (nstage == S) { nstage = 0; nphase ^= 1; }
        if (nbuf == G) { nbuf = 0; ++nround; }
#pragma unroll
        for (int ky = 0; ky < 7; ++ky) {
          if (ky == 5)                          // look ahead, non-blocking
            ready = !(p.dbg_skip & 16) && i + 1 < my_tiles && mbar_test(smem_u32(&full_bar[nstage]), nphase) &&
                    mbar_test(smem_u32(&tmem_empty[nbuf]), (nround & 1) ^ 1);
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            const uint64_t da = make_desc_noswz(a_addr + ky * kStemPatchPitch + m * 32, 16, 2 * kStemPatchPitch);
            const uint64_t db = make_desc_noswz(w_addr + ky * 4096 + m * 256, 128, 512);
            if (!(p.dbg_skip & 1)) umma_f16(tmem_d, da, db, idesc, (ky | m) != 0 ? 1u : 0u);
          }
        }
        if (tr) p.trace[i * 8 + 3] = clock64();
        // one commit per tile (each costs this thread ~95 cycles): the epilogue group that sees the accumulator complete
        // hands the patch slot back to the producer
        if (p.dbg_skip & 32) umma_commit(smem_u32(&empty_bar[stage]));
        umma_commit(smem_u32(&tmem_full[buf]));
        if (tr) p.trace[i * 8 + 4] = clock64();
        stage = nstage; phase = nphase; buf = nbuf; eround = nround;
      }
    }
  } else {
    // two groups of 8 epilogue warps on alternating tiles (group k owns accumulator k and staging buffer k): a tile's
    // epilogue is a ~2000-cycle latency chain against ~700 cycles of MMAs, so one chain at a time set the pace
    const int grp = (warp - 2) >> 3;
    const int q4 = warp & 3;
    const int g = ((warp - 2) & 7) >> 2;
    const int row = q4 * 32 + lane;            // = y*8 + x, the order the {64, 8, 16, 1} store box expects
    const int sw = row & 7;
    const bool store_thread = (((warp - 2) & 7) == 0 && lane == 0);
    const int bar_id = 1 + grp;
    TileWalk t;
    t.init(walk_first + grp * walk_step, G * walk_step, p.tiles_w, p.tiles_h);
    int pstage = grp % S;                      // patch slot of tile i = i % S
    uint32_t round = 0;                        // i / G
    for (int i = grp; i < my_tiles; i += G, t.advance(), ++round) {
      int x0, y0;
      tile_origin(t, x0, y0);
      const int n = t.n;
      const int buf = grp;
      if (store_thread && i > grp) tma_store_wait_read();   // this group's previous tile has left its staging buffer
      const bool etr = p.trace && blockIdx.x == 0 && store_thread && i < 64;
      if (etr) p.trace[i * 8 + 5] = clock64();
      asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
      mbar_wait(smem_u32(&tmem_full[buf]), round & 1);
      tc_fence_after();
      if (store_thread && !(p.dbg_skip & 32)) mbar_arrive(smem_u32(&empty_bar[pstage]));   // the tile's MMAs have read its patch
      pstage += G;
      while (pstage >= S) pstage -= S;
      if (etr) p.trace[i * 8 + 6] = clock64();
      uint32_t acc[32];
      if (!(p.dbg_skip & 4)) {
        tmem_ld_32x32b_x32(tmem_base + (uint32_t(q4 * 32) << 16) + uint32_t(buf * BLOCK_N + g * 32), acc);
        tmem_ld_wait();
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&tmem_empty[buf]));
      uint8_t* stage_out = outbuf + buf * kChunkBytes;
      uint4* out_row = reinterpret_cast<uint4*>(stage_out + row * 128);
#pragma unroll
      for (int jj = 0; jj < 4 && !(p.dbg_skip & 4); ++jj) {
        const int j = g * 4 + jj;
        const float4 b0 = *reinterpret_cast<const float4*>(s_bias + j * 8);
        const float4 b1 = *reinterpret_cast<const float4*>(s_bias + j * 8 + 4);
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
      fence_proxy_async_smem();
      asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
      if (p.dbg_skip & 2) {
      } else if (p.pool) {
        // 3 x 7 pooled pixels x eight 16-byte channel chunks = 168 work items for the group's 256 threads; the conv
        // tile sits in the staging buffer as 128 rows (y*8 + x) of 128 swizzled bytes
        const int tg = int(threadIdx.x) - 64 - grp * kEpiThreads;
        if (tg < 168) {
          const int pp = tg >> 3, j = tg & 7;
          const int pyl = pp / 3, pxl = pp - pyl * 3;
          const int px = (x0 + 1) / 2 + pxl, py = (y0 + 1) / 2 + pyl;
          if (px < p.pool_w && py < p.pool_h) {
            uint32_t best[4] = {lowest2_u32<T>(), lowest2_u32<T>(), lowest2_u32<T>(), lowest2_u32<T>()};
            const uint32_t st = smem_u32(stage_out);
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
              const int yl = 2 * pyl + dy, gy = y0 + yl;
              if (gy < 0 || gy >= p.conv_h) continue;
#pragma unroll
              for (int dx = 0; dx < 3; ++dx) {
                const int xl = 2 * pxl + dx, gx = x0 + xl;
                if (gx < 0 || gx >= p.conv_w) continue;
                const int r = yl * 8 + xl;
                const uint4 v = lds_u4(st + uint32_t(r) * 128u + uint32_t((j ^ (r & 7)) << 4));
                best[0] = max2_u32<T>(best[0], v.x); best[1] = max2_u32<T>(best[1], v.y);
                best[2] = max2_u32<T>(best[2], v.z); best[3] = max2_u32<T>(best[3], v.w);
              }
            }
            const size_t off = ((size_t(n) * p.pool_h + py) * p.pool_w + px) * size_t(p.cout) + size_t(n_tile * BLOCK_N + j * 8);
            if (p.pool_out) *reinterpret_cast<uint4*>(static_cast<T*>(p.pool_out) + off) = make_uint4(best[0], best[1], best[2], best[3]);
            if (p.pool_act) {
              uint32_t a[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float2 f = unpack2<T>(best[k]);
                float a0 = fmaf(s_scale[j * 8 + 2 * k], f.x, s_shift[j * 8 + 2 * k]);
                float a1 = fmaf(s_scale[j * 8 + 2 * k + 1], f.y, s_shift[j * 8 + 2 * k + 1]);
                if (p.act_relu) { a0 = fmaxf(a0, 0.f); a1 = fmaxf(a1, 0.f); }
                a[k] = pack2<T>(a0, a1);
              }
              *reinterpret_cast<uint4*>(static_cast<T*>(p.pool_act) + off) = make_uint4(a[0], a[1], a[2], a[3]);
            }
          }
        }
      } else if (store_thread) {
        tma_store_4d(&tmOut, smem_u32(stage_out), n_tile * BLOCK_N, x0, y0, n);   // box {64 ch, 8, 16, 1}
        tma_store_commit();
      }
      if (etr) p.trace[i * 8 + 7] = clock64();
    }
    if (store_thread) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, G > 2 ? 4 * BLOCK_N : 2 * BLOCK_N);
}

// ---- two M tiles per weight tile ("dual-M") ----------------------------------------------------------
// The large-K layers of stages 3 and 4 (1x1 1024->256, 3x3 256->256, 1x1 2048->512, ...) are bound by the
// L2 -> shared-memory feed: every 128-pixel tile streams the whole weight slab of its N tile again
// (measured ~7 TB/s aggregate, profiles/).  Here one CTA works on TWO M tiles at once: each pipeline stage holds
// A0 | A1 | B and the B tile is used by eight MMAs (four into accumulator D0, four into D1), so the bytes
// staged per output tile drop from A+B to A+B/2 (1.5x less at N = 256).  Both accumulators fill TMEM
// (2 x BLOCK_N columns), so the epilogue of an item does not overlap the next item's MMAs; with K >= 512 the
// main loop dominates.  Epilogue: bias + optional ReLU, one output (the only form these layers need).
struct DualParams {
  int num_k_iters, cblks, kw, stride, pad;
  int bw, bh, bn, tiles_w, tiles_h, m_tiles, n_tiles, total_items, num_stages, relu;
  uint32_t a_bytes, b_bytes;
  const float* bias;
};

template <typename T, int BLOCK_N>
__global__ void __launch_bounds__(kTwoGroupThreads, 1)
conv_dualm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                  const __grid_constant__ CUtensorMap tmOut, const DualParams p) {
  constexpr int kChunks = BLOCK_N / 64;
  constexpr uint32_t A_BYTES = kTileM * 128;
  constexpr uint32_t B_BYTES = BLOCK_N * 128;
  constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + B_BYTES;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int S = p.num_stages;
  uint8_t* pipe = smem;
  uint8_t* outbuf = pipe + S * STAGE_BYTES;        // two 16 KB chunk buffers, one per epilogue group
  float* s_bias = reinterpret_cast<float*>(outbuf + 2 * kChunkBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_bias + BLOCK_N);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + S;
  uint64_t* tmem_full = empty_bar + S;
  uint64_t* tmem_empty = tmem_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int my_items = (int(blockIdx.x) < p.total_items) ? (p.total_items - 1 - int(blockIdx.x)) / int(gridDim.x) + 1 : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmOut);
    for (int i = 0; i < S; ++i) { mbar_init(smem_u32(&full_bar[i]), 1); mbar_init(smem_u32(&empty_bar[i]), 1); }
    mbar_init(smem_u32(tmem_full), 1);
    mbar_init(smem_u32(tmem_empty), 2 * kEpiThreads);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 2 * BLOCK_N);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait_prior_grid();

  // M tile index -> patch origin; an index past the last tile lands beyond the batch, where TMA loads read
  // zeros and TMA stores are clipped
  auto m_coord = [&](int mt, int& w0, int& h0, int& n0) {
    const int tw = mt % p.tiles_w; mt /= p.tiles_w;
    const int th = mt % p.tiles_h;
    w0 = tw * p.bw; h0 = th * p.bh; n0 = (mt / p.tiles_h) * p.bn;
  };

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_items; ++i) {
        const int item = int(blockIdx.x) + i * int(gridDim.x);
        const int col0 = (item % p.n_tiles) * BLOCK_N;
        const int pm = item / p.n_tiles;
        int w0[2], h0[2], n0[2];
        m_coord(2 * pm, w0[0], h0[0], n0[0]);
        m_coord(2 * pm + 1, w0[1], h0[1], n0[1]);
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          const uint32_t dst = smem_u32(pipe + stage * STAGE_BYTES);
          mbar_expect_tx(fb, 2 * p.a_bytes + p.b_bytes);
          const int tap = it / p.cblks, cb = it - tap * p.cblks;
          const int ky = tap / p.kw, kx = tap - ky * p.kw;
#pragma unroll
          for (int h = 0; h < 2; ++h)
            tma_load_4d(dst + h * A_BYTES, &tmA, fb, cb * 64, w0[h] * p.stride - p.pad + kx, h0[h] * p.stride - p.pad + ky, n0[h]);
          tma_load_2d(dst + 2 * A_BYTES, &tmW, fb, it * 64, col0);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc<T, BLOCK_N>();
    constexpr uint32_t desc_hi = smem_desc_hi<128>();
    if (elect_one()) {
      const uint32_t pipe_lo = smem_desc_lo(smem_u32(pipe));
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_items; ++i) {
        mbar_wait(smem_u32(tmem_empty), (uint32_t(i) & 1) ^ 1);      // epilogue has drained both accumulators
        tc_fence_after();
        for (int it = 0; it < p.num_k_iters; ++it) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t a_lo = pipe_lo + uint32_t(stage) * (STAGE_BYTES >> 4);
          const uint32_t b_lo = a_lo + 2 * (A_BYTES >> 4);
#pragma unroll
          for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_f16_lo(tmem_base + uint32_t(h * BLOCK_N), a_lo + h * (A_BYTES >> 4) + 2 * k, b_lo + 2 * k, desc_hi, idesc,
                          (it | k) != 0 ? 1u : 0u);
          umma_commit(smem_u32(&empty_bar[stage]));
          if (it == p.num_k_iters - 1) umma_commit(smem_u32(tmem_full));
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // Two groups of 8 epilogue warps: group h drains accumulator h (M tile h of the item) through its own staging
    // buffer.  Nothing overlaps this epilogue (both accumulators are in use), so halving its length is a direct gain:
    // 8 chunks in lock step were ~12k cycles beside ~37k of MMAs on the 3x3 @14x14 layers.
    const int grp = (warp - 2) >> 3;
    const int q4 = warp & 3;
    const int g = ((warp - 2) & 7) >> 2;
    const int row = q4 * 32 + lane;
    const int sw = row & 7;
    const int et = threadIdx.x - 64;             // 0..511 over both groups
    const bool store_thread = (((warp - 2) & 7) == 0 && lane == 0);
    const int bar_id = 1 + grp;
    uint8_t* stage_out = outbuf + grp * kChunkBytes;
    int cur_ntile = -1;
    int cc = 0;                                  // this group's running chunk counter
    for (int i = 0; i < my_items; ++i) {
      const int item = int(blockIdx.x) + i * int(gridDim.x);
      const int n_tile = item % p.n_tiles;
      const int col0 = n_tile * BLOCK_N;
      const int pm = item / p.n_tiles;
      if (n_tile != cur_ntile) {
        cur_ntile = n_tile;
        // both groups walk the same item list, so they meet here: nobody still reads the previous bias set
        asm volatile("bar.sync 3, 512;" ::: "memory");
        for (int k = et; k < BLOCK_N; k += 2 * kEpiThreads) s_bias[k] = __ldg(p.bias + col0 + k);
        asm volatile("bar.sync 3, 512;" ::: "memory");
      }
      mbar_wait(smem_u32(tmem_full), uint32_t(i) & 1);
      tc_fence_after();
      {
        const int h = grp;
        int w0, h0, n0;
        m_coord(2 * pm + h, w0, h0, n0);
#pragma unroll 1
        for (int c = 0; c < kChunks; ++c, ++cc) {
          uint32_t acc[32];
          tmem_ld_32x32b_x32(tmem_base + (uint32_t(q4 * 32) << 16) + uint32_t(h * BLOCK_N + c * 64 + g * 32), acc);
          tmem_ld_wait();
          if (c == kChunks - 1) {
            tc_fence_before();
            mbar_arrive(smem_u32(tmem_empty));
          }
          if (store_thread && cc > 0) tma_store_wait_read();   // the previous chunk's store has read the staging buffer
          asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
          uint4* out_row = reinterpret_cast<uint4*>(stage_out + row * 128);
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
          fence_proxy_async_smem();
          asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
          if (store_thread) {
            tma_store_4d(&tmOut, smem_u32(stage_out), col0 + c * 64, w0, h0, n0);
            tma_store_commit();
          }
        }
      }
    }
    if (store_thread) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 2 * BLOCK_N);
}

#include "conv_halo2.cuh"
#include "conv_tail.cuh"
#include "conv_wide.cuh"
#include "conv_pair.cuh"
#include "conv_pairwide.cuh"
#include "conv_tf32.cuh"

// ---- host side --------------------------------------------------------------------------------
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// swz: 0 = SWIZZLE_128B, 1 = SWIZZLE_64B, 2 = none
bool encode(CUtensorMap* m, DType dt, int rank, const void* base, const uint64_t* dims, const uint64_t* strides_bytes,
            const uint32_t* box, const uint32_t* estr, int swz, std::string& err) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { err = "cuTensorMapEncodeTiled is not available from the driver"; return false; }
  cuuint64_t d[5]; cuuint64_t s[4]; cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = estr[i]; }
  for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
  CUresult r = fn(m, dt == DType::BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                  cuuint32_t(rank), const_cast<void*>(base), d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swz == 1 ? CU_TENSOR_MAP_SWIZZLE_64B : (swz == 2 ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B),
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    std::snprintf(buf, sizeof buf, "cuTensorMapEncodeTiled failed (CUresult %d) rank=%d dims=[%llu,%llu,%llu,%llu] box=[%u,%u,%u,%u]",
                  int(r), rank, (unsigned long long)dims[0], (unsigned long long)dims[1],
                  (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0), box[0], box[1],
                  rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    err = buf;
    return false;
  }
  return true;
}

// NHWC activation [B,H,W,C] as a 4-D map (C, W, H, N) with a {64, bw*es, bh*es, bn} box
bool encode_act(CUtensorMap* m, DType dt, const void* base, int B, int H, int W, int C, int bw, int bh, int bn, int es,
                std::string& err) {
  uint64_t dims[4] = {uint64_t(C), uint64_t(W), uint64_t(H), uint64_t(B)};
  uint64_t str[3] = {uint64_t(C) * 2, uint64_t(W) * C * 2, uint64_t(H) * W * C * 2};
  uint32_t box[4] = {64, uint32_t(bw * es), uint32_t(bh * es), uint32_t(bn)};
  uint32_t estr[4] = {1, uint32_t(es), uint32_t(es), 1};
  return encode(m, dt, 4, base, dims, str, box, estr, false, err);
}

}  // namespace

// Stem frame geometry (fast path): image at (3,3) inside [Hp, Wp] with zero borders.
static inline int stem_Hp(int H) { return H + 6; }
static inline int stem_Wp(int W) { return ((W + 3 + 5) + 7) / 8 * 8; }

struct UmmaConvPlan {
  DType dt;
  ConvGeom g;
  int stem;
  int block_n, kb;
  UmmaParams p;
  CUtensorMap tmA, tmW, tmRes, tmOut, tmAct;
  uint32_t smem_bytes;
  int grid;
  int halo = 0;           // 3x3/s1/p1 layer routed to conv3x3_halo_kernel (tmA = patch map, tmOut = row-store map)
  HaloParams hp;
  int halo2 = 0;          // ... to conv3x3_halo2_kernel (two M tiles per streamed weight tile) when the slab does not fit
  Halo2Params h2;
  int dual = 0;           // large-K layer routed to conv_dualm_kernel (two M tiles per weight tile)
  DualParams dp;
  int pair = 0;           // large-K layer routed to conv_pair_kernel (CTA pair, tcgen05.mma.cta_group::2, M = 256)
  PairParams pp;
  int stem2 = 0;          // stem routed to stem7x7_kernel (tmA = patch map, tmOut = {64, 8, 16, 1} store map)
  Stem2Params sp;
  int wide = 0;           // epilogue-bound 1x1 conv with a 256-wide N tile routed to conv_wide_kernel
  int pairwide = 0;       // ... and, when its K is large and its weights stream, run as a CTA pair (conv_pairwide_kernel)
  WideParams wp;
  int tail = 0;           // 1x1 conv + residual fused with the following 1x1 conv (conv_tail_kernel)
  TailParams tp;
  CUtensorMap tmW1, tmOut1;
  CUtensorMap tmA2;       // second input of a two-input conv (wide / tail kernels)
  int tp_n1 = 0;
};

bool umma_conv_supported(const ConvGeom& g, int stem_mode, char* why, size_t cap) {
  auto no = [&](const char* m) { if (why && cap) std::snprintf(why, cap, "%s", m); return false; };
  if (stem_mode) {
    if (!(g.kh == 7 && g.kw == 7 && g.stride == 2 && g.pad == 3 && g.Cin == 4 && g.Cout % 64 == 0))
      return no("stem path needs a 7x7/2 pad-3 conv on a 4-channel (3 padded) input with Cout % 64 == 0");
    return true;
  }
  if (g.Cin % 64 || g.Cout % 64) return no("tensor-core conv needs Cin % 64 == 0 and Cout % 64 == 0");
  if (g.stride != 1 && g.stride != 2) return no("tensor-core conv supports stride 1 or 2");
  if (g.kh != g.kw) return no("tensor-core conv needs a square filter");
  return true;
}

size_t umma_weight_bytes(const ConvGeom& g, int stem_mode) {
  if (stem_mode) return size_t(g.Cout) * 7 * 32 * 2;
  return size_t(g.Cout) * g.kh * g.kw * g.Cin * 2;
}

template <typename T> static T host_cvt(float f);
template <> __nv_bfloat16 host_cvt<__nv_bfloat16>(float f) { return __float2bfloat16_rn(f); }
template <> __half host_cvt<__half>(float f) { return __float2half_rn(f); }

template <typename T>
static void pack_weights_t(const float* w, const ConvGeom& g, int stem_mode, T* dst) {
  if (!stem_mode) {
    size_t n = size_t(g.Cout) * g.kh * g.kw * g.Cin;   // already [Cout][ky][kx][Cin] = K-major
    for (size_t i = 0; i < n; ++i) dst[i] = host_cvt<T>(w[i]);
    return;
  }
  if (stem_mode == 2) {
    // stem7x7_kernel: per 64-channel N tile, per filter row ky a 4 KB block of un-swizzled K-major core matrices
    // [n_group 0..7][k_core 0..3][8 rows x 16 bytes]; k = kx*4 + c (kx = 7 and c = 3 are zero)
    std::memset(dst, 0, size_t(g.Cout) * 7 * 32 * sizeof(T));
    for (int o = 0; o < g.Cout; ++o)
      for (int ky = 0; ky < 7; ++ky)
        for (int kx = 0; kx < 7; ++kx)
          for (int c = 0; c < 4; ++c) {
            const int k = kx * 4 + c;
            const size_t byte = size_t(o / 64) * (7 * 4096) + size_t(ky) * 4096 + size_t((o % 64) / 8) * 512 + size_t(k / 8) * 128 +
                                size_t(o % 8) * 16 + size_t(k % 8) * 2;
            dst[byte / 2] = host_cvt<T>(w[((size_t(o) * 7 + ky) * 7 + kx) * g.Cin + c]);
          }
    return;
  }
  // stem: [Cout][ky 0..6][kx 0..7][c 0..3]; kx = 7 is zero (8-pixel granule), c = 3 is the pad channel
  for (int o = 0; o < g.Cout; ++o)
    for (int ky = 0; ky < 7; ++ky)
      for (int kx = 0; kx < 8; ++kx)
        for (int c = 0; c < 4; ++c) {
          float v = (kx < 7) ? w[((size_t(o) * 7 + ky) * 7 + kx) * g.Cin + c] : 0.f;
          dst[((size_t(o) * 7 + ky) * 8 + kx) * 4 + c] = host_cvt<T>(v);
        }
}

void umma_pack_weights(const float* w, const ConvGeom& g, DType dt, int stem_mode, void* dst) {
  if (dt == DType::BF16) pack_weights_t<__nv_bfloat16>(w, g, stem_mode, static_cast<__nv_bfloat16*>(dst));
  else pack_weights_t<__half>(w, g, stem_mode, static_cast<__half*>(dst));
}

// pick the spatial patch (bw x bh of bn images) that wastes the fewest MMA rows
static void choose_patch(int B, int Ho, int Wo, int& bw, int& bh, int& bn) {
  double best = -1;
  for (int w = 1; w <= Wo && w <= kTileM; ++w)
    for (int h = 1; h <= Ho && w * h <= kTileM; ++h) {
      int n = kTileM / (w * h);
      if (n > B) n = B;
      if (n > 256) n = 256;
      long tiles = long((Wo + w - 1) / w) * ((Ho + h - 1) / h) * ((B + n - 1) / n);
      double eff = double(B) * Ho * Wo / (double(tiles) * kTileM);
      // prefer wider patches on ties (longer contiguous runs per TMA box)
      if (eff > best + 1e-9 || (eff > best - 1e-9 && w > bw)) { best = eff; bw = w; bh = h; bn = n; }
    }
}

static int sm_count() {
  static int n[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (!n[dev]) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    n[dev] = v;
  }
  return n[dev];
}

// Tile geometry, pipeline depth and launch shape of one layer at one batch size.  Needs pl.dt, pl.g,
// pl.stem and the has_* flags of pl.p; fills everything else except the tensor maps.
static void plan_geometry(UmmaConvPlan& pl) {
  const ConvGeom& g = pl.g;
  UmmaParams& p = pl.p;
  const int stem_mode = pl.stem;
  pl.kb = stem_mode ? 64 : 128;
  int bw = 0, bh = 0, bn = 0;
  choose_patch(g.B, g.Ho, g.Wo, bw, bh, bn);
  p.bw = bw; p.bh = bh; p.bn = bn;
  p.tiles_w = (g.Wo + bw - 1) / bw;
  p.tiles_h = (g.Ho + bh - 1) / bh;
  const int tiles_n = (g.B + bn - 1) / bn;
  const int m_tiles = p.tiles_w * p.tiles_h * tiles_n;
  const int sms = sm_count();
  // N-tile width: minimise (rounds over the SMs) x (operand bytes staged per K block), i.e. the
  // shared-memory fill traffic per SM, among the widths that divide Cout and leave >= 3 stages.
  const uint32_t budget = 227u * 1024;
  auto fixed_bytes = [&](int bn, int nb, int nab = 1) {
    const uint32_t raw_bytes = uint32_t(bn / 64) * kChunkBytes;
    return 1024u + ((p.has_out || p.has_res) ? uint32_t(nb) * raw_bytes : 0u) + (p.has_act ? uint32_t(nab) * raw_bytes : 0u) +
           3u * uint32_t(bn) * 4 + (p.has_pre ? 8u * uint32_t(g.Cin) : 0u) + 8u * (3 * 8 + 9) + 64;
  };
  auto stage_bytes_of = [&](int bn) { return uint32_t(kTileM * pl.kb + bn * pl.kb); };
  int best_bn = 64;
  long best_cost = -1;
  for (int bn_c : {64, 128, 256}) {
    if (g.Cout % bn_c) continue;
    if (stem_mode && bn_c != 64) continue;
    const int nb_c = p.has_res ? 2 : 1;
    const uint32_t fx = std::min(fixed_bytes(bn_c, nb_c), fixed_bytes(bn_c, 1));
    if (bn_c > 64 && (budget < fx || (budget - fx) / stage_bytes_of(bn_c) < 3)) continue;
    const long tiles = long(m_tiles) * (g.Cout / bn_c);
    const long cost = ((tiles + sms - 1) / sms) * long(kTileM + bn_c);
    if (best_cost < 0 || cost <= best_cost) { best_cost = cost; best_bn = bn_c; }
  }
  pl.block_n = best_bn;
  p.n_tiles = g.Cout / pl.block_n;
  p.total_tiles = m_tiles * p.n_tiles;
  p.rows = uint32_t(bw * bh * bn);
  p.stem = stem_mode; p.kw = g.kw; p.stride = g.stride; p.pad = g.pad;
  p.cblks = stem_mode ? 1 : g.Cin / 64;
  p.num_k_iters = stem_mode ? 7 : g.kh * g.kw * p.cblks;
  p.a_bytes = p.rows * uint32_t(pl.kb);
  p.b_bytes = uint32_t(pl.block_n * pl.kb);
  const uint32_t stage_bytes = stage_bytes_of(pl.block_n);
  // double-buffer the residual/raw staging when a residual is prefetched and it still leaves 3 stages
  p.raw_bufs = (p.has_res && budget > fixed_bytes(pl.block_n, 2) &&
                (budget - fixed_bytes(pl.block_n, 2)) / stage_bytes >= 3) ? 2 : 1;
  p.act_bufs = 1;
  p.relaxed_store_wait = 0;
  // no residual: double-buffer the output staging instead, so a tile's epilogue never waits for the previous
  // tile's TMA stores to finish reading (matters for the many-small-tiles layers: stem, 64-channel convs)
  if (!p.has_res && budget > fixed_bytes(pl.block_n, 2, 2) && (budget - fixed_bytes(pl.block_n, 2, 2)) / stage_bytes >= 3) {
    p.raw_bufs = 2; p.act_bufs = 2; p.relaxed_store_wait = 1;
  }
  uint32_t fixed = fixed_bytes(pl.block_n, p.raw_bufs, p.act_bufs);
  pl.grid = std::min(p.total_tiles, sms);
  // Stationary weights: when the N tile's whole weight slab is small, keep it in shared memory for the CTA's
  // lifetime instead of re-fetching a B tile with every A tile (the grid is trimmed to a multiple of n_tiles so
  // that a CTA always works on the same N tile).
  const uint32_t slab = uint32_t(p.num_k_iters) * p.b_bytes;
  const uint32_t a_only = uint32_t(kTileM * pl.kb);
  const int grid_bs = (pl.grid / p.n_tiles) * p.n_tiles;
  const char* bs_env = std::getenv("B200RN_BSTAT");
  p.b_stationary = 0;
  uint32_t stage_eff = stage_bytes;
  if (!(bs_env && bs_env[0] == '0') && slab <= 72u * 1024 && grid_bs >= 1 && grid_bs * 8 >= pl.grid * 7 &&
      p.total_tiles >= 4 * pl.grid && budget > fixed + slab && (budget - fixed - slab) / a_only >= 3) {
    p.b_stationary = 1;
    fixed += slab;
    stage_eff = a_only;
    pl.grid = grid_bs;
  }
  // K blocks per stage: the largest divisor of the K-block count that keeps a stage <= 64 KB and >= 3 stages
  p.kgroup = 1;
  const char* kg_env = std::getenv("B200RN_KGROUP");
  if (!p.has_pre && !(kg_env && kg_env[0] == '0'))
    for (int kg = std::min(8, p.num_k_iters); kg > 1; --kg)
      if (p.num_k_iters % kg == 0 && uint32_t(kg) * stage_eff <= 64u * 1024 && (budget - fixed) / (uint32_t(kg) * stage_eff) >= 3) {
        p.kgroup = kg;
        break;
      }
  stage_eff *= uint32_t(p.kgroup);
  int stages = int((budget - fixed) / stage_eff);
  stages = std::max(2, std::min(stages, 8));
  p.num_stages = stages;
  pl.smem_bytes = fixed + uint32_t(stages) * stage_eff;
}

// CTA-pair kernel switches: B200RN_PAIR=0 keeps the single-SM kernels; B200RN_PAIR_MINK = smallest K-block count routed there
static bool pair_enabled() {
  const char* v = std::getenv("B200RN_PAIR");
  return !(v && v[0] == '0');
}
// Measured at batch 256 (profiles/r2_pair_ab.md): with the 256-wide N tile the pair kernel wins on every large-K layer
// (3x3 @14x14 0.062 -> 0.049 ms, 1x1 1024->256 0.043 -> 0.038), with N = 128 (Cout = 128: a CTA stages a 64-row
// weight half per 128-row A tile, 96 B/clk/SM of L2 feed like the single-SM kernels) it loses 0.054 -> 0.067 ms, so
// Cout % 256 != 0 layers stay on the single-SM kernels unless B200RN_PAIR=2 (tests).
static bool pair_n_ok(int cout) {
  const char* v = std::getenv("B200RN_PAIR");
  return cout % 256 == 0 || (v && v[0] == '2' && cout % 128 == 0);
}
static int pair_min_k() {
  const char* v = std::getenv("B200RN_PAIR_MINK");
  return v ? std::max(1, std::atoi(v)) : 8;
}
// With few pair tiles most SMs idle while a handful of clusters walk the whole K: at batch 8 the single-SM kernels (twice
// the CTAs per layer) run the step in 0.560 ms against 0.593, at batch 32 it is a tie (0.791 / 0.782), from 64 on the pair
// kernel wins (1.120 / 1.194).  B200RN_PAIR_MINITEMS (default 24 pair tiles x N tiles); B200RN_PAIR=2 ignores it (tests).
static bool pair_items_ok(long m_tiles, int cout) {
  const char* v = std::getenv("B200RN_PAIR");
  if (v && v[0] == '2') return true;
  const char* t = std::getenv("B200RN_PAIR_MINITEMS");
  const long min_items = t ? std::max(1, std::atoi(t)) : 24;
  const int block_n = cout % 256 == 0 ? 256 : 128;
  return ((m_tiles + 1) / 2) * (cout / block_n) >= min_items;
}

static bool pdl_enabled() {
  static const bool on = !(std::getenv("B200RN_PDL") && std::getenv("B200RN_PDL")[0] == '0');
  return on;
}

template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), int grid, int block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(grid));
  cfg.blockDim = dim3(unsigned(block));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

template <typename T, int BLOCK_N, int KB, int EPI>
static cudaError_t launch_inst(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = conv_umma_kernel<T, BLOCK_N, KB, EPI>;
  static thread_local int configured_device = -1;  // per (instantiation, thread)
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  // the four operand-transform warps exist only for layers that need them
  return launch_pdl(kern, pl->grid, pl->p.has_pre ? kThreads : kBaseThreads, pl->smem_bytes, s, pl->tmA, pl->tmW, pl->tmRes, pl->tmOut, pl->tmAct, pl->p);
}

template <typename T, int BLOCK_N, int KB>
static cudaError_t launch_epi(const UmmaConvPlan* pl, cudaStream_t s) {
  const UmmaParams& p = pl->p;
  const int flags = (p.relu ? EPI_RELU : 0) | (p.has_res ? EPI_RES : 0) | (p.has_out ? EPI_OUT : 0) |
                    (p.has_act ? EPI_ACT : 0) | ((p.has_act && p.act_relu) ? EPI_ACT_RELU : 0);
  switch (flags) {   // the combinations a ResNet produces get their own code; anything else runs the generic one
    case EPI_OUT | EPI_RELU: return launch_inst<T, BLOCK_N, KB, EPI_OUT | EPI_RELU>(pl, s);
    case EPI_OUT: return launch_inst<T, BLOCK_N, KB, EPI_OUT>(pl, s);
    case EPI_OUT | EPI_RES: return launch_inst<T, BLOCK_N, KB, EPI_OUT | EPI_RES>(pl, s);
    case EPI_OUT | EPI_RES | EPI_ACT | EPI_ACT_RELU: return launch_inst<T, BLOCK_N, KB, EPI_OUT | EPI_RES | EPI_ACT | EPI_ACT_RELU>(pl, s);
    case EPI_RES | EPI_ACT | EPI_ACT_RELU: return launch_inst<T, BLOCK_N, KB, EPI_RES | EPI_ACT | EPI_ACT_RELU>(pl, s);
    default: return launch_inst<T, BLOCK_N, KB, -1>(pl, s);
  }
}

template <typename T, int BLOCK_N>
static cudaError_t launch_halo(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = conv3x3_halo_kernel<T, BLOCK_N>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  return launch_pdl(kern, pl->grid, kTwoGroupThreads, pl->smem_bytes, s, pl->tmA, pl->tmW, pl->tmOut, pl->hp);
}

template <typename T, int BLOCK_N>
static cudaError_t launch_halo2(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = conv3x3_halo2_kernel<T, BLOCK_N>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  return launch_pdl(kern, pl->grid, kTwoGroupThreads, pl->smem_bytes, s, pl->tmA, pl->tmW, pl->tmOut, pl->h2);
}

template <typename T>
static cudaError_t launch_stem2(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = stem7x7_kernel<T>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  return launch_pdl(kern, pl->grid, 64 + pl->sp.groups * kEpiThreads, pl->smem_bytes, s, pl->tmA, pl->tmOut, pl->sp);
}

template <typename T, int BLOCK_N>
static cudaError_t launch_dual(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = conv_dualm_kernel<T, BLOCK_N>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  return launch_pdl(kern, pl->grid, kTwoGroupThreads, pl->smem_bytes, s, pl->tmA, pl->tmW, pl->tmOut, pl->dp);
}

template <typename T, int BLOCK_N>
static cudaError_t launch_pair(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = conv_pair_kernel<T, BLOCK_N>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  // the cluster shape (2,1,1) is part of the kernel (__cluster_dims__); the grid is a multiple of 2
  return launch_pdl(kern, pl->grid, kPairThreads, pl->smem_bytes, s, pl->tmA, pl->tmW, pl->tmOut, pl->pp);
}

template <typename T, int N1>
static cudaError_t launch_tail(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = conv_tail_kernel<T, N1>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  return launch_pdl(kern, pl->grid, kTailThreads, pl->smem_bytes, s, pl->tmA, pl->tmA2, pl->tmW, pl->tmW1, pl->tmRes, pl->tmOut, pl->tmOut1, pl->tp);
}

// Two-input wide layers with this many K blocks or more (streamed operands, no residual to prefetch) take three
// operand stages and two slots per staging ring: their main loop, not the epilogue, sets the pace.  Measured:
// stage-2/3/4 projection blocks 0.165 -> 0.134, 0.132 -> 0.101, 0.126 -> 0.093 ms; the K = 512 residual layers of
// stage 4 lose 10 % with three stages and keep two.  (B200RN_WIDE_S3_MINK; 0 = never)
static int wide_s3_min_k() {
  static const int v = [] { const char* e = std::getenv("B200RN_WIDE_S3_MINK"); int k = e ? std::atoi(e) : 6; return k <= 0 ? 1 << 30 : k; }();
  return v;
}

template <typename T, int N>
static cudaError_t launch_wide(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = conv_wide_kernel<T, N>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  return launch_pdl(kern, pl->grid, kWideThreads, pl->smem_bytes, s, pl->tmA, pl->tmA2, pl->tmW, pl->tmRes, pl->tmOut, pl->tmAct, pl->wp);
}

template <typename T>
static cudaError_t launch_pairwide(const UmmaConvPlan* pl, cudaStream_t s) {
  auto kern = conv_pairwide_kernel<T, 256>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  // the cluster shape (2,1,1) is part of the kernel (__cluster_dims__); the grid is a multiple of 2 x n_tiles
  return launch_pdl(kern, pl->grid, kPairWideThreads, pl->smem_bytes, s, pl->tmA, pl->tmA2, pl->tmW, pl->tmRes, pl->tmOut, pl->tmAct, pl->wp);
}

cudaError_t launch_conv_umma(const UmmaConvPlan* pl, cudaStream_t s) {
  const bool bf = pl->dt == DType::BF16;
  if (pl->pairwide) return bf ? launch_pairwide<__nv_bfloat16>(pl, s) : launch_pairwide<__half>(pl, s);
  if (pl->wide) {
    if (pl->block_n == 128) return bf ? launch_wide<__nv_bfloat16, 128>(pl, s) : launch_wide<__half, 128>(pl, s);
    return bf ? launch_wide<__nv_bfloat16, 256>(pl, s) : launch_wide<__half, 256>(pl, s);
  }
  if (pl->tail) {
    if (pl->tp_n1 == 64) return bf ? launch_tail<__nv_bfloat16, 64>(pl, s) : launch_tail<__half, 64>(pl, s);
    return bf ? launch_tail<__nv_bfloat16, 128>(pl, s) : launch_tail<__half, 128>(pl, s);
  }
  if (pl->pair) {
    if (pl->block_n == 128) return bf ? launch_pair<__nv_bfloat16, 128>(pl, s) : launch_pair<__half, 128>(pl, s);
    return bf ? launch_pair<__nv_bfloat16, 256>(pl, s) : launch_pair<__half, 256>(pl, s);
  }
  if (pl->dual) {
    if (pl->block_n == 128) return bf ? launch_dual<__nv_bfloat16, 128>(pl, s) : launch_dual<__half, 128>(pl, s);
    return bf ? launch_dual<__nv_bfloat16, 256>(pl, s) : launch_dual<__half, 256>(pl, s);
  }
  if (pl->stem2) return bf ? launch_stem2<__nv_bfloat16>(pl, s) : launch_stem2<__half>(pl, s);
  if (pl->halo2) {
    if (pl->block_n == 64) return bf ? launch_halo2<__nv_bfloat16, 64>(pl, s) : launch_halo2<__half, 64>(pl, s);
    return bf ? launch_halo2<__nv_bfloat16, 128>(pl, s) : launch_halo2<__half, 128>(pl, s);
  }
  if (pl->halo) {
    if (pl->block_n == 64) return bf ? launch_halo<__nv_bfloat16, 64>(pl, s) : launch_halo<__half, 64>(pl, s);
    return bf ? launch_halo<__nv_bfloat16, 128>(pl, s) : launch_halo<__half, 128>(pl, s);
  }
  if (pl->kb == 64) return bf ? launch_epi<__nv_bfloat16, 64, 64>(pl, s) : launch_epi<__half, 64, 64>(pl, s);
  if (pl->block_n == 64) return bf ? launch_epi<__nv_bfloat16, 64, 128>(pl, s) : launch_epi<__half, 64, 128>(pl, s);
  if (pl->block_n == 128) return bf ? launch_epi<__nv_bfloat16, 128, 128>(pl, s) : launch_epi<__half, 128, 128>(pl, s);
  return bf ? launch_epi<__nv_bfloat16, 256, 128>(pl, s) : launch_epi<__half, 256, 128>(pl, s);
}

// A 1x1 / stride-1 / pad-0 conv is a plain GEMM over the B*H*W pixels, and NHWC keeps those pixels' channel rows at a
// constant pitch: such a layer is planned as ONE image of 1 x (B*H*W) pixels, so that an M tile is 128 CONSECUTIVE
// rows of every tensor it touches (input, second input, residual, outputs).  With the spatial patches choose_patch
// otherwise finds for 14x14 and 7x7 maps (2 x 2 pixels x 32 images, 1 pixel x 128 images) every 128- or 256-byte piece
// of a tile lies in a different image, 100-400 KB apart, and the residual / output traffic of stages 3 and 4 ran at
// half of the DRAM rate; flattened tiles also always fill all 128 rows.  B200RN_FLATTEN=0 keeps the spatial patches.
static bool flatten_1x1(ConvGeom& g, Epilogue& e, int stem_mode) {
  const char* fe = std::getenv("B200RN_FLATTEN");
  if (fe && fe[0] == '0') return false;
  if (stem_mode || g.kh != 1 || g.kw != 1 || g.stride != 1 || g.pad != 0 || g.H * g.W <= 1) return false;
  if (e.in_b && (e.in_b_stride != 1 || e.in_b_h != g.H || e.in_b_w != g.W)) return false;
  if (e.out_f32) return false;
  const long long px = 1LL * g.B * g.H * g.W;
  if (px > 0x7fffffffLL) return false;
  g.B = 1; g.H = 1; g.W = int(px); g.Ho = 1; g.Wo = int(px);
  if (e.in_b) { e.in_b_h = 1; e.in_b_w = int(px); }
  return true;
}

UmmaConvPlan* umma_conv_plan_create(DType dt, const ConvGeom& g_in, const void* in, const void* w_packed,
                                    const Epilogue& e_in, int stem_mode, char* err, size_t errcap) {
  auto failmsg = [&](const std::string& m) -> UmmaConvPlan* {
    if (err && errcap) std::snprintf(err, errcap, "%s", m.c_str());
    return nullptr;
  };
  char why[200];
  if (dt == DType::F32) return failmsg("tensor-core conv needs a 16-bit activation type");
  if (!umma_conv_supported(g_in, stem_mode, why, sizeof why)) return failmsg(why);
  ConvGeom g = g_in;
  Epilogue e = e_in;
  flatten_1x1(g, e, stem_mode);
  auto pl = new UmmaConvPlan();
  pl->dt = dt; pl->g = g; pl->stem = stem_mode;
  UmmaParams& p = pl->p;
  std::memset(&p, 0, sizeof p);
  p.has_res = e.residual != nullptr; p.has_out = e.out != nullptr; p.has_act = e.out_act != nullptr;
  p.relu = e.relu; p.act_relu = e.act_relu;
  p.bias = e.bias; p.act_scale = e.act_scale; p.act_shift = e.act_shift;
  p.has_pre = e.pre_scale != nullptr; p.pre_relu = e.pre_relu; p.cin = g.Cin;
  if (std::getenv("B200RN_DBG_TRACE")) {
    cudaMalloc(&p.trace, 64 * 8 * sizeof(unsigned long long));
    cudaMemset(p.trace, 0, 64 * 8 * sizeof(unsigned long long));
  }
  if (const char* v = std::getenv("B200RN_DBG_ASHIFT")) p.dbg_ashift = std::atoi(v);
  p.pre_scale = e.pre_scale; p.pre_shift = e.pre_shift;
  if (p.has_pre && (stem_mode || g.kh != 1 || g.kw != 1 || g.pad != 0)) { delete pl; return failmsg("operand prologue needs a 1x1, pad-0 conv"); }
  p.out_f32 = e.out_f32; p.f32_ld = e.out_f32_ld; p.f32_cols = e.out_f32_cols; p.f32_rows = g.B;
  if (p.out_f32 && (stem_mode || g.Ho != 1 || g.Wo != 1 || p.has_out || p.has_act || p.has_res || p.has_pre || e.tail_out)) {
    delete pl;
    return failmsg("fp32 row output needs a 1x1-spatial conv with no other output, residual, prologue or tail");
  }
  if (!p.has_out && !p.has_act && !p.out_f32) { delete pl; return failmsg("conv with no output"); }
  if (e.tail_out) {
    // 1x1 conv (Cout 256) + residual -> raw output, and the following 1x1 conv on relu(bn(.)) of it
    if (stem_mode || g.kh != 1 || g.kw != 1 || g.stride != 1 || g.pad != 0 || g.Cout != kTailN3 || (!p.has_res && !e.in_b) || !p.has_out ||
        p.has_act || p.has_pre || !e.act_scale || !e.act_shift || !e.tail_w || !e.tail_bias ||
        (e.tail_cout != 64 && e.tail_cout != 128)) {
      delete pl;
      return failmsg("fused tail conv needs a 1x1/s1/p0 conv to 256 channels with a residual or a second input and a raw output, and a 64- or 128-channel tail");
    }
    TailParams& tp = pl->tp;
    std::memset(&tp, 0, sizeof tp);
    pl->tail = 1; pl->tp_n1 = e.tail_cout; pl->kb = 128; pl->block_n = kTailN3;
    int bw = 1, bh = 1, bn = 1;
    choose_patch(g.B, g.Ho, g.Wo, bw, bh, bn);
    const int tKtot = g.Cin + (e.in_b ? e.in_b_cin : 0);
    tp.cblks = tKtot / 64; tp.k1_blocks = g.Cin / 64; tp.stride_b = e.in_b ? e.in_b_stride : 1;
    tp.has_res = p.has_res;
    tp.bw = bw; tp.bh = bh; tp.bn = bn;
    tp.tiles_w = (g.Wo + bw - 1) / bw; tp.tiles_h = (g.Ho + bh - 1) / bh;
    tp.total_tiles = tp.tiles_w * tp.tiles_h * ((g.B + bn - 1) / bn);
    tp.rows = uint32_t(bw * bh * bn);
    tp.relu = e.relu; tp.act_relu = e.act_relu; tp.tail_relu = e.tail_relu;
    {
      const char* penv = std::getenv("B200RN_TAIL_PF");
      tp.pf_dist = penv ? std::max(0, std::min(4, std::atoi(penv))) : 0;   // measured: 2 makes the layer 5 % slower
      const char* p2env = std::getenv("B200RN_TAIL_PF_INB");
      tp.pf_in_b = p2env ? std::max(0, std::min(4, std::atoi(p2env))) : 2;
    }
    tp.bias = e.bias; tp.act_scale = e.act_scale; tp.act_shift = e.act_shift; tp.tail_bias = e.tail_bias;
    if (std::getenv("B200RN_DBG_TRACE")) {
      cudaMalloc(&tp.trace, 8 * 32 * 8 * sizeof(unsigned long long));
      cudaMemset(tp.trace, 0, 8 * 32 * 8 * sizeof(unsigned long long));
    }
    const uint32_t fixed = 1024u + uint32_t(tp.cblks) * kTailN3 * 128 + 4u * uint32_t(e.tail_cout) * 128 +
                           uint32_t(e.tail_cout / 64) * kChunkBytes + (3u * kTailN3 + uint32_t(e.tail_cout)) * 4 + 8u * 48 + 64;
    const int slots = (int(227u * 1024) - int(fixed)) / int(kChunkBytes);    // 16 KB units left for the three rings
    tp.act_bufs = 2;
    tp.a_stages = 1;     // the next tile's operand is only needed a tile from now; the slots are worth more in the residual ring
    tp.r_slots = std::min(8, slots - 2 - tp.a_stages * tp.cblks);
    if (slots < 2 + tp.cblks + 2 || tp.r_slots < 2) { delete pl; return failmsg("fused tail conv: shared memory does not hold the weights plus staging"); }
    pl->smem_bytes = fixed + uint32_t(tp.act_bufs + tp.a_stages * tp.cblks + tp.r_slots) * kChunkBytes;
    pl->grid = std::min(tp.total_tiles, sm_count());
    std::string tmsg;
    bool tok = encode_act(&pl->tmA, dt, in, g.B, g.H, g.W, g.Cin, bw, bh, bn, 1, tmsg);
    if (tok && e.in_b) {
      if (e.in_b_cin % 64 || (e.in_b_h - 1) / e.in_b_stride + 1 != g.Ho || (e.in_b_w - 1) / e.in_b_stride + 1 != g.Wo) {
        delete pl;
        return failmsg("two-input conv: the second input must have 64-multiple channels and map onto the output grid at its stride");
      }
      tok = encode_act(&pl->tmA2, dt, e.in_b, g.B, e.in_b_h, e.in_b_w, e.in_b_cin, bw, bh, bn, e.in_b_stride, tmsg);
    } else {
      pl->tmA2 = pl->tmA;
    }
    if (tok) {
      uint64_t wd[2] = {uint64_t(tKtot), uint64_t(kTailN3)};
      uint64_t ws[1] = {uint64_t(tKtot) * 2};
      uint32_t wb[2] = {64, uint32_t(kTailN3)};
      uint32_t we[2] = {1, 1};
      tok = encode(&pl->tmW, dt, 2, w_packed, wd, ws, wb, we, false, tmsg);
    }
    if (tok) {
      uint64_t wd[2] = {uint64_t(kTailN3), uint64_t(e.tail_cout)};
      uint64_t ws[1] = {uint64_t(kTailN3) * 2};
      uint32_t wb[2] = {64, uint32_t(e.tail_cout)};
      uint32_t we[2] = {1, 1};
      tok = encode(&pl->tmW1, dt, 2, e.tail_w, wd, ws, wb, we, false, tmsg);
    }
    if (tok) tok = encode_act(&pl->tmRes, dt, e.residual ? e.residual : e.out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, tmsg);
    if (tok) tok = encode_act(&pl->tmOut, dt, e.out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, tmsg);
    if (tok) tok = encode_act(&pl->tmOut1, dt, e.tail_out, g.B, g.Ho, g.Wo, e.tail_cout, bw, bh, bn, 1, tmsg);
    if (!tok) { delete pl; return failmsg(tmsg); }
    pl->tmAct = pl->tmOut;
    if (std::getenv("B200RN_VERBOSE"))
      std::fprintf(stderr, "[b200rn] conv 1x1 %d->%d + tail ->%d @%dx%d B%d: patch=%dx%dx%d tiles=%d grid=%d a_stages=%d r_slots=%d smem=%u\n",
                   g.Cin, g.Cout, e.tail_cout, g.H, g.W, g.B, bw, bh, bn, tp.total_tiles, pl->grid, tp.a_stages, tp.r_slots, pl->smem_bytes);
    return pl;
  }
  if (stem_mode == 2) {
    if (p.has_res || p.has_pre || (!e.pool_fused && (!p.has_out || p.has_act))) {
      delete pl;
      return failmsg("stem kernel supports bias + ReLU + one output, or a fused 3x3/2 max-pool with raw and/or activated pooled outputs");
    }
    if (e.pool_fused && (e.pool_h != (g.Ho + 2 - 3) / 2 + 1 || e.pool_w != (g.Wo + 2 - 3) / 2 + 1)) {
      delete pl;
      return failmsg("stem kernel: the fused pool must be 3x3 / stride 2 / pad 1 over the conv output");
    }
    Stem2Params& sp = pl->sp;
    std::memset(&sp, 0, sizeof sp);
    pl->stem2 = 1;
    pl->kb = 64; pl->block_n = 64;
    sp.tiles_w = (g.Wo + kStemTW - 1) / kStemTW;
    sp.tiles_h = (g.Ho + kStemTH - 1) / kStemTH;
    if (e.pool_fused) {
      sp.pool = 1; sp.conv_w = g.Wo; sp.conv_h = g.Ho; sp.pool_w = e.pool_w; sp.pool_h = e.pool_h; sp.cout = g.Cout;
      sp.pool_out = e.out; sp.pool_act = e.out_act; sp.act_scale = e.act_scale; sp.act_shift = e.act_shift; sp.act_relu = e.act_relu;
      sp.tiles_w = (e.pool_w + 2) / 3;          // 3 x 7 pooled pixels per tile
      sp.tiles_h = (e.pool_h + 6) / 7;
    }
    sp.n_tiles = g.Cout / 64;
    sp.total_tiles = sp.tiles_w * sp.tiles_h * g.B * sp.n_tiles;
    sp.relu = e.relu;
    sp.bias = e.bias;
    sp.w_packed = w_packed;
    if (const char* v = std::getenv("B200RN_STEM_DBG")) sp.dbg_skip = std::atoi(v);
    if (std::getenv("B200RN_DBG_TRACE")) {
      cudaMalloc(&sp.trace, 64 * 8 * sizeof(unsigned long long));
      cudaMemset(sp.trace, 0, 64 * 8 * sizeof(unsigned long long));
    }
    const int sms = sm_count();
    int grid = std::min(sp.total_tiles, sms);
    grid = std::max(sp.n_tiles, (grid / sp.n_tiles) * sp.n_tiles);
    {
      // a tile's epilogue (TMEM -> staging -> store: ~2300 cycles; with the fused pool ~4000) is a longer chain than its
      // 14 MMAs (~630 cycles): three groups keep three tiles in flight (0.132 -> 0.119 ms at batch 256 against two;
      // B200RN_STEM_GROUPS overrides)
      const char* ge = std::getenv("B200RN_STEM_GROUPS");
      sp.groups = ge ? std::max(2, std::min(3, std::atoi(ge))) : 3;
    }
    const uint32_t fixed = 1024u + kStemWBytes + uint32_t(sp.groups) * kChunkBytes + 3u * 64u * 4 + 8u * (2 * 16 + 9) + 64;
    sp.stages = std::max(2, std::min(16, int((227u * 1024 - fixed) / kStemStageBytes)));
    if (const char* v = std::getenv("B200RN_STEM_STAGES")) sp.stages = std::max(2, std::min(sp.stages, std::atoi(v)));
    pl->smem_bytes = fixed + uint32_t(sp.stages) * kStemStageBytes;
    pl->grid = grid;
    std::string smsg;
    const int Hp = stem_Hp(g.H), Wp = stem_Wp(g.W);
    uint64_t dims[3] = {uint64_t(Wp) * 4, uint64_t(Hp), uint64_t(g.B)};
    uint64_t str[2] = {uint64_t(Wp) * 8, uint64_t(Hp) * Wp * 8};
    uint32_t box[3] = {uint32_t(kStemPatchPx * 4), uint32_t(kStemPatchRows), 1};
    uint32_t es[4] = {1, 1, 1, 1};
    bool sok = encode(&pl->tmA, dt, 3, in, dims, str, box, es, 2, smsg);
    if (sok && !e.pool_fused) {
      uint64_t od[4] = {uint64_t(g.Cout), uint64_t(g.Wo), uint64_t(g.Ho), uint64_t(g.B)};
      uint64_t os[3] = {uint64_t(g.Cout) * 2, uint64_t(g.Wo) * g.Cout * 2, uint64_t(g.Ho) * g.Wo * g.Cout * 2};
      uint32_t ob[4] = {64, uint32_t(kStemTW), uint32_t(kStemTH), 1};
      sok = encode(&pl->tmOut, dt, 4, e.out, od, os, ob, es, 0, smsg);
    } else if (sok) {
      pl->tmOut = pl->tmA;                      // pooled outputs are written with plain stores: no store map
    }
    if (!sok) { delete pl; return failmsg(smsg); }
    pl->tmW = pl->tmOut; pl->tmAct = pl->tmOut; pl->tmRes = pl->tmOut;
    return pl;
  }
  {
    // 3x3 / stride 1 / pad 1 on 64 or 128 input channels with a plain (bias, ReLU) epilogue: halo kernel
    const char* henv = std::getenv("B200RN_HALO");
    if (!(henv && henv[0] == '0') && !stem_mode && g.kh == 3 && g.kw == 3 && g.stride == 1 && g.pad == 1 &&
        (g.Cin == 64 || g.Cin == 128) && g.Wo >= kHaloBW && p.has_out && !p.has_res && !p.has_act && !p.has_pre) {
      HaloParams& hp = pl->hp;
      std::memset(&hp, 0, sizeof hp);
      pl->halo = 1;
      pl->kb = 128;
      pl->block_n = (g.Cout % 128 == 0) ? 128 : 64;
      hp.cblks = g.Cin / 64;
      hp.n_tiles = g.Cout / pl->block_n;
      hp.tiles_w = (g.Wo + kHaloBW - 1) / kHaloBW;
      hp.tiles_h = (g.Ho + kHaloBH - 1) / kHaloBH;
      hp.total_tiles = hp.tiles_w * hp.tiles_h * g.B * hp.n_tiles;
      hp.relu = e.relu;
      hp.bias = e.bias;
      if (const char* v = std::getenv("B200RN_DBG_SKIP")) hp.dbg_skip = std::atoi(v);
      if (std::getenv("B200RN_DBG_TRACE")) {
        cudaMalloc(&hp.trace, 64 * 8 * sizeof(unsigned long long));
        cudaMemset(hp.trace, 0, 64 * 8 * sizeof(unsigned long long));
      }
      const int sms = sm_count();
      int grid = std::min(hp.total_tiles, sms);
      grid = std::max(hp.n_tiles, (grid / hp.n_tiles) * hp.n_tiles);
      const uint32_t budget = 227u * 1024;
      const uint32_t b_bytes = uint32_t(pl->block_n) * 128;
      const uint32_t slab = 9u * uint32_t(hp.cblks) * b_bytes;
      const uint32_t out_bytes = uint32_t(pl->block_n / 64) * kChunkBytes;
      hp.b_stationary = (slab <= 80u * 1024 && hp.total_tiles >= 2 * grid) ? 1 : 0;
      hp.b_stages = hp.b_stationary ? 1 : 4;
      const uint32_t fixed = 1024u + (hp.b_stationary ? slab : uint32_t(hp.b_stages) * b_bytes) + 2 * out_bytes +
                             2u * uint32_t(pl->block_n) * 4 + 8u * (2 * 8 + 2 * 4 + 4) + 64;
      hp.halo_stages = std::max(2, std::min(8, int((budget - fixed) / kHaloStageBytes)));
      pl->smem_bytes = fixed + uint32_t(hp.halo_stages) * kHaloStageBytes;
      pl->grid = grid;
      {
        // The weight slab does not fit (it is streamed per tile): two M tiles per streamed weight tile instead
        // (conv3x3_halo2_kernel).  Measured with the MMAs switched off these layers ran at the pace of that stream.
        const char* h2env = std::getenv("B200RN_HALO2");
        const int m_tiles = hp.tiles_w * hp.tiles_h * g.B;
        if (!(h2env && h2env[0] == '0') && !hp.b_stationary && m_tiles >= 2 && !hp.trace && !hp.dbg_skip) {
          Halo2Params& h2 = pl->h2;
          std::memset(&h2, 0, sizeof h2);
          pl->halo2 = 1;
          h2.cblks = hp.cblks; h2.n_tiles = hp.n_tiles; h2.m_tiles = m_tiles; h2.tiles_w = hp.tiles_w; h2.tiles_h = hp.tiles_h;
          h2.total_items = ((m_tiles + 1) / 2) * hp.n_tiles;
          h2.relu = hp.relu; h2.bias = hp.bias;
          h2.b_stages = 4;
          const uint32_t fixed2 = 1024u + 4u * b_bytes + 2 * out_bytes + 2u * uint32_t(pl->block_n) * 4 + 8u * (2 * 8 + 2 * 4 + 4) + 64;
          h2.halo_stages = std::max(2, std::min(6, int((budget - fixed2) / (2 * kHaloStageBytes))));
          pl->smem_bytes = fixed2 + uint32_t(h2.halo_stages) * 2 * kHaloStageBytes;
          pl->grid = std::min(h2.total_items, sms);
        }
      }
      std::string hmsg;
      uint64_t dims[4] = {uint64_t(g.Cin), uint64_t(g.W), uint64_t(g.H), uint64_t(g.B)};
      uint64_t str[3] = {uint64_t(g.Cin) * 2, uint64_t(g.W) * g.Cin * 2, uint64_t(g.H) * g.W * g.Cin * 2};
      uint32_t box[4] = {64, uint32_t(kHaloPW), uint32_t(kHaloPH), 1};
      uint32_t es[4] = {1, 1, 1, 1};
      bool hok = encode(&pl->tmA, dt, 4, in, dims, str, box, es, false, hmsg);
      if (hok) {
        const uint64_t K = 9ull * g.Cin;
        uint64_t wd[2] = {K, uint64_t(g.Cout)};
        uint64_t ws[1] = {K * 2};
        uint32_t wb[2] = {64, uint32_t(pl->block_n)};
        uint32_t we[2] = {1, 1};
        hok = encode(&pl->tmW, dt, 2, w_packed, wd, ws, wb, we, false, hmsg);
      }
      if (hok) {
        uint64_t od[4] = {uint64_t(g.Cout), uint64_t(g.Wo), uint64_t(g.Ho), uint64_t(g.B)};
        uint64_t os[3] = {uint64_t(g.Cout) * 2, uint64_t(g.Wo) * g.Cout * 2, uint64_t(g.Ho) * g.Wo * g.Cout * 2};
        uint32_t ob[4] = {64, uint32_t(kHaloBW), uint32_t(kHaloBH), 1};
        hok = encode(&pl->tmOut, dt, 4, e.out, od, os, ob, es, false, hmsg);
      }
      if (!hok) { delete pl; return failmsg(hmsg); }
      pl->tmAct = pl->tmOut; pl->tmRes = pl->tmOut;
      return pl;
    }
  }
  {
    // Block-ending 1x1 convs (wide output + residual, raw and/or activated copy written): conv_wide_kernel.
    // Measured per layer at batch 256 (B200RN_WIDE=0/1): the residual layers gain up to 21 % (most where only the
    // activated copy is written), the projection shortcuts (no residual, one output) lose 15-20 % against the
    // generic kernel's stationary-weight path, so those and the plain 1x1 convs stay where they were.
    // B200RN_WIDE=2 routes every wide-output 1x1 conv here (tests).
    const char* wenv = std::getenv("B200RN_WIDE");
    const bool wide_all = wenv && wenv[0] == '2';
    // "Deep" configuration for the large-K 1x1 convs with a plain epilogue: their main loop is bound by how many
    // operand bytes are in flight (L2 latency ~2500 cycles x 41 B/clk with three 48 KB stages in the general
    // kernel, whose whole-tile staging buffer takes the room a fourth would need).  Chunk-ring staging needs only
    // two 16 KB slots, which leaves four stages.
    const char* denv2 = std::getenv("B200RN_DEEP");
    // Measured per layer at batch 256: 1x1 1024->256 @14x14 0.0455 -> 0.0422 ms, 512->256 @28x28 0.076 -> 0.068;
    // with fewer than two waves of tiles (2048->512 @7x7) the dual-M kernel stays ahead (0.0355 vs 0.0397).
    int pbw = 1, pbh = 1, pbn = 1;
    choose_patch(g.B, g.Ho, g.Wo, pbw, pbh, pbn);
    const long deep_tiles = long((g.Wo + pbw - 1) / pbw) * ((g.Ho + pbh - 1) / pbh) * ((g.B + pbn - 1) / pbn) * (g.Cout / kWideN);
    // layers the CTA-pair kernel takes (decided after plan_geometry below) do not go the "deep" way
    const bool pair_takes = pair_enabled() && !stem_mode && p.has_out && !p.has_res && !p.has_act && !p.has_pre && !p.out_f32 &&
                            g.kh * g.kw * (g.Cin / 64) >= pair_min_k() && pair_n_ok(g.Cout) &&
                            long((g.Wo + pbw - 1) / pbw) * ((g.Ho + pbh - 1) / pbh) * ((g.B + pbn - 1) / pbn) >= 2 && sm_count() >= 2 &&
                            pair_items_ok(long((g.Wo + pbw - 1) / pbw) * ((g.Ho + pbh - 1) / pbh) * ((g.B + pbn - 1) / pbn), g.Cout);
    const bool deep = !pair_takes && !(denv2 && denv2[0] == '0') && p.has_out && !p.has_res && !p.has_act && g.Cin >= 512 &&
                      g.Cout % kWideN == 0 && deep_tiles >= 2L * sm_count();
    if (!(wenv && wenv[0] == '0') && !stem_mode && !p.has_pre && g.kh == 1 && g.kw == 1 && g.pad == 0 && g.Cout % kWideN == 0 &&
        (p.has_res || deep || e.in_b || (wide_all && (p.has_act || g.Cout > g.Cin)))) {
      WideParams& wp = pl->wp;
      std::memset(&wp, 0, sizeof wp);
      pl->wide = 1; pl->kb = 128;
      int bw = 1, bh = 1, bn = 1;
      choose_patch(g.B, g.Ho, g.Wo, bw, bh, bn);
      const int Ktot = g.Cin + (e.in_b ? e.in_b_cin : 0);
      wp.cblks = g.Cin / 64; wp.k1_iters = wp.cblks; wp.num_k_iters = Ktot / 64; wp.kw = 1; wp.stride = g.stride; wp.pad = 0;
      wp.stride_b = e.in_b ? e.in_b_stride : 1;
      wp.bw = bw; wp.bh = bh; wp.bn = bn;
      wp.tiles_w = (g.Wo + bw - 1) / bw; wp.tiles_h = (g.Ho + bh - 1) / bh;
      wp.m_tiles = wp.tiles_w * wp.tiles_h * ((g.B + bn - 1) / bn);
      wp.has_res = p.has_res; wp.has_out = p.has_out; wp.has_act = p.has_act; wp.relu = p.relu; wp.act_relu = p.act_relu;
      wp.rows = uint32_t(bw * bh * bn); wp.a_bytes = wp.rows * 128u;
      wp.bias = e.bias; wp.act_scale = e.act_scale; wp.act_shift = e.act_shift;
      if (const char* v = std::getenv("B200RN_WIDE_DBG")) wp.dbg_skip = std::atoi(v);
      if (std::getenv("B200RN_DBG_TRACE")) {
        cudaMalloc(&wp.trace, 96 * 8 * sizeof(unsigned long long));
        cudaMemset(wp.trace, 0, 96 * 8 * sizeof(unsigned long long));
      }
      const char* penv = std::getenv("B200RN_WIDE_PF");
      wp.pf_dist = penv ? std::max(0, std::min(4, std::atoi(penv))) : 0;   // measured: 2 costs 4 % of the whole step
      // Stationary weights: the CTA keeps one N tile, so when that tile's whole slab (N x Cin) fits beside the staging
      // rings it is loaded once and only A streams -- the operand ring was what bounded the raw + activated layers
      // (two 48 KB stages, B re-fetched for every tile).  Cin <= 128: N = 256 (slab <= 64 KB); Cin = 256: N = 128
      // (64 KB) with four A stages = one tile ahead; larger Cin streams both operands as before.
      const char* benv = std::getenv("B200RN_WIDE_BSTAT");
      const bool bstat_ok = !(benv && benv[0] == '0') && !deep;
      int N = kWideN;
      wp.b_stationary = 0;
      // CTA pair (conv_pairwide_kernel): K >= 512 with streamed weights (half the weight bytes per SM and stage, so three
      // or four stages fit beside the rings).  Measured at batch 256 (one CTA -> pair): projection blocks of stages 3 / 4
      // 0.097 -> 0.089 / 0.087 -> 0.075 ms, 512 -> 2048 @7x7 + residual 0.059 -> 0.054 and 0.051 -> 0.046.
      // Opt-in only (B200RN_PAIRWIDE_BSTAT=1): K = 256 with the pair's 256-wide slab stationary, half in each CTA -- it
      // halves the A re-reads of the single-CTA form (which drops to N = 128 to keep its slab) but ran the 256 -> 1024
      // @14x14 residual layers at 0.083-0.089 ms against 0.070: they are bound by the epilogue rings, not the feed.
      // B200RN_PAIRWIDE=0 turns the pair route off, B200RN_PAIRWIDE_MINK (default 8) is the smallest streamed K-block
      // count, B200RN_PAIRWIDE_STAGES the stage count.
      const char* pwe = std::getenv("B200RN_PAIRWIDE");
      const char* pwk = std::getenv("B200RN_PAIRWIDE_MINK");
      const char* pwb = std::getenv("B200RN_PAIRWIDE_BSTAT");
      const int pw_min_k = pwk ? std::max(2, std::atoi(pwk)) : 8;
      const bool pw_ok = pair_enabled() && !(pwe && pwe[0] == '0') && !deep && (p.has_res || e.in_b) && wp.m_tiles >= 2 &&
                         sm_count() >= 2 * (g.Cout / kWideN);
      if (bstat_ok && Ktot <= 128) wp.b_stationary = 1;
      else if (bstat_ok && Ktot == 256 && pw_ok && pwb && pwb[0] == '1') { wp.b_stationary = 1; pl->pairwide = 1; }
      else if (bstat_ok && Ktot == 256 && g.Cout % 128 == 0) { wp.b_stationary = 1; N = 128; }
      else if (pw_ok && wp.num_k_iters >= pw_min_k) pl->pairwide = 1;
      pl->block_n = N;
      wp.n_tiles = g.Cout / N;
      const char* senv = std::getenv(pl->pairwide ? "B200RN_PAIRWIDE_STAGES" : "B200RN_WIDE_STAGES");
      const uint32_t slab = wp.b_stationary ? uint32_t(wp.num_k_iters) * uint32_t(pl->pairwide ? N / 2 : N) * 128u : 0u;
      const uint32_t stage_bytes = wp.b_stationary ? uint32_t(kTileM) * 128u
                                                   : uint32_t(kTileM + (pl->pairwide ? N / 2 : N)) * 128u;
      wp.stages = senv ? std::max(2, std::min(pl->pairwide ? 5 : 4, std::atoi(senv)))
                       : (pl->pairwide && !wp.b_stationary) ? (p.has_res ? 3 : 4)    // residual layers: the rings need the room more (measured)
                       : (wp.b_stationary ? std::max(2, std::min(4, wp.num_k_iters + 1))
                                           : ((deep && !p.has_res) ? 4 : ((e.in_b && wp.num_k_iters >= wide_s3_min_k()) ? 3 : 2)));
      const uint32_t fixed = 1024u + (p.has_act ? 3u : 1u) * uint32_t(N) * 4 + 8u * (2 * 5 + 4 + 33) + 64;
      const int slots = (int(227u * 1024) - int(fixed) - int(slab) - int(uint32_t(wp.stages) * stage_bytes)) / int(kChunkBytes);
      const bool use_r = p.has_res || p.has_out;
      if (use_r && p.has_act) {
        wp.c_slots = std::max(2, slots * 3 / 7);
        if (const char* ce = std::getenv("B200RN_WIDE_CSLOTS")) wp.c_slots = std::max(2, std::min(slots - 2, std::atoi(ce)));
        wp.r_slots = slots - wp.c_slots;
      }
      else if (use_r) wp.r_slots = slots;
      else wp.c_slots = slots;
      wp.r_slots = std::min(8, wp.r_slots); wp.c_slots = std::min(8, wp.c_slots);
      if ((use_r && wp.r_slots < 2) || (p.has_act && wp.c_slots < 2)) { delete pl; return failmsg("wide-epilogue conv: shared memory does not hold the staging rings"); }
      const int sms = sm_count();
      const long total = long(wp.m_tiles) * wp.n_tiles;
      int grid = int(std::min<long>(total, sms));
      grid = std::max(wp.n_tiles, (grid / wp.n_tiles) * wp.n_tiles);
      if (pl->pairwide) {                       // clusters of two CTAs; the cluster count is a multiple of the N tile count
        const long pair_total = long((wp.m_tiles + 1) / 2) * wp.n_tiles;
        int clusters = int(std::min<long>(pair_total, sms / 2));
        clusters = std::max(wp.n_tiles, (clusters / wp.n_tiles) * wp.n_tiles);
        grid = 2 * clusters;
      }
      pl->grid = grid;
      pl->smem_bytes = fixed + slab + uint32_t(wp.stages) * stage_bytes +
                       uint32_t((use_r ? wp.r_slots : 0) + (p.has_act ? wp.c_slots : 0)) * kChunkBytes;
      std::string wmsg;
      bool wok = encode_act(&pl->tmA, dt, in, g.B, g.H, g.W, g.Cin, bw, bh, bn, g.stride, wmsg);
      if (wok && e.in_b) {
        if (e.in_b_cin % 64 || e.in_b_stride < 1 || (e.in_b_h - 1) / e.in_b_stride + 1 != g.Ho || (e.in_b_w - 1) / e.in_b_stride + 1 != g.Wo) {
          delete pl;
          return failmsg("two-input conv: the second input must have 64-multiple channels and map onto the output grid at its stride");
        }
        wok = encode_act(&pl->tmA2, dt, e.in_b, g.B, e.in_b_h, e.in_b_w, e.in_b_cin, bw, bh, bn, e.in_b_stride, wmsg);
      } else {
        pl->tmA2 = pl->tmA;
      }
      if (wok) {
        uint64_t wd[2] = {uint64_t(Ktot), uint64_t(g.Cout)};
        uint64_t ws[1] = {uint64_t(Ktot) * 2};
        uint32_t wb[2] = {64, uint32_t(pl->pairwide ? N / 2 : N)};     // a CTA of a pair stages half of the weight tile
        uint32_t we[2] = {1, 1};
        wok = encode(&pl->tmW, dt, 2, w_packed, wd, ws, wb, we, false, wmsg);
      }
      const void* any_out = e.out ? e.out : e.out_act;
      if (wok) wok = encode_act(&pl->tmOut, dt, any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, wmsg);
      if (wok) wok = encode_act(&pl->tmAct, dt, e.out_act ? e.out_act : any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, wmsg);
      if (wok) wok = encode_act(&pl->tmRes, dt, e.residual ? e.residual : any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, wmsg);
      if (!wok) { delete pl; return failmsg(wmsg); }
      if (std::getenv("B200RN_VERBOSE"))
        std::fprintf(stderr, "[b200rn] conv 1x1 s%d %d->%d @%dx%d B%d: %swide-epilogue N=%d bstat=%d patch=%dx%dx%d m_tiles=%d n_tiles=%d grid=%d stages=%d r_slots=%d c_slots=%d smem=%u\n",
                     g.stride, g.Cin, g.Cout, g.H, g.W, g.B, pl->pairwide ? "pair-" : "", N, wp.b_stationary, bw, bh, bn, wp.m_tiles, wp.n_tiles, grid, wp.stages,
                     wp.r_slots, wp.c_slots, pl->smem_bytes);
      return pl;
    }
  }
  if (e.in_b) { delete pl; return failmsg("two-input conv needs a 1x1 / pad-0 conv with 256-multiple output channels (conv_wide_kernel / conv_tail_kernel)"); }
  plan_geometry(*pl);
  const int bw = p.bw, bh = p.bh, bn = p.bn;
  {
    // Large-K layer with a plain epilogue: one 256 x N tile per CTA pair (conv_pair_kernel)
    const int m_tiles = p.tiles_w * p.tiles_h * ((g.B + bn - 1) / bn);
    const int sms = sm_count();
    if (pair_enabled() && !stem_mode && p.has_out && !p.has_res && !p.has_act && !p.has_pre && !p.out_f32 &&
        p.num_k_iters >= pair_min_k() && pair_n_ok(g.Cout) && m_tiles >= 2 && sms >= 2 && pair_items_ok(m_tiles, g.Cout)) {
      PairParams& pp = pl->pp;
      std::memset(&pp, 0, sizeof pp);
      pl->pair = 1;
      pl->block_n = (g.Cout % 256 == 0) ? 256 : 128;
      pp.num_k_iters = p.num_k_iters; pp.cblks = p.cblks; pp.kw = p.kw; pp.stride = p.stride; pp.pad = p.pad;
      pp.bw = bw; pp.bh = bh; pp.bn = bn; pp.tiles_w = p.tiles_w; pp.tiles_h = p.tiles_h;
      pp.m_tiles = m_tiles; pp.n_tiles = g.Cout / pl->block_n;
      pp.total_items = ((m_tiles + 1) / 2) * pp.n_tiles;
      pp.relu = p.relu; pp.a_bytes = p.a_bytes; pp.bias = p.bias;
      const uint32_t stage_bytes = uint32_t(kTileM * 128 + (pl->block_n / 2) * 128);
      const uint32_t fixed = 1024u + 2 * kChunkBytes + 8u * (2 * 8 + 4) + 64;
      pp.num_stages = std::max(2, std::min(8, int((227u * 1024 - fixed) / stage_bytes)));
      pl->smem_bytes = fixed + uint32_t(pp.num_stages) * stage_bytes;
      pl->grid = 2 * std::min(pp.total_items, sms / 2);
      p.n_tiles = pp.n_tiles;
      p.b_bytes = uint32_t(pl->block_n) * 128;
    }
  }
  if (!pl->pair) {
    // Large-K layer with a plain epilogue: consider two M tiles per weight tile.  Same cost model as the N-tile
    // choice: (rounds over the SMs) x (bytes staged per K block) must come out lower than the single-M plan's.
    const char* denv = std::getenv("B200RN_DUALM");
    const int m_tiles = p.tiles_w * p.tiles_h * ((g.B + bn - 1) / bn);
    if (!(denv && denv[0] == '0') && !stem_mode && p.has_out && !p.has_res && !p.has_act && !p.has_pre &&
        p.num_k_iters >= 8 && (pl->block_n == 256 || pl->block_n == 128) && m_tiles >= 2) {
      const int sms = sm_count();
      const long items = long((m_tiles + 1) / 2) * p.n_tiles;
      const long cost_dual = ((items + sms - 1) / sms) * long(2 * kTileM + pl->block_n);
      const long cost_single = ((long(p.total_tiles) + sms - 1) / sms) * long(kTileM + pl->block_n);
      const uint32_t stage_bytes = uint32_t(2 * kTileM * 128 + pl->block_n * 128);
      const uint32_t fixed = 1024u + 2 * kChunkBytes + uint32_t(pl->block_n) * 4 + 8u * (2 * 8 + 2) + 64;
      const int stages = int((227u * 1024 - fixed) / stage_bytes);
      if (cost_dual < cost_single && stages >= 2) {
        DualParams& dp = pl->dp;
        std::memset(&dp, 0, sizeof dp);
        pl->dual = 1;
        dp.num_k_iters = p.num_k_iters; dp.cblks = p.cblks; dp.kw = p.kw; dp.stride = p.stride; dp.pad = p.pad;
        dp.bw = bw; dp.bh = bh; dp.bn = bn; dp.tiles_w = p.tiles_w; dp.tiles_h = p.tiles_h;
        dp.m_tiles = m_tiles; dp.n_tiles = p.n_tiles; dp.total_items = int(items);
        dp.num_stages = std::min(stages, 6);
        dp.relu = p.relu; dp.a_bytes = p.a_bytes; dp.b_bytes = p.b_bytes; dp.bias = p.bias;
        pl->smem_bytes = fixed + uint32_t(dp.num_stages) * stage_bytes;
        pl->grid = int(std::min<long>(items, sms));
      }
    }
  }

  std::string msg;
  bool ok = true;
  if (stem_mode) {
    // 5-D overlapping view of the zero-framed input [B, Hp, Wp, 4]:
    //   d0 = 32 elements (8 px x 4 ch), d1 = output column (stride 2 px = 16 B), d2 = row parity,
    //   d3 = row pair, d4 = image
    const int Hp = stem_Hp(g.H), Wp = stem_Wp(g.W);
    const uint64_t rowpitch = uint64_t(Wp) * 4 * 2;
    uint64_t dims[5] = {32, uint64_t(g.Wo), 2, uint64_t(Hp / 2), uint64_t(g.B)};
    uint64_t str[4] = {16, rowpitch, 2 * rowpitch, uint64_t(Hp) * rowpitch};
    uint32_t box[5] = {32, uint32_t(bw), 1, uint32_t(bh), uint32_t(bn)};
    uint32_t es[5] = {1, 1, 1, 1, 1};
    ok = encode(&pl->tmA, dt, 5, in, dims, str, box, es, true, msg);
    if (ok) {
      uint64_t wd[2] = {uint64_t(7 * 32), uint64_t(g.Cout)};
      uint64_t ws[1] = {uint64_t(7 * 32) * 2};
      uint32_t wb[2] = {32, uint32_t(pl->block_n)};
      uint32_t we[2] = {1, 1};
      ok = encode(&pl->tmW, dt, 2, w_packed, wd, ws, wb, we, true, msg);
    }
  } else {
    ok = encode_act(&pl->tmA, dt, in, g.B, g.H, g.W, g.Cin, bw, bh, bn, g.stride, msg);
    if (ok) {
      const uint64_t K = uint64_t(g.kh) * g.kw * g.Cin;
      uint64_t wd[2] = {K, uint64_t(g.Cout)};
      uint64_t ws[1] = {K * 2};
      uint32_t wb[2] = {64, uint32_t(pl->pair ? pl->block_n / 2 : pl->block_n)};   // a CTA of a pair stages half the N tile
      uint32_t we[2] = {1, 1};
      ok = encode(&pl->tmW, dt, 2, w_packed, wd, ws, wb, we, false, msg);
    }
  }
  // unused maps still need to be valid objects to pass by value: alias the output map
  const void* any_out = e.out ? e.out : (e.out_act ? e.out_act : in);   // fp32-row convs have no 16-bit output: any valid address
  if (ok) ok = encode_act(&pl->tmOut, dt, e.out ? e.out : any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, msg);
  if (ok) ok = encode_act(&pl->tmAct, dt, e.out_act ? e.out_act : any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, msg);
  if (ok) ok = encode_act(&pl->tmRes, dt, e.residual ? e.residual : any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, msg);
  if (!ok) { delete pl; return failmsg(msg); }
  if (std::getenv("B200RN_VERBOSE"))
    std::fprintf(stderr, "[b200rn] conv %dx%d s%d %d->%d @%dx%d B%d: %s bn=%d patch=%dx%dx%d tiles=%d grid=%d stages=%d kgroup=%d bstat=%d smem=%u\n",
                 g.kh, g.kw, g.stride, g.Cin, g.Cout, g.H, g.W, g.B, pl->pair ? "pair (2-CTA)" : (pl->dual ? "dual-M" : "generic"), pl->block_n, bw, bh, bn,
                 pl->pair ? pl->pp.total_items : (pl->dual ? pl->dp.total_items : p.total_tiles), pl->grid,
                 pl->pair ? pl->pp.num_stages : (pl->dual ? pl->dp.num_stages : p.num_stages), p.kgroup,
                 p.b_stationary, pl->smem_bytes);
  return pl;
}

void umma_conv_plan_set_out_f32(UmmaConvPlan* p, float* out_f32) {
  if (p && p->p.out_f32) p->p.out_f32 = out_f32;
}

void umma_conv_plan_destroy(UmmaConvPlan* p) {
  if (p && p->stem2 && p->sp.trace) {
    std::vector<unsigned long long> h(64 * 8);
    cudaDeviceSynchronize();
    cudaMemcpy(h.data(), p->sp.trace, h.size() * 8, cudaMemcpyDeviceToHost);
    const unsigned long long t0 = h[0];
    std::fprintf(stderr, "stem kernel CTA 0 (groups %d stages %d), per tile: producer slot free | MMA top | waits done | MMAs issued | committed | "
                         "epilogue top | accumulator full | epilogue end (cycles)\n", p->sp.groups, p->sp.stages);
    for (int i = 0; i < 40; ++i) {
      std::fprintf(stderr, "tile %2d:", i);
      for (int k = 0; k < 8; ++k) std::fprintf(stderr, " %7lld", h[size_t(i) * 8 + k] ? (long long)(h[size_t(i) * 8 + k] - t0) : -1LL);
      std::fprintf(stderr, "\n");
    }
    cudaFree(p->sp.trace);
    delete p;
    return;
  }
  if (p && p->wide && p->wp.trace) {
    std::vector<unsigned long long> h(96 * 8);
    cudaDeviceSynchronize();
    cudaMemcpy(h.data(), p->wp.trace, h.size() * 8, cudaMemcpyDeviceToHost);
    const unsigned long long t0 = h[0];
    auto rel = [&](unsigned long long v) { return v ? (long long)(v - t0) : -1LL; };
    std::fprintf(stderr, "wide kernel CTA 0 (k_iters %d stages %d r_slots %d c_slots %d)\n", p->wp.num_k_iters, p->wp.stages, p->wp.r_slots, p->wp.c_slots);
    std::fprintf(stderr, "MMA thread, per tile: top | accumulator free | K block 0..4 landed (cycles)\n");
    for (int i = 0; i < 8; ++i) {
      std::fprintf(stderr, "tile %2d:", i);
      for (int k = 0; k < 7; ++k) std::fprintf(stderr, " %7lld", rel(h[size_t(i) * 8 + k]));
      std::fprintf(stderr, "\n");
    }
    std::fprintf(stderr, "epilogue (first warp of the chunk's group), per chunk: top | slots ready | math + smem writes done | signalled ; tile top\n");
    for (int cc = 0; cc < 32; ++cc) {
      std::fprintf(stderr, "chunk %2d:", cc);
      for (int k = 0; k < 4; ++k) std::fprintf(stderr, " %7lld", rel(h[size_t(32 + cc) * 8 + k]));
      std::fprintf(stderr, " ; %7lld\n", rel(h[size_t(32 + cc) * 8 + 7]));
    }
    cudaFree(p->wp.trace);
    delete p;
    return;
  }
  if (p && p->tail && p->tp.trace) {
    std::vector<unsigned long long> h(8 * 32 * 8);
    cudaDeviceSynchronize();
    cudaMemcpy(h.data(), p->tp.trace, h.size() * 8, cudaMemcpyDeviceToHost);
    unsigned long long t0 = std::min(h[0], h[size_t(32 + 1) * 8]);
    std::fprintf(stderr, "tail kernel, CTA 0, first warp of each epilogue group, per chunk: top | buffers ready | - | - | tmem + residual read | "
                         "math + smem writes | signalled | tail tile done (cycles)\n");
    for (int i = 8; i < 28; ++i) {
      const int w = i & 1;
      std::fprintf(stderr, "chunk %2d group %d:", i, w);
      for (int k = 0; k < 8; ++k) {
        unsigned long long v = h[(size_t(w) * 32 + i) * 8 + k];
        std::fprintf(stderr, " %7lld", v ? (long long)(v - t0) : -1LL);
      }
      std::fprintf(stderr, "\n");
    }
    cudaFree(p->tp.trace);
    delete p;
    return;
  }
  if (p && !p->halo && p->p.trace) {
    std::vector<unsigned long long> h(64 * 8);
    cudaDeviceSynchronize();
    cudaMemcpy(h.data(), p->p.trace, h.size() * 8, cudaMemcpyDeviceToHost);
    unsigned long long t0 = h[0];
    std::fprintf(stderr, "generic kernel, k_iters %d kgroup %d stages %d bn %d bstat %d grid %d tiles %d\n", p->p.num_k_iters, p->p.kgroup,
                 p->p.num_stages, p->block_n, p->p.b_stationary, p->grid, p->p.total_tiles);
    std::fprintf(stderr, "tile: first_load_issue | mma_top got_tmem got_first_stage all_issued | epi_top epi_got_acc epi_end (cycles)\n");
    for (int i = 0; i < 20; ++i) {
      std::fprintf(stderr, "%2d:", i);
      for (int k = 0; k < 8; ++k) std::fprintf(stderr, " %8lld", h[i * 8 + k] ? (long long)(h[i * 8 + k] - t0) : -1LL);
      std::fprintf(stderr, "\n");
    }
    cudaFree(p->p.trace);
  }
  if (p && p->halo && p->hp.trace) {
    // timeline experiment: print CTA 0's per-tile stamps (cycles relative to the first)
    std::vector<unsigned long long> h(64 * 8);
    cudaDeviceSynchronize();
    cudaMemcpy(h.data(), p->hp.trace, h.size() * 8, cudaMemcpyDeviceToHost);
    unsigned long long t0 = h[0];
    std::fprintf(stderr, "unit: load_issue | lookahead_wait_begin unit_top mma_issued after_commit1 after_commit2 | - lookahead_wait_end (cycles)\n");
    for (int i = 0; i < 24; ++i) {
      std::fprintf(stderr, "%2d:", i);
      for (int k = 0; k < 8; ++k) std::fprintf(stderr, " %8lld", h[i * 8 + k] ? (long long)(h[i * 8 + k] - t0) : -1LL);
      std::fprintf(stderr, "\n");
    }
    cudaFree(p->hp.trace);
  }
  delete p;
}


// ---- exact mode on the tensor cores: 3 x TF32 (conv_tf32.cuh) --------------------------------------------------
struct Tf32ConvPlan {
  ConvGeom g;
  int block_n;
  Tf32Params p;
  CUtensorMap tmA, tmWhi, tmWlo, tmOut, tmAct, tmRes;
  uint32_t smem_bytes;
  int grid;
};

bool tf32_conv_supported(const ConvGeom& g) {
  return g.Cin % 32 == 0 && g.Cout % 64 == 0 && g.kh == g.kw && (g.stride == 1 || g.stride == 2);
}
// the 7x7 / stride-2 / pad-3 stem on a 4-channel (3 padded) input of even height, read from a zero-framed copy
bool tf32_stem_supported(const ConvGeom& g) {
  return g.kh == 7 && g.kw == 7 && g.stride == 2 && g.pad == 3 && g.Cin == 4 && g.Cout % 64 == 0 && g.H % 2 == 0;
}
size_t tf32_stem_weight_elems(int Cout) { return size_t(Cout) * 7 * 32; }
// [Cout][7][7][4] -> [Cout][ky 0..6][kx 0..7][c 0..3] with kx = 7 zero
void tf32_pack_stem_weights(const float* w, int Cout, float* dst) {
  for (int o = 0; o < Cout; ++o)
    for (int ky = 0; ky < 7; ++ky)
      for (int kx = 0; kx < 8; ++kx)
        for (int c = 0; c < 4; ++c)
          dst[((size_t(o) * 7 + ky) * 8 + kx) * 4 + c] = kx < 7 ? w[((size_t(o) * 7 + ky) * 7 + kx) * 4 + c] : 0.f;
}

void tf32_split_weights(const float* w, size_t n, float* hi, float* lo) {
  for (size_t i = 0; i < n; ++i) {
    uint32_t u;
    std::memcpy(&u, &w[i], 4);
    u &= 0xFFFFE000u;
    float h;
    std::memcpy(&h, &u, 4);
    float l = w[i] - h;               // exact: the 13 bits that were cut
    std::memcpy(&u, &l, 4);
    u &= 0xFFFFE000u;
    std::memcpy(&l, &u, 4);
    hi[i] = h; lo[i] = l;
  }
}

static bool encode_f32(CUtensorMap* m, int rank, const void* base, const uint64_t* dims, const uint64_t* strides_bytes,
                       const uint32_t* box, const uint32_t* estr, std::string& err) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { err = "cuTensorMapEncodeTiled is not available from the driver"; return false; }
  cuuint64_t d[5]; cuuint64_t st[4]; cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = estr[i]; }
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, cuuint32_t(rank), const_cast<void*>(base), d, st, b, e,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[200];
    std::snprintf(buf, sizeof buf, "cuTensorMapEncodeTiled (fp32) failed (CUresult %d) rank=%d box=[%u,%u,%u,%u]", int(r), rank, box[0], box[1],
                  rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    err = buf;
    return false;
  }
  return true;
}

// fp32 NHWC activation [B,H,W,C] as a 4-D map (C, W, H, N) with a {32, bw*es, bh*es, bn} box
static bool encode_act_f32(CUtensorMap* m, const void* base, int B, int H, int W, int C, int bw, int bh, int bn, int es, std::string& err) {
  uint64_t dims[4] = {uint64_t(C), uint64_t(W), uint64_t(H), uint64_t(B)};
  uint64_t str[3] = {uint64_t(C) * 4, uint64_t(W) * C * 4, uint64_t(H) * W * C * 4};
  uint32_t box[4] = {32, uint32_t(bw * es), uint32_t(bh * es), uint32_t(bn)};
  uint32_t estr[4] = {1, uint32_t(es), uint32_t(es), 1};
  return encode_f32(m, 4, base, dims, str, box, estr, err);
}

Tf32ConvPlan* tf32_conv_plan_create(const ConvGeom& g_in, const float* in, const float* w_hi, const float* w_lo, const Epilogue& e_in,
                                    char* err, size_t errcap, int stem) {
  auto failmsg = [&](const std::string& m) -> Tf32ConvPlan* {
    if (err && errcap) std::snprintf(err, errcap, "%s", m.c_str());
    return nullptr;
  };
  ConvGeom g = g_in;
  Epilogue e = e_in;
  if (!e.in_b && !e.out_f32) flatten_1x1(g, e, stem);       // 1x1 / stride-1 convs: M tile = 128 consecutive pixel rows
  if (stem ? !tf32_stem_supported(g) : !tf32_conv_supported(g))
    return failmsg(stem ? "3xTF32 stem needs a 7x7/2 pad-3 conv on a 4-channel input of even height with Cout % 64 == 0"
                        : "3xTF32 conv needs Cin % 32 == 0, Cout % 64 == 0, a square filter and stride 1 or 2");
  if (e.pre_scale || e.tail_out || e.in_b || e.out_f32 || e.pool_fused) return failmsg("3xTF32 conv has no prologue / tail / second input / fp32-row / pool form");
  if (!e.out && !e.out_act) return failmsg("conv with no output");
  auto pl = new Tf32ConvPlan();
  pl->g = g;
  Tf32Params& p = pl->p;
  std::memset(&p, 0, sizeof p);
  pl->block_n = (g.Cout % 128 == 0) ? 128 : 64;
  int bw = 1, bh = 1, bn = 1;
  choose_patch(g.B, g.Ho, g.Wo, bw, bh, bn);
  p.bw = bw; p.bh = bh; p.bn = bn;
  p.tiles_w = (g.Wo + bw - 1) / bw; p.tiles_h = (g.Ho + bh - 1) / bh;
  p.n_tiles = g.Cout / pl->block_n;
  p.total_tiles = p.tiles_w * p.tiles_h * ((g.B + bn - 1) / bn) * p.n_tiles;
  p.stem = stem;
  p.cblks = stem ? 1 : g.Cin / 32; p.kw = g.kw; p.stride = g.stride; p.pad = g.pad;
  p.num_k_iters = stem ? 7 : g.kh * g.kw * p.cblks;
  p.rows = uint32_t(bw * bh * bn); p.a_bytes = p.rows * 128u;
  p.has_res = e.residual != nullptr; p.has_out = e.out != nullptr; p.has_act = e.out_act != nullptr;
  p.relu = e.relu; p.act_relu = e.act_relu;
  p.cout = g.Cout; p.Ho = g.Ho; p.Wo = g.Wo; p.B = g.B;
  p.bias = e.bias; p.act_scale = e.act_scale; p.act_shift = e.act_shift;
  p.residual = static_cast<const float*>(e.residual);
  if (p.has_act && (!e.act_scale || !e.act_shift)) { delete pl; return failmsg("activated output needs scale and shift"); }
  const uint32_t stage_bytes = 2u * kTileM * 128 + 2u * uint32_t(pl->block_n) * 128;
  // staging: one 16 KB chunk per output and epilogue group
  const uint32_t fixed = 1024u + 2u * uint32_t(p.has_out + p.has_act) * kChunkBytes + 8u * (3 * 8 + 10) + 64;
  p.num_stages = std::max(2, std::min(8, int((227u * 1024 - fixed) / stage_bytes)));
  pl->smem_bytes = fixed + uint32_t(p.num_stages) * stage_bytes;
  {
    const char* rh = std::getenv("B200RN_TF32_RAW_HI");
    p.raw_hi = !(rh && rh[0] == '0');
    // Both A_hi products as one instruction of width 2N: for the 3x3 convs (18-144 K blocks).  Measured at batch 256:
    // 3x3 layers 0.305 -> 0.293 (N = 128), 0.446 -> 0.424 ms (N = 64); the 1x1 layers with 16-32 K blocks lose 5-10 %
    // (the next tile's first MMA waits for the epilogue to read the single LO pair), so they keep the three-instruction
    // form.  B200RN_TF32_FUSED=0 off, =N smallest K-block count of a 3x3 layer (default 16), negative: any filter size.
    const char* fe = std::getenv("B200RN_TF32_FUSED");
    const int fmin = fe ? std::atoi(fe) : 16;
    p.fused = !stem && fmin != 0 && (fmin < 0 || g.kh * g.kw >= 9) && p.num_k_iters >= std::max(2 * kT32ChunkBlocks, std::abs(fmin));
  }
  {
    // Residual by TMA into a 16 KB landing buffer per epilogue group, one chunk ahead, instead of per-thread global loads
    // of 64-byte row pieces -- also when the two buffers cost an operand stage.  Measured at batch 256 (exact mode):
    // residual + raw + activated layers 0.79 -> 0.72 (56x56), 0.44 -> 0.385 (28x28), 0.26 -> 0.24 ms (14x14); the
    // activated-only block ends, which give up their third stage, 0.78 -> 0.51, 0.44 -> 0.30, 0.26 -> 0.22 ms.
    // B200RN_TF32_RES_TMA: 0 = never, 1 = only where no stage is lost.
    const char* re = std::getenv("B200RN_TF32_RES_TMA");
    const int mode = re ? std::atoi(re) : 2;
    const uint32_t with_res = pl->smem_bytes + 2u * kChunkBytes;
    if (p.has_res && mode >= 1 && with_res <= 227u * 1024) { p.res_tma = 1; pl->smem_bytes = with_res; }
    else if (p.has_res && mode >= 2 && p.num_stages > 2) { p.res_tma = 1; --p.num_stages; pl->smem_bytes = fixed + 2u * kChunkBytes + uint32_t(p.num_stages) * stage_bytes; }
  }
  pl->grid = std::min(p.total_tiles, sm_count());
  std::string msg;
  bool ok;
  if (stem) {
    // 5-D overlapping view of the zero-framed input [B, Hp, Wp, 4] fp32 (image at (3,3)):
    //   d0 = 32 floats (8 px x 4 ch), d1 = output column (2 px = 32 B apart), d2 = row parity, d3 = row pair, d4 = image
    const int Hp = stem_Hp(g.H), Wp = stem_Wp(g.W);
    const uint64_t rowpitch = uint64_t(Wp) * 4 * 4;
    uint64_t dims[5] = {32, uint64_t(g.Wo), 2, uint64_t(Hp / 2), uint64_t(g.B)};
    uint64_t str[4] = {32, rowpitch, 2 * rowpitch, uint64_t(Hp) * rowpitch};
    uint32_t box[5] = {32, uint32_t(bw), 1, uint32_t(bh), uint32_t(bn)};
    uint32_t es[5] = {1, 1, 1, 1, 1};
    ok = encode_f32(&pl->tmA, 5, in, dims, str, box, es, msg);
  } else {
    ok = encode_act_f32(&pl->tmA, in, g.B, g.H, g.W, g.Cin, bw, bh, bn, g.stride, msg);
  }
  const uint64_t K = stem ? uint64_t(7 * 32) : uint64_t(g.kh) * g.kw * g.Cin;
  uint64_t wd[2] = {K, uint64_t(g.Cout)};
  uint64_t ws[1] = {K * 4};
  uint32_t wb[2] = {32, uint32_t(pl->block_n)};
  uint32_t we[2] = {1, 1};
  if (ok) ok = encode_f32(&pl->tmWhi, 2, w_hi, wd, ws, wb, we, msg);
  if (ok) ok = encode_f32(&pl->tmWlo, 2, w_lo, wd, ws, wb, we, msg);
  const void* any_out = e.out ? e.out : e.out_act;
  if (ok) ok = encode_act_f32(&pl->tmOut, e.out ? e.out : any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, msg);
  if (ok) ok = encode_act_f32(&pl->tmAct, e.out_act ? e.out_act : any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, msg);
  if (ok) ok = encode_act_f32(&pl->tmRes, p.res_tma ? e.residual : any_out, g.B, g.Ho, g.Wo, g.Cout, bw, bh, bn, 1, msg);
  if (!ok) { delete pl; return failmsg(msg); }
  if (std::getenv("B200RN_VERBOSE"))
    std::fprintf(stderr, "[b200rn] conv %dx%d s%d %d->%d @%dx%d B%d: 3xTF32 bn=%d patch=%dx%dx%d tiles=%d grid=%d stages=%d res_tma=%d smem=%u\n", g.kh, g.kw,
                 g.stride, g.Cin, g.Cout, g.H, g.W, g.B, pl->block_n, bw, bh, bn, p.total_tiles, pl->grid, p.num_stages, p.res_tma, pl->smem_bytes);
  return pl;
}

void tf32_conv_plan_destroy(Tf32ConvPlan* p) { delete p; }

template <int BLOCK_N, bool RES_TMA>
static cudaError_t launch_tf32_inst(const Tf32ConvPlan* pl, cudaStream_t s) {
  auto kern = conv_tf32x3_kernel<BLOCK_N, RES_TMA>;
  static thread_local int configured_device = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (configured_device != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    configured_device = dev;
  }
  return launch_pdl(kern, pl->grid, kT32Threads, pl->smem_bytes, s, pl->tmA, pl->tmWhi, pl->tmWlo, pl->tmOut, pl->tmAct, pl->tmRes, pl->p);
}

cudaError_t launch_conv_tf32(const Tf32ConvPlan* pl, cudaStream_t s) {
  // the residual-by-TMA form is a separate instantiation: it does not keep a chunk's residual in registers across the K loop
  if (pl->p.res_tma) return pl->block_n == 128 ? launch_tf32_inst<128, true>(pl, s) : launch_tf32_inst<64, true>(pl, s);
  return pl->block_n == 128 ? launch_tf32_inst<128, false>(pl, s) : launch_tf32_inst<64, false>(pl, s);
}

cudaError_t umma_take_timeout_flag(unsigned int* flag) {
  *flag = 0;
  cudaError_t e = cudaMemcpyFromSymbol(flag, g_umma_timeout, sizeof(unsigned int));
  if (e != cudaSuccess) return e;
  if (*flag) {
    unsigned int zero = 0;
    e = cudaMemcpyToSymbol(g_umma_timeout, &zero, sizeof zero);
  }
  return e;
}

}  // namespace rn

// =================================================================================================
// b200rn_stem_conv_static -- the stem convolution for callers that must not allocate
// =================================================================================================
// Some hosts embed this library where device memory cannot be requested at call time: a worker
// whose memory pool was sized and handed out at start-up, a latency-critical request thread for
// which cudaMalloc's implicit device synchronisation is unacceptable, or a plugin slot whose
// contract forbids the plugin to own device memory.  b200rn_stem_conv needs device scratch (the
// zero-framed re-laid input, the packed filter bank, bias and input scale/shift) and takes it from
// cudaMalloc.  This variant takes the same scratch from the fixed array below instead, so it calls
// no cudaMalloc / cudaFree and no allocating routine of any library, on the host side either.
//
// The array costs kStaticScratchBytes (4 MiB + 58 KiB) of device memory on every GPU that loads
// libb200rn.so, from load until the process exits, whether or not the entry point is ever called.
// All calls on a device share it and it keeps whatever the previous call left behind, so every call
// first zeroes the part of the frame it is about to use (the frame's borders are the convolution's
// zero padding).  Two calls whose GPU work overlaps -- different host threads are serialised here by
// a mutex and each call synchronises before returning, but another process-external stream or a
// captured graph replay could still overlap -- would race on the array.
namespace rn {
namespace {
constexpr size_t kStaticFrameBytes = size_t(4) << 20;                     // framed input images
constexpr int kStaticMaxCout = 128;
constexpr size_t kStaticWeightBytes = size_t(kStaticMaxCout) * 7 * 32 * 2;  // packed 16-bit filters
constexpr size_t kStaticParamBytes = 1024;                                 // bias[128] + scale[4] + shift[4] (fp32)
constexpr size_t kStaticScratchBytes = kStaticFrameBytes + kStaticWeightBytes + kStaticParamBytes;
constexpr int kStaticMaxDevices = 64;
}  // namespace
}  // namespace rn

// In host code this name is only a handle for the runtime, not a device address: passing it to a
// kernel would fault (or silently touch host memory).  Host code below resolves the real address
// with cudaGetSymbolAddress, once per device.
static __device__ __align__(256) unsigned char g_static_scratch[rn::kStaticScratchBytes];

namespace rn {
namespace {

std::mutex g_static_mu;
unsigned char* g_static_dev_addr[kStaticMaxDevices] = {nullptr};   // resolved address per device ordinal
// host-side staging, also static: nothing is allocated per call
unsigned char g_static_host_weights[kStaticWeightBytes];
float g_static_host_params[kStaticParamBytes / 4];

CUresult encode_plain(CUtensorMap* m, DType dt, int rank, const void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box, bool sw64) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return CUDA_ERROR_NOT_SUPPORTED;
  cuuint64_t d[5]; cuuint64_t s[4]; cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; e[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
  return fn(m, dt == DType::BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
            cuuint32_t(rank), const_cast<void*>(base), d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
            sw64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

// Same tile geometry, pipeline depth and tensor maps that umma_conv_plan_create builds for a stem
// layer, written into a caller-provided (stack) object.
bool fill_stem_plan(UmmaConvPlan& pl, DType dt, const ConvGeom& g, const void* frame, const void* w_packed,
                    const float* d_bias, void* d_out, int relu) {
  pl.dt = dt; pl.g = g; pl.stem = 1;
  UmmaParams& p = pl.p;
  std::memset(&p, 0, sizeof p);
  p.has_out = 1; p.relu = relu; p.bias = d_bias;
  plan_geometry(pl);
  const int bw = p.bw, bh = p.bh, bn = p.bn;

  const int Hp = stem_Hp(g.H), Wp = stem_Wp(g.W);
  const uint64_t rowpitch = uint64_t(Wp) * 4 * 2;
  uint64_t dims[5] = {32, uint64_t(g.Wo), 2, uint64_t(Hp / 2), uint64_t(g.B)};
  uint64_t str[4] = {16, rowpitch, 2 * rowpitch, uint64_t(Hp) * rowpitch};
  uint32_t box[5] = {32, uint32_t(bw), 1, uint32_t(bh), uint32_t(bn)};
  if (encode_plain(&pl.tmA, dt, 5, frame, dims, str, box, true) != CUDA_SUCCESS) return false;
  uint64_t wd[2] = {uint64_t(7 * 32), uint64_t(g.Cout)};
  uint64_t ws[1] = {uint64_t(7 * 32) * 2};
  uint32_t wb[2] = {32, 64};
  if (encode_plain(&pl.tmW, dt, 2, w_packed, wd, ws, wb, true) != CUDA_SUCCESS) return false;
  uint64_t od[4] = {uint64_t(g.Cout), uint64_t(g.Wo), uint64_t(g.Ho), uint64_t(g.B)};
  uint64_t os[3] = {uint64_t(g.Cout) * 2, uint64_t(g.Wo) * g.Cout * 2, uint64_t(g.Ho) * g.Wo * g.Cout * 2};
  uint32_t ob[4] = {64, uint32_t(bw), uint32_t(bh), uint32_t(bn)};
  if (encode_plain(&pl.tmOut, dt, 4, d_out, od, os, ob, false) != CUDA_SUCCESS) return false;
  pl.tmAct = pl.tmOut;    // unused by this launch, but passed by value: must be valid objects
  pl.tmRes = pl.tmOut;
  return true;
}

}  // namespace
}  // namespace rn

// Declared in include/b200rn.h.  Same arguments and same result as b200rn_stem_conv, plus an
// optional buffer for the error text.  The FIRST call on each device resolves the scratch array's
// device address (cudaGetSymbolAddress); make that first call outside any stream capture.
extern "C" int b200rn_stem_conv_static(int dtype, const float* d_in_nchw, int B, int H, int W, const float* in_scale,
                                       const float* in_shift, const float* w, const float* bias, int Cout, int relu,
                                       void* d_out, char* err, size_t err_cap) {
  using namespace rn;
  auto fail = [&](int code, const char* fmt, long long a = 0, long long b = 0) {
    if (err && err_cap) std::snprintf(err, err_cap, fmt, a, b);
    return code;
  };
  if (!d_in_nchw || !w || !d_out || dtype < 1 || dtype > 2 || B <= 0 || H <= 0 || W <= 0)
    return fail(1, "b200rn_stem_conv_static: bad argument");
  if (H % 2) return fail(1, "b200rn_stem_conv_static: image height %lld must be even", H);
  if (Cout <= 0 || Cout % 64 || Cout > kStaticMaxCout)
    return fail(1, "b200rn_stem_conv_static: Cout %lld must be a multiple of 64 and at most %lld", Cout, kStaticMaxCout);
  const int Hp = stem_Hp(H), Wp = stem_Wp(W);
  const size_t frame_bytes = size_t(B) * Hp * Wp * 4 * 2;
  if (frame_bytes > kStaticFrameBytes)
    return fail(1, "b200rn_stem_conv_static: this batch needs %lld bytes of frame scratch, the static array holds %lld "
                   "(use b200rn_stem_conv, which allocates, for larger batches)", (long long)frame_bytes, (long long)kStaticFrameBytes);

  std::lock_guard<std::mutex> lk(g_static_mu);
  int dev = 0;
  cudaError_t ce = cudaGetDevice(&dev);
  if (ce != cudaSuccess) return fail(3, "b200rn_stem_conv_static: cudaGetDevice failed (%lld)", ce);
  if (dev < 0 || dev >= kStaticMaxDevices) return fail(3, "b200rn_stem_conv_static: device ordinal %lld not supported", dev);
  if (!g_static_dev_addr[dev]) {                 // one-time set-up for this device, with it current
    void* p = nullptr;
    ce = cudaGetSymbolAddress(&p, g_static_scratch);
    if (ce != cudaSuccess || !p) return fail(3, "b200rn_stem_conv_static: cudaGetSymbolAddress failed (%lld)", ce);
    g_static_dev_addr[dev] = static_cast<unsigned char*>(p);
  }
  unsigned char* scratch = g_static_dev_addr[dev];
  unsigned char* d_frame = scratch;
  unsigned char* d_weights = scratch + kStaticFrameBytes;
  float* d_params = reinterpret_cast<float*>(scratch + kStaticFrameBytes + kStaticWeightBytes);
  float* d_bias = d_params;                       // [Cout]
  float* d_scale = d_params + kStaticMaxCout;     // [4]
  float* d_shift = d_scale + 4;                   // [4]

  const DType dt = DType(dtype);
  const int Ho = (H + 6 - 7) / 2 + 1, Wo = (W + 6 - 7) / 2 + 1;
  ConvGeom g{B, H, W, 4, Ho, Wo, Cout, 7, 7, 2, 3};
  // host staging (static): packed filters, then bias | scale | shift
  umma_pack_weights(w, g, dt, 1, g_static_host_weights);
  for (int c = 0; c < Cout; ++c) g_static_host_params[c] = bias ? bias[c] : 0.f;
  for (int c = 0; c < 4; ++c) {
    g_static_host_params[kStaticMaxCout + c] = (c < 3) ? (in_scale ? in_scale[c] : 1.f) : 0.f;
    g_static_host_params[kStaticMaxCout + 4 + c] = (c < 3) ? (in_shift ? in_shift[c] : 0.f) : 0.f;
  }
  // the previous call's frame contents are still there: the borders must read as zero
  if ((ce = cudaMemsetAsync(d_frame, 0, frame_bytes, 0)) != cudaSuccess)
    return fail(3, "b200rn_stem_conv_static: cudaMemsetAsync failed (%lld)", ce);
  if ((ce = cudaMemcpyAsync(d_weights, g_static_host_weights, umma_weight_bytes(g, 1), cudaMemcpyHostToDevice, 0)) != cudaSuccess)
    return fail(3, "b200rn_stem_conv_static: filter upload failed (%lld)", ce);
  if ((ce = cudaMemcpyAsync(d_params, g_static_host_params, (kStaticMaxCout + 8) * sizeof(float), cudaMemcpyHostToDevice, 0)) != cudaSuccess)
    return fail(3, "b200rn_stem_conv_static: parameter upload failed (%lld)", ce);
  if ((ce = launch_input_xform(d_in_nchw, d_frame, dt, B, 3, H, W, 4, Hp, Wp, 3, 3, d_scale, d_shift, 0, 0)) != cudaSuccess)
    return fail(3, "b200rn_stem_conv_static: input re-layout launch failed (%lld)", ce);
  UmmaConvPlan pl;
  if (!fill_stem_plan(pl, dt, g, d_frame, d_weights, d_bias, d_out, relu))
    return fail(3, "b200rn_stem_conv_static: tensor-map encoding failed");
  const cudaError_t le = launch_conv_umma(&pl, 0);
  const cudaError_t se = cudaDeviceSynchronize();
  unsigned int flag = 0;
  const cudaError_t fe = (se == cudaSuccess) ? umma_take_timeout_flag(&flag) : se;
  if (le != cudaSuccess || se != cudaSuccess || fe != cudaSuccess || flag)
    return fail(3, "b200rn_stem_conv_static: launch/sync error %lld, barrier timeout %lld", (long long)(le ? le : (se ? se : fe)), flag);
  return 0;
}
