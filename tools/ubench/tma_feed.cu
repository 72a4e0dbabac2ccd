// Microbenchmark: how fast can one SM's TMA pull L2-resident GEMM operands into shared memory, as a function of
// ring depth, and does cluster multicast of the shared operand (B) raise it?   (sm_100a, B200)
//   A: [M rows][K=256] fp16, one 128-row x 64-col box per K block (16 KB), distinct per tile
//   B: [N=1024][K=256] fp16, one 256-row x 64-col box per K block (32 KB), same for every tile of an N tile
// One producer thread per CTA keeps a ring of `stages` (A+B) slots full; one consumer thread frees each slot the
// moment it lands (an infinitely fast MMA).  With clusters of C CTAs (same N tile, different M tiles) each CTA
// issues 1/C of every B box with .multicast::cluster, and a slot is reused only after all C consumers released it.
// Single-GPU tool: A and B are cudaMalloc'd on the current device.  TMA multicast loads here read only this GPU's
// memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/ubench/tma_feed.cu -o tools/ubench/tma_feed
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(n)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar, uint32_t cta) {
  asm volatile("{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\tmbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}"
               ::"r"(bar), "r"(cta) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"((uint64_t)m), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_2d_mc(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, uint16_t mask) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
               ::"r"(dst), "l"((uint64_t)m), "r"(bar), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

struct Params { int stages, k_iters, m_tiles, n_tiles, passes, cluster, load_a, load_b; unsigned long long* cycles; };

__global__ void __launch_bounds__(64, 1)
feed_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr uint32_t A_BYTES = 128 * 128, B_BYTES = 256 * 128, STAGE = A_BYTES + B_BYTES;
  __shared__ uint64_t full_bar[8], empty_bar[8];
  const int C = p.cluster;
  const uint32_t rank = C > 1 ? cluster_rank() : 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < p.stages; ++i) { mbar_init(smem_u32(&full_bar[i]), 1); mbar_init(smem_u32(&empty_bar[i]), C); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (C > 1) cluster_sync();
  // cluster = C consecutive CTAs: same N tile, consecutive M tiles
  const int cl = blockIdx.x / C, n_cl = gridDim.x / C;
  const int n_tile = cl % p.n_tiles;
  const int m_first = (cl / p.n_tiles) * C + rank, m_step = (n_cl / p.n_tiles) * C;
  const int my_tiles = (p.m_tiles - (cl / p.n_tiles) * C + m_step - 1) / m_step;    // same count for every CTA of a cluster
  const uint32_t stage_bytes = (p.load_a ? A_BYTES : 0) + (p.load_b ? B_BYTES : 0);
  const long long t0 = clock64();
  if (threadIdx.x == 0) {
    int stage = 0; uint32_t phase = 0;
    for (int pass = 0; pass < p.passes; ++pass)
      for (int i = 0; i < my_tiles; ++i) {
        int m = m_first + i * m_step;
        if (m >= p.m_tiles) m = p.m_tiles - 1;               // keep the cluster in lock step on the ragged end
        for (int it = 0; it < p.k_iters; ++it) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_expect_tx(fb, stage_bytes);
          const uint32_t dst = smem_u32(smem + stage * STAGE);
          if (p.load_a) tma_load_2d(dst, &tmA, fb, it * 64, m * 128);
          if (p.load_b) {
            if (C == 1) tma_load_2d(dst + A_BYTES, &tmB, fb, it * 64, n_tile * 256);
            else {
              const uint32_t rows = 256 / C;     // this CTA's slice of the B box, delivered to every CTA of the cluster
              tma_load_2d_mc(dst + A_BYTES + rank * rows * 128, &tmB, fb, it * 64, n_tile * 256 + rank * rows, (uint16_t)((1u << C) - 1));
            }
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
  } else if (threadIdx.x == 32) {
    int stage = 0; uint32_t phase = 0;
    for (int pass = 0; pass < p.passes; ++pass)
      for (int i = 0; i < my_tiles; ++i)
        for (int it = 0; it < p.k_iters; ++it) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          if (C == 1) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty_bar[stage])) : "memory");
          else for (int r = 0; r < C; ++r) mbar_arrive_remote(smem_u32(&empty_bar[stage]), r);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
  }
  __syncthreads();
  if (C > 1) cluster_sync();
  if (threadIdx.x == 0) p.cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(EncodeFn fn, void* base, uint64_t cols, uint64_t rows, uint32_t box_rows) {
  CUtensorMap m;
  cuuint64_t d[2] = {cols, rows}, s[1] = {cols * 2};
  cuuint32_t b[2] = {64, box_rows}, e[2] = {1, 1};
  CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

int main() {
  void* fnp = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &q));
  EncodeFn fn = (EncodeFn)fnp;
  const int K = 256, M = 50176, N = 1024;
  __half *A, *B;
  CK(cudaMalloc(&A, size_t(M) * K * 2));
  CK(cudaMalloc(&B, size_t(N) * K * 2));
  CK(cudaMemset(A, 0, size_t(M) * K * 2));
  CK(cudaMemset(B, 0, size_t(N) * K * 2));
  unsigned long long* cyc;
  CK(cudaMalloc(&cyc, 256 * 8));
  CK(cudaFuncSetAttribute(feed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  CK(cudaFuncSetAttribute(feed_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  printf("%-8s %-7s %-7s %-5s %10s %12s %12s\n", "mode", "cluster", "stages", "grid", "ms", "B/clk/SM", "TB/s chip");
  for (int mode = 0; mode < 3; ++mode)
    for (int C : {1, 2, 4})
      for (int stages : {2, 3, 4}) {
        if (mode == 2 && C > 1) continue;       // A only: nothing to multicast
        Params p;
        p.stages = stages; p.k_iters = K / 64; p.m_tiles = M / 128; p.n_tiles = N / 256; p.passes = 4; p.cluster = C;
        p.load_a = mode != 1; p.load_b = mode != 2; p.cycles = cyc;
        const int grid = (148 / (C * p.n_tiles)) * (C * p.n_tiles);
        CUtensorMap tmA = make_map(fn, A, K, M, 128), tmB = make_map(fn, B, K, N, 256 / C);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(64); cfg.dynamicSmemBytes = 1024 + stages * 48 * 1024;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        CK(cudaLaunchKernelEx(&cfg, feed_kernel, tmA, tmB, p));      // warm: operands into L2
        CK(cudaDeviceSynchronize());
        cudaEventRecord(e0);
        CK(cudaLaunchKernelEx(&cfg, feed_kernel, tmA, tmB, p));
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        std::vector<unsigned long long> h(grid);
        CK(cudaMemcpy(h.data(), cyc, grid * 8, cudaMemcpyDeviceToHost));
        unsigned long long mx = 0;
        for (auto v : h) mx = v > mx ? v : mx;
        // bytes landed in one SM's shared memory
        const int n_cl = grid / C, m_step = (n_cl / p.n_tiles) * C;
        const int my_tiles = (p.m_tiles + m_step - 1) / m_step;
        const double per_sm = double(p.passes) * my_tiles * p.k_iters * ((p.load_a ? 16384.0 : 0) + (p.load_b ? 32768.0 : 0));
        printf("%-8s %-7d %-7d %-5d %10.4f %12.1f %12.2f\n", mode == 0 ? "A+B" : (mode == 1 ? "B" : "A"), C, stages, grid, ms,
               per_sm / double(mx), per_sm * grid / (ms * 1e-3) / 1e12);
      }
  return 0;
}
