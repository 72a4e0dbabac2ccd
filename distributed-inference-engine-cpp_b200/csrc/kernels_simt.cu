// CUDA-core kernels of the B200 ResNet backend: layout transform, fp32 implicit-GEMM conv (exact
// mode + FC), max-pool, global average pool, element-wise fallbacks, softmax.
// All memory-bound kernels use 16-byte vector accesses with consecutive threads on consecutive
// channel vectors (NHWC => fully coalesced) and grids of at least a few waves over 148 SMs.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <algorithm>

#include "kernels.h"

namespace rn {
namespace {

constexpr int kSMs = 148;

// ---- 16-byte vectors of T <-> float -------------------------------------------------------
template <typename T> struct Vec;
template <> struct Vec<float> {
  static constexpr int N = 4;
  __device__ static void load(const float* p, float (&v)[4]) {
    float4 x = *reinterpret_cast<const float4*>(p);
    v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
  }
  __device__ static void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
  __device__ static float lowest() { return -3.402823466e+38f; }
};
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 x = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&x);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
  }
  __device__ static void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 x;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&x);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = x;
  }
};
template <> struct Vec<__half> {
  static constexpr int N = 8;
  __device__ static void load(const __half* p, float (&v)[8]) {
    uint4 x = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&x);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __half22float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
  }
  __device__ static void store(__half* p, const float (&v)[8]) {
    uint4 x;
    __half2* h = reinterpret_cast<__half2*>(&x);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = x;
  }
};

template <typename T> __device__ inline T from_float(float f);
template <> __device__ inline float from_float<float>(float f) { return f; }
template <> __device__ inline __nv_bfloat16 from_float<__nv_bfloat16>(float f) { return __float2bfloat16_rn(f); }
template <> __device__ inline __half from_float<__half>(float f) { return __float2half_rn(f); }

// one pixel = 4 channels: 8 bytes (16-bit types) or 16 bytes (fp32), naturally aligned
template <typename T> __device__ inline void store_pixel4(T* dst, const float (&v)[4]);
template <> __device__ inline void store_pixel4<float>(float* dst, const float (&v)[4]) {
  *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
}
template <> __device__ inline void store_pixel4<__nv_bfloat16>(__nv_bfloat16* dst, const float (&v)[4]) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(dst) = u;
}
template <> __device__ inline void store_pixel4<__half>(__half* dst, const float (&v)[4]) {
  __half2 a = __floats2half2_rn(v[0], v[1]), b = __floats2half2_rn(v[2], v[3]);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(dst) = u;
}

inline int grid_for(int64_t threads, int block) {
  int64_t g = (threads + block - 1) / block;
  if (g < 1) g = 1;
  return int(g);
}

// ---- input re-layout ------------------------------------------------------------------------
// One thread per 4 consecutive pixels of a row: three float4 plane reads (coalesced along W), one
// 32-byte (16-bit types) or 64-byte (fp32) store of 4 pixels x 4 channels.  Requires W % 4 == 0,
// Cpad == 4 and C <= 4; anything else takes the scalar kernel below.
template <typename T>
__global__ void input_xform_vec4_kernel(const float* __restrict__ in, T* __restrict__ out, int B, int C, int H, int W,
                                        int Hp, int Wp, int off_h, int off_w, const float* __restrict__ scale,
                                        const float* __restrict__ shift, int relu) {
  const int W4 = W >> 2;
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  int64_t total = int64_t(B) * H * W4;
  if (idx >= total) return;
  int w = int(idx % W4) * 4;
  int h = int((idx / W4) % H);
  int b = int(idx / (int64_t(W4) * H));
  float v[4][4];   // [pixel][channel]
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    float sc = 0.f, sh = 0.f;
    if (c < C) {
      x = __ldg(reinterpret_cast<const float4*>(in + ((int64_t(b) * C + c) * H + h) * W + w));
      sc = __ldg(scale + c); sh = __ldg(shift + c);
    }
    v[0][c] = fmaf(sc, x.x, sh); v[1][c] = fmaf(sc, x.y, sh); v[2][c] = fmaf(sc, x.z, sh); v[3][c] = fmaf(sc, x.w, sh);
    if (relu) {
#pragma unroll
      for (int px = 0; px < 4; ++px) v[px][c] = fmaxf(v[px][c], 0.f);
    }
    if (c >= C) { v[0][c] = v[1][c] = v[2][c] = v[3][c] = 0.f; }
  }
  T* dst = out + ((int64_t(b) * Hp + h + off_h) * Wp + (w + off_w)) * 4;
#pragma unroll
  for (int px = 0; px < 4; ++px) store_pixel4<T>(dst + px * 4, v[px]);
}

// Scalar fallback: one thread per pixel, any C / Cpad / W.
template <typename T>
__global__ void input_xform_kernel(const float* __restrict__ in, T* __restrict__ out, int B, int C, int H, int W,
                                   int Cpad, int Hp, int Wp, int off_h, int off_w, const float* __restrict__ scale,
                                   const float* __restrict__ shift, int relu) {
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  int64_t total = int64_t(B) * H * W;
  if (idx >= total) return;
  int w = int(idx % W);
  int h = int((idx / W) % H);
  int b = int(idx / (int64_t(W) * H));
  const float* src = in + (int64_t(b) * C * H + h) * W + w;
  T* dst = out + ((int64_t(b) * Hp + h + off_h) * Wp + (w + off_w)) * Cpad;
  for (int c = 0; c < Cpad; ++c) {
    float v = 0.f;
    if (c < C) {
      v = fmaf(__ldg(scale + c), __ldg(src + int64_t(c) * H * W), __ldg(shift + c));
      if (relu) v = fmaxf(v, 0.f);
    }
    dst[c] = from_float<T>(v);
  }
}

// ---- fp32 implicit-GEMM conv ----------------------------------------------------------------
// C[M = B*Ho*Wo][N = Cout] = A[M][K] * Wt[K][N], K ordered (ky, kx, cin).  BM x 64 CTA tile (BM = 128
// for convolutions, 32 for the skinny FC so that a 256-row problem still covers the chip), BK = 16,
// 256 threads, (BM/16) x 4 register tile, double-buffered shared memory, global->register prefetch of
// the next K slice during the FMAs.  A is gathered straight from the NHWC input in float4 granules
// (4 consecutive channels of one tap), zero for padding taps.
constexpr int BN = 64, BK = 16, SIMT_THREADS = 256;

struct SimtConvParams {
  const float* in; const float* w;
  Epilogue e;
  ConvGeom g;
  int M, K;
  // split-K (skinny problems only): blockIdx.z walks K slices of k_per_slice; partial sums go to
  // `partial` [slices][M][Cout]; the last CTA to finish a tile (ticket from `counters`) adds the
  // slices in index order -- a fixed order, so the result is bit-reproducible -- and runs the epilogue.
  int slices, k_per_slice;
  float* partial;
  unsigned int* counters;
};

template <int BM>
__global__ void __launch_bounds__(SIMT_THREADS) conv_simt_f32_kernel(SimtConvParams p) {
  constexpr int TM = BM / 16;                       // rows per thread
  constexpr int A_VECS = BM * BK / 4;               // float4 granules of one A tile
  constexpr int A_PER_THREAD = (A_VECS + SIMT_THREADS - 1) / SIMT_THREADS;
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Bs[2][BK][BN];
  const ConvGeom& g = p.g;
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;

  // A-load role: granule v = tid + j*256 -> row v/4, k offset (v%4)*4
  int a_row[A_PER_THREAD], a_k[A_PER_THREAD], ih0[A_PER_THREAD], iw0[A_PER_THREAD];
  const float* a_img[A_PER_THREAD];
  bool a_ok[A_PER_THREAD];
#pragma unroll
  for (int j = 0; j < A_PER_THREAD; ++j) {
    const int v = tid + j * SIMT_THREADS;
    a_row[j] = v >> 2;
    a_k[j] = (v & 3) * 4;
    const int am = m0 + a_row[j];
    a_ok[j] = (v < A_VECS) && (am < p.M);
    ih0[j] = iw0[j] = 0;
    a_img[j] = p.in;
    if (a_ok[j]) {
      int wo = am % g.Wo, t = am / g.Wo;
      int ho = t % g.Ho, b = t / g.Ho;
      ih0[j] = ho * g.stride - g.pad;
      iw0[j] = wo * g.stride - g.pad;
      a_img[j] = p.in + int64_t(b) * g.H * g.W * g.Cin;
    }
  }
  // B-load role: k row bk (0..15), 4 consecutive n starting at bn
  const int bk = tid >> 4, bn = (tid & 15) * 4;
  const bool b_col_ok = (n0 + bn) < g.Cout;

  float4 ra[A_PER_THREAD], rb;
  auto load_global = [&](int k0) {
#pragma unroll
    for (int j = 0; j < A_PER_THREAD; ++j) {
      int k = k0 + a_k[j];
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (a_ok[j] && k < p.K) {
        int tap = k / g.Cin, c = k - tap * g.Cin;
        int ky = tap / g.kw, kx = tap - ky * g.kw;
        int ih = ih0[j] + ky, iw = iw0[j] + kx;
        if (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W) {
          v = __ldg(reinterpret_cast<const float4*>(a_img[j] + (int64_t(ih) * g.W + iw) * g.Cin + c));
          if (p.e.pre_scale) {   // BN -> ReLU of the producer side, applied on the way in
            const float4 ps = __ldg(reinterpret_cast<const float4*>(p.e.pre_scale + c));
            const float4 pt = __ldg(reinterpret_cast<const float4*>(p.e.pre_shift + c));
            v.x = fmaf(ps.x, v.x, pt.x); v.y = fmaf(ps.y, v.y, pt.y); v.z = fmaf(ps.z, v.z, pt.z); v.w = fmaf(ps.w, v.w, pt.w);
            if (p.e.pre_relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
          }
        }
      }
      ra[j] = v;
    }
    int k = k0 + bk;
    rb = (b_col_ok && k < p.K) ? __ldg(reinterpret_cast<const float4*>(p.w + int64_t(k) * g.Cout + n0 + bn))
                               : make_float4(0.f, 0.f, 0.f, 0.f);
  };
  auto store_smem = [&](int buf) {
#pragma unroll
    for (int j = 0; j < A_PER_THREAD; ++j) {
      if (tid + j * SIMT_THREADS < A_VECS) {
        As[buf][a_k[j] + 0][a_row[j]] = ra[j].x;
        As[buf][a_k[j] + 1][a_row[j]] = ra[j].y;
        As[buf][a_k[j] + 2][a_row[j]] = ra[j].z;
        As[buf][a_k[j] + 3][a_row[j]] = ra[j].w;
      }
    }
    *reinterpret_cast<float4*>(&Bs[buf][bk][bn]) = rb;
  };

  const int ty = tid >> 4, tx = tid & 15;  // 16 x 16 threads, TM rows x 4 cols each
  float acc[TM][4];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int k_begin = (p.slices > 1) ? blockIdx.z * p.k_per_slice : 0;
  const int k_end = (p.slices > 1) ? min(p.K, k_begin + p.k_per_slice) : p.K;
  const int KT = (k_end - k_begin + BK - 1) / BK;
  load_global(k_begin);
  store_smem(0);
  __syncthreads();
  for (int kt = 0; kt < KT; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < KT) load_global(k_begin + (kt + 1) * BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM];
      if constexpr (TM % 4 == 0) {
#pragma unroll
        for (int i = 0; i < TM; i += 4) {
          float4 t4 = *reinterpret_cast<const float4*>(&As[buf][k][ty * TM + i]);
          a[i] = t4.x; a[i + 1] = t4.y; a[i + 2] = t4.z; a[i + 3] = t4.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < TM; i += 2) {
          float2 t2 = *reinterpret_cast<const float2*>(&As[buf][k][ty * TM + i]);
          a[i] = t2.x; a[i + 1] = t2.y;
        }
      }
      float4 b = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
      const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    if (kt + 1 < KT) store_smem(buf ^ 1);
    __syncthreads();
  }

  const int n = n0 + tx * 4;
  if (p.slices > 1) {
    // publish this slice, then let the last arrival of the tile reduce all slices in order
    __shared__ unsigned int s_ticket;
    if (n < g.Cout) {
#pragma unroll
      for (int i = 0; i < TM; ++i) {
        int m = m0 + ty * TM + i;
        if (m < p.M)
          *reinterpret_cast<float4*>(p.partial + (int64_t(blockIdx.z) * p.M + m) * g.Cout + n) =
              make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_ticket = atomicAdd(p.counters + blockIdx.y * gridDim.x + blockIdx.x, 1u);
    __syncthreads();
    if (s_ticket != unsigned(p.slices - 1)) return;
    __threadfence();
    if (tid == 0) p.counters[blockIdx.y * gridDim.x + blockIdx.x] = 0;   // ready for the next launch
    if (n < g.Cout) {
#pragma unroll
      for (int i = 0; i < TM; ++i) {
        int m = m0 + ty * TM + i;
        if (m >= p.M) break;
        float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int z = 0; z < p.slices; ++z) {
          float4 v = __ldcg(reinterpret_cast<const float4*>(p.partial + (int64_t(z) * p.M + m) * g.Cout + n));
          sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
        }
        acc[i][0] = sum.x; acc[i][1] = sum.y; acc[i][2] = sum.z; acc[i][3] = sum.w;
      }
    }
  }
  // fused epilogue
  if (n >= g.Cout) return;
  const Epilogue& e = p.e;
  const float4 bias = __ldg(reinterpret_cast<const float4*>(e.bias + n));
  float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e.out_act) {
    sc = __ldg(reinterpret_cast<const float4*>(e.act_scale + n));
    sh = __ldg(reinterpret_cast<const float4*>(e.act_shift + n));
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int m = m0 + ty * TM + i;
    if (m >= p.M) break;
    int64_t off = int64_t(m) * g.Cout + n;
    float4 v = make_float4(acc[i][0] + bias.x, acc[i][1] + bias.y, acc[i][2] + bias.z, acc[i][3] + bias.w);
    if (e.relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    if (e.residual) {
      float4 r = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(e.residual) + off));
      v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w;
    }
    if (e.out) *reinterpret_cast<float4*>(static_cast<float*>(e.out) + off) = v;
    if (e.out_act) {
      float4 a = make_float4(fmaf(sc.x, v.x, sh.x), fmaf(sc.y, v.y, sh.y), fmaf(sc.z, v.z, sh.z), fmaf(sc.w, v.w, sh.w));
      if (e.act_relu) { a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); a.z = fmaxf(a.z, 0.f); a.w = fmaxf(a.w, 0.f); }
      *reinterpret_cast<float4*>(static_cast<float*>(e.out_act) + off) = a;
    }
  }
}

// ---- max-pool -------------------------------------------------------------------------------
template <typename T>
__global__ void maxpool_kernel(const T* __restrict__ in, int B, int H, int W, int C, int Ho, int Wo, int k,
                               int stride, int pad, Epilogue e) {
  constexpr int V = Vec<T>::N;
  const int cv = C / V;
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  int64_t total = int64_t(B) * Ho * Wo * cv;
  if (idx >= total) return;
  int c = int(idx % cv) * V;
  int64_t pix = idx / cv;
  int wo = int(pix % Wo);
  int ho = int((pix / Wo) % Ho);
  int b = int(pix / (int64_t(Wo) * Ho));
  float best[V];
#pragma unroll
  for (int i = 0; i < V; ++i) best[i] = -3.402823466e+38f;
  for (int ky = 0; ky < k; ++ky) {
    int ih = ho * stride - pad + ky;
    if (ih < 0 || ih >= H) continue;
    for (int kx = 0; kx < k; ++kx) {
      int iw = wo * stride - pad + kx;
      if (iw < 0 || iw >= W) continue;
      float v[V];
      Vec<T>::load(in + ((int64_t(b) * H + ih) * W + iw) * C + c, v);
#pragma unroll
      for (int i = 0; i < V; ++i) best[i] = fmaxf(best[i], v[i]);
    }
  }
  int64_t off = pix * C + c;
  if (e.out) Vec<T>::store(static_cast<T*>(e.out) + off, best);
  if (e.out_act) {
    float a[V];
#pragma unroll
    for (int i = 0; i < V; ++i) {
      a[i] = fmaf(__ldg(e.act_scale + c + i), best[i], __ldg(e.act_shift + c + i));
      if (e.act_relu) a[i] = fmaxf(a[i], 0.f);
    }
    Vec<T>::store(static_cast<T*>(e.out_act) + off, a);
  }
}

// ---- global average pool (+ affine + ReLU on the way in) ------------------------------------
template <typename T>
__global__ void gap_kernel(const T* __restrict__ in, int B, int HW, int C, int Ctrue, const float* __restrict__ scale,
                           const float* __restrict__ shift, int relu, float* __restrict__ out, T* __restrict__ out16) {
  constexpr int V = Vec<T>::N;
  const int cv = C / V;
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * cv) return;
  int c = (idx % cv) * V, b = idx / cv;
  float s[V], t[V], acc[V];
#pragma unroll
  for (int i = 0; i < V; ++i) { s[i] = __ldg(scale + c + i); t[i] = __ldg(shift + c + i); acc[i] = 0.f; }
  const T* p = in + int64_t(b) * HW * C + c;
  // the pixels are summed in order (the result does not depend on the unrolling); eight loads are kept in flight --
  // a thread's 49 loads one after the other made this kernel latency-bound at 1.6 TB/s
#pragma unroll 8
  for (int i = 0; i < HW; ++i) {
    float v[V];
    Vec<T>::load(p + int64_t(i) * C, v);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float x = fmaf(s[j], v[j], t[j]);
      acc[j] += relu ? fmaxf(x, 0.f) : x;
    }
  }
  const float inv = 1.f / float(HW);
#pragma unroll
  for (int j = 0; j < V; ++j)
    if (c + j < Ctrue) out[int64_t(b) * Ctrue + c + j] = acc[j] * inv;
  if (out16) {
    float m[V];
#pragma unroll
    for (int j = 0; j < V; ++j) m[j] = acc[j] * inv;
    Vec<T>::store(out16 + int64_t(b) * C + c, m);
  }
}

// ---- element-wise fallbacks -----------------------------------------------------------------
template <typename T>
__global__ void affine_kernel(const T* __restrict__ in, T* __restrict__ out, int64_t pixels, int C,
                              const float* __restrict__ scale, const float* __restrict__ shift, int relu) {
  constexpr int V = Vec<T>::N;
  const int cv = C / V;
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= pixels * cv) return;
  int c = int(idx % cv) * V;
  int64_t off = (idx / cv) * C + c;
  float v[V];
  Vec<T>::load(in + off, v);
#pragma unroll
  for (int i = 0; i < V; ++i) {
    v[i] = fmaf(__ldg(scale + c + i), v[i], __ldg(shift + c + i));
    if (relu) v[i] = fmaxf(v[i], 0.f);
  }
  Vec<T>::store(out + off, v);
}

template <typename T>
__global__ void add_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, int64_t nvec) {
  constexpr int V = Vec<T>::N;
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= nvec) return;
  float x[V], y[V];
  Vec<T>::load(a + idx * V, x);
  Vec<T>::load(b + idx * V, y);
#pragma unroll
  for (int i = 0; i < V; ++i) x[i] += y[i];
  Vec<T>::store(out + idx * V, x);
}

// ---- softmax over rows (one CTA per row) ----------------------------------------------------
__global__ void softmax_rows_kernel(const float* __restrict__ in, float* __restrict__ out, int N) {
  __shared__ float red[32];
  const float* x = in + int64_t(blockIdx.x) * N;
  float* y = out + int64_t(blockIdx.x) * N;
  float mx = -3.402823466e+38f;
  for (int i = threadIdx.x; i < N; i += blockDim.x) mx = fmaxf(mx, x[i]);
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int i = 1; i < int(blockDim.x >> 5); ++i) mx = fmaxf(mx, red[i]);
  __syncthreads();
  float sum = 0.f;
  for (int i = threadIdx.x; i < N; i += blockDim.x) sum += expf(x[i] - mx);
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
  __syncthreads();
  sum = 0.f;
  for (int i = 0; i < int(blockDim.x >> 5); ++i) sum += red[i];
  const float inv = 1.f / sum;
  for (int i = threadIdx.x; i < N; i += blockDim.x) y[i] = expf(x[i] - mx) * inv;
}

}  // namespace

// ---- launchers --------------------------------------------------------------------------------

cudaError_t launch_input_xform(const float* nchw, void* out, DType dt, int B, int C, int H, int W, int Cpad,
                               int Hp, int Wp, int off_h, int off_w, const float* scale, const float* shift,
                               int relu, cudaStream_t s) {
  const int block = 256;
  const bool vec = (W % 4 == 0) && Cpad == 4 && C <= 4 && (reinterpret_cast<uintptr_t>(nchw) % 16 == 0);
  if (vec) {
    const int grid = grid_for(int64_t(B) * H * (W / 4), block);
    switch (dt) {
      case DType::F32:
        input_xform_vec4_kernel<float><<<grid, block, 0, s>>>(nchw, static_cast<float*>(out), B, C, H, W, Hp, Wp, off_h, off_w, scale, shift, relu);
        break;
      case DType::BF16:
        input_xform_vec4_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(nchw, static_cast<__nv_bfloat16*>(out), B, C, H, W, Hp, Wp, off_h, off_w, scale, shift, relu);
        break;
      case DType::F16:
        input_xform_vec4_kernel<__half><<<grid, block, 0, s>>>(nchw, static_cast<__half*>(out), B, C, H, W, Hp, Wp, off_h, off_w, scale, shift, relu);
        break;
    }
    return cudaGetLastError();
  }
  const int grid = grid_for(int64_t(B) * H * W, block);
  switch (dt) {
    case DType::F32:
      input_xform_kernel<float><<<grid, block, 0, s>>>(nchw, static_cast<float*>(out), B, C, H, W, Cpad, Hp, Wp, off_h, off_w, scale, shift, relu);
      break;
    case DType::BF16:
      input_xform_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(nchw, static_cast<__nv_bfloat16*>(out), B, C, H, W, Cpad, Hp, Wp, off_h, off_w, scale, shift, relu);
      break;
    case DType::F16:
      input_xform_kernel<__half><<<grid, block, 0, s>>>(nchw, static_cast<__half*>(out), B, C, H, W, Cpad, Hp, Wp, off_h, off_w, scale, shift, relu);
      break;
  }
  return cudaGetLastError();
}

cudaError_t launch_conv_simt_f32(const float* in, const float* w, const ConvGeom& g, const Epilogue& e, cudaStream_t s,
                                 float* splitk_scratch, size_t splitk_scratch_bytes) {
  if (g.Cin % 4 || g.Cout % 4) return cudaErrorInvalidValue;
  SimtConvParams p;
  p.in = in; p.w = w; p.e = e; p.g = g;
  p.M = g.B * g.Ho * g.Wo;
  p.K = g.kh * g.kw * g.Cin;
  p.slices = 1; p.k_per_slice = p.K; p.partial = nullptr; p.counters = nullptr;
  // a skinny problem (the FC: M = batch) takes 32-row tiles and, given scratch, split-K, so that it
  // still spreads over the SMs
  const long tiles128 = long((p.M + 127) / 128) * ((g.Cout + BN - 1) / BN);
  if (tiles128 < 2 * kSMs) {
    dim3 grid((p.M + 31) / 32, (g.Cout + BN - 1) / BN, 1);
    const long tiles = long(grid.x) * grid.y;
    if (splitk_scratch && tiles < 4 * kSMs) {
      int slices = int(std::min<long>(16, (4 * kSMs + tiles - 1) / tiles));
      int per = ((p.K + slices - 1) / slices + BK - 1) / BK * BK;
      slices = (p.K + per - 1) / per;
      const size_t counter_bytes = (size_t(tiles) * 4 + 255) / 256 * 256;
      const size_t need = counter_bytes + size_t(slices) * p.M * g.Cout * 4;
      if (slices > 1 && need <= splitk_scratch_bytes) {
        p.slices = slices; p.k_per_slice = per;
        p.counters = reinterpret_cast<unsigned int*>(splitk_scratch);   // zeroed at allocation, self-resetting
        p.partial = reinterpret_cast<float*>(reinterpret_cast<char*>(splitk_scratch) + counter_bytes);
        grid.z = unsigned(slices);
      }
    }
    conv_simt_f32_kernel<32><<<grid, SIMT_THREADS, 0, s>>>(p);
  } else {
    dim3 grid((p.M + 127) / 128, (g.Cout + BN - 1) / BN);
    conv_simt_f32_kernel<128><<<grid, SIMT_THREADS, 0, s>>>(p);
  }
  return cudaGetLastError();
}

size_t simt_splitk_scratch_bytes(int max_rows, int Cout) {
  // 16 slices of partial sums + one ticket counter per 32x64 tile
  const size_t tiles = size_t((max_rows + 31) / 32) * ((Cout + BN - 1) / BN);
  return (tiles * 4 + 255) / 256 * 256 + size_t(16) * max_rows * Cout * 4;
}

// 3x3 / stride 2 / pad 1 max-pool on 16-bit NHWC (the ResNet stem's pool), two horizontally adjacent outputs per
// thread: 15 window loads for two outputs instead of 18, and the max stays in the packed 16-bit type (max is exact,
// so no conversion to fp32 and back is needed; NaNs are ignored like fmaxf does).  Padding taps are skipped.
template <typename T> struct Pk;
template <> struct Pk<__half> {
  using V2 = __half2;
  __device__ static V2 lowest() { return __half2half2(__ushort_as_half(0xFBFFu)); }       // -65504
  __device__ static float2 to_float2(V2 v) { return __half22float2(v); }
};
template <> struct Pk<__nv_bfloat16> {
  using V2 = __nv_bfloat162;
  __device__ static V2 lowest() { return __bfloat162bfloat162(__ushort_as_bfloat16(0xFF7Fu)); }   // -3.39e38
  __device__ static float2 to_float2(V2 v) { return __bfloat1622float2(v); }
};

template <typename T>
__global__ void maxpool3x3s2_pair_kernel(const T* __restrict__ in, int B, int H, int W, int C, int Ho, int Wo, Epilogue e) {
  using V2 = typename Pk<T>::V2;
  const int cv = C / 8;
  const int Wp = (Wo + 1) / 2;                                  // output pairs per row
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t total = int64_t(B) * Ho * Wp * cv;
  if (idx >= total) return;
  const int c = int(idx % cv) * 8;
  int64_t t = idx / cv;
  const int wp = int(t % Wp); t /= Wp;
  const int ho = int(t % Ho);
  const int b = int(t / Ho);
  const int wo = 2 * wp;
  V2 best[2][4];
#pragma unroll
  for (int o = 0; o < 2; ++o)
#pragma unroll
    for (int i = 0; i < 4; ++i) best[o][i] = Pk<T>::lowest();
  const int iw0 = wo * 2 - 1;                                   // columns iw0 .. iw0+4 cover both windows
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    const int ih = ho * 2 - 1 + ky;
    if (ih < 0 || ih >= H) continue;
    const T* rowp = in + (int64_t(b) * H + ih) * W * C + c;
#pragma unroll
    for (int kx = 0; kx < 5; ++kx) {
      const int iw = iw0 + kx;
      if (iw < 0 || iw >= W) continue;
      const uint4 x = __ldg(reinterpret_cast<const uint4*>(rowp + int64_t(iw) * C));
      const V2* h = reinterpret_cast<const V2*>(&x);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (kx <= 2) best[0][i] = __hmax2(best[0][i], h[i]);
        if (kx >= 2) best[1][i] = __hmax2(best[1][i], h[i]);
      }
    }
  }
#pragma unroll
  for (int o = 0; o < 2; ++o) {
    if (wo + o >= Wo) break;
    const int64_t off = ((int64_t(b) * Ho + ho) * Wo + wo + o) * C + c;
    if (e.out) *reinterpret_cast<uint4*>(static_cast<T*>(e.out) + off) = *reinterpret_cast<const uint4*>(best[o]);
    if (e.out_act) {
      float a[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = Pk<T>::to_float2(best[o][i]);
        a[2 * i] = fmaf(__ldg(e.act_scale + c + 2 * i), f.x, __ldg(e.act_shift + c + 2 * i));
        a[2 * i + 1] = fmaf(__ldg(e.act_scale + c + 2 * i + 1), f.y, __ldg(e.act_shift + c + 2 * i + 1));
        if (e.act_relu) { a[2 * i] = fmaxf(a[2 * i], 0.f); a[2 * i + 1] = fmaxf(a[2 * i + 1], 0.f); }
      }
      Vec<T>::store(static_cast<T*>(e.out_act) + off, a);
    }
  }
}

cudaError_t launch_maxpool(const void* in, DType dt, int B, int H, int W, int C, int Ho, int Wo, int k, int stride,
                           int pad, const Epilogue& e, cudaStream_t s) {
  int V = dt == DType::F32 ? 4 : 8;
  if (C % V) return cudaErrorInvalidValue;
  if (dt != DType::F32 && k == 3 && stride == 2 && pad == 1 && H >= 2 && W >= 2) {
    const int64_t pairs = int64_t(B) * Ho * ((Wo + 1) / 2) * (C / 8);
    const int pblock = 256, pgrid = grid_for(pairs, pblock);
    if (dt == DType::BF16)
      maxpool3x3s2_pair_kernel<__nv_bfloat16><<<pgrid, pblock, 0, s>>>(static_cast<const __nv_bfloat16*>(in), B, H, W, C, Ho, Wo, e);
    else
      maxpool3x3s2_pair_kernel<__half><<<pgrid, pblock, 0, s>>>(static_cast<const __half*>(in), B, H, W, C, Ho, Wo, e);
    return cudaGetLastError();
  }
  int64_t total = int64_t(B) * Ho * Wo * (C / V);
  int block = 256, grid = grid_for(total, block);
  switch (dt) {
    case DType::F32: maxpool_kernel<float><<<grid, block, 0, s>>>(static_cast<const float*>(in), B, H, W, C, Ho, Wo, k, stride, pad, e); break;
    case DType::BF16: maxpool_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(static_cast<const __nv_bfloat16*>(in), B, H, W, C, Ho, Wo, k, stride, pad, e); break;
    case DType::F16: maxpool_kernel<__half><<<grid, block, 0, s>>>(static_cast<const __half*>(in), B, H, W, C, Ho, Wo, k, stride, pad, e); break;
  }
  return cudaGetLastError();
}

cudaError_t launch_gap(const void* in, DType dt, int B, int HW, int C, int Ctrue, const float* scale,
                       const float* shift, int relu, float* out, cudaStream_t s, void* out16) {
  int V = dt == DType::F32 ? 4 : 8;
  if (C % V) return cudaErrorInvalidValue;
  int total = B * (C / V);
  int block = 128, grid = grid_for(total, block);
  switch (dt) {
    case DType::F32: gap_kernel<float><<<grid, block, 0, s>>>(static_cast<const float*>(in), B, HW, C, Ctrue, scale, shift, relu, out, nullptr); break;
    case DType::BF16: gap_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(static_cast<const __nv_bfloat16*>(in), B, HW, C, Ctrue, scale, shift, relu, out, static_cast<__nv_bfloat16*>(out16)); break;
    case DType::F16: gap_kernel<__half><<<grid, block, 0, s>>>(static_cast<const __half*>(in), B, HW, C, Ctrue, scale, shift, relu, out, static_cast<__half*>(out16)); break;
  }
  return cudaGetLastError();
}

cudaError_t launch_affine(const void* in, void* out, DType dt, int64_t pixels, int C, const float* scale,
                          const float* shift, int relu, cudaStream_t s) {
  int V = dt == DType::F32 ? 4 : 8;
  if (C % V) return cudaErrorInvalidValue;
  int block = 256, grid = grid_for(pixels * (C / V), block);
  switch (dt) {
    case DType::F32: affine_kernel<float><<<grid, block, 0, s>>>(static_cast<const float*>(in), static_cast<float*>(out), pixels, C, scale, shift, relu); break;
    case DType::BF16: affine_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(static_cast<const __nv_bfloat16*>(in), static_cast<__nv_bfloat16*>(out), pixels, C, scale, shift, relu); break;
    case DType::F16: affine_kernel<__half><<<grid, block, 0, s>>>(static_cast<const __half*>(in), static_cast<__half*>(out), pixels, C, scale, shift, relu); break;
  }
  return cudaGetLastError();
}

cudaError_t launch_add(const void* a, const void* b, void* out, DType dt, int64_t elems, cudaStream_t s) {
  int V = dt == DType::F32 ? 4 : 8;
  if (elems % V) return cudaErrorInvalidValue;
  int block = 256, grid = grid_for(elems / V, block);
  switch (dt) {
    case DType::F32: add_kernel<float><<<grid, block, 0, s>>>(static_cast<const float*>(a), static_cast<const float*>(b), static_cast<float*>(out), elems / V); break;
    case DType::BF16: add_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b), static_cast<__nv_bfloat16*>(out), elems / V); break;
    case DType::F16: add_kernel<__half><<<grid, block, 0, s>>>(static_cast<const __half*>(a), static_cast<const __half*>(b), static_cast<__half*>(out), elems / V); break;
  }
  return cudaGetLastError();
}

cudaError_t launch_softmax_rows(const float* in, float* out, int B, int N, cudaStream_t s) {
  softmax_rows_kernel<<<B, 256, 0, s>>>(in, out, N);
  return cudaGetLastError();
}

}  // namespace rn
