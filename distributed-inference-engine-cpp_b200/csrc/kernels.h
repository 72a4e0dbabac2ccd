// Launch interface of the sm_100a kernels (host side).  Activations are NHWC; element type is
// float (exact mode) or bf16 / fp16 (tensor-core modes).  All launchers are asynchronous on the
// given stream and return the cudaError_t of the launch.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

namespace rn {

enum class DType : int { F32 = 0, BF16 = 1, F16 = 2 };
inline size_t dtype_size(DType t) { return t == DType::F32 ? 4 : 2; }

// Fused epilogue shared by both conv kernels and (partly) max-pool:
//   v = acc + bias[c];  if (relu) v = max(v, 0);  if (residual) v += residual[m][c];
//   if (out)     out[m][c]     = v;
//   if (out_act) out_act[m][c] = act_relu ? max(act_scale[c]*v + act_shift[c], 0) : act_scale[c]*v + act_shift[c];
// Optional prologue (1x1, pad-0 convs only): the input is read as
//   x' = pre_relu ? max(pre_scale[c]*x + pre_shift[c], 0) : pre_scale[c]*x + pre_shift[c]
struct Epilogue {
  const float* pre_scale = nullptr;
  const float* pre_shift = nullptr;
  int pre_relu = 0;
  const float* bias = nullptr;
  const void* residual = nullptr;
  void* out = nullptr;
  void* out_act = nullptr;
  const float* act_scale = nullptr;
  const float* act_shift = nullptr;
  int relu = 0;
  int act_relu = 0;
  // Fused tail (tensor-core path only): a following 1x1 / stride-1 conv that is the only reader of out_act.
  //   tail_out[m][n] = tail_relu ? max(sum_c act(v)[m][c] * tail_w[n][c] + tail_bias[n], 0) : ...
  // where act(v) is the out_act expression above, kept in shared memory; out_act itself must be null.
  const void* tail_w = nullptr;       // packed like any 1x1 conv's weights (umma_pack_weights)
  const float* tail_bias = nullptr;
  void* tail_out = nullptr;
  int tail_cout = 0;
  int tail_relu = 0;
  // Second input (tensor-core path, 1x1 / pad-0 convs on conv_wide_kernel / conv_tail_kernel only): the GEMM's K runs
  // over g.Cin channels of `in` and then in_b_cin channels of in_b, an NHWC tensor [B, in_b_h, in_b_w, in_b_cin] read
  // at stride in_b_stride (output pixel (y, x) <- in_b pixel (y*s, x*s)); the weights are [Cout][g.Cin + in_b_cin].
  const void* in_b = nullptr;
  int in_b_cin = 0, in_b_h = 0, in_b_w = 0, in_b_stride = 1;
  // Fused max-pool (stem kernel only, stem_mode 2): `out` / `out_act` are then the POOLED tensors
  // [B, pool_h, pool_w, Cout] of a 3x3 / stride-2 / pad-1 max-pool over relu?(conv + bias); act_* apply to the pooled
  // value.  The conv's own output is never written.
  int pool_fused = 0, pool_h = 0, pool_w = 0;
  // fp32 row output (tensor-core path, 1x1 spatial outputs only: the FC as a 1x1 conv over pooled rows):
  //   out_f32[image][c] = acc + bias[c] for c < out_f32_cols, row pitch out_f32_ld floats; no 16-bit output then.
  float* out_f32 = nullptr;
  int out_f32_ld = 0, out_f32_cols = 0;
};

struct ConvGeom {
  int B, H, W, Cin;        // input NHWC (Cin as stored)
  int Ho, Wo, Cout;        // output NHWC (Cout as stored)
  int kh, kw, stride, pad;
};

// NCHW fp32 [B,C,H,W] -> NHWC (C padded to Cpad with zeros), y = relu?(scale[c]*x + shift[c]).
// The destination may be a spatially padded frame [B,Hp,Wp,Cpad] with the image at (off_h, off_w);
// only the interior is written (borders are zeroed once at allocation).
cudaError_t launch_input_xform(const float* nchw, void* out, DType dt, int B, int C, int H, int W, int Cpad,
                               int Hp, int Wp, int off_h, int off_w, const float* scale, const float* shift,
                               int relu, cudaStream_t s);

// fp32 implicit-GEMM convolution on CUDA cores (exact mode, and the tiny FC in every mode).
// w: [K = kh*kw*Cin][Cout] fp32.
// splitk_scratch (optional, zero-initialised device memory of simt_splitk_scratch_bytes()) lets a skinny
// problem split K across CTAs with a deterministic in-kernel reduction.
cudaError_t launch_conv_simt_f32(const float* in, const float* w, const ConvGeom& g, const Epilogue& e, cudaStream_t s,
                                 float* splitk_scratch = nullptr, size_t splitk_scratch_bytes = 0);
size_t simt_splitk_scratch_bytes(int max_rows, int Cout);

// max-pool k x k / stride / pad (padding excluded from the max), NHWC, with optional second output.
cudaError_t launch_maxpool(const void* in, DType dt, int B, int H, int W, int C, int Ho, int Wo, int k, int stride,
                           int pad, const Epilogue& e, cudaStream_t s);

// out[b][c] = mean over H*W of relu?(scale[c]*x + shift[c]);  fp32 rows [B, Ctrue]; C = stored channels
// out16 (optional, 16-bit types only): the same means as [B, C] rows in the activation type (the FC's operand)
cudaError_t launch_gap(const void* in, DType dt, int B, int HW, int C, int Ctrue, const float* scale,
                       const float* shift, int relu, float* out, cudaStream_t s, void* out16 = nullptr);

// stand-alone element-wise fallbacks (only used when the graph does not match a fusion)
cudaError_t launch_affine(const void* in, void* out, DType dt, int64_t pixels, int C, const float* scale,
                          const float* shift, int relu, cudaStream_t s);
cudaError_t launch_add(const void* a, const void* b, void* out, DType dt, int64_t elems, cudaStream_t s);

cudaError_t launch_softmax_rows(const float* in, float* out, int B, int N, cudaStream_t s);

// ---- tensor-core path (conv_umma.cu) -------------------------------------------------------
struct UmmaConvPlan;  // opaque: TMA descriptors + tile geometry for one conv layer at one batch size
UmmaConvPlan* umma_conv_plan_create(DType dt, const ConvGeom& g, const void* in, const void* w_packed,
                                    const Epilogue& e, int stem_mode, char* err, size_t errcap);
void umma_conv_plan_destroy(UmmaConvPlan*);
// fp32-row plans only: the rows' destination may change between launches (it is the caller's output buffer)
void umma_conv_plan_set_out_f32(UmmaConvPlan*, float* out_f32);
cudaError_t launch_conv_umma(const UmmaConvPlan* p, cudaStream_t s);
// bytes of the packed weight buffer and the host-side packer ([Cout][kh][kw][Cin] fp32 -> kernel layout)
size_t umma_weight_bytes(const ConvGeom& g, int stem_mode);
void umma_pack_weights(const float* w, const ConvGeom& g, DType dt, int stem_mode, void* dst);
// Reads (and clears) the device flag a tensor-core conv sets when one of its barrier waits gave up.
// Call after synchronising; a set flag means the results of that launch are invalid.
cudaError_t umma_take_timeout_flag(unsigned int* flag);
bool umma_conv_supported(const ConvGeom& g, int stem_mode, char* why, size_t cap);


// ---- exact mode on the tensor cores: fp32 storage, 3 x TF32 split products (conv_tf32.cuh) ------------------
struct Tf32ConvPlan;
bool tf32_conv_supported(const ConvGeom& g);
// w: [Cout][kh][kw][Cin] fp32 -> hi (low 13 mantissa bits cleared) and lo (w - hi, cut the same way)
void tf32_split_weights(const float* w, size_t n, float* hi, float* lo);
// stem != 0: the 7x7/2 stem; `in` is then the zero-framed fp32 input [B, H + 6, Wp, 4] (image at (3,3)) and the weights are
// packed by tf32_pack_stem_weights ([Cout][7][8][4]) before the split
Tf32ConvPlan* tf32_conv_plan_create(const ConvGeom& g, const float* in, const float* w_hi, const float* w_lo, const Epilogue& e,
                                    char* err, size_t errcap, int stem = 0);
bool tf32_stem_supported(const ConvGeom& g);
size_t tf32_stem_weight_elems(int Cout);
void tf32_pack_stem_weights(const float* w, int Cout, float* dst);
void tf32_conv_plan_destroy(Tf32ConvPlan*);
cudaError_t launch_conv_tf32(const Tf32ConvPlan* p, cudaStream_t s);

}  // namespace rn
