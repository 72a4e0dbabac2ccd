// Graph -> fused layer plan for the B200 ResNet backend.
//
// This is the host-side "graph optimiser" the reference gets from ONNX Runtime
// (src/inference_engine.cpp:20 `SetGraphOptimizationLevel(ORT_ENABLE_EXTENDED)`), rebuilt for one
// job: turn the parsed ONNX node list into the smallest sequence of device launches.
// Fusions are decided from the graph (consumer counts), never from layer names:
//   Conv -> BN            (BN the only consumer)      : folded into weights + bias
//   ... -> ReLU           (only consumer)             : epilogue ReLU
//   Conv1x1 + Conv1x1 -> Add (projection block)        : (dual_input) one conv over both inputs' channels
//   ... -> Add(x, other)  (only consumer)             : epilogue residual add; when both operands come from such
//                                                       convs (projection block) the one with fewer MACs takes it
//   T -> BN -> ReLU -> 1x1 convs only              : applied to those convs' A operand while it is
//                                                       staged in shared memory; no activated copy exists
//   T -> BN -> ReLU       (T made by Conv/MaxPool)    : otherwise a second epilogue output relu(s*T+t);
//                                                       the raw T is not written if nothing else reads it
//   input -> BN           (BN the only consumer)      : applied while re-laying NCHW fp32 -> NHWC
//   wide 1x1 Conv(+Add) -> BN -> ReLU -> 1x1 Conv     : (tail_fusion) both convs in one launch, see PlanLayer::tail_*
//   BN -> ReLU -> GlobalAveragePool                    : affine+ReLU applied inside the pooling kernel
// Anything else falls back to a stand-alone element-wise layer; an operator outside the
// supported set is an error (no CPU fallback).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "onnx_reader.h"

namespace rn {

enum class LayerKind : int { InputXform = 0, Conv, MaxPool, Affine, Add, Gap, Gemm, Softmax };
const char* layer_kind_name(LayerKind k);

struct PlanTensor {
  int id = -1;
  std::string name;        // ONNX value name it holds
  int C = 0, H = 0, W = 0; // logical per-sample dims
  int Cpad = 0;            // channels as stored in NHWC (>= C, zero-filled)
  bool vec_f32 = false;    // [B, C] fp32 row (pool / Gemm / Softmax outputs) instead of NHWC activations
  int def_layer = -1, last_use = -1;
  bool is_output = false;
  int64_t elems_per_sample() const { return vec_f32 ? int64_t(C) : int64_t(Cpad) * H * W; }
};

struct PlanLayer {
  LayerKind kind = LayerKind::Conv;
  std::string name;        // ONNX node names fused into this launch, '+'-joined
  int in = -1;             // main input tensor
  int in2 = -1;            // residual operand (Conv) / second operand (Add)
  int in_b = -1;           // Conv: a second 1x1 conv summed into the same accumulator (projection shortcut), see below
  int out = -1;            // primary output (-1 when only the activated copy is kept)
  int out_act = -1;        // relu?(act_scale * primary + act_shift)
  // Conv / Gemm
  int Cin = 0, Cout = 0, kh = 1, kw = 1, stride = 1, pad = 0;
  std::vector<float> weight;   // Conv: [Cout][kh][kw][Cin_pad] (BN folded); Gemm: [N][K] (alpha folded)
  std::vector<float> bias;     // [Cout] (always present; zeros if the graph has none)
  bool relu = false;           // ReLU on (acc + bias), before the residual add
  // second output / stand-alone affine / pooling prologue
  std::vector<float> act_scale, act_shift;
  bool act_relu = false;
  // Conv prologue: the input is read as relu?(pre_scale[c] * x + pre_shift[c]) (a BN -> ReLU that sits
  // between a residual sum and the 1x1 convs that read it, applied while the operand is staged, so the
  // activated copy of the tensor is never written to memory)
  std::vector<float> pre_scale, pre_shift;
  bool pre_relu = false;
  // Two-input conv (dual_input fusion): out = conv1x1(in, W[:, :Cin]) + conv1x1_stride_b(in_b, W[:, Cin:]) + bias.
  // A projection block's "main-branch 1x1 conv + shortcut 1x1 conv -> Add" is one GEMM whose K runs over both
  // inputs' channels, so the shortcut tensor is never written or read.  weight is [Cout][Cin + Cin_b], kh = kw = 1.
  int Cin_b = 0, stride_b = 1;
  // Stem conv with the 3x3 / stride-2 / pad-1 max-pool behind it in the same launch (stem_pool fusion): out / out_act
  // are then the POOLED tensors and act_* apply to the pooled value; the conv's own output is never materialised.
  bool fused_pool = false;
  // Conv tail: the 1x1 / stride-1 / pad-0 conv that was the only reader of this layer's activated output, run
  // inside the same launch on the activated tile while it is still in shared memory.  out_act is then -1
  // (act_scale / act_shift / act_relu still describe the activation between the two convs) and
  //   tail_out = tail_relu ? max(conv1x1(act, tail_weight) + tail_bias, 0) : ...
  int tail_out = -1, tail_Cout = 0;
  std::vector<float> tail_weight;  // [tail_Cout][Cout]
  std::vector<float> tail_bias;
  bool tail_relu = false;
  // MaxPool
  int pool_k = 0, pool_stride = 0, pool_pad = 0;
  double macs_per_sample = 0;  // algorithmic multiply-accumulates (Conv/Gemm), for rooflines
};

struct Plan {
  std::vector<PlanTensor> tensors;
  std::vector<PlanLayer> layers;
  std::string input_name, output_name;
  std::vector<int64_t> input_shape, output_shape;   // as the reference reports them (-1 -> 1)
  int64_t in_elems_per_sample = 0;                  // product of input dims after dim 0
  int64_t out_elems_per_sample = 0;                 // product of output dims after dim 0
  int in_C = 0, in_H = 0, in_W = 0;
  int output_tensor = -1;
  double total_macs_per_sample = 0;
  std::string describe() const;                     // one line per layer (tests / DESIGN.md tables)
};

// channel_align: activations are stored with channels rounded up to this (4 for the fp32 path,
// 8 for 16-bit paths; the 3-channel network input is always stored as 4).
// consumer_prologue: fuse "T -> BN -> ReLU -> {1x1 pad-0 convs}" into those convs' operand load instead
// of a second output of T's producer (saves writing and re-reading the activated copy).
// tail_fusion: fuse "Conv(1x1, Cin 64 -> 256 channels, + residual) -> BN -> ReLU -> Conv 1x1 (<= 128 channels)"
// pairs -- the end of one bottleneck block and the start of the next -- into one launch when that second conv
// is the activated tensor's only reader (tensor-core modes only: the shapes are what conv_tail_kernel takes).
// dual_input: fuse "Conv1x1(a) + Conv1x1(b) -> Add" (both linear, both only read by the Add, 64-multiple input
// channels, 256-multiple output channels) into one two-input conv (tensor-core modes only).
// stem_pool: fuse "7x7 / stride-2 / pad-3 conv -> 3x3 / stride-2 / pad-1 MaxPool" (the pool the conv's only reader)
// into the conv's launch (tensor-core modes with the dedicated stem kernel only).
Plan build_plan(const OnnxModel& model, int channel_align, bool consumer_prologue = false, bool tail_fusion = false,
                bool dual_input = false, bool stem_pool = false);

}  // namespace rn
