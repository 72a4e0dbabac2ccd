// See plan.h.
#include "plan.h"

#include <algorithm>
#include <cmath>
#include <map>
#include <set>
#include <sstream>
#include <stdexcept>

namespace rn {

const char* layer_kind_name(LayerKind k) {
  switch (k) {
    case LayerKind::InputXform: return "InputXform";
    case LayerKind::Conv: return "Conv";
    case LayerKind::MaxPool: return "MaxPool";
    case LayerKind::Affine: return "Affine";
    case LayerKind::Add: return "Add";
    case LayerKind::Gap: return "GlobalAvgPool";
    case LayerKind::Gemm: return "Gemm";
    case LayerKind::Softmax: return "Softmax";
  }
  return "?";
}

namespace {

[[noreturn]] void fail(const std::string& m) { throw std::runtime_error("plan: " + m); }

int round_up(int v, int a) { return (v + a - 1) / a * a; }

struct PendingPrologue {
  int tensor = -1;
  std::vector<float> scale, shift;
  bool relu = false;
  std::string names;
};

struct Builder {
  const OnnxModel& m;
  int align;
  bool consumer_prologue = true;
  bool tail_fusion = false;
  bool dual_input = false;
  bool stem_pool = false;
  std::map<std::string, PendingPrologue> prologues;   // activated value name -> how to make it on the fly
  Plan plan;
  std::map<std::string, std::vector<int>> consumers;  // value name -> node indices
  std::set<std::string> graph_outputs;
  std::vector<char> absorbed;
  std::map<std::string, int> value;       // ONNX value name -> tensor id
  std::map<int, int> producer_layer;      // tensor id -> index into `order`
  std::vector<PlanLayer> layers;          // creation order; `order` is the execution order
  std::vector<int> order;

  Builder(const OnnxModel& model, int a, bool cp) : m(model), align(a), consumer_prologue(cp), absorbed(model.nodes.size(), 0) {
    for (size_t i = 0; i < m.nodes.size(); ++i)
      for (auto& in : m.nodes[i].inputs) consumers[in].push_back(int(i));
    for (auto& o : m.outputs) graph_outputs.insert(o.name);
  }

  int cpad(int C) const { return C <= 4 ? 4 : round_up(C, align); }

  int new_tensor(const std::string& name, int C, int H, int W, bool vec) {
    PlanTensor t;
    t.id = int(plan.tensors.size());
    t.name = name;
    t.C = C; t.H = H; t.W = W;
    t.vec_f32 = vec;
    t.Cpad = vec ? C : cpad(C);
    plan.tensors.push_back(t);
    return t.id;
  }

  const OnnxTensor& init(const std::string& name, const char* what) const {
    auto it = m.initializers.find(name);
    if (it == m.initializers.end()) fail(std::string(what) + " '" + name + "' is not an initializer (dynamic weights are not supported)");
    if (it->second.data_type != 1) fail(std::string(what) + " '" + name + "' is not float32");
    return it->second;
  }

  // the single consumer of `v` if it has exactly one, is not a graph output and that consumer has op `op`
  int sole_consumer(const std::string& v, const char* op) const {
    if (graph_outputs.count(v)) return -1;
    auto it = consumers.find(v);
    if (it == consumers.end() || it->second.size() != 1) return -1;
    int n = it->second[0];
    if (absorbed[size_t(n)] || m.nodes[size_t(n)].op != op) return -1;
    return n;
  }

  int tensor_of(const std::string& v, const OnnxNode& n) const {
    auto it = value.find(v);
    if (it == value.end()) fail("node '" + n.name + "' (" + n.op + ") reads '" + v + "' which no supported node produced");
    return it->second;
  }

  int emit(PlanLayer&& L) {
    layers.push_back(std::move(L));
    int idx = int(layers.size()) - 1;
    order.push_back(idx);
    const PlanLayer& l = layers.back();
    if (l.out >= 0) producer_layer[l.out] = idx;
    if (l.out_act >= 0) producer_layer[l.out_act] = idx;
    return idx;
  }

  // BatchNormalization (inference): y = scale*(x-mean)/sqrt(var+eps)+bias  ->  y = s*x + t
  void bn_affine(const OnnxNode& n, std::vector<float>& s, std::vector<float>& t) const {
    if (n.inputs.size() < 5) fail("BatchNormalization '" + n.name + "' needs 5 inputs");
    const auto& g = init(n.inputs[1], "BN scale");
    const auto& b = init(n.inputs[2], "BN bias");
    const auto& mu = init(n.inputs[3], "BN mean");
    const auto& var = init(n.inputs[4], "BN var");
    float eps = n.attr_f("epsilon", 1e-5f);
    size_t C = g.f32.size();
    if (b.f32.size() != C || mu.f32.size() != C || var.f32.size() != C) fail("BN '" + n.name + "': parameter sizes differ");
    s.resize(C); t.resize(C);
    for (size_t c = 0; c < C; ++c) {
      // computed in double, rounded once: keeps the fold within 1 ulp of the unfused fp32 form
      double inv = double(g.f32[c]) / std::sqrt(double(var.f32[c]) + double(eps));
      s[c] = float(inv);
      t[c] = float(double(b.f32[c]) - double(mu.f32[c]) * inv);
    }
  }

  static void pad_to(std::vector<float>& v, size_t n, float fill) { v.resize(n, fill); }

  // "T -> BN [-> ReLU] -> only 1x1, pad-0, group-1 convs": let those convs apply the affine+ReLU to
  // their operand as it is staged; nothing is materialised.  (With padding this would be wrong: the
  // conv pads the activated tensor with zeros, not with relu(shift).)
  bool defer_act_to_consumers(int node_idx) {
    if (!consumer_prologue) return false;
    const OnnxNode& bn = m.nodes[size_t(node_idx)];
    int tid = tensor_of(bn.inputs[0], bn);
    const PlanTensor src = plan.tensors[size_t(tid)];
    if (src.vec_f32) return false;
    // only worth it when the raw tensor is written anyway (someone else reads it); otherwise the
    // producer can simply write the activated copy instead of the raw one, at the same cost
    if (!graph_outputs.count(bn.inputs[0]) && consumers[bn.inputs[0]].size() <= 1) return false;
    std::string outname = bn.outputs[0];
    int r = sole_consumer(outname, "Relu");
    if (r >= 0) outname = m.nodes[size_t(r)].outputs[0];
    if (graph_outputs.count(outname)) return false;
    auto cit = consumers.find(outname);
    if (cit == consumers.end() || cit->second.empty()) return false;
    for (int c : cit->second) {
      const OnnxNode& n = m.nodes[size_t(c)];
      if (n.op != "Conv" || n.inputs[0] != outname) return false;
      for (size_t k = 1; k < n.inputs.size(); ++k)
        if (n.inputs[k] == outname) return false;
      auto ks = n.attr_ints("kernel_shape", {});
      auto pd = n.attr_ints("pads", {0, 0, 0, 0});
      if (ks.size() != 2 || ks[0] != 1 || ks[1] != 1) return false;
      for (auto v : pd) if (v != 0) return false;
      if (n.attr_i("group", 1) != 1) return false;
    }
    PendingPrologue pp;
    pp.tensor = tid;
    bn_affine(bn, pp.scale, pp.shift);
    if (int(pp.scale.size()) != src.C) fail("BN '" + bn.name + "': channel count mismatch");
    pad_to(pp.scale, size_t(src.Cpad), 0.f);
    pad_to(pp.shift, size_t(src.Cpad), 0.f);
    pp.relu = r >= 0;
    pp.names = bn.name + (r >= 0 ? "+" + m.nodes[size_t(r)].name : std::string());
    absorbed[size_t(node_idx)] = 1;
    if (r >= 0) absorbed[size_t(r)] = 1;
    prologues[outname] = std::move(pp);
    return true;
  }

  // Try to hang "-> BN [-> ReLU]" on tensor `tid` as the producing layer's second output.
  // Returns true (and registers the value names) when done.
  bool fuse_act_into_producer(int node_idx) {
    const OnnxNode& bn = m.nodes[size_t(node_idx)];
    int tid = tensor_of(bn.inputs[0], bn);
    auto pit = producer_layer.find(tid);
    if (pit == producer_layer.end()) return false;
    PlanLayer& P = layers[size_t(pit->second)];
    if (!(P.kind == LayerKind::Conv || P.kind == LayerKind::MaxPool)) return false;
    if (P.out != tid || P.out_act >= 0) return false;
    const PlanTensor src = plan.tensors[size_t(tid)];
    std::vector<float> s, t;
    bn_affine(bn, s, t);
    if (int(s.size()) != src.C) fail("BN '" + bn.name + "': channel count mismatch");
    std::string outname = bn.outputs[0];
    bool relu = false;
    int r = sole_consumer(outname, "Relu");
    if (r >= 0) { relu = true; absorbed[size_t(r)] = 1; outname = m.nodes[size_t(r)].outputs[0]; P.name += "+" + bn.name + "+" + m.nodes[size_t(r)].name; }
    else P.name += "+" + bn.name;
    pad_to(s, size_t(src.Cpad), 0.f);
    pad_to(t, size_t(src.Cpad), 0.f);
    P.act_scale = std::move(s); P.act_shift = std::move(t); P.act_relu = relu;
    P.out_act = new_tensor(outname, src.C, src.H, src.W, false);
    value[outname] = P.out_act;
    producer_layer[P.out_act] = pit->second;
    // drop the raw copy if this BN was its only reader
    const std::string& rawname = bn.inputs[0];
    if (!graph_outputs.count(rawname) && consumers[rawname].size() == 1) P.out = -1;
    absorbed[size_t(node_idx)] = 1;
    return true;
  }

  void do_input() {
    const auto& in = m.inputs[0];
    if (in.dims.size() != 4) fail("model input must be 4-D NCHW, got rank " + std::to_string(in.dims.size()));
    for (size_t d = 1; d < 4; ++d)
      if (in.dims[d] <= 0) fail("model input has a dynamic non-batch dimension");
    plan.input_name = in.name;
    plan.in_C = int(in.dims[1]); plan.in_H = int(in.dims[2]); plan.in_W = int(in.dims[3]);
    PlanLayer L;
    L.kind = LayerKind::InputXform;
    L.name = "input:" + in.name;
    std::string outname = in.name;
    std::vector<float> s(size_t(plan.in_C), 1.f), t(size_t(plan.in_C), 0.f);
    int b = sole_consumer(in.name, "BatchNormalization");
    if (b >= 0 && m.nodes[size_t(b)].inputs[0] == in.name) {
      bn_affine(m.nodes[size_t(b)], s, t);
      if (int(s.size()) != plan.in_C) fail("input BN channel mismatch");
      absorbed[size_t(b)] = 1;
      outname = m.nodes[size_t(b)].outputs[0];
      L.name += "+" + m.nodes[size_t(b)].name;
      int r = sole_consumer(outname, "Relu");
      if (r >= 0) { absorbed[size_t(r)] = 1; outname = m.nodes[size_t(r)].outputs[0]; L.act_relu = true; L.name += "+" + m.nodes[size_t(r)].name; }
    }
    L.out = new_tensor(outname, plan.in_C, plan.in_H, plan.in_W, false);
    pad_to(s, size_t(plan.tensors[size_t(L.out)].Cpad), 0.f);
    pad_to(t, size_t(plan.tensors[size_t(L.out)].Cpad), 0.f);
    L.act_scale = std::move(s); L.act_shift = std::move(t);
    value[outname] = L.out;
    emit(std::move(L));
  }

  void do_conv(int idx) {
    const OnnxNode& n = m.nodes[size_t(idx)];
    const PendingPrologue* pro = nullptr;
    int tin;
    auto pit = prologues.find(n.inputs[0]);
    if (pit != prologues.end()) { pro = &pit->second; tin = pro->tensor; }
    else tin = tensor_of(n.inputs[0], n);
    const PlanTensor in = plan.tensors[size_t(tin)];
    if (in.vec_f32) fail("Conv '" + n.name + "' on a flattened tensor");
    const auto& w = init(n.inputs[1], "Conv weight");
    if (w.dims.size() != 4) fail("Conv '" + n.name + "': weight must be 4-D");
    int Cout = int(w.dims[0]), Cin = int(w.dims[1]), kh = int(w.dims[2]), kw = int(w.dims[3]);
    if (n.attr_i("group", 1) != 1) fail("Conv '" + n.name + "': group != 1 is not supported");
    auto dil = n.attr_ints("dilations", {1, 1});
    if (dil[0] != 1 || dil[1] != 1) fail("Conv '" + n.name + "': dilation != 1 is not supported");
    auto st = n.attr_ints("strides", {1, 1});
    auto pd = n.attr_ints("pads", {0, 0, 0, 0});
    if (st[0] != st[1]) fail("Conv '" + n.name + "': anisotropic stride");
    if (!(pd[0] == pd[1] && pd[1] == pd[2] && pd[2] == pd[3])) fail("Conv '" + n.name + "': asymmetric padding");
    if (Cin != in.C) fail("Conv '" + n.name + "': weight expects " + std::to_string(Cin) + " input channels, tensor has " + std::to_string(in.C));
    int stride = int(st[0]), pad = int(pd[0]);
    int Ho = (in.H + 2 * pad - kh) / stride + 1, Wo = (in.W + 2 * pad - kw) / stride + 1;
    if (Ho <= 0 || Wo <= 0) fail("Conv '" + n.name + "': empty output");

    PlanLayer L;
    L.kind = LayerKind::Conv;
    L.name = n.name;
    L.in = tin;
    if (pro) {
      L.pre_scale = pro->scale; L.pre_shift = pro->shift; L.pre_relu = pro->relu;
      L.name = pro->names + "+" + n.name;
    }
    L.Cin = in.Cpad; L.kh = kh; L.kw = kw; L.stride = stride; L.pad = pad;
    L.macs_per_sample = double(Cout) * Cin * kh * kw * Ho * Wo;
    std::vector<double> scale(size_t(Cout), 1.0);
    std::vector<double> bias(size_t(Cout), 0.0);
    if (n.inputs.size() > 2 && !n.inputs[2].empty()) {
      const auto& b = init(n.inputs[2], "Conv bias");
      for (int c = 0; c < Cout; ++c) bias[size_t(c)] = b.f32[size_t(c)];
    }
    std::string cur = n.outputs[0];
    int b = sole_consumer(cur, "BatchNormalization");
    if (b >= 0) {
      std::vector<float> s, t;
      bn_affine(m.nodes[size_t(b)], s, t);
      if (int(s.size()) != Cout) fail("BN after Conv '" + n.name + "': channel mismatch");
      for (int c = 0; c < Cout; ++c) { scale[size_t(c)] = s[size_t(c)]; bias[size_t(c)] = bias[size_t(c)] * s[size_t(c)] + t[size_t(c)]; }
      absorbed[size_t(b)] = 1;
      cur = m.nodes[size_t(b)].outputs[0];
      L.name += "+" + m.nodes[size_t(b)].name;
    }
    int r = sole_consumer(cur, "Relu");
    if (r >= 0) { L.relu = true; absorbed[size_t(r)] = 1; cur = m.nodes[size_t(r)].outputs[0]; L.name += "+" + m.nodes[size_t(r)].name; }

    int CoutP = cpad(Cout);
    L.Cout = CoutP;
    L.weight.assign(size_t(CoutP) * kh * kw * in.Cpad, 0.f);
    for (int o = 0; o < Cout; ++o)
      for (int c = 0; c < Cin; ++c)
        for (int y = 0; y < kh; ++y)
          for (int x = 0; x < kw; ++x) {
            double v = double(w.f32[((size_t(o) * Cin + c) * kh + y) * kw + x]) * scale[size_t(o)];
            L.weight[((size_t(o) * kh + y) * kw + x) * in.Cpad + c] = float(v);
          }
    L.bias.assign(size_t(CoutP), 0.f);
    for (int o = 0; o < Cout; ++o) L.bias[size_t(o)] = float(bias[size_t(o)]);
    L.out = new_tensor(cur, Cout, Ho, Wo, false);
    value[cur] = L.out;
    emit(std::move(L));
  }

  void do_add(int idx) {
    const OnnxNode& n = m.nodes[size_t(idx)];
    int ta = tensor_of(n.inputs[0], n), tb = tensor_of(n.inputs[1], n);
    const PlanTensor A = plan.tensors[size_t(ta)], B = plan.tensors[size_t(tb)];
    if (A.C != B.C || A.H != B.H || A.W != B.W || A.vec_f32 != B.vec_f32) fail("Add '" + n.name + "': shapes differ (broadcast is not supported)");
    // Both operands are linear 1x1 convs read only by this Add: one GEMM over both inputs' channels.
    if (dual_input && ta != tb) {
      int li[2] = {-1, -1};
      for (int side = 0; side < 2; ++side) {
        const int t = side ? tb : ta;
        const std::string& vname = n.inputs[size_t(side)];
        auto pit = producer_layer.find(t);
        if (pit == producer_layer.end()) continue;
        const PlanLayer& P = layers[size_t(pit->second)];
        if (P.kind != LayerKind::Conv || P.out != t || P.in2 >= 0 || P.in_b >= 0 || P.out_act >= 0 || P.relu) continue;
        if (P.kh != 1 || P.kw != 1 || P.pad != 0 || !P.pre_scale.empty() || P.Cin % 64 || P.Cout % 256) continue;
        if (graph_outputs.count(vname) || consumers[vname].size() != 1) continue;
        li[side] = pit->second;
      }
      if (li[0] >= 0 && li[1] >= 0 && layers[size_t(li[0])].Cout == layers[size_t(li[1])].Cout) {
        // the carrier is a stride-1 conv (its geometry is the output's); prefer the one with fewer MACs
        int carrier = -1;
        for (int side = 0; side < 2; ++side) {
          const PlanLayer& P = layers[size_t(li[side])];
          if (P.stride != 1) continue;
          if (carrier < 0 || P.macs_per_sample < layers[size_t(carrier)].macs_per_sample) carrier = li[side];
        }
        if (carrier >= 0) {
          const int other_li = carrier == li[0] ? li[1] : li[0];
          PlanLayer& Cc = layers[size_t(carrier)];
          PlanLayer& O = layers[size_t(other_li)];
          const int K1 = Cc.Cin, K2 = O.Cin, Co = Cc.Cout;
          std::vector<float> wcat(size_t(Co) * size_t(K1 + K2));
          for (int o = 0; o < Co; ++o) {
            std::copy(Cc.weight.begin() + size_t(o) * K1, Cc.weight.begin() + size_t(o + 1) * K1, wcat.begin() + size_t(o) * (K1 + K2));
            std::copy(O.weight.begin() + size_t(o) * K2, O.weight.begin() + size_t(o + 1) * K2, wcat.begin() + size_t(o) * (K1 + K2) + K1);
          }
          Cc.weight = std::move(wcat);
          for (int o = 0; o < Co; ++o) Cc.bias[size_t(o)] += O.bias[size_t(o)];
          Cc.in_b = O.in; Cc.Cin_b = K2; Cc.stride_b = O.stride;
          Cc.macs_per_sample += O.macs_per_sample;
          Cc.name += "+" + O.name + "+" + n.name;
          plan.tensors[size_t(Cc.out)].name = n.outputs[0];
          value[n.outputs[0]] = Cc.out;
          producer_layer.erase(O.out);
          // drop the absorbed conv from the execution order and run the carrier last (both inputs exist by then)
          order.erase(std::find(order.begin(), order.end(), other_li));
          order.erase(std::find(order.begin(), order.end(), carrier));
          order.push_back(carrier);
          O.weight.clear();
          return;
        }
      }
    }
    // fuse into the Conv that produced one operand, if this Add is that operand's only reader.  When both sides
    // qualify (first block of a stage: main-branch 1x1 conv vs projection shortcut) the one with fewer MACs carries
    // the residual epilogue -- that epilogue is the expensive part of such a layer, and the smaller reduction keeps
    // its weights resident in shared memory (conv_wide_kernel); on a tie the later one in graph order.
    int best = -1, best_pos = -1, other = -1;
    double best_macs = 0;
    for (int side = 0; side < 2; ++side) {
      int t = side ? tb : ta;
      const std::string& vname = n.inputs[size_t(side)];
      auto pit = producer_layer.find(t);
      if (pit == producer_layer.end()) continue;
      const PlanLayer& P = layers[size_t(pit->second)];
      if (P.kind != LayerKind::Conv || P.out != t || P.in2 >= 0 || P.out_act >= 0) continue;
      if (graph_outputs.count(vname) || consumers[vname].size() != 1) continue;
      int pos = int(std::find(order.begin(), order.end(), pit->second) - order.begin());
      const bool better = best < 0 || P.macs_per_sample < best_macs || (P.macs_per_sample == best_macs && pos > best_pos);
      if (better) { best = pit->second; best_pos = pos; best_macs = P.macs_per_sample; other = side ? ta : tb; }
    }
    if (best >= 0 && ta != tb) {
      PlanLayer& P = layers[size_t(best)];
      P.in2 = other;
      P.name += "+" + n.name;
      plan.tensors[size_t(P.out)].name = n.outputs[0];
      value[n.outputs[0]] = P.out;
      // run it after the other operand exists: move to the end of the execution order
      order.erase(order.begin() + best_pos);
      order.push_back(best);
      return;
    }
    PlanLayer L;
    L.kind = LayerKind::Add;
    L.name = n.name;
    L.in = ta; L.in2 = tb;
    L.out = new_tensor(n.outputs[0], A.C, A.H, A.W, A.vec_f32);
    value[n.outputs[0]] = L.out;
    emit(std::move(L));
  }

  void do_bn_or_relu(int idx) {
    const OnnxNode& n = m.nodes[size_t(idx)];
    bool is_bn = n.op == "BatchNormalization";
    if (is_bn && defer_act_to_consumers(idx)) return;
    if (is_bn && fuse_act_into_producer(idx)) return;
    int tin = tensor_of(n.inputs[0], n);
    const PlanTensor in = plan.tensors[size_t(tin)];
    if (in.vec_f32) fail(n.op + " '" + n.name + "' on a flattened tensor is not supported");
    std::vector<float> s(size_t(in.C), 1.f), t(size_t(in.C), 0.f);
    bool relu = !is_bn;
    std::string cur = n.outputs[0], name = n.name;
    if (is_bn) {
      bn_affine(n, s, t);
      if (int(s.size()) != in.C) fail("BN '" + n.name + "': channel mismatch");
      int r = sole_consumer(cur, "Relu");
      if (r >= 0) { relu = true; absorbed[size_t(r)] = 1; cur = m.nodes[size_t(r)].outputs[0]; name += "+" + m.nodes[size_t(r)].name; }
    }
    pad_to(s, size_t(in.Cpad), 0.f);
    pad_to(t, size_t(in.Cpad), 0.f);
    PlanLayer L;
    L.in = tin;
    L.act_scale = std::move(s); L.act_shift = std::move(t); L.act_relu = relu;
    int g = sole_consumer(cur, "GlobalAveragePool");
    if (g >= 0) {
      L.kind = LayerKind::Gap;
      absorbed[size_t(g)] = 1;
      cur = m.nodes[size_t(g)].outputs[0];
      L.name = name + "+" + m.nodes[size_t(g)].name;
      L.out = new_tensor(cur, in.C, 1, 1, true);
    } else {
      L.kind = LayerKind::Affine;
      L.name = name;
      L.out = new_tensor(cur, in.C, in.H, in.W, false);
    }
    value[cur] = L.out;
    emit(std::move(L));
  }

  void do_maxpool(int idx) {
    const OnnxNode& n = m.nodes[size_t(idx)];
    int tin = tensor_of(n.inputs[0], n);
    const PlanTensor in = plan.tensors[size_t(tin)];
    auto ks = n.attr_ints("kernel_shape", {});
    auto st = n.attr_ints("strides", {1, 1});
    auto pd = n.attr_ints("pads", {0, 0, 0, 0});
    if (ks.size() != 2 || ks[0] != ks[1] || st[0] != st[1] || !(pd[0] == pd[1] && pd[1] == pd[2] && pd[2] == pd[3]))
      fail("MaxPool '" + n.name + "': only square kernels with symmetric stride/pads are supported");
    PlanLayer L;
    L.kind = LayerKind::MaxPool;
    L.name = n.name;
    L.in = tin;
    L.pool_k = int(ks[0]); L.pool_stride = int(st[0]); L.pool_pad = int(pd[0]);
    int Ho = (in.H + 2 * L.pool_pad - L.pool_k) / L.pool_stride + 1;
    int Wo = (in.W + 2 * L.pool_pad - L.pool_k) / L.pool_stride + 1;
    L.out = new_tensor(n.outputs[0], in.C, Ho, Wo, false);
    value[n.outputs[0]] = L.out;
    emit(std::move(L));
  }

  void do_gap(int idx) {
    const OnnxNode& n = m.nodes[size_t(idx)];
    int tin = tensor_of(n.inputs[0], n);
    const PlanTensor in = plan.tensors[size_t(tin)];
    PlanLayer L;
    L.kind = LayerKind::Gap;
    L.name = n.name;
    L.in = tin;
    L.act_scale.assign(size_t(in.Cpad), 1.f);
    L.act_shift.assign(size_t(in.Cpad), 0.f);
    L.out = new_tensor(n.outputs[0], in.C, 1, 1, true);
    value[n.outputs[0]] = L.out;
    emit(std::move(L));
  }

  void do_flatten(int idx) {
    const OnnxNode& n = m.nodes[size_t(idx)];
    int tin = tensor_of(n.inputs[0], n);
    const PlanTensor& in = plan.tensors[size_t(tin)];
    // NHWC storage equals the NCHW flattening only when H = W = 1 (pooled rows)
    if (!in.vec_f32) fail(n.op + " '" + n.name + "' of a tensor with spatial extent is not supported (would need an NHWC->NCHW re-layout)");
    if (n.op == "Flatten" && n.attr_i("axis", 1) != 1) fail("Flatten '" + n.name + "': axis != 1");
    value[n.outputs[0]] = tin;  // alias
    plan.tensors[size_t(tin)].name = n.outputs[0];
  }

  void do_gemm(int idx) {
    const OnnxNode& n = m.nodes[size_t(idx)];
    int tin = tensor_of(n.inputs[0], n);
    const PlanTensor in = plan.tensors[size_t(tin)];
    if (!in.vec_f32) fail("Gemm '" + n.name + "' input must be a flattened [B, K] tensor");
    if (n.attr_i("transA", 0) != 0) fail("Gemm '" + n.name + "': transA is not supported");
    const auto& w = init(n.inputs[1], "Gemm weight");
    if (w.dims.size() != 2) fail("Gemm '" + n.name + "': weight must be 2-D");
    bool transB = n.attr_i("transB", 0) != 0;
    int N = int(transB ? w.dims[0] : w.dims[1]), K = int(transB ? w.dims[1] : w.dims[0]);
    if (K != in.C) fail("Gemm '" + n.name + "': K mismatch");
    double alpha = n.attr_f("alpha", 1.f), beta = n.attr_f("beta", 1.f);
    PlanLayer L;
    L.kind = LayerKind::Gemm;
    L.name = n.name;
    L.in = tin;
    L.Cin = K; L.Cout = N;
    L.weight.resize(size_t(N) * K);
    for (int o = 0; o < N; ++o)
      for (int k = 0; k < K; ++k)
        L.weight[size_t(o) * K + k] = float(alpha * double(transB ? w.f32[size_t(o) * K + k] : w.f32[size_t(k) * N + o]));
    L.bias.assign(size_t(N), 0.f);
    if (n.inputs.size() > 2 && !n.inputs[2].empty()) {
      const auto& b = init(n.inputs[2], "Gemm bias");
      if (int(b.f32.size()) != N) fail("Gemm '" + n.name + "': bias must have N elements");
      for (int o = 0; o < N; ++o) L.bias[size_t(o)] = float(beta * double(b.f32[size_t(o)]));
    }
    L.macs_per_sample = double(N) * K;
    L.out = new_tensor(n.outputs[0], N, 1, 1, true);
    value[n.outputs[0]] = L.out;
    emit(std::move(L));
  }

  void do_softmax(int idx) {
    const OnnxNode& n = m.nodes[size_t(idx)];
    int tin = tensor_of(n.inputs[0], n);
    const PlanTensor in = plan.tensors[size_t(tin)];
    if (!in.vec_f32 || n.attr_i("axis", 1) != 1) fail("Softmax '" + n.name + "': only [B, N] rows with axis=1 are supported");
    PlanLayer L;
    L.kind = LayerKind::Softmax;
    L.name = n.name;
    L.in = tin;
    L.out = new_tensor(n.outputs[0], in.C, 1, 1, true);
    value[n.outputs[0]] = L.out;
    emit(std::move(L));
  }

  // End of one bottleneck block + start of the next: the wide 1x1 conv's activated output feeds exactly one
  // narrow 1x1 conv.  Decided from reader counts and shapes only.
  void fuse_tails() {
    auto& Ls = plan.layers;
    for (size_t i = 0; i < Ls.size(); ++i) {
      PlanLayer& L = Ls[i];
      if (L.kind != LayerKind::Conv || L.out < 0 || L.out_act < 0 || (L.in2 < 0 && L.in_b < 0) || L.tail_out >= 0) continue;
      if (L.kh != 1 || L.kw != 1 || L.stride != 1 || L.pad != 0 || !L.pre_scale.empty()) continue;
      if (L.Cout != 256 || L.Cin != 64 || (L.in_b >= 0 && (L.Cin_b != 64 || L.in2 >= 0))) continue;
      if (L.out_act == value_or(plan.output_name)) continue;
      int readers = 0, j = -1;
      for (size_t k = 0; k < Ls.size(); ++k)
        for (int t : {Ls[k].in, Ls[k].in2, Ls[k].in_b})
          if (t == L.out_act) { ++readers; j = int(k); }
      if (readers != 1 || j <= int(i)) continue;
      PlanLayer& M = Ls[size_t(j)];
      if (M.kind != LayerKind::Conv || M.in != L.out_act || M.in2 >= 0 || M.out < 0 || M.out_act >= 0) continue;
      if (M.kh != 1 || M.kw != 1 || M.stride != 1 || M.pad != 0 || !M.pre_scale.empty() || M.tail_out >= 0) continue;
      if (M.Cin != L.Cout || (M.Cout != 64 && M.Cout != 128)) continue;
      L.tail_out = M.out; L.tail_Cout = M.Cout; L.tail_relu = M.relu;
      L.tail_weight = std::move(M.weight); L.tail_bias = std::move(M.bias);
      L.macs_per_sample += M.macs_per_sample;
      L.name += "+" + M.name;
      L.out_act = -1;
      Ls.erase(Ls.begin() + j);
    }
  }
  // The stem conv's 112x112x64 output is the network's largest tensor and only the max-pool reads it.
  void fuse_stem_pool() {
    auto& Ls = plan.layers;
    for (size_t i = 0; i < Ls.size(); ++i) {
      PlanLayer& L = Ls[i];
      if (L.kind != LayerKind::Conv || L.kh != 7 || L.kw != 7 || L.stride != 2 || L.pad != 3 || L.Cin != 4) continue;
      if (L.out < 0 || L.out_act >= 0 || L.in2 >= 0 || L.in_b >= 0 || L.tail_out >= 0 || !L.pre_scale.empty()) continue;
      int readers = 0, j = -1;
      for (size_t k = 0; k < Ls.size(); ++k)
        for (int t : {Ls[k].in, Ls[k].in2, Ls[k].in_b})
          if (t == L.out) { ++readers; j = int(k); }
      if (readers != 1 || j <= int(i)) continue;
      PlanLayer& P = Ls[size_t(j)];
      if (P.kind != LayerKind::MaxPool || P.in != L.out || P.pool_k != 3 || P.pool_stride != 2 || P.pool_pad != 1) continue;
      if (L.out == value_or(plan.output_name)) continue;
      L.out = P.out; L.out_act = P.out_act;
      L.act_scale = std::move(P.act_scale); L.act_shift = std::move(P.act_shift); L.act_relu = P.act_relu;
      L.fused_pool = true;
      L.name += "+" + P.name;
      Ls.erase(Ls.begin() + j);
      return;
    }
  }
  int value_or(const std::string& name) const {
    auto it = value.find(name);
    return it == value.end() ? -2 : it->second;
  }

  Plan finish() {
    for (int li : order) plan.layers.push_back(std::move(layers[size_t(li)]));
    plan.output_name = m.outputs[0].name;
    if (tail_fusion) fuse_tails();
    if (stem_pool) fuse_stem_pool();
    const auto& out = m.outputs[0];
    auto it = value.find(out.name);
    if (it == value.end()) fail("graph output '" + out.name + "' is not produced by a supported node");
    plan.output_name = out.name;
    plan.output_tensor = it->second;
    PlanTensor& ot = plan.tensors[size_t(plan.output_tensor)];
    if (!ot.vec_f32) fail("graph output must be a flattened [B, N] tensor");
    ot.is_output = true;
    // shapes as the reference reports them: dynamic dims (-1) become 1 (src/inference_engine.cpp:47-51,64-68)
    for (auto d : m.inputs[0].dims) plan.input_shape.push_back(d == -1 ? 1 : d);
    if (out.dims.empty()) { plan.output_shape = {1, ot.C}; }
    else for (auto d : out.dims) plan.output_shape.push_back(d == -1 ? 1 : d);
    plan.in_elems_per_sample = int64_t(plan.in_C) * plan.in_H * plan.in_W;
    plan.out_elems_per_sample = ot.C;
    // liveness + dependency check
    for (size_t i = 0; i < plan.layers.size(); ++i) {
      PlanLayer& L = plan.layers[i];
      for (int t : {L.in, L.in2, L.in_b}) {
        if (t < 0) continue;
        PlanTensor& T = plan.tensors[size_t(t)];
        if (T.def_layer < 0) fail("internal: layer '" + L.name + "' reads '" + T.name + "' before it is produced");
        T.last_use = int(i);
      }
      for (int t : {L.out, L.out_act, L.tail_out}) {
        if (t < 0) continue;
        PlanTensor& T = plan.tensors[size_t(t)];
        if (T.def_layer < 0) { T.def_layer = int(i); T.last_use = std::max(T.last_use, int(i)); }
      }
      plan.total_macs_per_sample += L.macs_per_sample;
    }
    ot.last_use = int(plan.layers.size());
    return std::move(plan);
  }

  Plan run() {
    do_input();
    for (size_t i = 0; i < m.nodes.size(); ++i) {
      if (absorbed[i]) continue;
      const std::string& op = m.nodes[i].op;
      if (op == "Conv") do_conv(int(i));
      else if (op == "BatchNormalization" || op == "Relu") do_bn_or_relu(int(i));
      else if (op == "Add") do_add(int(i));
      else if (op == "MaxPool") do_maxpool(int(i));
      else if (op == "GlobalAveragePool") do_gap(int(i));
      else if (op == "Flatten" || op == "Reshape") do_flatten(int(i));
      else if (op == "Gemm") do_gemm(int(i));
      else if (op == "Softmax") do_softmax(int(i));
      else fail("operator '" + op + "' (node '" + m.nodes[i].name + "') is not supported by the B200 backend");
    }
    return finish();
  }
};

}  // namespace

Plan build_plan(const OnnxModel& model, int channel_align, bool consumer_prologue, bool tail_fusion, bool dual_input,
                bool stem_pool) {
  Builder b(model, channel_align, consumer_prologue);
  b.tail_fusion = tail_fusion;
  b.dual_input = dual_input;
  b.stem_pool = stem_pool;
  return b.run();
}

std::string Plan::describe() const {
  std::ostringstream os;
  os << "input " << input_name << " [" << in_C << "," << in_H << "," << in_W << "] -> output " << output_name
     << " [" << out_elems_per_sample << "]; layers=" << layers.size() << " macs/sample=" << (long long)total_macs_per_sample << "\n";
  auto tn = [&](int t) {
    if (t < 0) return std::string("-");
    const PlanTensor& T = tensors[size_t(t)];
    std::ostringstream s;
    s << "t" << t << "[" << T.C << "x" << T.H << "x" << T.W << (T.vec_f32 ? " f32row" : "") << "]";
    return s.str();
  };
  for (size_t i = 0; i < layers.size(); ++i) {
    const PlanLayer& L = layers[i];
    os << i << " " << layer_kind_name(L.kind) << " in=" << tn(L.in) << " in2=" << tn(L.in2) << " out=" << tn(L.out)
       << " out_act=" << tn(L.out_act);
    if (L.in_b >= 0) os << " in_b=" << tn(L.in_b) << " cin_b=" << L.Cin_b << " stride_b=" << L.stride_b;
    if (L.fused_pool) os << " fused_pool=1";
    if (L.tail_out >= 0) os << " tail_out=" << tn(L.tail_out) << " tail_cout=" << L.tail_Cout;
    if (L.kind == LayerKind::Conv)
      os << " k=" << L.kh << "x" << L.kw << " s=" << L.stride << " p=" << L.pad << " cin=" << L.Cin << " cout=" << L.Cout;
    if (L.kind == LayerKind::MaxPool) os << " k=" << L.pool_k << " s=" << L.pool_stride << " p=" << L.pool_pad;
    os << " relu=" << int(L.relu) << " act_relu=" << int(L.act_relu) << " pre=" << int(!L.pre_scale.empty()) << " | " << L.name << "\n";
  }
  return os.str();
}

}  // namespace rn
