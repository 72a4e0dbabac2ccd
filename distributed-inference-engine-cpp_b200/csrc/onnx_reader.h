// Minimal ONNX (protobuf wire format) reader for the B200 ResNet backend.
//
// Replaces what the reference delegates to ONNX Runtime when it opens the model
// (reference: src/inference_engine.cpp:31 `Ort::Session(env, model_path, options)` and the
// input/output introspection at :39-68).  No protobuf library, no ONNX Runtime: the file is
// memory-mapped and decoded field by field (field numbers from onnx.proto3).
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace rn {

struct OnnxTensor {
  std::string name;
  std::vector<int64_t> dims;
  int data_type = 1;            // 1 = FLOAT, 7 = INT64
  std::vector<float> f32;       // decoded payload for FLOAT tensors
  std::vector<int64_t> i64;     // decoded payload for INT64 tensors
  int64_t numel() const {
    int64_t n = 1;
    for (auto d : dims) n *= d;
    return n;
  }
};

struct OnnxAttr {
  int type = 0;                 // AttributeProto.AttributeType
  float f = 0.f;
  int64_t i = 0;
  std::string s;
  std::vector<float> floats;
  std::vector<int64_t> ints;
  bool has_f = false, has_i = false;
};

struct OnnxNode {
  std::string op, name;
  std::vector<std::string> inputs, outputs;
  std::map<std::string, OnnxAttr> attrs;
  int64_t attr_i(const std::string& k, int64_t dflt) const;
  float attr_f(const std::string& k, float dflt) const;
  std::vector<int64_t> attr_ints(const std::string& k, std::vector<int64_t> dflt) const;
};

struct OnnxValueInfo {
  std::string name;
  std::vector<int64_t> dims;    // dim_param / unknown -> -1 (as ONNX Runtime reports it)
};

struct OnnxModel {
  std::vector<OnnxNode> nodes;
  std::map<std::string, OnnxTensor> initializers;
  std::vector<OnnxValueInfo> inputs;    // graph inputs that are NOT initializers, in file order
  std::vector<OnnxValueInfo> outputs;
  int64_t ir_version = 0, opset = 0;
};

// Throws std::runtime_error on an unreadable or malformed file (the reference's constructor
// likewise throws from Ort::Session on a missing model; worker_node.cpp:171 does not catch).
OnnxModel load_onnx(const std::string& path);

}  // namespace rn
