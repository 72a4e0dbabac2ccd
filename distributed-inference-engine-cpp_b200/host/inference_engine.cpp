#include "inference_engine.h"

#include <cstdlib>
#include <cstring>
#include <iostream>
#include <stdexcept>

#include "../../include/b200rn.h"

namespace {
int env_int(const char* name, int dflt) {
  const char* v = std::getenv(name);
  return (v && *v) ? std::atoi(v) : dflt;
}
int env_mode() {
  const char* v = std::getenv("B200RN_MODE");
  if (!v || !*v) return B200RN_MODE_FP16;
  if (!std::strcmp(v, "exact") || !std::strcmp(v, "fp32")) return B200RN_MODE_EXACT;
  if (!std::strcmp(v, "bf16")) return B200RN_MODE_BF16;
  if (!std::strcmp(v, "fp16")) return B200RN_MODE_FP16;
  throw std::runtime_error(std::string("B200RN_MODE: unknown mode '") + v + "'");
}
void print_shape(const char* label, const std::vector<int64_t>& s) {
  std::cout << "  " << label << ": [";
  for (size_t i = 0; i < s.size(); ++i) std::cout << s[i] << (i + 1 < s.size() ? ", " : "");
  std::cout << "]" << std::endl;
}
}  // namespace

InferenceEngine::InferenceEngine(const std::string& model_path, int shard_id)
    : model_path_(model_path), shard_id_(shard_id) {
  const int rc = b200rn_create(model_path.c_str(), env_int("B200RN_DEVICE", 0), env_mode(),
                               env_int("B200RN_MAX_BATCH", 256), &handle_);
  if (rc != B200RN_OK)  // the reference throws out of its constructor too (Ort::Session on a bad path)
    throw std::runtime_error(std::string("InferenceEngine: ") + b200rn_global_error());
  int64_t d[8];
  int r = b200rn_input_shape(handle_, d, 8);
  input_shape_.assign(d, d + r);
  r = b200rn_output_shape(handle_, d, 8);
  output_shape_.assign(d, d + r);
  std::cout << "ONNX model loaded: " << model_path_ << std::endl;
  print_shape("Input shape", input_shape_);
  print_shape("Output shape", output_shape_);
}

InferenceEngine::~InferenceEngine() { b200rn_destroy(handle_); }

std::vector<float> InferenceEngine::predict(const std::vector<float>& input) {
  std::vector<float> out(static_cast<size_t>(b200rn_output_elems(handle_)));
  if (b200rn_predict(handle_, input.data(), input.size(), out.data()) != B200RN_OK)
    throw std::runtime_error(std::string("InferenceEngine::predict: ") + b200rn_last_error(handle_));
  return out;
}

std::vector<std::vector<float>> InferenceEngine::batchPredict(const std::vector<std::vector<float>>& inputs) {
  if (inputs.empty()) return {};
  const size_t n = inputs.size(), per = static_cast<size_t>(b200rn_output_elems(handle_));
  std::vector<const float*> ptrs(n);
  std::vector<size_t> lens(n);
  for (size_t i = 0; i < n; ++i) {
    ptrs[i] = inputs[i].data();
    lens[i] = inputs[i].size();
  }
  std::vector<float> flat(n * per);
  if (b200rn_infer(handle_, ptrs.data(), lens.data(), n, flat.data()) != B200RN_OK)
    throw std::runtime_error(std::string("InferenceEngine::batchPredict: ") + b200rn_last_error(handle_));
  std::vector<std::vector<float>> results;
  results.reserve(n);
  for (size_t i = 0; i < n; ++i) results.emplace_back(flat.begin() + long(i * per), flat.begin() + long((i + 1) * per));
  return results;
}

std::vector<int64_t> InferenceEngine::getInputShape() const { return input_shape_; }
std::vector<int64_t> InferenceEngine::getOutputShape() const { return output_shape_; }
