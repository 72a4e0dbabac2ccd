// Drop-in replacement for the reference's InferenceEngine with the identical public surface
// (reference: include/inference_engine.h:10-21) and no ONNX Runtime anywhere: the body calls the
// C ABI of libb200rn.so (include/b200rn.h), which parses the ONNX file itself and runs the
// sm_100a kernels.  worker_node.cpp needs no source change beyond including this header
// (its only use is `engine_.batchPredict(inputs)`, src/worker_node.cpp:116).
//
// Environment (optional; the constructor signature is fixed by the reference):
//   B200RN_MODE      exact | bf16 | fp16   arithmetic mode          (default fp16)
//   B200RN_DEVICE    CUDA device ordinal                            (default 0, as the reference)
//   B200RN_MAX_BATCH activation arena size in samples               (default 256)
#ifndef INFERENCE_ENGINE_H
#define INFERENCE_ENGINE_H

#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

struct b200rn;

class InferenceEngine {
 public:
  explicit InferenceEngine(const std::string& model_path, int shard_id = 0);
  ~InferenceEngine();
  InferenceEngine(const InferenceEngine&) = delete;
  InferenceEngine& operator=(const InferenceEngine&) = delete;

  std::vector<float> predict(const std::vector<float>& input);
  std::vector<std::vector<float>> batchPredict(const std::vector<std::vector<float>>& inputs);
  const std::string& getModelPath() const { return model_path_; }
  int getShardId() const { return shard_id_; }
  std::vector<int64_t> getInputShape() const;
  std::vector<int64_t> getOutputShape() const;

 private:
  std::string model_path_;
  int shard_id_;
  b200rn* handle_ = nullptr;
  std::vector<int64_t> input_shape_, output_shape_;
};

#endif
