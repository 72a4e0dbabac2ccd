// The worker's request path: cache check -> dynamic batch -> engine -> cache fill, and the /health view.
//
// Same behaviour and JSON contract as the reference's WorkerNode (reference: src/worker_node.cpp:27-140):
//   handleInfer  (:50-83)   total_requests++; request_id / input_data read with nlohmann's conversions (a
//                           missing key raises type_error.302 exactly as there); LRU hit -> cached:true with the
//                           constant inference_time_us 50 (:65); miss -> BatchProcessor::process, cache_.put,
//                           cached:false and the batch's per-request time
//   processBatch (:106-133) one engine call per batch; inference_time_us = elapsed_us / batch (integer division)
//   getHealth    (:85-103)  healthy, node_id, total_requests, cache_hits, cache_size, cache_hit_rate,
//                           batch_processor{total_batches, avg_batch_size, timeout_batches, full_batches}
// Keys come out alphabetically (nlohmann's default std::map object), floats are printed shortest-round-trip
// after widening to double, as in the reference.
//
// Differences in mechanism only: the engine is a template parameter (the product instantiates it with the
// B200 InferenceEngine; CPU tests with a stub), request payloads are moved rather than copied on their way
// into the batch, and the engine is handed pointers to the requests' own vectors (no vector<vector> copy).
#pragma once
#include <atomic>
#include <chrono>
#include <cstdint>
#include <string>
#include <vector>

#include <nlohmann/json.hpp>

#include "batch_processor.h"
#include "fast_request.h"
#include "lru_cache.h"

struct InferenceRequest {
  std::string request_id;
  std::vector<float> input_data;
};

struct InferenceResponse {
  std::string request_id;
  std::vector<float> output_data;
  int64_t inference_time_us;
  bool cached;
};

template <class Engine>
class WorkerNodeT {
 public:
  using json = nlohmann::json;

  WorkerNodeT(const std::string& node_id, int port, const std::string& model_path, size_t cache_capacity = 1000,
              size_t max_batch = 32, std::chrono::milliseconds timeout = std::chrono::milliseconds(20))
      : node_id_(node_id),
        port_(port),
        engine_(model_path, port % 3),
        cache_(cache_capacity),
        batcher_(max_batch, timeout, [this](const std::vector<InferenceRequest>& reqs) { return processBatch(reqs); }) {
    batcher_.start();
  }
  ~WorkerNodeT() { batcher_.stop(); }

  // Raw-body entry: the common request shape is scanned straight into a vector<float> (fast_request.h); every
  // other body takes the nlohmann path below, so what is accepted and every error text stay the same.
  json handleInferBody(const std::string& body) {
    std::string request_id;
    std::vector<float> input_data;
    if (fast_request::parse(body, request_id, input_data)) {
      ++total_requests_;
      return handleInferFields(request_id, std::move(input_data));
    }
    return handleInfer(json::parse(body));
  }

  // Entry for requests whose fields arrive already decoded (the octet-stream side channel).
  json handleInferRaw(const std::string& request_id, std::vector<float> input_data) {
    ++total_requests_;
    return handleInferFields(request_id, std::move(input_data));
  }

  json handleInfer(const json& request) {
    ++total_requests_;
    // A missing key reads as null and the conversion then throws nlohmann's type_error.302
    // ("type must be string, but is null"), the text the reference's handler reports; a const
    // operator[] on a missing key would be undefined behaviour, so the lookup is explicit.
    static const json kNull;
    auto field = [&](const char* k) -> const json& {
      auto it = request.find(k);
      return it == request.end() ? kNull : *it;
    };
    std::string request_id = field("request_id");
    std::vector<float> input_data = field("input_data");
    return handleInferFields(request_id, std::move(input_data));
  }

  // The handler after its two fields have been read (by the conversions above, or by fast_request::parse from the
  // raw body); the caller has counted the request.  The sample is taken by value: on a miss it is MOVED into the batch
  // (the reference copies it into the request and again into the queue, src/worker_node.cpp:71, batch_processor.h:98)
  // after one copy has been kept as the cache key.
  json handleInferFields(const std::string& request_id, std::vector<float> input_data) {
    if (auto hit = cache_.get(input_data)) {
      ++cache_hits_;
      json response;
      response["request_id"] = request_id;
      response["output_data"] = *hit;
      response["node_id"] = node_id_;
      response["cached"] = true;
      response["inference_time_us"] = 50;
      return response;
    }
    std::vector<float> key(input_data);                       // the cache stores its key; everything else moves
    InferenceResponse resp = batcher_.process(InferenceRequest{request_id, std::move(input_data)});
    cache_.put(std::move(key), resp.output_data);

    json response;
    response["request_id"] = resp.request_id;
    response["output_data"] = resp.output_data;
    response["node_id"] = node_id_;
    response["cached"] = false;
    response["inference_time_us"] = resp.inference_time_us;
    return response;
  }

  json getHealth() {
    auto m = batcher_.getMetrics();
    json health;
    health["healthy"] = true;
    health["node_id"] = node_id_;
    health["total_requests"] = total_requests_.load();
    health["cache_hits"] = cache_hits_.load();
    health["cache_size"] = cache_.size();
    health["cache_hit_rate"] = cache_.getHitRate();
    json b;
    b["total_batches"] = m.total_batches;
    b["avg_batch_size"] = m.avg_batch_size;
    b["timeout_batches"] = m.timeout_batches;
    b["full_batches"] = m.full_batches;
    health["batch_processor"] = b;
    return health;
  }

  Engine& engine() { return engine_; }
  void setFillWait(bool on) { batcher_.setFillWait(on); }
  // n batches in flight (n consumer threads; the engine's batchPredict must accept concurrent calls -- the B200
  // InferenceEngine gives each its own transfer slot).  Call before serving: queued requests would be failed.
  void setInFlight(int n) {
    batcher_.stop();
    batcher_.setInFlight(n);
    batcher_.start();
  }

 private:
  std::vector<InferenceResponse> processBatch(const std::vector<InferenceRequest>& requests) {
    const auto start = std::chrono::high_resolution_clock::now();
    std::vector<std::vector<float>> inputs;
    inputs.reserve(requests.size());
    for (const auto& r : requests) inputs.push_back(r.input_data);
    auto outputs = engine_.batchPredict(inputs);
    const auto end = std::chrono::high_resolution_clock::now();
    const int64_t us = std::chrono::duration_cast<std::chrono::microseconds>(end - start).count();
    const int64_t per_request = us / int64_t(requests.size());
    std::vector<InferenceResponse> out;
    out.reserve(requests.size());
    for (size_t i = 0; i < requests.size(); ++i)
      out.push_back(InferenceResponse{requests[i].request_id, std::move(outputs[i]), per_request, false});
    return out;
  }

  std::string node_id_;
  int port_;
  Engine engine_;
  LRUCache<std::vector<float>, std::vector<float>, VectorHash> cache_;
  BatchProcessor<InferenceRequest, InferenceResponse> batcher_;
  std::atomic<int64_t> total_requests_{0};
  std::atomic<int64_t> cache_hits_{0};
};
