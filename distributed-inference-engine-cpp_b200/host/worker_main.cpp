// worker_node: one replica of the B200 ResNet backend behind the reference's worker API.
//   usage: worker_node <port> <node_id> [model_path]      (or MODEL_PATH in the environment)
// Same CLI, routes, JSON and error behaviour as the reference's main (src/worker_node.cpp:145-204):
// POST /infer -> handleInfer, exceptions -> HTTP 500 {"error": what}; GET /health -> getHealth.
// Build with -DWORKER_FAKE_ENGINE for a CPU-only binary whose "engine" returns a checksum row per sample
// (used by the host-side tests of the JSON contract; never by the product).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "http_server.h"
#include "worker_node.h"

#ifdef WORKER_FAKE_ENGINE
#include <numeric>
struct FakeEngine {
  FakeEngine(const std::string&, int) {}
  std::vector<std::vector<float>> batchPredict(const std::vector<std::vector<float>>& in) {
    std::vector<std::vector<float>> out;
    for (auto& v : in) {
      float s = std::accumulate(v.begin(), v.end(), 0.0f);
      out.push_back({s, float(v.size()), 0.1f, 1e-8f});
    }
    return out;
  }
};
using Worker = WorkerNodeT<FakeEngine>;
#else
#include "inference_engine.h"
using Worker = WorkerNodeT<InferenceEngine>;
#endif

using json = nlohmann::json;

int main(int argc, char** argv) {
  if (argc < 3) {
    std::cerr << "Usage: " << argv[0] << " <port> <node_id> [model_path]" << std::endl;
    std::cerr << "  Or set MODEL_PATH environment variable" << std::endl;
    return 1;
  }
  const int port = std::stoi(argv[1]);
  const std::string node_id = argv[2];
  std::string model_path;
  if (argc >= 4) model_path = argv[3];
  else if (const char* env = std::getenv("MODEL_PATH")) model_path = env;
  else {
    std::cerr << "Error: No model path provided!" << std::endl;
    return 1;
  }
  std::cout << "Using model: " << model_path << std::endl;
  Worker worker(node_id, port, model_path);
  if (const char* fw = std::getenv("B200RN_FILL_WAIT"))
    if (fw[0] == '1') worker.setFillWait(true);      // optional fill-or-timeout batching (changes /health counters)
  if (const char* nf = std::getenv("B200RN_INFLIGHT"))
    if (std::atoi(nf) > 1) worker.setInFlight(std::atoi(nf));   // optional: batch i+1 is packed and transferred while batch i computes

  mini_http::Server server(std::getenv("B200RN_HTTP_THREADS") ? std::atoi(std::getenv("B200RN_HTTP_THREADS")) : 0);
  if (const char* to = std::getenv("B200RN_HTTP_TIMEOUTS")) {      // "<read seconds>,<idle seconds>" (defaults 5, 30)
    int rd = 5, idle = 30;
    if (std::sscanf(to, "%d,%d", &rd, &idle) >= 1) server.set_timeouts(rd > 0 ? rd : 5, idle > 0 ? idle : 30);
  }
  server.Post("/infer", [&worker](const mini_http::Request& req, mini_http::Response& res) {
    try {
      // B200RN_FAST_JSON=0 forces the reference's parse-then-convert path for every body (A/B and tests)
      static const bool fast = !(std::getenv("B200RN_FAST_JSON") && std::getenv("B200RN_FAST_JSON")[0] == '0');
      // Additive side channel: Content-Type application/octet-stream carries the sample as raw
      // little-endian float32 and the id in X-Request-Id; the response is the usual JSON.  JSON stays the default.
      auto ct = req.headers.find("content-type");
      if (ct != req.headers.end() && ct->second.compare(0, 24, "application/octet-stream") == 0) {
        auto id = req.headers.find("x-request-id");
        if (id == req.headers.end()) throw std::runtime_error("octet-stream request without an X-Request-Id header");
        if (req.body.size() % sizeof(float)) throw std::runtime_error("octet-stream body is not a whole number of float32 values");
        const float* first = reinterpret_cast<const float*>(req.body.data());
        std::vector<float> data(first, first + req.body.size() / sizeof(float));
        res.set_content(worker.handleInferRaw(id->second, std::move(data)).dump(), "application/json");
        return;
      }
      if (fast) res.set_content(worker.handleInferBody(req.body).dump(), "application/json");
      else res.set_content(worker.handleInfer(json::parse(req.body)).dump(), "application/json");
    } catch (const std::exception& e) {
      json error;
      error["error"] = e.what();
      res.status = 500;
      res.set_content(error.dump(), "application/json");
    }
  });
  server.Get("/health", [&worker](const mini_http::Request&, mini_http::Response& res) {
    res.set_content(worker.getHealth().dump(), "application/json");
  });
  std::cout << "Worker Node: " << node_id << "  port " << port << "  (cache 1000 entries, batch 32, timeout 20ms)" << std::endl;
  std::cout << "Ready to accept requests!" << std::endl;
  if (!server.listen("0.0.0.0", port)) {
    std::cerr << "listen failed on port " << port << std::endl;
    return 1;
  }
  return 0;
}
