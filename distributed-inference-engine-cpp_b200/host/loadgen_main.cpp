// loadgen: closed-loop HTTP load on a worker or the gateway, in the manner of the reference's benchmark.py
// (threads x requests, request_id "req_{i}", percentiles by sorted index int(n * p): benchmark.py:49-52,113-116) but
// native, so that the client is not what bounds the measurement (the Python generator saturates one core's GIL at
// ~600 full-image requests/s).  One keep-alive connection per thread; prints one JSON object.
//   usage: loadgen <host> <port> <threads> <requests> <payload> [floats]
//   payload: small  = {"request_id":"req_i","input_data":[i%10, i%10+1, i%10+2]}     (benchmark.py's body: cache path)
//            json   = a full sample as JSON text, first value unique per request      (inference path, ~2 MB bodies)
//            octet  = the same sample as raw float32, id in X-Request-Id              (additive side channel)
//   floats: sample length for json / octet (default 150528 = 3 x 224 x 224)
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "http_client.h"

int main(int argc, char** argv) {
  if (argc < 6) {
    std::fprintf(stderr, "usage: %s <host> <port> <threads> <requests> small|json|octet [floats]\n", argv[0]);
    return 1;
  }
  const std::string host = argv[1];
  const int port = std::atoi(argv[2]), threads = std::max(1, std::atoi(argv[3]));
  const long total = std::max(1L, std::atol(argv[4]));
  const std::string kind = argv[5];
  const size_t floats = argc > 6 ? size_t(std::atol(argv[6])) : 150528;
  if (kind != "small" && kind != "json" && kind != "octet") { std::fprintf(stderr, "unknown payload '%s'\n", kind.c_str()); return 1; }
  const long per = total / threads;

  // one pseudo-random sample shared by all requests; element 0 is replaced per request so that no two keys are equal
  std::vector<float> base(floats);
  uint32_t s = 12345u;
  for (auto& v : base) { s = s * 1664525u + 1013904223u; v = float(int32_t(s >> 8) % 2000) / 1000.0f - 1.0f; }
  std::string json_tail;                      // ",v1,v2,...]}" of the json payload, built once
  if (kind == "json") {
    json_tail.reserve(floats * 8);
    char buf[32];
    for (size_t i = 1; i < floats; ++i) { int n = std::snprintf(buf, sizeof buf, ",%.3f", double(base[i])); json_tail.append(buf, size_t(n)); }
    json_tail += "]}";
  }

  std::vector<std::vector<double>> lat{size_t(threads)};
  std::atomic<long> failed{0};
  size_t bytes_one = 0;
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> ts;
  for (int t = 0; t < threads; ++t)
    ts.emplace_back([&, t] {
      mini_http::ClientPool pool(host, port, 30);
      std::vector<float> mine(kind == "octet" ? base : std::vector<float>());
      std::string body;
      lat[size_t(t)].reserve(size_t(per));
      for (long k = 0; k < per; ++k) {
        const long i = long(t) * per + k;
        std::string ctype = "application/json", extra;
        char head[128];
        if (kind == "small") {
          std::snprintf(head, sizeof head, "{\"request_id\":\"req_%ld\",\"input_data\":[%ld.0,%ld.0,%ld.0]}", i, i % 10, i % 10 + 1, i % 10 + 2);
          body = head;
        } else if (kind == "json") {
          std::snprintf(head, sizeof head, "{\"request_id\":\"req_%ld\",\"input_data\":[%ld.5", i, i);
          body.assign(head);
          body += json_tail;
        } else {
          mine[0] = float(i) + 0.25f;
          body.assign(reinterpret_cast<const char*>(mine.data()), mine.size() * sizeof(float));
          ctype = "application/octet-stream";
          std::snprintf(head, sizeof head, "X-Request-Id: raw_%ld\r\n", i);
          extra = head;
        }
        if (t == 0 && k == 0) bytes_one = body.size();
        const auto a = std::chrono::steady_clock::now();
        mini_http::ClientResult r = pool.Post("/infer", body, ctype, extra);
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - a).count();
        if (r.ok && r.status == 200) lat[size_t(t)].push_back(ms);
        else ++failed;
      }
    });
  for (auto& th : ts) th.join();
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::vector<double> all;
  for (auto& v : lat) all.insert(all.end(), v.begin(), v.end());
  std::sort(all.begin(), all.end());
  const size_t n = all.size();
  auto pct = [&](double p) { return n ? all[std::min(n - 1, size_t(double(n) * p))] : 0.0; };
  double mean = 0;
  for (double v : all) mean += v;
  std::printf("{\"payload\":\"%s\",\"threads\":%d,\"requests\":%ld,\"ok\":%zu,\"failed\":%ld,\"seconds\":%.3f,\"req_per_s\":%.1f,"
              "\"p50_ms\":%.3f,\"p90_ms\":%.3f,\"p99_ms\":%.3f,\"mean_ms\":%.3f,\"body_bytes\":%zu}\n",
              kind.c_str(), threads, per * threads, n, failed.load(), secs, n / secs, pct(0.50), pct(0.90), pct(0.99), n ? mean / n : 0.0,
              bytes_one);
  return 0;
}
