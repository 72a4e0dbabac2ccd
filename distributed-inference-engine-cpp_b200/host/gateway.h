// Gateway: routes /infer by request_id over the consistent-hash ring, fails over across the other workers, one
// circuit breaker and one connection pool per worker, /stats.  Behaviour follows the reference's Gateway
// (src/gateway.cpp:12-163): ring of the worker addresses as given on the command line (150 virtual nodes each),
// breaker 5 failures / 2 successes / 30 s, 5 s connect and read timeouts, primary node first and then every other
// node in ring order, "No workers available" / "All workers failed or circuit breakers open" as the error texts,
// a worker counts as failed unless it answers 200 with a JSON body.
//
// Two deliberate differences, both invisible on the wire for well-formed traffic:
//  * the request body is forwarded as received instead of parse -> dump -> parse again on the worker
//    (gateway.cpp:101,107): only request_id is extracted, by a SAX pass that builds no tree (a 150,528-float body
//    is 1-3 MB); a worker's 200 body is passed through after validation (the reference re-dumps the parsed tree,
//    which for a nlohmann-produced body gives the same bytes);
//  * concurrent requests to one worker use separate pooled connections instead of one serialised client.
// A request without "request_id" is undefined behaviour in the reference (const operator[] on a missing key);
// here it is HTTP 500 with the type_error.302 text the worker gives for the same mistake.
#pragma once
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include <nlohmann/json.hpp>

#include "circuit_breaker.h"
#include "consistent_hash.h"
#include "fast_request.h"
#include "http_client.h"

class Gateway {
 public:
  using json = nlohmann::json;

  explicit Gateway(const std::vector<std::string>& workers,
                   CircuitBreaker::Clock::duration breaker_timeout = std::chrono::seconds(30), int io_timeout_s = 5) {
    for (const auto& w : workers) {
      ring_.addNode(w);
      breakers_[w] = std::make_unique<CircuitBreaker>(5, 2, breaker_timeout);
      auto hp = parseUrl(w);
      clients_[w] = std::make_unique<mini_http::ClientPool>(hp.first, hp.second, io_timeout_s);
    }
  }

  // host:port out of "http://host:port/path" | "host:port" | "host" (port 8080 when absent or unparsable,
  // gateway.cpp:128-156)
  static std::pair<std::string, int> parseUrl(const std::string& url) {
    std::string s = url;
    const size_t proto = s.find("://");
    if (proto != std::string::npos) s = s.substr(proto + 3);
    const size_t colon = s.find_last_of(':');
    if (colon == std::string::npos) return {s, 8080};
    std::string port = s.substr(colon + 1);
    const size_t slash = port.find('/');
    if (slash != std::string::npos) port = port.substr(0, slash);
    int p = 8080;
    try { p = std::stoi(port); } catch (...) {}
    return {s.substr(0, colon), p};
  }

  // The request_id of a request body, with the reference's error texts when it cannot be had
  // (json::parse errors; type errors of `std::string request_id = request["request_id"]`).
  static std::string extractRequestId(const std::string& body);

  // Forwards `body` and returns the answering worker's response body.
  std::string route(const std::string& body) { return routeAs(extractRequestId(body), body, "application/json", ""); }

  // The octet-stream side channel (raw float32 sample, id in X-Request-Id): same placement and fail-over, the
  // header travels with the body.
  std::string routeRaw(const std::string& request_id, const std::string& body) {
    return routeAs(request_id, body, "application/octet-stream", "X-Request-Id: " + request_id + "\r\n");
  }

  std::string routeAs(const std::string& request_id, const std::string& body, const std::string& content_type,
                      const std::string& extra_headers) {
    const std::string target = ring_.getNode(request_id);
    if (target.empty()) throw std::runtime_error("No workers available");
    if (auto r = tryNode(target, body, content_type, extra_headers)) return *r;
    for (const auto& node : ring_.getAllNodes())
      if (node != target)
        if (auto r = tryNode(node, body, content_type, extra_headers)) return *r;
    throw std::runtime_error("All workers failed or circuit breakers open");
  }

  json getStats() const {
    json stats;
    stats["total_workers"] = ring_.getAllNodes().size();
    json states = json::array();
    for (const auto& kv : breakers_) {
      json s;
      s["node"] = kv.first;
      s["state"] = kv.second->getStateString();
      s["failures"] = kv.second->getFailureCount();
      s["successes"] = kv.second->getSuccessCount();
      states.push_back(s);
    }
    stats["circuit_breakers"] = states;
    return stats;
  }

  std::string nodeFor(const std::string& request_id) const { return ring_.getNode(request_id); }

 private:
  std::optional<std::string> tryNode(const std::string& node, const std::string& body, const std::string& content_type,
                                     const std::string& extra_headers) {
    auto b = breakers_.find(node);
    if (b == breakers_.end()) return std::nullopt;
    if (!b->second->allowRequest()) return std::nullopt;
    auto c = clients_.find(node);
    if (c == clients_.end()) { b->second->recordFailure(); return std::nullopt; }
    mini_http::ClientResult r = c->second->Post("/infer", body, content_type, extra_headers);
    if (r.ok && r.status == 200 && json::accept(r.body)) {
      b->second->recordSuccess();
      return std::move(r.body);
    }
    b->second->recordFailure();
    return std::nullopt;
  }

  ConsistentHash ring_;
  std::map<std::string, std::unique_ptr<CircuitBreaker>> breakers_;
  std::map<std::string, std::unique_ptr<mini_http::ClientPool>> clients_;
};

// ---- request_id extraction -------------------------------------------------------------------------------
namespace gateway_detail {
using json = nlohmann::json;

// SAX handler: validates the whole document, remembers the value stored under the top-level key "request_id"
// (last occurrence wins, as in a parsed tree) and the type of the top-level value; builds nothing else.
struct IdSax {
  int depth = 0;
  bool top_is_object = false, top_seen = false, want_value = false, found = false;
  json top_scalar;          // a stand-in of the top-level value's type when it is not an object
  json id_value;            // the value under "request_id" (containers are recorded as empty ones: only the type matters)
  std::string error;

  bool scalar(json v) {
    if (depth == 0) { top_seen = true; top_scalar = std::move(v); }
    else if (depth == 1 && want_value) { id_value = std::move(v); found = true; want_value = false; }
    return true;
  }
  bool null() { return scalar(nullptr); }
  bool boolean(bool v) { return scalar(v); }
  bool number_integer(json::number_integer_t v) { return scalar(v); }
  bool number_unsigned(json::number_unsigned_t v) { return scalar(v); }
  bool number_float(json::number_float_t v, const std::string&) { return scalar(v); }
  bool string(std::string& v) { return scalar(v); }
  bool binary(json::binary_t&) { return scalar(json::binary({})); }
  bool open(json empty) {
    if (depth == 0) { top_seen = true; top_is_object = empty.is_object(); top_scalar = empty; }
    else if (depth == 1 && want_value) { id_value = std::move(empty); found = true; want_value = false; }
    ++depth;
    return true;
  }
  bool start_object(std::size_t) { return open(json::object()); }
  bool start_array(std::size_t) { return open(json::array()); }
  bool end_object() { --depth; return true; }
  bool end_array() { --depth; return true; }
  bool key(std::string& k) {
    if (depth == 1 && top_is_object) want_value = (k == "request_id");
    return true;
  }
  bool parse_error(std::size_t, const std::string&, const nlohmann::detail::exception& ex) {
    error = ex.what();
    return false;
  }
};
}  // namespace gateway_detail

inline std::string Gateway::extractRequestId(const std::string& body) {
  {
    // The one shape every client sends is validated and its id taken by the worker's request scanner (fast_request.h:
    // ~6x quicker than a SAX pass over a full-image body); any other body is not judged by it and takes the SAX
    // path below, so acceptance and error texts stay what they were.
    static thread_local std::vector<float> scratch;
    std::string id;
    if (fast_request::parse(body, id, scratch)) return id;
  }
  gateway_detail::IdSax sax;
  if (!json::sax_parse(body, &sax)) throw std::runtime_error(sax.error.empty() ? "parse error" : sax.error);
  if (!sax.top_is_object) {
    // what `request["request_id"]` does on a const non-object: type_error.305 with the value's type name
    const json& top = sax.top_scalar;
    (void)top["request_id"];
  }
  const json value = sax.found ? sax.id_value : json(nullptr);
  return value.get<std::string>();      // type_error.302 "type must be string, but is ..." unless it is a string
}
