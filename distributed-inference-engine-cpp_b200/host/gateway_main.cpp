// gateway: the reference's L4 router (src/gateway.cpp:165-200) over this repository's ring, breaker and HTTP pieces.
//   usage: gateway <worker1:port> [worker2:port] ...
// POST /infer -> routed to a worker (HTTP 500 {"error": what} on any failure), GET /stats.  Listens on port 8000
// like the reference; GATEWAY_PORT overrides it, GATEWAY_BREAKER_TIMEOUT_MS and GATEWAY_IO_TIMEOUT_S shorten the
// 30 s / 5 s constants (tests).
#include <cstdlib>
#include <iostream>

#include "gateway.h"
#include "http_server.h"

using json = nlohmann::json;

int main(int argc, char** argv) {
  if (argc < 2) {
    std::cerr << "Usage: " << argv[0] << " <worker1:port> [worker2:port] ..." << std::endl;
    std::cerr << "Example: " << argv[0] << " localhost:8001 localhost:8002 localhost:8003" << std::endl;
    return 1;
  }
  std::vector<std::string> workers(argv + 1, argv + argc);
  int port = 8000;
  if (const char* p = std::getenv("GATEWAY_PORT")) port = std::atoi(p);
  CircuitBreaker::Clock::duration breaker_timeout = std::chrono::seconds(30);
  if (const char* t = std::getenv("GATEWAY_BREAKER_TIMEOUT_MS")) breaker_timeout = std::chrono::milliseconds(std::atol(t));
  int io_timeout = 5;
  if (const char* t = std::getenv("GATEWAY_IO_TIMEOUT_S")) io_timeout = std::max(1, std::atoi(t));

  Gateway gateway(workers, breaker_timeout, io_timeout);
  mini_http::Server server;
  server.Post("/infer", [&gateway](const mini_http::Request& req, mini_http::Response& res) {
    try {
      auto ct = req.headers.find("content-type");
      if (ct != req.headers.end() && ct->second.compare(0, 24, "application/octet-stream") == 0) {
        auto id = req.headers.find("x-request-id");
        if (id == req.headers.end()) throw std::runtime_error("octet-stream request without an X-Request-Id header");
        if (id->second.find_first_of("\r\n") != std::string::npos) throw std::runtime_error("bad X-Request-Id header");
        res.set_content(gateway.routeRaw(id->second, req.body), "application/json");
        return;
      }
      res.set_content(gateway.route(req.body), "application/json");
    } catch (const std::exception& e) {
      json error;
      error["error"] = e.what();
      res.status = 500;
      res.set_content(error.dump(), "application/json");
    }
  });
  server.Get("/stats", [&gateway](const mini_http::Request&, mini_http::Response& res) {
    res.set_content(gateway.getStats().dump(), "application/json");
  });
  std::cout << "Gateway listening on port " << port << std::endl;
  std::cout << "Workers: " << workers.size() << std::endl;
  std::cout << "Circuit breakers enabled" << std::endl;
  std::cout << "Ready!" << std::endl;
  if (!server.listen("0.0.0.0", port)) {
    std::cerr << "Error: cannot listen on port " << port << std::endl;
    return 1;
  }
  return 0;
}
