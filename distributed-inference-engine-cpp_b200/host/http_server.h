// Minimal HTTP/1.1 server for the worker's two routes (POST /infer, GET /health).
//
// The reference uses cpp-httplib (un-vendored; not on this image): `server.Post("/infer", ...)`,
// `server.Get("/health", ...)`, `server.listen("0.0.0.0", port)` (src/worker_node.cpp:173-203).  This is a small
// stand-in with the same shape: a blocking accept loop, one thread per connection out of a fixed pool,
// keep-alive, Content-Length bodies of any size (a 150,528-float JSON body is 1-3 MB), no chunked encoding,
// no TLS.  Handlers get the body and fill status / content / content type, like httplib's Request/Response.
#pragma once
#include <atomic>
#include <functional>
#include <map>
#include <string>
#include <thread>
#include <vector>

namespace mini_http {

struct Request {
  std::string method, path, body;
  std::map<std::string, std::string> headers;   // lower-cased names
};

struct Response {
  int status = 200;
  std::string body, content_type = "text/plain";
  void set_content(const std::string& b, const std::string& type) { body = b; content_type = type; }
};

class Server {
 public:
  using Handler = std::function<void(const Request&, Response&)>;
  explicit Server(int threads = 0);
  ~Server();
  void Post(const std::string& path, Handler h) { routes_["POST " + path] = std::move(h); }
  void Get(const std::string& path, Handler h) { routes_["GET " + path] = std::move(h); }
  bool listen(const std::string& host, int port);   // blocks until stop()
  void stop();
  int bound_port() const { return port_; }

 private:
  void serve(int fd);
  std::map<std::string, Handler> routes_;
  std::atomic<bool> running_{false};
  std::atomic<int> active_{0};
  int listen_fd_ = -1, port_ = 0, max_threads_;
};

}  // namespace mini_http
