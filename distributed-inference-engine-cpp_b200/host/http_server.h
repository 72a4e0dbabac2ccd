// Minimal HTTP/1.1 server for the worker's two routes (POST /infer, GET /health) and the gateway's.
//
// The reference uses cpp-httplib (un-vendored; not on this image): `server.Post("/infer", ...)`,
// `server.Get("/health", ...)`, `server.listen("0.0.0.0", port)` (src/worker_node.cpp:173-203).  This is a small
// stand-in with the same shape: keep-alive, Content-Length bodies of any size (a 150,528-float JSON body is 1-3 MB),
// no chunked encoding, no TLS.  Handlers get the body and fill status / content / content type, like httplib's
// Request/Response.
//
// Threads bound concurrent REQUESTS, not connections: one thread runs an epoll loop over the listening socket and
// every idle keep-alive connection; a connection with bytes to read is handed to a pool thread for exactly one
// request (plus any already-pipelined ones) and goes back to the epoll set afterwards.  An idle connection therefore
// holds no thread (a gateway's pooled sockets or 50 keep-alive clients cannot starve a health probe), a peer that
// stalls mid-request is dropped after `read_timeout_s`, and connections idle for `idle_timeout_s` are closed
// (cpp-httplib recycles them after 5 s).
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_set>
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
  // threads: requests served at once (0 = max(64, 2 x hardware threads): handlers block in BatchProcessor::process,
  // and a batch can only be as large as the number of requests waiting there)
  explicit Server(int threads = 0);
  ~Server();
  void Post(const std::string& path, Handler h) { routes_["POST " + path] = std::move(h); }
  void Get(const std::string& path, Handler h) { routes_["GET " + path] = std::move(h); }
  bool listen(const std::string& host, int port);   // blocks until stop()
  void stop();
  int bound_port() const { return port_; }
  void set_timeouts(int read_timeout_s, int idle_timeout_s) { read_timeout_s_ = read_timeout_s; idle_timeout_s_ = idle_timeout_s; }

 private:
  struct Conn;
  void worker_loop();
  bool serve_one(Conn* c);          // false: close the connection
  void park(Conn* c);               // back to the epoll set
  void drop(Conn* c);
  std::map<std::string, Handler> routes_;
  std::atomic<bool> running_{false};
  int listen_fd_ = -1, epoll_fd_ = -1, wake_fd_ = -1, port_ = 0, max_threads_;
  int read_timeout_s_ = 5, idle_timeout_s_ = 30;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<Conn*> ready_;
  std::unordered_set<Conn*> idle_;
  std::vector<std::thread> pool_;
};

}  // namespace mini_http
