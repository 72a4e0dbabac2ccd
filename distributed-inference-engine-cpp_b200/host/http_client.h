// Minimal blocking HTTP/1.1 client for the gateway -> worker hop (the reference uses httplib::Client with 5 s
// connect / read timeouts, src/gateway.cpp:29-33).  One ClientPool per worker keeps idle keep-alive connections
// and hands one to each request, so concurrent gateway requests to the same worker are not serialised on a
// single socket ("per-worker connection pool instead of one serialised client").
#pragma once
#include <mutex>
#include <string>
#include <vector>

namespace mini_http {

struct ClientResult {
  bool ok = false;          // a complete response was received
  int status = 0;
  std::string body;
  std::string error;        // when !ok: what went wrong (connect / send / timeout / malformed response)
};

class ClientPool {
 public:
  ClientPool(std::string host, int port, int timeout_s = 5) : host_(std::move(host)), port_(port), timeout_s_(timeout_s) {}
  ~ClientPool();
  ClientPool(const ClientPool&) = delete;
  ClientPool& operator=(const ClientPool&) = delete;
  // POST `body` to `path`; retries once on a fresh connection if a pooled keep-alive connection turns out dead.
  // extra_headers: complete "Name: value\r\n" lines appended to the request head
  ClientResult Post(const std::string& path, const std::string& body, const std::string& content_type,
                    const std::string& extra_headers = "");
  ClientResult Get(const std::string& path);

 private:
  ClientResult request(const std::string& method, const std::string& path, const std::string& body, const std::string& content_type,
                       const std::string& extra_headers);
  int take();               // an idle connection or -1
  void give(int fd);
  int connect_new(std::string& err) const;
  std::string host_;
  int port_, timeout_s_;
  std::mutex mu_;
  std::vector<int> idle_;
};

}  // namespace mini_http
