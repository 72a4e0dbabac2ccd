#include "http_client.h"

#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <cstring>

namespace mini_http {
namespace {

bool send_all(int fd, const char* p, size_t n) {
  while (n) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k <= 0) {
      if (k < 0 && errno == EINTR) continue;
      return false;
    }
    p += k; n -= size_t(k);
  }
  return true;
}

// Reads one response (status line, headers, Content-Length body).  `got_bytes` says whether the request may NOT be
// retried: a keep-alive connection the peer closed while idle fails with EOF / reset before the first byte (false,
// retry once on a fresh socket); bytes received, or a read timeout, are final (true).
bool read_response(int fd, ClientResult& r, bool& got_bytes, bool& keep_alive) {
  std::string buf;
  size_t header_end = std::string::npos;
  char tmp[65536];
  got_bytes = false;
  while (header_end == std::string::npos) {
    ssize_t k = ::recv(fd, tmp, sizeof tmp, 0);
    if (k < 0 && errno == EINTR) continue;
    if (k <= 0) {
      const bool timed_out = k < 0 && (errno == EAGAIN || errno == EWOULDBLOCK);
      r.error = k == 0 ? "connection closed by worker" : (timed_out ? "read timeout" : std::strerror(errno));
      // a read timeout is a final answer (the worker HAS the request: sending it again would run the inference twice
      // and delay the breaker's failure count); only EOF / reset before the first byte marks a stale pooled socket
      if (timed_out) got_bytes = true;
      return false;
    }
    got_bytes = true;
    buf.append(tmp, size_t(k));
    header_end = buf.find("\r\n\r\n");
    if (header_end == std::string::npos && buf.size() > (1u << 20)) { r.error = "response headers too large"; return false; }
  }
  if (buf.compare(0, 5, "HTTP/") != 0) { r.error = "malformed status line"; return false; }
  size_t sp = buf.find(' ');
  r.status = sp == std::string::npos ? 0 : std::atoi(buf.c_str() + sp + 1);
  std::string head = buf.substr(0, header_end);
  std::transform(head.begin(), head.end(), head.begin(), [](unsigned char c) { return char(std::tolower(c)); });
  size_t want = 0;
  size_t cl = head.find("\r\ncontent-length:");
  if (cl != std::string::npos) want = size_t(std::strtoull(head.c_str() + cl + 17, nullptr, 10));
  keep_alive = head.find("connection: close") == std::string::npos;
  r.body = buf.substr(header_end + 4);
  while (r.body.size() < want) {
    ssize_t k = ::recv(fd, tmp, std::min(sizeof tmp, want - r.body.size()), 0);
    if (k < 0 && errno == EINTR) continue;
    if (k <= 0) { r.error = k == 0 ? "connection closed mid-body" : "read timeout"; return false; }
    r.body.append(tmp, size_t(k));
  }
  return true;
}

}  // namespace

ClientPool::~ClientPool() {
  for (int fd : idle_) ::close(fd);
}

int ClientPool::take() {
  std::lock_guard<std::mutex> g(mu_);
  if (idle_.empty()) return -1;
  int fd = idle_.back();
  idle_.pop_back();
  return fd;
}

void ClientPool::give(int fd) {
  std::lock_guard<std::mutex> g(mu_);
  if (idle_.size() < 64) idle_.push_back(fd);
  else ::close(fd);
}

int ClientPool::connect_new(std::string& err) const {
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_UNSPEC;
  hints.ai_socktype = SOCK_STREAM;
  const std::string port = std::to_string(port_);
  if (int rc = ::getaddrinfo(host_.c_str(), port.c_str(), &hints, &res)) { err = std::string("resolve: ") + gai_strerror(rc); return -1; }
  int fd = -1;
  for (addrinfo* a = res; a; a = a->ai_next) {
    fd = ::socket(a->ai_family, a->ai_socktype, a->ai_protocol);
    if (fd < 0) continue;
    // connect with a timeout: non-blocking connect + poll
    int fl = ::fcntl(fd, F_GETFL, 0);
    ::fcntl(fd, F_SETFL, fl | O_NONBLOCK);
    int rc = ::connect(fd, a->ai_addr, a->ai_addrlen);
    if (rc < 0 && errno == EINPROGRESS) {
      pollfd pf{fd, POLLOUT, 0};
      rc = ::poll(&pf, 1, timeout_s_ * 1000);
      int soerr = 0; socklen_t sl = sizeof soerr;
      if (rc == 1 && ::getsockopt(fd, SOL_SOCKET, SO_ERROR, &soerr, &sl) == 0 && soerr == 0) rc = 0;
      else { if (rc == 0) err = "connect timeout"; else err = std::string("connect: ") + std::strerror(soerr ? soerr : errno); rc = -1; }
    } else if (rc < 0) {
      err = std::string("connect: ") + std::strerror(errno);
    }
    if (rc == 0) {
      ::fcntl(fd, F_SETFL, fl);
      timeval tv{timeout_s_, 0};
      ::setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
      ::setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof tv);
      int one = 1;
      ::setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
      break;
    }
    ::close(fd);
    fd = -1;
  }
  ::freeaddrinfo(res);
  if (fd < 0 && err.empty()) err = "connect failed";
  return fd;
}

ClientResult ClientPool::request(const std::string& method, const std::string& path, const std::string& body,
                                 const std::string& content_type, const std::string& extra_headers) {
  std::string head = method + " " + path + " HTTP/1.1\r\nHost: " + host_ + ":" + std::to_string(port_) + "\r\nConnection: keep-alive\r\n";
  if (method == "POST") head += "Content-Type: " + content_type + "\r\nContent-Length: " + std::to_string(body.size()) + "\r\n";
  head += extra_headers;
  head += "\r\n";
  for (int attempt = 0; attempt < 2; ++attempt) {
    ClientResult r;
    int fd = attempt == 0 ? take() : -1;
    const bool pooled = fd >= 0;
    if (fd < 0) fd = connect_new(r.error);
    if (fd < 0) return r;
    bool got = false, keep = false;
    const bool sent = send_all(fd, head.data(), head.size()) && (body.empty() || send_all(fd, body.data(), body.size()));
    if (sent && read_response(fd, r, got, keep)) {
      r.ok = true;
      if (keep) give(fd); else ::close(fd);
      return r;
    }
    if (!sent && r.error.empty()) r.error = "send failed";
    ::close(fd);
    if (!(pooled && !got)) return r;     // only a stale pooled connection (nothing received) is worth a second try
  }
  ClientResult r;
  r.error = "request failed";
  return r;
}

ClientResult ClientPool::Post(const std::string& path, const std::string& body, const std::string& content_type,
                              const std::string& extra_headers) {
  return request("POST", path, body, content_type, extra_headers);
}
ClientResult ClientPool::Get(const std::string& path) { return request("GET", path, "", "", ""); }

}  // namespace mini_http
