#include "http_server.h"

#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <cctype>
#include <cstring>

namespace mini_http {
namespace {

bool send_all(int fd, const char* p, size_t n) {
  while (n) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k <= 0) return false;
    p += k;
    n -= size_t(k);
  }
  return true;
}

const char* reason(int status) {
  switch (status) {
    case 200: return "OK";
    case 400: return "Bad Request";
    case 404: return "Not Found";
    case 413: return "Payload Too Large";
    case 500: return "Internal Server Error";
    default: return "Status";
  }
}

}  // namespace

Server::Server(int threads) {
  unsigned hc = std::thread::hardware_concurrency();
  // cpp-httplib v0.14.3's default pool is max(8, hardware_concurrency - 1); keep the same bound on concurrency
  max_threads_ = threads > 0 ? threads : int(std::max(8u, hc > 0 ? hc - 1 : 8u));
}

Server::~Server() { stop(); }

void Server::stop() {
  running_ = false;
  if (listen_fd_ >= 0) {
    ::shutdown(listen_fd_, SHUT_RDWR);
    ::close(listen_fd_);
    listen_fd_ = -1;
  }
}

bool Server::listen(const std::string& host, int port) {
  listen_fd_ = ::socket(AF_INET, SOCK_STREAM, 0);
  if (listen_fd_ < 0) return false;
  int one = 1;
  ::setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(uint16_t(port));
  if (host.empty() || host == "0.0.0.0") addr.sin_addr.s_addr = htonl(INADDR_ANY);
  else if (::inet_pton(AF_INET, host.c_str(), &addr.sin_addr) != 1) return false;
  if (::bind(listen_fd_, reinterpret_cast<sockaddr*>(&addr), sizeof addr) != 0) return false;
  if (::listen(listen_fd_, 512) != 0) return false;
  socklen_t len = sizeof addr;
  ::getsockname(listen_fd_, reinterpret_cast<sockaddr*>(&addr), &len);
  port_ = ntohs(addr.sin_port);
  running_ = true;
  while (running_) {
    int fd = ::accept(listen_fd_, nullptr, nullptr);
    if (fd < 0) {
      if (!running_) break;
      continue;
    }
    ::setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    while (running_ && active_.load() >= max_threads_) std::this_thread::sleep_for(std::chrono::microseconds(200));
    ++active_;
    std::thread([this, fd] {
      serve(fd);
      ::close(fd);
      --active_;
    }).detach();
  }
  while (active_.load() > 0) std::this_thread::sleep_for(std::chrono::milliseconds(1));
  return true;
}

void Server::serve(int fd) {
  std::string buf;
  buf.reserve(1 << 16);
  char tmp[1 << 16];
  for (;;) {
    // ---- header block ----
    size_t hdr_end;
    while ((hdr_end = buf.find("\r\n\r\n")) == std::string::npos) {
      ssize_t k = ::recv(fd, tmp, sizeof tmp, 0);
      if (k <= 0) return;
      buf.append(tmp, size_t(k));
      if (buf.size() > (1 << 20) && buf.find("\r\n\r\n") == std::string::npos) return;   // header flood
    }
    Request req;
    {
      size_t line_end = buf.find("\r\n");
      std::string line = buf.substr(0, line_end);
      size_t a = line.find(' '), b = line.rfind(' ');
      if (a == std::string::npos || b == a) return;
      req.method = line.substr(0, a);
      req.path = line.substr(a + 1, b - a - 1);
      size_t q = req.path.find('?');
      if (q != std::string::npos) req.path.resize(q);
      size_t pos = line_end + 2;
      while (pos < hdr_end) {
        size_t e = buf.find("\r\n", pos);
        std::string h = buf.substr(pos, e - pos);
        size_t c = h.find(':');
        if (c != std::string::npos) {
          std::string name = h.substr(0, c), val = h.substr(c + 1);
          std::transform(name.begin(), name.end(), name.begin(), [](unsigned char ch) { return char(std::tolower(ch)); });
          size_t s = val.find_first_not_of(" \t");
          req.headers[name] = s == std::string::npos ? "" : val.substr(s);
        }
        pos = e + 2;
      }
    }
    size_t need = 0;
    auto cl = req.headers.find("content-length");
    if (cl != req.headers.end()) need = size_t(std::strtoull(cl->second.c_str(), nullptr, 10));
    if (need > (size_t(1) << 30)) return;
    auto ex = req.headers.find("expect");
    if (ex != req.headers.end() && ex->second.find("100-continue") != std::string::npos)
      send_all(fd, "HTTP/1.1 100 Continue\r\n\r\n", 25);
    // ---- body ----
    size_t body_at = hdr_end + 4;
    while (buf.size() - body_at < need) {
      ssize_t k = ::recv(fd, tmp, sizeof tmp, 0);
      if (k <= 0) return;
      buf.append(tmp, size_t(k));
    }
    req.body.assign(buf, body_at, need);
    buf.erase(0, body_at + need);

    Response res;
    auto it = routes_.find(req.method + " " + req.path);
    if (it == routes_.end()) {
      res.status = 404;
      res.set_content("Not Found", "text/plain");
    } else {
      try {
        it->second(req, res);
      } catch (const std::exception& e) {
        res.status = 500;
        res.set_content(e.what(), "text/plain");
      }
    }
    bool close_conn = false;
    auto cn = req.headers.find("connection");
    if (cn != req.headers.end()) {
      std::string v = cn->second;
      std::transform(v.begin(), v.end(), v.begin(), [](unsigned char ch) { return char(std::tolower(ch)); });
      close_conn = v.find("close") != std::string::npos;
    }
    std::string head = "HTTP/1.1 " + std::to_string(res.status) + " " + reason(res.status) + "\r\nContent-Type: " + res.content_type +
                       "\r\nContent-Length: " + std::to_string(res.body.size()) + "\r\nConnection: " +
                       (close_conn ? "close" : "keep-alive") + "\r\n\r\n";
    if (!send_all(fd, head.data(), head.size()) || !send_all(fd, res.body.data(), res.body.size())) return;
    if (close_conn || !running_) return;
  }
}

}  // namespace mini_http
