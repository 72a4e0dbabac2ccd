#include "http_server.h"

#include <arpa/inet.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/epoll.h>
#include <sys/eventfd.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <unistd.h>

#include <algorithm>
#include <cctype>
#include <chrono>
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

struct Server::Conn {
  int fd = -1;
  std::string buf;                                    // bytes read beyond the last request (pipelining)
  std::chrono::steady_clock::time_point parked;       // when it last went idle
};

Server::Server(int threads) {
  unsigned hc = std::thread::hardware_concurrency();
  max_threads_ = threads > 0 ? threads : int(std::max(64u, 2 * hc));
}

Server::~Server() { stop(); }

void Server::stop() {
  running_ = false;
  if (wake_fd_ >= 0) {
    uint64_t one = 1;
    ssize_t k = ::write(wake_fd_, &one, sizeof one);
    (void)k;
  }
  cv_.notify_all();
}

void Server::drop(Conn* c) {
  ::close(c->fd);
  delete c;
}

void Server::park(Conn* c) {
  c->parked = std::chrono::steady_clock::now();
  {
    std::lock_guard<std::mutex> lk(mu_);
    idle_.insert(c);
  }
  epoll_event ev{};
  ev.events = EPOLLIN | EPOLLRDHUP | EPOLLONESHOT;
  ev.data.ptr = c;
  if (::epoll_ctl(epoll_fd_, EPOLL_CTL_MOD, c->fd, &ev) != 0 && ::epoll_ctl(epoll_fd_, EPOLL_CTL_ADD, c->fd, &ev) != 0) {
    std::lock_guard<std::mutex> lk(mu_);
    if (idle_.erase(c)) drop(c);
  }
}

void Server::worker_loop() {
  for (;;) {
    Conn* c = nullptr;
    {
      std::unique_lock<std::mutex> lk(mu_);
      cv_.wait(lk, [&] { return !ready_.empty() || !running_; });
      if (ready_.empty()) return;
      c = ready_.front();
      ready_.pop_front();
    }
    bool keep = true;
    do {
      keep = serve_one(c);
    } while (keep && running_ && c->buf.find("\r\n\r\n") != std::string::npos);   // requests already in the buffer
    if (keep && running_) park(c);
    else drop(c);
  }
}

bool Server::listen(const std::string& host, int port) {
  listen_fd_ = ::socket(AF_INET, SOCK_STREAM | SOCK_NONBLOCK, 0);
  if (listen_fd_ < 0) return false;
  int one = 1;
  ::setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(uint16_t(port));
  if (host.empty() || host == "0.0.0.0") addr.sin_addr.s_addr = htonl(INADDR_ANY);
  else if (::inet_pton(AF_INET, host.c_str(), &addr.sin_addr) != 1) return false;
  if (::bind(listen_fd_, reinterpret_cast<sockaddr*>(&addr), sizeof addr) != 0) return false;
  if (::listen(listen_fd_, 1024) != 0) return false;
  socklen_t len = sizeof addr;
  ::getsockname(listen_fd_, reinterpret_cast<sockaddr*>(&addr), &len);
  port_ = ntohs(addr.sin_port);
  epoll_fd_ = ::epoll_create1(0);
  wake_fd_ = ::eventfd(0, EFD_NONBLOCK);
  if (epoll_fd_ < 0 || wake_fd_ < 0) return false;
  epoll_event ev{};
  ev.events = EPOLLIN;
  ev.data.ptr = nullptr;                    // the listening socket
  ::epoll_ctl(epoll_fd_, EPOLL_CTL_ADD, listen_fd_, &ev);
  ev.data.ptr = &wake_fd_;
  ::epoll_ctl(epoll_fd_, EPOLL_CTL_ADD, wake_fd_, &ev);
  running_ = true;
  for (int i = 0; i < max_threads_; ++i) pool_.emplace_back([this] { worker_loop(); });

  auto last_sweep = std::chrono::steady_clock::now();
  std::vector<epoll_event> evs(256);
  while (running_) {
    int n = ::epoll_wait(epoll_fd_, evs.data(), int(evs.size()), 500);
    for (int i = 0; i < n && running_; ++i) {
      void* tag = evs[size_t(i)].data.ptr;
      if (tag == &wake_fd_) continue;
      if (tag == nullptr) {
        for (;;) {                            // accept everything that is waiting
          int fd = ::accept4(listen_fd_, nullptr, nullptr, 0);
          if (fd < 0) break;
          ::setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
          timeval tv{read_timeout_s_, 0};     // a peer that stalls mid-request does not pin a thread for ever
          ::setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
          ::setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof tv);
          Conn* c = new Conn();
          c->fd = fd;
          park(c);
        }
        continue;
      }
      Conn* c = static_cast<Conn*>(tag);      // readable (or closed by the peer): one pool thread takes it
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (!idle_.erase(c)) continue;        // already swept
        ready_.push_back(c);
      }
      cv_.notify_one();
    }
    const auto now = std::chrono::steady_clock::now();
    if (now - last_sweep > std::chrono::seconds(1)) {
      last_sweep = now;
      std::vector<Conn*> dead;
      {
        std::lock_guard<std::mutex> lk(mu_);
        for (auto it = idle_.begin(); it != idle_.end();)
          if (now - (*it)->parked > std::chrono::seconds(idle_timeout_s_)) { dead.push_back(*it); it = idle_.erase(it); }
          else ++it;
      }
      for (Conn* c : dead) {
        ::epoll_ctl(epoll_fd_, EPOLL_CTL_DEL, c->fd, nullptr);
        drop(c);
      }
    }
  }
  cv_.notify_all();
  for (auto& t : pool_) t.join();
  pool_.clear();
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (Conn* c : idle_) drop(c);
    idle_.clear();
    for (Conn* c : ready_) drop(c);
    ready_.clear();
  }
  ::close(listen_fd_); listen_fd_ = -1;
  ::close(epoll_fd_); epoll_fd_ = -1;
  ::close(wake_fd_); wake_fd_ = -1;
  return true;
}

// One request of connection c: read (bounded by the socket's receive timeout), dispatch, answer.
bool Server::serve_one(Conn* c) {
  const int fd = c->fd;
  std::string& buf = c->buf;
  char tmp[1 << 16];
  {
    // ---- header block ----
    size_t hdr_end;
    while ((hdr_end = buf.find("\r\n\r\n")) == std::string::npos) {
      ssize_t k = ::recv(fd, tmp, sizeof tmp, 0);
      if (k <= 0) return false;
      buf.append(tmp, size_t(k));
      if (buf.size() > (1 << 20) && buf.find("\r\n\r\n") == std::string::npos) return false;   // header flood
    }
    Request req;
    {
      size_t line_end = buf.find("\r\n");
      std::string line = buf.substr(0, line_end);
      size_t a = line.find(' '), b = line.rfind(' ');
      if (a == std::string::npos || b == a) return false;
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
    if (need > (size_t(1) << 30)) return false;
    auto ex = req.headers.find("expect");
    if (ex != req.headers.end() && ex->second.find("100-continue") != std::string::npos)
      send_all(fd, "HTTP/1.1 100 Continue\r\n\r\n", 25);
    // ---- body ----
    size_t body_at = hdr_end + 4;
    while (buf.size() - body_at < need) {
      ssize_t k = ::recv(fd, tmp, sizeof tmp, 0);
      if (k <= 0) return false;
      buf.append(tmp, size_t(k));
    }
    // the body is handed over without a second buffer: the connection's buffer becomes the request's body (bytes of a
    // pipelined next request, rare, are kept aside) and its storage goes back to the connection after the handler
    std::string rest;
    if (buf.size() > body_at + need) rest.assign(buf, body_at + need, std::string::npos);
    buf.resize(body_at + need);
    buf.erase(0, body_at);
    req.body.swap(buf);
    buf = std::move(rest);

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
    if (buf.empty()) {                       // keep the large allocation with the connection for its next request
      req.body.clear();
      buf.swap(req.body);
    }
    if (!send_all(fd, head.data(), head.size()) || !send_all(fd, res.body.data(), res.body.size())) return false;
    return !close_conn;
  }
}

}  // namespace mini_http
