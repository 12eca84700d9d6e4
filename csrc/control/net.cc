#include "net.h"

#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/types.h>
#include <unistd.h>

#include <chrono>
#include <thread>

namespace mp4j {
namespace net {

static void set_common_opts(int fd) {
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  int sz = 4 << 20;
  setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &sz, sizeof(sz));
  setsockopt(fd, SOL_SOCKET, SO_RCVBUF, &sz, sizeof(sz));
}

int tcp_listen(int port, int* bound_port, int backlog) {
  int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) throw Error(std::string("socket: ") + strerror(errno));
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_addr.s_addr = htonl(INADDR_ANY);
  addr.sin_port = htons((uint16_t)port);
  if (::bind(fd, (sockaddr*)&addr, sizeof(addr)) < 0) {
    int e = errno;
    ::close(fd);
    throw Error("bind port " + std::to_string(port) + ": " + strerror(e));
  }
  if (::listen(fd, backlog) < 0) {
    int e = errno;
    ::close(fd);
    throw Error(std::string("listen: ") + strerror(e));
  }
  socklen_t len = sizeof(addr);
  getsockname(fd, (sockaddr*)&addr, &len);
  if (bound_port) *bound_port = ntohs(addr.sin_port);
  return fd;
}

int tcp_connect(const std::string& host, int port, int retries, int delay_ms) {
  std::string last = "unknown";
  for (int attempt = 0; attempt < retries; ++attempt) {
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    int rc = getaddrinfo(host.c_str(), std::to_string(port).c_str(), &hints, &res);
    if (rc != 0) {
      last = gai_strerror(rc);
    } else {
      for (addrinfo* ai = res; ai; ai = ai->ai_next) {
        int fd = ::socket(ai->ai_family, ai->ai_socktype, ai->ai_protocol);
        if (fd < 0) continue;
        if (::connect(fd, ai->ai_addr, ai->ai_addrlen) == 0) {
          freeaddrinfo(res);
          set_common_opts(fd);
          return fd;
        }
        last = strerror(errno);
        ::close(fd);
      }
      freeaddrinfo(res);
    }
    std::this_thread::sleep_for(std::chrono::milliseconds(delay_ms));
  }
  throw Error("connect " + host + ":" + std::to_string(port) + " failed: " + last);
}

int tcp_accept(int listen_fd) {
  for (;;) {
    int fd = ::accept(listen_fd, nullptr, nullptr);
    if (fd >= 0) {
      set_common_opts(fd);
      return fd;
    }
    if (errno == EINTR) continue;
    return -1;
  }
}

void close_fd(int fd) {
  if (fd >= 0) ::close(fd);
}
void shutdown_fd(int fd) {
  if (fd >= 0) ::shutdown(fd, SHUT_RDWR);
}

void send_all(int fd, const void* buf, size_t n) {
  const char* p = (const char*)buf;
  while (n > 0) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k < 0) {
      if (errno == EINTR) continue;
      throw Error(std::string("send: ") + strerror(errno));
    }
    p += k;
    n -= (size_t)k;
  }
}

bool recv_all(int fd, void* buf, size_t n) {
  char* p = (char*)buf;
  size_t got = 0;
  while (got < n) {
    ssize_t k = ::recv(fd, p + got, n - got, 0);
    if (k == 0) {
      if (got == 0) return false;
      throw Error("connection closed mid-frame");
    }
    if (k < 0) {
      if (errno == EINTR) continue;
      throw Error(std::string("recv: ") + strerror(errno));
    }
    got += (size_t)k;
  }
  return true;
}

std::string local_hostname() {
  char buf[256];
  if (gethostname(buf, sizeof(buf)) != 0) return "127.0.0.1";
  buf[sizeof(buf) - 1] = 0;
  // container hostnames frequently do not resolve; peers then could not connect back
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  if (getaddrinfo(buf, nullptr, &hints, &res) != 0 || !res) return "127.0.0.1";
  freeaddrinfo(res);
  return buf;
}

void sendrecv(int send_fd, const void* sbuf, size_t slen, int recv_fd, void* rbuf, size_t rlen) {
  const char* sp = (const char*)sbuf;
  char* rp = (char*)rbuf;
  size_t sent = 0, got = 0;
  while (sent < slen || got < rlen) {
    pollfd fds[2];
    int nf = 0, si = -1, ri = -1;
    if (sent < slen) { fds[nf] = {send_fd, POLLOUT, 0}; si = nf++; }
    if (got < rlen) { fds[nf] = {recv_fd, POLLIN, 0}; ri = nf++; }
    int rc = ::poll(fds, nf, 1000);
    if (rc < 0) {
      if (errno == EINTR) continue;
      throw Error(std::string("poll: ") + strerror(errno));
    }
    if (si >= 0 && (fds[si].revents & (POLLOUT | POLLERR | POLLHUP))) {
      ssize_t k = ::send(send_fd, sp + sent, slen - sent, MSG_NOSIGNAL | MSG_DONTWAIT);
      if (k < 0) {
        if (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR)
          throw Error(std::string("send: ") + strerror(errno));
      } else sent += (size_t)k;
    }
    if (ri >= 0 && (fds[ri].revents & (POLLIN | POLLERR | POLLHUP))) {
      ssize_t k = ::recv(recv_fd, rp + got, rlen - got, MSG_DONTWAIT);
      if (k == 0) throw Error("peer closed during sendrecv");
      if (k < 0) {
        if (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR)
          throw Error(std::string("recv: ") + strerror(errno));
      } else got += (size_t)k;
    }
  }
}

void send_frame(int fd, uint8_t tag, const void* payload, size_t n) {
  char hdr[5];
  uint32_t len = (uint32_t)n;
  memcpy(hdr, &len, 4);
  hdr[4] = (char)tag;
  if (n <= 1024) {  // one syscall for small frames
    char tmp[1024 + 5];
    memcpy(tmp, hdr, 5);
    if (n) memcpy(tmp + 5, payload, n);
    send_all(fd, tmp, n + 5);
  } else {
    send_all(fd, hdr, 5);
    send_all(fd, payload, n);
  }
}

void send_frame(int fd, uint8_t tag, const std::vector<char>& payload) {
  send_frame(fd, tag, payload.data(), payload.size());
}

bool recv_frame(int fd, uint8_t* tag, std::vector<char>* payload) {
  char hdr[5];
  if (!recv_all(fd, hdr, 5)) return false;
  uint32_t len;
  memcpy(&len, hdr, 4);
  *tag = (uint8_t)hdr[4];
  payload->resize(len);
  if (len && !recv_all(fd, payload->data(), len)) throw Error("connection closed mid-frame");
  return true;
}

}  // namespace net
}  // namespace mp4j
