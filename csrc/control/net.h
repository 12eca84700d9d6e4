// Small blocking-socket toolkit used by the control plane and the host data plane.
// Replaces the reference's Hadoop IPC + per-message java.net sockets (PCS:151,391-425)
// with persistent connections carrying length-prefixed binary frames.
#pragma once
#include <stdint.h>
#include <string.h>
#include <string>
#include <vector>
#include "../core/error.h"

namespace mp4j {
namespace net {

// Listens on `port` (0 = ephemeral) on all interfaces; returns fd, stores the bound port.
int tcp_listen(int port, int* bound_port, int backlog = 128);
// Connects with retry (parity with the 50 x 30 ms retry of PCS:62-63,395-414).
int tcp_connect(const std::string& host, int port, int retries = 50, int delay_ms = 30);
int tcp_accept(int listen_fd);  // returns -1 when the listener was closed
void close_fd(int fd);
void shutdown_fd(int fd);
void send_all(int fd, const void* buf, size_t n);
// Returns false on orderly EOF at a frame boundary (n bytes not yet started), throws on error.
bool recv_all(int fd, void* buf, size_t n);
std::string local_hostname();

// Full-duplex exchange on two (possibly identical) sockets without deadlock:
// sends `slen` bytes to send_fd while receiving `rlen` bytes from recv_fd (poll-driven).
void sendrecv(int send_fd, const void* sbuf, size_t slen, int recv_fd, void* rbuf, size_t rlen);

class Writer {
 public:
  void u8(uint8_t v) { buf_.push_back((char)v); }
  void i32(int32_t v) { raw(&v, 4); }
  void i64(int64_t v) { raw(&v, 8); }
  void str(const std::string& s) { i32((int32_t)s.size()); raw(s.data(), s.size()); }
  void bytes(const void* p, size_t n) { i64((int64_t)n); raw(p, n); }
  void raw(const void* p, size_t n) { buf_.insert(buf_.end(), (const char*)p, (const char*)p + n); }
  const std::vector<char>& data() const { return buf_; }
 private:
  std::vector<char> buf_;
};

class Reader {
 public:
  Reader(const char* p, size_t n) : p_(p), end_(p + n) {}
  uint8_t u8() { need(1); return (uint8_t)*p_++; }
  int32_t i32() { int32_t v; need(4); memcpy(&v, p_, 4); p_ += 4; return v; }
  int64_t i64() { int64_t v; need(8); memcpy(&v, p_, 8); p_ += 8; return v; }
  std::string str() { int32_t n = i32(); need(n); std::string s(p_, p_ + n); p_ += n; return s; }
  std::string bytes() { int64_t n = i64(); need(n); std::string s(p_, p_ + n); p_ += n; return s; }
  size_t left() const { return end_ - p_; }
 private:
  void need(size_t n) { if ((size_t)(end_ - p_) < n) throw Error("malformed frame"); }
  const char* p_;
  const char* end_;
};

// Frame = u32 payload_len | u8 tag | payload. tag is an opcode (request) or status (response).
void send_frame(int fd, uint8_t tag, const std::vector<char>& payload);
void send_frame(int fd, uint8_t tag, const void* payload, size_t n);
// Returns false on EOF before a frame starts.
bool recv_frame(int fd, uint8_t* tag, std::vector<char>* payload);

}  // namespace net
}  // namespace mp4j
