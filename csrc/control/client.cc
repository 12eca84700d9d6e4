// MasterClient: synchronous request/response proxy to the Master (parity: the IServer proxy
// obtained at PCS:151 and used throughout ProcessCommSlave).
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>

#include "control.h"

namespace mp4j {

MasterClient::MasterClient(const std::string& host, int port) : host_(host), port_(port) {
  fd_ = net::tcp_connect(host, port, 200, 50);
}

MasterClient::~MasterClient() {
  stop_heartbeat();
  net::close_fd(fd_);
}

void MasterClient::call(int fd, uint8_t op, const net::Writer& req, std::vector<char>* resp) {
  net::send_frame(fd, op, req.data());
  uint8_t st;
  std::vector<char> buf;
  if (!net::recv_frame(fd, &st, &buf)) throw Error("master closed the connection");
  if (st != ST_OK) {
    net::Reader rd(buf.data(), buf.size());
    throw Error("master: " + rd.str());
  }
  if (resp) resp->swap(buf);
}

void MasterClient::call_main(uint8_t op, const net::Writer& req, std::vector<char>* resp) {
  std::lock_guard<std::mutex> lk(mu_);
  call(fd_, op, req, resp);
}

void MasterClient::call_side(uint8_t op, const net::Writer& req, std::vector<char>* resp) {
  std::unique_lock<std::mutex> lk(mu_, std::try_to_lock);
  if (lk.owns_lock()) {
    call(fd_, op, req, resp);
    return;
  }
  int fd = net::tcp_connect(host_, port_, 20, 50);
  try {
    call(fd, op, req, resp);
  } catch (...) {
    net::close_fd(fd);
    throw;
  }
  net::close_fd(fd);
}

int MasterClient::rendezvous(const std::string& my_addr, int rank_hint, std::vector<std::string>* all) {
  net::Writer w;
  w.str(my_addr);
  w.i32(rank_hint);
  std::vector<char> resp;
  call_main(OP_RENDEZVOUS, w, &resp);
  net::Reader rd(resp.data(), resp.size());
  int rank = rd.i32();
  int n = rd.i32();
  all->clear();
  for (int i = 0; i < n; ++i) all->push_back(rd.str());
  return rank;
}

void MasterClient::barrier() {
  net::Writer w;
  call_main(OP_BARRIER, w, nullptr);
}

int MasterClient::exchange(int rank) {
  net::Writer w;
  w.i32(rank);
  std::vector<char> resp;
  call_main(OP_EXCHANGE, w, &resp);
  net::Reader rd(resp.data(), resp.size());
  return rd.i32();
}

void MasterClient::open_local_log(int rank) {
  LogSettings ls = load_log_settings("slave");
  const char* dir = getenv("MP4J_SLAVE_LOG_DIR");
  if (dir && *dir) ls.log_dir = dir;
  if (!ls.log_dir.empty()) local_log_.open(ls, "slave_" + std::to_string(rank));
}

void MasterClient::log(LogLevel lvl, const std::string& msg) {
  if (local_log_.active()) {
    local_log_.write(lvl, msg);
    if (local_log_.echo(lvl)) fprintf(stderr, "%s\n", msg.c_str());
  }
  net::Writer w;
  w.u8(lvl);
  w.str(msg);
  call_side(OP_LOG, w, nullptr);
}

void MasterClient::close(int rank, int code) {
  net::Writer w;
  w.i32(rank);
  w.i32(code);
  call_side(OP_CLOSE, w, nullptr);
}

void MasterClient::kill_me(int rank, const std::string& cmd) {
  net::Writer w;
  w.i32(rank);
  w.str(cmd);
  call_main(OP_KILLME, w, nullptr);
}

void MasterClient::write_file(const std::string& content, const std::string& fname) {
  net::Writer w;
  w.str(content);
  w.str(fname);
  call_main(OP_WRITEFILE, w, nullptr);
}

void MasterClient::shutdown_master() {
  net::Writer w;
  w.i32(0);
  w.str("shutdown requested by a slave");
  call_side(OP_SHUTDOWN, w, nullptr);
}

std::vector<std::string> MasterClient::gather_all(int rank, const void* data, size_t n) {
  net::Writer w;
  w.i32(rank);
  w.bytes(data, n);
  std::vector<char> resp;
  call_main(OP_GATHER_ALL, w, &resp);
  net::Reader rd(resp.data(), resp.size());
  std::vector<std::string> out;
  while (rd.left() > 0) out.push_back(rd.bytes());
  return out;
}

void MasterClient::start_heartbeat(int rank, int first_ms, int period_ms, int max_fail,
                                   void (*on_dead)(void*), void* arg) {
  stop_heartbeat();
  hb_stop_ = false;
  hb_thr_ = std::thread([=] {
    int fd = -1;
    int fails = 0;
    int wait_ms = first_ms;
    for (;;) {
      {
        std::unique_lock<std::mutex> lk(hb_mu_);
        if (hb_cv_.wait_for(lk, std::chrono::milliseconds(wait_ms), [this] { return hb_stop_; })) break;
      }
      wait_ms = period_ms;
      try {
        if (fd < 0) fd = net::tcp_connect(host_, port_, 3, 50);
        net::Writer w;
        w.i32(rank);
        call(fd, OP_HEARTBEAT, w, nullptr);
        fails = 0;
      } catch (const std::exception&) {
        net::close_fd(fd);
        fd = -1;
        if (++fails > max_fail) {  // PCS:219-224
          if (on_dead) on_dead(arg);
          else _exit(4);
          break;
        }
      }
    }
    net::close_fd(fd);
  });
}

void MasterClient::stop_heartbeat() {
  {
    std::lock_guard<std::mutex> lk(hb_mu_);
    hb_stop_ = true;
  }
  hb_cv_.notify_all();
  if (hb_thr_.joinable()) hb_thr_.join();
}

}  // namespace mp4j
