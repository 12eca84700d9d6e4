// Master: one thread per connection, all shared state under one mutex + condvar.
// See control.h for the parity map.
#include <limits.h>
#include <stdio.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <set>

#include "control.h"

namespace mp4j {

struct Master::Conn {
  int fd = -1;
  std::thread thr;
  std::atomic<bool> done{false};
};

namespace {
struct Finished {};  // thrown inside handlers when the job ended while they were blocked
const char* level_name(LogLevel l) {
  return l == LOG_DEBUG ? "DEBUG" : (l == LOG_ERROR ? "ERROR" : "INFO");
}
}  // namespace

Master::Master(int nslaves, int port, MasterOptions opt) : n_(nslaves), port_(port), opt_(opt) {
  if (nslaves <= 0) throw Error("slave number must be > 0");
  last_beat_.resize(n_);
  close_codes_.assign(n_, INT_MIN);
  ga_parts_.assign(n_, std::string());
  log_settings_ = load_log_settings("master");
  if (!opt_.log_dir.empty()) log_settings_.log_dir = opt_.log_dir;  // explicit argument wins over the file
  log_sink_.open(log_settings_, "master");
}

Master::~Master() {
  shutdown();
}

void Master::start() {
  {
    std::lock_guard<std::mutex> lk(mu_);
    if (started_) return;
    started_ = true;
    start_time_ = std::chrono::steady_clock::now();
  }
  listen_fd_ = net::tcp_listen(port_, &port_);
  accept_thr_ = std::thread([this, lfd = listen_fd_] { accept_loop(lfd); });
  watchdog_thr_ = std::thread([this] { watchdog_loop(); });
  log_line(LOG_INFO, "master started, slave num:" + std::to_string(n_) + ", port:" + std::to_string(port_));
}

void Master::log_line(LogLevel lvl, const std::string& msg) {
  std::string line = std::string("[") + level_name(lvl) + "] " + msg;
  std::lock_guard<std::mutex> lk(log_mu_);
  log_lines_.push_back(line);
  if (log_lines_.size() > std::max<size_t>(log_settings_.memory_lines, 1)) log_lines_.pop_front();
  if (!opt_.quiet && log_sink_.echo(lvl)) fprintf(stderr, "[mp4j-master] %s\n", line.c_str());
  log_sink_.write(lvl, msg);
}

std::vector<std::string> Master::logs() {
  std::lock_guard<std::mutex> lk(log_mu_);
  return std::vector<std::string>(log_lines_.begin(), log_lines_.end());
}

void Master::finish(int code) {
  std::lock_guard<std::mutex> lk(mu_);
  if (!finished_) {
    finished_ = true;
    final_code_ = code;
  }
  cv_.notify_all();
}

int Master::wait() {
  std::unique_lock<std::mutex> lk(mu_);
  cv_.wait(lk, [this] { return finished_; });
  return final_code_;
}

bool Master::wait_for(int timeout_ms, int* code) {
  std::unique_lock<std::mutex> lk(mu_);
  bool ok = cv_.wait_for(lk, std::chrono::milliseconds(timeout_ms), [this] { return finished_; });
  if (ok && code) *code = final_code_;
  return ok;
}

void Master::shutdown() {
  {
    std::lock_guard<std::mutex> lk(mu_);
    if (!started_) return;
    if (!finished_) {
      finished_ = true;
      // a master torn down before every slave closed reports failure, like a killed job
      final_code_ = (closed_cnt_ >= n_) ? final_code_ : 1;
    }
    cv_.notify_all();
  }
  if (listen_fd_ >= 0) net::shutdown_fd(listen_fd_);  // wakes the acceptor (it owns a copy of the number)
  if (accept_thr_.joinable()) accept_thr_.join();
  if (listen_fd_ >= 0) {
    net::close_fd(listen_fd_);
    listen_fd_ = -1;
  }
  if (watchdog_thr_.joinable()) watchdog_thr_.join();
  std::vector<std::shared_ptr<Conn>> conns;
  {
    std::lock_guard<std::mutex> lk(conns_mu_);
    conns.swap(conns_);
  }
  for (auto& c : conns) net::shutdown_fd(c->fd);
  for (auto& c : conns)
    if (c->thr.joinable()) c->thr.join();
  for (auto& c : conns) net::close_fd(c->fd);
}

void Master::accept_loop(int listen_fd) {
  for (;;) {
    int fd = net::tcp_accept(listen_fd);
    if (fd < 0) return;
    auto c = std::make_shared<Conn>();
    c->fd = fd;
    {
      std::lock_guard<std::mutex> lk(conns_mu_);
      // reap finished side connections so a long job does not accumulate threads
      for (size_t i = 0; i < conns_.size();) {
        if (conns_[i]->done.load()) {
          if (conns_[i]->thr.joinable()) conns_[i]->thr.join();
          net::close_fd(conns_[i]->fd);
          conns_.erase(conns_.begin() + i);
        } else ++i;
      }
      conns_.push_back(c);
    }
    c->thr = std::thread([this, c] { serve(c); });
  }
}

void Master::watchdog_loop() {
  std::unique_lock<std::mutex> lk(mu_);
  int first = std::min(opt_.watchdog_period_ms, 10000);
  cv_.wait_for(lk, std::chrono::milliseconds(first), [this] { return finished_; });
  while (!finished_) {
    auto now = std::chrono::steady_clock::now();
    if (!rendezvous_done_) {
      if (now - start_time_ > std::chrono::seconds(opt_.connect_timeout_s)) {
        lk.unlock();
        log_line(LOG_ERROR, "waiting for connecting check timeout > " +
                                std::to_string(opt_.connect_timeout_s) + "s, master will be shut down!");
        finish(2);
        return;
      }
    } else {
      for (int i = 0; i < n_; ++i) {
        if (close_codes_[i] != INT_MIN) continue;  // closed slaves stop beating
        if (now - last_beat_[i] > std::chrono::seconds(opt_.heartbeat_timeout_s)) {
          lk.unlock();
          log_line(LOG_ERROR, "waiting for heartbeat of rank " + std::to_string(i) + " timeout > " +
                                  std::to_string(opt_.heartbeat_timeout_s) + "s, master will be shut down!");
          finish(3);
          return;
        }
      }
    }
    cv_.wait_for(lk, std::chrono::milliseconds(opt_.watchdog_period_ms), [this] { return finished_; });
  }
}

void Master::serve(std::shared_ptr<Conn> c) {
  std::vector<char> req;
  uint8_t op;
  try {
    while (net::recv_frame(c->fd, &op, &req)) {
      net::Reader rd(req.data(), req.size());
      net::Writer wr;
      uint8_t st = ST_OK;
      try {
        switch (op) {
          case OP_RENDEZVOUS: h_rendezvous(rd, wr); break;
          case OP_BARRIER: h_barrier(); break;
          case OP_EXCHANGE: h_exchange(rd, wr); break;
          case OP_HEARTBEAT: {
            int r = rd.i32();
            std::lock_guard<std::mutex> lk(mu_);
            if (r >= 0 && r < n_) last_beat_[r] = std::chrono::steady_clock::now();
            break;
          }
          case OP_LOG: {
            LogLevel lvl = (LogLevel)rd.u8();
            log_line(lvl, rd.str());
            break;
          }
          case OP_CLOSE: h_close(rd); break;
          case OP_KILLME: h_killme(rd); break;
          case OP_WRITEFILE: h_writefile(rd); break;
          case OP_SHUTDOWN: {
            int code = rd.i32();
            log_line(LOG_INFO, rd.str());
            log_line(LOG_INFO, "shutdown(" + std::to_string(code) + ") invoked");
            finish(code);
            break;
          }
          case OP_GATHER_ALL: h_gather_all(rd, wr); break;
          default: throw Error("unknown opcode " + std::to_string(op));
        }
      } catch (const Finished&) {
        st = ST_ERR;
        wr = net::Writer();
        wr.str("master finished (job ended) while the request was pending");
      } catch (const Error& e) {
        st = ST_ERR;
        wr = net::Writer();
        wr.str(e.what());
      }
      net::send_frame(c->fd, st, wr.data());
    }
  } catch (const std::exception&) {
    // connection dropped: the heartbeat watchdog decides whether that is fatal
  }
  c->done.store(true);
}

void Master::h_rendezvous(net::Reader& rd, net::Writer& wr) {
  std::string addr = rd.str();
  int hint = rd.i32();
  std::unique_lock<std::mutex> lk(mu_);
  if (rendezvous_done_ || (int)pending_.size() >= n_)
    throw Error("task failed: all " + std::to_string(n_) + " slaves are already connected");  // SRV:173-176
  pending_.emplace_back(addr, hint);
  if ((int)pending_.size() == n_) {
    std::set<int> hints;
    bool use_hints = true;
    for (auto& p : pending_) {
      if (p.second < 0 || p.second >= n_ || !hints.insert(p.second).second) use_hints = false;
    }
    sorted_.assign(n_, std::string());
    if (use_hints) {
      for (auto& p : pending_) sorted_[p.second] = p.first;
    } else {  // reference rule: lexicographic order of "host###port" (SRV:188)
      for (int i = 0; i < n_; ++i) sorted_[i] = pending_[i].first;
      std::sort(sorted_.begin(), sorted_.end());
    }
    auto now = std::chrono::steady_clock::now();
    for (auto& t : last_beat_) t = now;
    rendezvous_done_ = true;
    cv_.notify_all();
  } else {
    cv_.wait(lk, [this] { return rendezvous_done_ || finished_; });
    if (!rendezvous_done_) throw Finished();
  }
  int rank = (int)(std::find(sorted_.begin(), sorted_.end(), addr) - sorted_.begin());
  if (rank >= n_) throw Error("rendezvous: address not found: " + addr);
  wr.i32(rank);
  wr.i32(n_);
  for (auto& s : sorted_) wr.str(s);
}

void Master::h_barrier() {
  std::unique_lock<std::mutex> lk(mu_);
  uint64_t gen = bar_gen_;
  if (++bar_count_ == n_) {
    bar_count_ = 0;
    ++bar_gen_;
    cv_.notify_all();
    return;
  }
  cv_.wait(lk, [&] { return bar_gen_ != gen || finished_; });
  if (bar_gen_ == gen) throw Finished();
}

void Master::h_exchange(net::Reader& rd, net::Writer& wr) {
  int rank = rd.i32();
  std::unique_lock<std::mutex> lk(mu_);
  if (exch_waiting_ < 0) {
    // park; each pairing gets its own slot so a later pairing cannot overwrite the answer
    auto slot = std::make_shared<int>(-1);
    exch_waiting_ = rank;
    exch_slot_ = slot;
    bool ok = cv_.wait_for(lk, std::chrono::seconds(opt_.exchange_timeout_s),
                           [&] { return *slot >= 0 || finished_; });
    if (!ok) {  // SRV:361-366
      lk.unlock();
      log_line(LOG_ERROR, "exchange timeout, master will be shut down!");
      finish(201);
      throw Finished();
    }
    if (*slot < 0) throw Finished();
    wr.i32(*slot);
  } else {
    int other = exch_waiting_;
    exch_waiting_ = -1;
    *exch_slot_ = rank;
    exch_slot_.reset();
    cv_.notify_all();
    wr.i32(other);
  }
}

void Master::h_close(net::Reader& rd) {
  int rank = rd.i32();
  int code = rd.i32();
  log_line(LOG_INFO, "recv close message from the slave rank:" + std::to_string(rank) +
                         ", code:" + std::to_string(code));
  bool all = false;
  {
    std::lock_guard<std::mutex> lk(mu_);
    if (rank >= 0 && rank < n_ && close_codes_[rank] == INT_MIN) {
      close_codes_[rank] = code;
      ++closed_cnt_;
    }
    if (code != 0) {
      if (!finished_) {
        finished_ = true;
        final_code_ = 1;
      }
      cv_.notify_all();
    } else if (closed_cnt_ >= n_) {
      all = true;
      if (!finished_) {
        finished_ = true;
        final_code_ = 0;
        for (int c : close_codes_)
          if (c != 0) final_code_ = 1;
      }
      cv_.notify_all();
    }
  }
  if (code != 0)
    log_line(LOG_INFO, "unnormally closed, rank=" + std::to_string(rank) + ", code:" + std::to_string(code));
  else if (all)
    log_line(LOG_INFO, "all slaves have sent close messages! server will be closed right now!");
}

void Master::h_killme(net::Reader& rd) {
  int rank = rd.i32();
  std::string cmd = rd.str();
  bool write = false;
  std::vector<std::string> cmds;
  {
    std::lock_guard<std::mutex> lk(mu_);
    if (kill_cmds_.empty()) kill_cmds_.assign(n_, std::string());
    if (rank >= 0 && rank < n_) kill_cmds_[rank] = cmd;
    if (++kill_cnt_ == n_) {
      write = true;
      cmds = kill_cmds_;
    }
  }
  if (write) {  // last arrival writes (the reference has rank 0 write after a barrier, SRV:327-348)
    std::string path = opt_.work_dir + "/kill_" + std::to_string(port_) + ".sh";
    FILE* f = fopen(path.c_str(), "w");
    if (!f) throw Error("cannot write " + path);
    for (auto& c : cmds) fprintf(f, "%s\n", c.c_str());
    fprintf(f, "kill -9 %d\n", (int)getpid());
    fclose(f);
  }
}

void Master::h_writefile(net::Reader& rd) {
  std::string content = rd.str();
  std::string fname = rd.str();
  std::string path = (fname.size() && fname[0] == '/') ? fname : opt_.work_dir + "/" + fname;
  FILE* f = fopen(path.c_str(), "w");
  if (!f) throw Error("cannot write " + path);
  fprintf(f, "%s\n", content.c_str());
  fclose(f);
}

void Master::h_gather_all(net::Reader& rd, net::Writer& wr) {
  int rank = rd.i32();
  std::string blob = rd.bytes();
  if (rank < 0 || rank >= n_) throw Error("gather_all: bad rank");
  std::shared_ptr<std::vector<std::string>> res;
  {
    std::unique_lock<std::mutex> lk(mu_);
    uint64_t gen = ga_gen_;
    ga_parts_[rank] = std::move(blob);
    if (++ga_count_ == n_) {
      ga_last_ = std::make_shared<std::vector<std::string>>(std::move(ga_parts_));
      ga_parts_.assign(n_, std::string());
      ga_count_ = 0;
      ++ga_gen_;
      cv_.notify_all();
    } else {
      cv_.wait(lk, [&] { return ga_gen_ != gen || finished_; });
      if (ga_gen_ == gen) throw Finished();
    }
    res = ga_last_;
  }
  // Rank order, not arrival order (fixes the slot hazard noted in SURVEY §5.2 for SRV:377).
  for (auto& s : *res) wr.bytes(s.data(), s.size());
}

}  // namespace mp4j
