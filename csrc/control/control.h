// Control plane: Master (CommMaster + rpc/Server equivalent) and MasterClient (the slave's
// proxy). Parity map:
//   rendezvous + sorted-address rank assignment   SRV:167-222
//   barrier                                        SRV:131
//   exchange (pairwise matchmaker)                 SRV:358-370
//   heartbeat + watchdog, exit codes 2/3/201       SRV:124,140-164,361-366
//   info/debug/error log sink                      SRV:225-239
//   close code aggregation / stop                  SRV:243-289
//   killMe kill-script writer                      SRV:319-355
//   writeFile                                      SRV:302-316
//   RPC allreduce mailbox (as rank-ordered gather) SRV:373-514
// New duty: the same gather mailbox brokers GPU memory handles at communicator set-up.
#pragma once
#include "../core/logconf.h"
#include <stdint.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "net.h"

namespace mp4j {

enum Opcode : uint8_t {
  OP_RENDEZVOUS = 1,
  OP_BARRIER = 2,
  OP_EXCHANGE = 3,
  OP_HEARTBEAT = 4,
  OP_LOG = 5,
  OP_CLOSE = 6,
  OP_KILLME = 7,
  OP_WRITEFILE = 8,
  OP_SHUTDOWN = 9,
  OP_GATHER_ALL = 10,
};
enum LogLevel : uint8_t { LOG_DEBUG = 0, LOG_INFO = 1, LOG_ERROR = 2 };
enum Status : uint8_t { ST_OK = 0, ST_ERR = 1 };

constexpr const char* kAddrDelim = "###";  // PCS:61

struct MasterOptions {
  int connect_timeout_s = 7200;   // SRV:82 -> exit 2
  int heartbeat_timeout_s = 600;  // SRV:83 -> exit 3
  int exchange_timeout_s = 3600;  // SRV:52 -> exit 201
  int watchdog_period_ms = 30000; // SRV:108-109 (first check after 10 s)
  std::string log_dir;            // "" = stderr only; else <dir>/master.log
  std::string work_dir = ".";     // where kill_<port>.sh and writeFile() outputs go
  bool quiet = false;             // suppress stderr echo (tests)
};

class Master {
 public:
  Master(int nslaves, int port, MasterOptions opt = MasterOptions());
  ~Master();
  void start();            // bind + serve in background threads
  int port() const { return port_; }
  // Blocks until every slave closed (or a failure ended the job); returns the job exit code:
  // 0 iff all slaves closed with 0 (J/comm/CommMaster.java:127-130).
  int wait();
  bool wait_for(int timeout_ms, int* code);
  void shutdown();         // stop serving, wake everyone
  std::vector<std::string> logs();  // in-memory copy of everything the sink received

 private:
  struct Conn;
  void accept_loop(int listen_fd);
  void serve(std::shared_ptr<Conn> c);
  void watchdog_loop();
  void finish(int code);
  void log_line(LogLevel lvl, const std::string& msg);
  // handlers
  void h_rendezvous(net::Reader& rd, net::Writer& wr);
  void h_barrier();
  void h_exchange(net::Reader& rd, net::Writer& wr);
  void h_close(net::Reader& rd);
  void h_killme(net::Reader& rd);
  void h_writefile(net::Reader& rd);
  void h_gather_all(net::Reader& rd, net::Writer& wr);

  const int n_;
  int port_;
  MasterOptions opt_;
  int listen_fd_ = -1;
  std::thread accept_thr_, watchdog_thr_;
  std::mutex conns_mu_;
  std::vector<std::shared_ptr<Conn>> conns_;

  std::mutex mu_;
  std::condition_variable cv_;
  bool started_ = false;
  bool finished_ = false;
  int final_code_ = 0;
  std::chrono::steady_clock::time_point start_time_;
  // rendezvous
  std::vector<std::pair<std::string, int>> pending_;  // (address, rank hint)
  std::vector<std::string> sorted_;
  bool rendezvous_done_ = false;
  // barrier
  int bar_count_ = 0;
  uint64_t bar_gen_ = 0;
  // exchange
  int exch_waiting_ = -1;      // rank parked in the matchmaker, or -1
  std::shared_ptr<int> exch_slot_;  // answer slot of the parked rank
  // heartbeat
  std::vector<std::chrono::steady_clock::time_point> last_beat_;
  // close
  std::vector<int> close_codes_;  // INT32_MIN = not closed
  int closed_cnt_ = 0;
  // killMe
  std::vector<std::string> kill_cmds_;
  int kill_cnt_ = 0;
  // gather-all mailbox
  std::vector<std::string> ga_parts_;
  int ga_count_ = 0;
  uint64_t ga_gen_ = 0;
  std::shared_ptr<std::vector<std::string>> ga_last_;
  // logs
  std::mutex log_mu_;
  std::deque<std::string> log_lines_;
  LogSettings log_settings_;
  LogSink log_sink_;  // <log_dir>/master.log (+ master_error.log), thresholds and rolling from the properties file
};

// Slave-side proxy. One mutex-guarded request connection + a dedicated heartbeat connection
// (a blocking barrier must not starve the heartbeat). Log/close requests that find the main
// connection busy go out on a short-lived side connection so an error report is never stuck
// behind a collective that will not complete.
class MasterClient {
 public:
  MasterClient(const std::string& host, int port);
  ~MasterClient();
  // Returns this slave's rank; fills the rank-ordered "host###port" list.
  int rendezvous(const std::string& my_addr, int rank_hint, std::vector<std::string>* all);
  void barrier();
  int exchange(int rank);
  void log(LogLevel lvl, const std::string& msg);
  // slave-side log files (slave_<rank>.log, slave_<rank>_error.log): settings "slave.*" of the properties
  // file, directory overridable with $MP4J_SLAVE_LOG_DIR. Called once the rank is known.
  void open_local_log(int rank);
  void close(int rank, int code);
  void kill_me(int rank, const std::string& cmd);
  void write_file(const std::string& content, const std::string& fname);
  void shutdown_master();
  std::vector<std::string> gather_all(int rank, const void* data, size_t n);
  // Heartbeat thread: first beat after first_ms, then every period_ms (PCS:199-200).
  // After more than `max_fail` consecutive failures calls on_dead (default: _exit(4), PCS:219-224).
  void start_heartbeat(int rank, int first_ms, int period_ms, int max_fail, void (*on_dead)(void*),
                       void* arg);
  void stop_heartbeat();

 private:
  void call(int fd, uint8_t op, const net::Writer& req, std::vector<char>* resp);
  void call_main(uint8_t op, const net::Writer& req, std::vector<char>* resp);
  void call_side(uint8_t op, const net::Writer& req, std::vector<char>* resp);
  std::string host_;
  int port_;
  int fd_ = -1;
  std::mutex mu_;
  std::thread hb_thr_;
  std::mutex hb_mu_;
  std::condition_variable hb_cv_;
  bool hb_stop_ = false;
  LogSink local_log_;
};

}  // namespace mp4j
