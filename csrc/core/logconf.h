// File logging of master and slaves, configured like the reference's log4j property files
// (/root/reference/config/log4j_master.properties:15-35, log4j_slave.properties:10-36): a threshold
// per sink, one file for everything at or above the file level plus one for errors only, appended,
// rolled over when the date changes (DailyRollingFileAppender: <file>-YYYY-MM-DD), and a console echo.
//
// Source of the settings: the file named by $MP4J_LOG_CONFIG (the stand-alone master binary and mp4j-launch set
// it to ./config/logging.properties when that file exists). Keys (role = master | slave):
//   <role>.log_dir        directory of the files ("" or absent: no files)
//   <role>.file_level     DEBUG | INFO | ERROR   threshold of <role>.log          (default DEBUG)
//   <role>.error_file     true | false           also write <role>_error.log      (default true)
//   <role>.echo_levels    comma list of levels echoed on stderr                   (default INFO,ERROR)
//   <role>.rolling        daily | none                                            (default daily)
//   master.memory_lines   lines kept in memory for CommMaster.logs()              (default 100000)
// Explicit arguments (CommMaster(log_dir=...), MP4J_SLAVE_LOG_DIR) override the file.
#pragma once
#include <stdio.h>

#include <map>
#include <mutex>
#include <string>

namespace mp4j {

struct LogSettings {
  std::string log_dir;
  int file_level = 0;          // LOG_DEBUG
  bool error_file = true;
  bool echo[3] = {false, true, true};
  bool rolling_daily = true;
  size_t memory_lines = 100000;
};

// Parses the property file (missing file = defaults). role: "master" or "slave".
LogSettings load_log_settings(const char* role);
std::map<std::string, std::string> parse_properties(const std::string& path);

class LogSink {
 public:
  LogSink() {}
  ~LogSink();
  // base: file name stem inside settings.log_dir, e.g. "master" or "slave_3"
  void open(const LogSettings& s, const std::string& base);
  bool active() const { return !dir_.empty(); }
  bool echo(int level) const { return level >= 0 && level < 3 && s_.echo[level]; }
  // "2026-11-07 04:15:01  INFO - msg" into <base>.log (level >= threshold) and <base>_error.log (errors)
  void write(int level, const std::string& msg);

 private:
  void roll_if_needed(const char* today);
  LogSettings s_;
  std::string dir_, base_, day_;
  FILE* all_ = nullptr;
  FILE* err_ = nullptr;
  std::mutex mu_;
};

}  // namespace mp4j
