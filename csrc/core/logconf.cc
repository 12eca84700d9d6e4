#include "logconf.h"

#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <time.h>

#include <fstream>

namespace mp4j {

namespace {
std::string trim(const std::string& s) {
  size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
  return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
}
int level_of(const std::string& name, int dflt) {
  if (name == "DEBUG" || name == "debug") return 0;
  if (name == "INFO" || name == "info") return 1;
  if (name == "ERROR" || name == "error" || name == "WARN" || name == "warn") return 2;
  return dflt;
}
const char* kLevelName[3] = {"DEBUG", " INFO", "ERROR"};
}  // namespace

std::map<std::string, std::string> parse_properties(const std::string& path) {
  std::map<std::string, std::string> kv;
  std::ifstream in(path);
  std::string line;
  while (std::getline(in, line)) {
    line = trim(line);
    if (line.empty() || line[0] == '#' || line[0] == '!') continue;
    size_t eq = line.find('=');
    if (eq == std::string::npos) continue;
    kv[trim(line.substr(0, eq))] = trim(line.substr(eq + 1));
  }
  return kv;
}

LogSettings load_log_settings(const char* role) {
  LogSettings s;
  // Only an explicitly named file is read: a library must not start writing into the working directory
  // of whoever links it. The stand-alone master and the launcher name ./config/logging.properties when present.
  const char* env = getenv("MP4J_LOG_CONFIG");
  if (!env || !*env) return s;
  std::string path = env;
  struct stat st;
  if (stat(path.c_str(), &st) != 0) return s;
  auto kv = parse_properties(path);
  const std::string r = std::string(role) + ".";
  auto get = [&](const char* k) -> const std::string* {
    auto it = kv.find(r + k);
    return it == kv.end() ? nullptr : &it->second;
  };
  if (auto v = get("log_dir")) s.log_dir = *v;
  if (auto v = get("file_level")) s.file_level = level_of(*v, s.file_level);
  if (auto v = get("error_file")) s.error_file = !(*v == "false" || *v == "0" || *v == "no");
  if (auto v = get("rolling")) s.rolling_daily = (*v != "none");
  if (auto v = get("memory_lines")) s.memory_lines = (size_t)strtoull(v->c_str(), nullptr, 10);
  if (auto v = get("echo_levels")) {
    s.echo[0] = s.echo[1] = s.echo[2] = false;
    size_t pos = 0;
    while (pos <= v->size()) {
      size_t c = v->find(',', pos);
      if (c == std::string::npos) c = v->size();
      int l = level_of(trim(v->substr(pos, c - pos)), -1);
      if (l >= 0) s.echo[l] = true;
      pos = c + 1;
    }
  }
  return s;
}

LogSink::~LogSink() {
  if (all_) fclose(all_);
  if (err_) fclose(err_);
}

void LogSink::open(const LogSettings& s, const std::string& base) {
  std::lock_guard<std::mutex> lk(mu_);
  s_ = s;
  base_ = base;
  dir_ = s.log_dir;
  if (dir_.empty()) return;
  mkdir(dir_.c_str(), 0755);
  all_ = fopen((dir_ + "/" + base_ + ".log").c_str(), "a");
  if (s_.error_file) err_ = fopen((dir_ + "/" + base_ + "_error.log").c_str(), "a");
  if (!all_) dir_.clear();
}

// DailyRollingFileAppender: when the date changes, yesterday's file becomes <file>-YYYY-MM-DD
void LogSink::roll_if_needed(const char* today) {
  if (day_.empty()) {
    day_ = today;
    return;
  }
  if (!s_.rolling_daily || day_ == today) return;
  const std::string a = dir_ + "/" + base_ + ".log", e = dir_ + "/" + base_ + "_error.log";
  if (all_) {
    fclose(all_);
    rename(a.c_str(), (a + "-" + day_).c_str());
    all_ = fopen(a.c_str(), "a");
  }
  if (err_) {
    fclose(err_);
    rename(e.c_str(), (e + "-" + day_).c_str());
    err_ = fopen(e.c_str(), "a");
  }
  day_ = today;
}

void LogSink::write(int level, const std::string& msg) {
  if (dir_.empty() || level < 0 || level > 2) return;
  time_t now = time(nullptr);
  struct tm tmv;
  localtime_r(&now, &tmv);
  char day[16], stamp[32];
  strftime(day, sizeof day, "%Y-%m-%d", &tmv);
  strftime(stamp, sizeof stamp, "%Y-%m-%d %H:%M:%S", &tmv);
  std::lock_guard<std::mutex> lk(mu_);
  roll_if_needed(day);
  if (all_ && level >= s_.file_level) {
    fprintf(all_, "%s %s - %s\n", stamp, kLevelName[level], msg.c_str());
    fflush(all_);
  }
  if (err_ && level == 2) {
    fprintf(err_, "%s %s - %s\n", stamp, kLevelName[level], msg.c_str());
    fflush(err_);
  }
}

}  // namespace mp4j
