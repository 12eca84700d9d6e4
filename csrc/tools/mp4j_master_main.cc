// mp4j-master: stand-alone master process (parity: CommMaster.main, J/comm/CommMaster.java:106-132).
// Usage: mp4j-master <slave_num> <port>
// Exit code: 0 iff every slave closed with code 0; 1 otherwise; 2/3/201 for watchdog timeouts.
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>

#include "../control/control.h"

int main(int argc, char** argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: %s <slave_num> <port> [log_dir]\n", argv[0]);
    return 64;
  }
  int code = 1;
  try {
    // like the reference's master, which finds config/log4j_master.properties next to its start script
    if (!getenv("MP4J_LOG_CONFIG") && access("config/logging.properties", R_OK) == 0)
      setenv("MP4J_LOG_CONFIG", "config/logging.properties", 0);
    mp4j::MasterOptions opt;
    if (argc > 3) opt.log_dir = argv[3];
    else if (!getenv("MP4J_LOG_CONFIG")) opt.log_dir = "log";
    if (getenv("MP4J_CONNECT_TIMEOUT_S")) opt.connect_timeout_s = atoi(getenv("MP4J_CONNECT_TIMEOUT_S"));
    if (getenv("MP4J_HEARTBEAT_TIMEOUT_S")) opt.heartbeat_timeout_s = atoi(getenv("MP4J_HEARTBEAT_TIMEOUT_S"));
    mp4j::Master master(atoi(argv[1]), atoi(argv[2]), opt);
    master.start();
    code = master.wait();
    master.shutdown();
  } catch (const std::exception& e) {
    fprintf(stderr, "mp4j-master: %s\n", e.what());
    return 1;
  }
  fprintf(stderr, "mp4j-master: exit code %d\n", code);
  return code;
}
