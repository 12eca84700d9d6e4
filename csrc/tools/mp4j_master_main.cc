// mp4j-master: stand-alone master process (parity: CommMaster.main, J/comm/CommMaster.java:106-132).
// Usage: mp4j-master <slave_num> <port>
// Exit code: 0 iff every slave closed with code 0; 1 otherwise; 2/3/201 for watchdog timeouts.
#include <stdio.h>
#include <stdlib.h>

#include "../control/control.h"

int main(int argc, char** argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: %s <slave_num> <port> [log_dir]\n", argv[0]);
    return 64;
  }
  int code = 1;
  try {
    mp4j::MasterOptions opt;
    opt.log_dir = argc > 3 ? argv[3] : "log";
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
