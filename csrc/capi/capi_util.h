// Shared plumbing of the extern "C" layer: error capture and the communicator handle.
#pragma once
#include <exception>
#include <string>

namespace mp4j {
class ProcessComm;
class ThreadComm;
extern thread_local std::string g_last_error;
}  // namespace mp4j

struct mp4j_comm {
  mp4j::ProcessComm* pc = nullptr;
  void* gpu = nullptr;                      // mp4j::GpuPlane*, owned; set by mp4j_gpu_init
  void (*gpu_dtor)(mp4j_comm*) = nullptr;   // installed by the GPU half so the host half stays CUDA-free
};

struct mp4j_tgroup {
  mp4j::ThreadComm* tc = nullptr;
};

#define MP4J_API_BEGIN try {
#define MP4J_API_END                     \
  }                                      \
  catch (const std::exception& e) {      \
    mp4j::g_last_error = e.what();       \
    return -1;                           \
  }                                      \
  catch (...) {                          \
    mp4j::g_last_error = "unknown error"; \
    return -1;                           \
  }                                      \
  return 0;
