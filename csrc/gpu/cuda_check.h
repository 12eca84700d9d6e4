// CUDA runtime error -> mp4j::Error.
#pragma once
#include <cuda_runtime.h>

#include <string>

#include "../core/error.h"

#define MP4J_CUDA(expr)                                                                          \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess)                                                                       \
      throw mp4j::Error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " (" + __FILE__ + ":" + \
                        std::to_string(__LINE__) + ")");                                         \
  } while (0)
