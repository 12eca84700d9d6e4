// mp4j::Error - the one exception type of the core (parity: J/exception/Mp4jException.java:30-44).
// The C ABI converts it to a negative return code + thread-local message; the Python
// binding re-raises it as Mp4jException.
#pragma once
#include <stdexcept>
#include <string>

namespace mp4j {

class Error : public std::runtime_error {
 public:
  explicit Error(const std::string& what) : std::runtime_error(what) {}
};

}  // namespace mp4j
