// --track-memory support with the reference's interface (src/memory.h:27-100:
// qflex::memory::get_memory_usage / print_memory_usage, driven by SIGALRM from
// src/main.cpp:108-112).  What a user of this engine wants to watch is the DEVICE
// memory, so the line reports the process' resident host memory and the GPU's used
// memory (qfx_device_memory, include/qflex_b200.h).
#ifndef QFLEX_HOST_MEMORY_H_
#define QFLEX_HOST_MEMORY_H_

#include <signal.h>
#include <unistd.h>

#include <array>
#include <fstream>
#include <iostream>
#include <string>

#include "errors.h"
#include "qflex_b200.h"

namespace qflex::memory {

// {resident host memory, peak resident host memory}
inline std::array<std::string, 2> get_memory_usage() noexcept {
  std::array<std::string, 2> out{std::string("n/a"), std::string("n/a")};
  std::ifstream status("/proc/self/status");
  std::string key;
  while (status >> key) {
    double kb = 0;
    if (key == "VmRSS:" && (status >> kb)) out[0] = utils::readable_memory_string(kb * 1024);
    else if (key == "VmHWM:" && (status >> kb)) out[1] = utils::readable_memory_string(kb * 1024);
    status.ignore(4096, '\n');
  }
  return out;
}

// {device memory in use, device memory total} of the current device
inline std::array<std::string, 2> get_device_memory_usage() noexcept {
  uint64_t free_b = 0, total_b = 0;
  if (qfx_device_memory(-1, &free_b, &total_b) != QFX_OK)
    return {std::string("n/a"), std::string("n/a")};
  return {utils::readable_memory_string((double)(total_b - free_b)),
          utils::readable_memory_string((double)total_b)};
}

inline void print_memory_usage(int /*signal_number*/ = 0) noexcept {
  const auto host = get_memory_usage();
  const auto dev = get_device_memory_usage();
  std::cerr << WARN_MSG("Memory usage: ", host[0], " (Peak: ", host[1], "), device: ", dev[0],
                        " of ", dev[1])
            << std::endl;
  if (global::track_memory_seconds > 0) alarm(global::track_memory_seconds);
}

}  // namespace qflex::memory

#endif  // QFLEX_HOST_MEMORY_H_
