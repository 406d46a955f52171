// Error/diagnostic conventions of the host layer.
// Mirrors the reference contract (src/errors.h:7-12, src/utils.h:14-27): every
// failure is a thrown std::string built by ERROR_MSG with file:function:line,
// re-wrapped with context at each level; callers and tests match substrings.
#ifndef QFLEX_HOST_ERRORS_H_
#define QFLEX_HOST_ERRORS_H_

#include <ctime>
#include <sstream>
#include <string>
#include <type_traits>

namespace qflex::global {
// Same process-wide knobs as the reference (src/global.h:9-15).
inline int verbose = 0;
inline std::size_t memory_limit = 1L << 30;
inline std::size_t track_memory_seconds = 0;
}  // namespace qflex::global

namespace qflex::utils {

// Appends one part of a message.  Strings, characters and integers take the direct
// route (index names are built tens of thousands of times per plan; a stringstream per
// name was a third of the planning time); everything else is formatted by operator<<
// exactly as the reference's concat does (src/utils.h:14-27).
namespace detail {
inline void append(std::string& out, const std::string& v) { out += v; }
inline void append(std::string& out, const char* v) { out += v; }
inline void append(std::string& out, char v) { out += v; }
template <typename T>
void append(std::string& out, const T& v) {
  if constexpr (std::is_integral_v<T> && !std::is_same_v<T, bool>) {
    out += std::to_string(v);
  } else {
    std::stringstream ss;
    ss << v;
    out += ss.str();
  }
}
}  // namespace detail

template <typename... Ts>
std::string concat(const Ts&... parts) {
  std::string out;
  (detail::append(out, parts), ...);
  return out;
}

inline std::string get_date_time() {
  std::time_t t = std::time(nullptr);
  char buf[100];
  if (std::strftime(buf, sizeof(buf), "%c", std::localtime(&t))) return buf;
  return "[time-readout error]";
}

// "4GB" / "512 MB" / "1024" -> bytes; bytes -> "52.76 MB"
// (behaviour of src/utils.cpp:22-72).
std::size_t from_readable_memory_string(std::string memory);
std::string readable_memory_string(double memory);

}  // namespace qflex::utils

#define ERROR_MSG(...)                                                        \
  qflex::utils::concat("ERROR (" __FILE__ ": ", __FUNCTION__, ":", __LINE__, \
                       ") --> ", __VA_ARGS__)
#define WARN_MSG(...) \
  qflex::utils::concat("[", qflex::utils::get_date_time(), "] ", __VA_ARGS__)

#endif  // QFLEX_HOST_ERRORS_H_
