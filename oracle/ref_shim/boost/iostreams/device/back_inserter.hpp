// TEST INFRASTRUCTURE (oracle/_ref build only): boost is not in this image.  common/port.h uses
// these in FastGzipString / FastGunzipString only, which the scan-matcher sources never call;
// the stand-ins abort if they are ever reached.
#ifndef ORACLE_REF_SHIM_BOOST_IOSTREAMS_
#define ORACLE_REF_SHIM_BOOST_IOSTREAMS_
#include <cstddef>
#include <cstdlib>
#include <string>
namespace boost {
namespace iostreams {
namespace zlib {
constexpr int best_speed = 1;
}
struct gzip_compressor {
  explicit gzip_compressor(int) {}
};
struct gzip_decompressor {};
struct string_sink {};
inline string_sink back_inserter(std::string&) { return string_sink(); }
class filtering_ostream {
 public:
  template <typename T>
  void push(const T&) {}
};
inline void write(filtering_ostream&, const char*, std::size_t) { std::abort(); }
}  // namespace iostreams
}  // namespace boost
#endif
