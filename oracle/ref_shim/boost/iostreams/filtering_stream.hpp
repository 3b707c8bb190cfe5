// TEST INFRASTRUCTURE (oracle/_ref build only): boost is not in this image; zlib, which boost's
// gzip filters wrap, is.  common/port.h's FastGzipString / FastGunzipString push one gzip filter
// and one back_inserter sink on a filtering_ostream and write once; this stand-in keeps that
// shape: bytes written are buffered and, when the stream is destroyed (where boost flushes),
// deflated into one gzip member at the requested level / inflated, and appended to the sink.
#ifndef ORACLE_REF_SHIM_BOOST_IOSTREAMS_
#define ORACLE_REF_SHIM_BOOST_IOSTREAMS_
#include <zlib.h>

#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <string>
namespace boost {
namespace iostreams {
namespace zlib {
constexpr int best_speed = 1;
}
struct gzip_compressor {
  explicit gzip_compressor(int level) : level(level) {}
  int level;
};
struct gzip_decompressor {};
struct string_sink {
  std::string* target;
};
inline string_sink back_inserter(std::string& s) { return string_sink{&s}; }
class filtering_ostream {
 public:
  void push(const gzip_compressor& c) { mode_ = 1, level_ = c.level; }
  void push(const gzip_decompressor&) { mode_ = 2; }
  void push(const string_sink& s) { sink_ = s.target; }
  void append(const char* data, std::size_t n) { pending_.append(data, n); }
  ~filtering_ostream() {
    if (sink_ == nullptr || mode_ == 0) return;
    z_stream z;
    std::memset(&z, 0, sizeof(z));
    if ((mode_ == 1 ? deflateInit2(&z, level_, Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY)
                    : inflateInit2(&z, 15 + 16)) != Z_OK)
      std::abort();
    z.next_in = reinterpret_cast<Bytef*>(const_cast<char*>(pending_.data()));
    z.avail_in = static_cast<uInt>(pending_.size());
    unsigned char chunk[1 << 15];
    int status = Z_OK;
    while (status != Z_STREAM_END) {
      z.next_out = chunk;
      z.avail_out = sizeof(chunk);
      status = mode_ == 1 ? deflate(&z, Z_FINISH) : inflate(&z, Z_NO_FLUSH);
      if (status != Z_OK && status != Z_STREAM_END) std::abort();  // boost throws gzip_error
      sink_->append(reinterpret_cast<const char*>(chunk), sizeof(chunk) - z.avail_out);
      if (status == Z_OK && z.avail_in == 0 && z.avail_out != 0) std::abort();  // truncated input
    }
    mode_ == 1 ? deflateEnd(&z) : inflateEnd(&z);
  }

 private:
  int mode_ = 0, level_ = 1;
  std::string* sink_ = nullptr;
  std::string pending_;
};
inline void write(filtering_ostream& out, const char* data, std::size_t n) { out.append(data, n); }
}  // namespace iostreams
}  // namespace boost
#endif
