// The .pbstream container: mirrors
//   io::ProtoStreamWriter / io::ProtoStreamReader                     io/proto_stream.{h,cc}
//   common::FastGzipString / FastGunzipString                         common/port.h:48-66
// File layout (proto_stream.cc:27-45,52-57,68-80): the eight bytes of kMagic
// 0x7b1d1f7b5bf501db, little endian, then per message an eight-byte little-endian size and
// that many bytes of a gzip member holding the serialized message.  The reference compresses
// with boost::iostreams::gzip_compressor(best_speed); boost is a wrapper around zlib, which
// this image has (zlib.h, libz), so the same deflate streams are made and read here with zlib
// itself.  libprotobuf is absent: messages cross this interface as serialized strings
// (`Read` / `Write` are the reference's private string methods made public) and are taken
// apart by io/proto_wire.h.
#ifndef CARTOGRAPHER_IO_PROTO_STREAM_H_
#define CARTOGRAPHER_IO_PROTO_STREAM_H_

#include <cstdint>
#include <fstream>
#include <string>

namespace cartographer {
namespace common {

// gzip (RFC 1952) at zlib level 1 / any gzip or zlib stream back to bytes.  A malformed stream
// aborts, as boost's filter throws in the reference.
void FastGzipString(const std::string& uncompressed, std::string* compressed);
void FastGunzipString(const std::string& compressed, std::string* decompressed);

}  // namespace common

namespace io {

class ProtoStreamWriter {
 public:
  explicit ProtoStreamWriter(const std::string& filename);
  ProtoStreamWriter(const ProtoStreamWriter&) = delete;
  ProtoStreamWriter& operator=(const ProtoStreamWriter&) = delete;

  // One serialized message (what WriteProto hands on after SerializeToString).
  void Write(const std::string& uncompressed_data);
  bool Close();

 private:
  std::ofstream out_;
};

class ProtoStreamReader {
 public:
  // Aborts when the file cannot be opened or does not start with the magic (the reference CHECKs).
  explicit ProtoStreamReader(const std::string& filename);
  ProtoStreamReader(const ProtoStreamReader&) = delete;
  ProtoStreamReader& operator=(const ProtoStreamReader&) = delete;

  // The next message's serialized bytes; false at the end of the stream.
  bool Read(std::string* decompressed_data);
  bool eof() const;

 private:
  std::ifstream in_;
};

}  // namespace io
}  // namespace cartographer

#endif  // CARTOGRAPHER_IO_PROTO_STREAM_H_
