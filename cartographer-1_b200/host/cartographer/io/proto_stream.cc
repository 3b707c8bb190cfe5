#include "cartographer/io/proto_stream.h"

#include <zlib.h>

#include <cstring>

#include "cartographer_compat.h"

namespace cartographer {
namespace common {

// common/port.h:48-57.  boost's gzip_compressor(best_speed) is deflate level 1 in a gzip
// wrapper with the default window (15 bits) and memory level.
void FastGzipString(const std::string& uncompressed, std::string* compressed) {
  z_stream z;
  std::memset(&z, 0, sizeof(z));
  CSM_CHECK(deflateInit2(&z, Z_BEST_SPEED, Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY) == Z_OK);
  z.next_in = reinterpret_cast<Bytef*>(const_cast<char*>(uncompressed.data()));
  z.avail_in = static_cast<uInt>(uncompressed.size());
  CSM_CHECK(uncompressed.size() <= 0xFFFFFFFFu);  // one protobuf message: < 2 GiB by construction
  unsigned char chunk[1 << 15];
  int status = Z_OK;
  while (status != Z_STREAM_END) {
    z.next_out = chunk;
    z.avail_out = sizeof(chunk);
    status = deflate(&z, Z_FINISH);
    CSM_CHECK(status == Z_OK || status == Z_STREAM_END || status == Z_BUF_ERROR);
    compressed->append(reinterpret_cast<const char*>(chunk), sizeof(chunk) - z.avail_out);
  }
  deflateEnd(&z);
}

// common/port.h:59-66.  Window bits 15 + 32: gzip or zlib header, detected by zlib.
void FastGunzipString(const std::string& compressed, std::string* decompressed) {
  z_stream z;
  std::memset(&z, 0, sizeof(z));
  CSM_CHECK(inflateInit2(&z, 15 + 32) == Z_OK);
  CSM_CHECK(compressed.size() <= 0xFFFFFFFFu);
  z.next_in = reinterpret_cast<Bytef*>(const_cast<char*>(compressed.data()));
  z.avail_in = static_cast<uInt>(compressed.size());
  unsigned char chunk[1 << 15];
  int status = Z_OK;
  while (status != Z_STREAM_END) {
    z.next_out = chunk;
    z.avail_out = sizeof(chunk);
    status = inflate(&z, Z_NO_FLUSH);
    // truncated or corrupt input: boost's filter throws, which ends the reference's process too
    CSM_CHECK(status == Z_OK || status == Z_STREAM_END);
    decompressed->append(reinterpret_cast<const char*>(chunk), sizeof(chunk) - z.avail_out);
    CSM_CHECK(status == Z_STREAM_END || z.avail_in != 0 || z.avail_out == 0);
  }
  inflateEnd(&z);
}

}  // namespace common

namespace io {
namespace {

// proto_stream.cc:27: the eight bytes every stream starts with, and the unit of its framing: an
// unsigned 64-bit integer, least significant byte first (proto_stream.cc:29-43).
constexpr uint64_t kMagic = 0x7b1d1f7b5bf501dbull;

void PutFrameWord(std::ofstream* file, const uint64_t value) {
  char bytes[8];
  for (int k = 0; k < 8; ++k) bytes[k] = static_cast<char>((value >> (8 * k)) & 0xffu);
  file->write(bytes, sizeof(bytes));
}

bool GetFrameWord(std::ifstream* file, uint64_t* value) {
  unsigned char bytes[8];
  file->read(reinterpret_cast<char*>(bytes), sizeof(bytes));  // a short read sets failbit (and eofbit)
  if (file->gcount() != static_cast<std::streamsize>(sizeof(bytes))) return false;
  uint64_t v = 0;
  for (int k = 7; k >= 0; --k) v = (v << 8) | bytes[k];
  *value = v;
  return true;
}

}  // namespace

ProtoStreamWriter::ProtoStreamWriter(const std::string& filename) : out_(filename, std::ios::binary | std::ios::out) {
  PutFrameWord(&out_, kMagic);
}

// proto_stream.cc:52-57: one gzip member per message behind its compressed size.
void ProtoStreamWriter::Write(const std::string& uncompressed_data) {
  std::string member;
  common::FastGzipString(uncompressed_data, &member);
  PutFrameWord(&out_, member.size());
  out_.write(member.data(), static_cast<std::streamsize>(member.size()));
}

bool ProtoStreamWriter::Close() {
  out_.close();
  return !out_.fail();
}

// proto_stream.cc:68-76: a file that cannot be opened or starts with anything but the magic is fatal.
ProtoStreamReader::ProtoStreamReader(const std::string& filename) : in_(filename, std::ios::binary | std::ios::in) {
  uint64_t first = 0;
  if (!in_.is_open() || !GetFrameWord(&in_, &first) || first != kMagic) {
    std::fprintf(stderr, "Failed to open proto stream '%s'.\n", filename.c_str());
    std::abort();
  }
}

// proto_stream.cc:78-90
bool ProtoStreamReader::Read(std::string* decompressed_data) {
  uint64_t member_size = 0;
  if (!GetFrameWord(&in_, &member_size)) return false;  // the regular end: eofbit is set now
  // a size that points beyond the file is a truncated stream: fail the read instead of allocating it
  const std::streampos at = in_.tellg();
  in_.seekg(0, std::ios::end);
  const std::streampos file_end = in_.tellg();
  in_.seekg(at);
  if (at < 0 || member_size > static_cast<uint64_t>(file_end - at)) {
    in_.setstate(std::ios::failbit);
    return false;
  }
  std::string member(static_cast<size_t>(member_size), '\0');
  if (member_size != 0 && !in_.read(&member[0], static_cast<std::streamsize>(member_size))) return false;
  decompressed_data->clear();
  common::FastGunzipString(member, decompressed_data);
  return true;
}

bool ProtoStreamReader::eof() const { return in_.eof(); }

}  // namespace io
}  // namespace cartographer
