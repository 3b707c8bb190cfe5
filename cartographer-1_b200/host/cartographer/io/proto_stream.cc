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

// proto_stream.cc:27
const uint64_t kMagic = 0x7b1d1f7b5bf501dbull;

// proto_stream.cc:29-34
void WriteSizeAsLittleEndian(uint64_t size, std::ostream* out) {
  for (int i = 0; i != 8; ++i) {
    out->put(static_cast<char>(size & 0xff));
    size >>= 8;
  }
}

// proto_stream.cc:36-43
bool ReadSizeAsLittleEndian(std::istream* in, uint64_t* size) {
  *size = 0;
  for (int i = 0; i != 8; ++i) {
    *size >>= 8;
    *size += static_cast<uint64_t>(in->get()) << 56;
  }
  return !in->fail();
}

}  // namespace

ProtoStreamWriter::ProtoStreamWriter(const std::string& filename)
    : out_(filename, std::ios::out | std::ios::binary) {
  WriteSizeAsLittleEndian(kMagic, &out_);
}

void ProtoStreamWriter::Write(const std::string& uncompressed_data) {
  std::string compressed_data;
  common::FastGzipString(uncompressed_data, &compressed_data);
  WriteSizeAsLittleEndian(compressed_data.size(), &out_);
  out_.write(compressed_data.data(), static_cast<std::streamsize>(compressed_data.size()));
}

bool ProtoStreamWriter::Close() {
  out_.close();
  return !out_.fail();
}

ProtoStreamReader::ProtoStreamReader(const std::string& filename)
    : in_(filename, std::ios::in | std::ios::binary) {
  uint64_t magic;
  if (!ReadSizeAsLittleEndian(&in_, &magic) || magic != kMagic) in_.setstate(std::ios::failbit);
  if (!in_.good()) {
    std::fprintf(stderr, "Failed to open proto stream '%s'.\n", filename.c_str());
    std::abort();
  }
}

bool ProtoStreamReader::Read(std::string* decompressed_data) {
  uint64_t compressed_size;
  if (!ReadSizeAsLittleEndian(&in_, &compressed_size)) return false;
  // a size beyond the file is a truncated stream: fail the read instead of allocating it
  const std::streampos here = in_.tellg();
  in_.seekg(0, std::ios::end);
  const std::streampos end = in_.tellg();
  in_.seekg(here);
  if (here < 0 || compressed_size > static_cast<uint64_t>(end - here)) {
    in_.setstate(std::ios::failbit);
    return false;
  }
  std::string compressed_data(compressed_size, '\0');
  if (compressed_size != 0 &&
      !in_.read(&compressed_data.front(), static_cast<std::streamsize>(compressed_size)))
    return false;
  decompressed_data->clear();
  common::FastGunzipString(compressed_data, decompressed_data);
  return true;
}

bool ProtoStreamReader::eof() const { return in_.eof(); }

}  // namespace io
}  // namespace cartographer
