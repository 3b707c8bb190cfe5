// TEST INFRASTRUCTURE (oracle/_ref build only): libprotobuf is not in this image.  The
// reference's io/proto_stream.cc needs exactly two members of google::protobuf::Message; the
// test surface (ref_capi_io.cc) passes messages that carry already-serialized bytes, so the
// container code that runs (magic, size prefixes, gzip members) is the reference's own.
#ifndef ORACLE_REF_SHIM_GOOGLE_PROTOBUF_MESSAGE_H_
#define ORACLE_REF_SHIM_GOOGLE_PROTOBUF_MESSAGE_H_
#include <string>
namespace google {
namespace protobuf {
class Message {
 public:
  virtual ~Message() {}
  virtual bool SerializeToString(std::string* output) const = 0;
  virtual bool ParseFromString(const std::string& data) = 0;
};
}  // namespace protobuf
}  // namespace google
#endif
