// A reader for the protocol-buffers wire format, enough to take serialized proto::Grid2D and
// proto::HybridGrid messages apart (libprotobuf and protoc are not in this image; inside a
// Cartographer tree the generated classes do this).  Wire format (public specification,
// protobuf.dev/programming-guides/encoding): a message is a sequence of (tag, value) records,
// tag = field_number << 3 | wire_type; wire types 0 varint, 1 fixed64, 2 length-delimited,
// 5 fixed32.  Repeated scalar fields of proto3 are packed (one length-delimited record) but a
// parser must also accept the unpacked form; sint32 is zig-zag encoded.
#ifndef CARTOGRAPHER_IO_PROTO_WIRE_H_
#define CARTOGRAPHER_IO_PROTO_WIRE_H_

#include <cstdint>
#include <cstring>
#include <string>

namespace cartographer {
namespace io {
namespace wire {

struct Span {
  const uint8_t* data = nullptr;
  size_t size = 0;
};

class Reader {
 public:
  Reader(const void* data, size_t size) : p_(static_cast<const uint8_t*>(data)), end_(p_ + size) {}
  explicit Reader(const Span& s) : Reader(s.data, s.size) {}
  bool ok() const { return ok_; }
  bool done() const { return !ok_ || p_ >= end_; }

  // Next record's tag; false at the end or on malformed input.
  bool Next(int* field, int* wire_type) {
    if (done()) return false;
    const uint64_t tag = Varint();
    *field = static_cast<int>(tag >> 3);
    *wire_type = static_cast<int>(tag & 7);
    return ok_ && *field > 0;
  }
  uint64_t Varint() {
    uint64_t v = 0;
    for (int shift = 0; shift < 64 && p_ < end_; shift += 7) {
      const uint8_t b = *p_++;
      v |= static_cast<uint64_t>(b & 0x7F) << shift;
      if ((b & 0x80) == 0) return v;
    }
    ok_ = false;
    return 0;
  }
  static int32_t ZigZag32(uint64_t v) {
    const uint32_t n = static_cast<uint32_t>(v);
    return static_cast<int32_t>((n >> 1) ^ (~(n & 1) + 1));
  }
  float Float() {
    uint32_t bits = 0;
    Fixed(&bits, 4);
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
  }
  double Double() {
    uint64_t bits = 0;
    Fixed(&bits, 8);
    double d;
    std::memcpy(&d, &bits, 8);
    return d;
  }
  Span Bytes() {
    const uint64_t n = Varint();
    Span s;
    if (!ok_ || n > static_cast<uint64_t>(end_ - p_)) {
      ok_ = false;
      return s;
    }
    s.data = p_;
    s.size = static_cast<size_t>(n);
    p_ += n;
    return s;
  }
  // Skips the value of a record whose field the caller does not know.
  void Skip(int wire_type) {
    switch (wire_type) {
      case 0: Varint(); break;
      case 1: Advance(8); break;
      case 2: Bytes(); break;
      case 5: Advance(4); break;
      default: ok_ = false;  // groups (3, 4) are not used by proto3
    }
  }

 private:
  void Fixed(void* out, size_t n) {
    if (static_cast<size_t>(end_ - p_) < n) {
      ok_ = false;
      return;
    }
    std::memcpy(out, p_, n);  // little endian on the wire and on this machine
    p_ += n;
  }
  void Advance(size_t n) {
    if (static_cast<size_t>(end_ - p_) < n) ok_ = false;
    else p_ += n;
  }
  const uint8_t* p_;
  const uint8_t* end_;
  bool ok_ = true;
};

// A repeated varint field: `fn(value)` for every element of a packed record (wire type 2) or of
// one unpacked element (wire type 0).
template <typename F>
void RepeatedVarint(Reader* r, int wire_type, const F& fn) {
  if (wire_type == 2) {
    Reader packed(r->Bytes());
    while (!packed.done()) fn(packed.Varint());
    if (!packed.ok()) r->Skip(-1);
  } else if (wire_type == 0) {
    fn(r->Varint());
  } else {
    r->Skip(-1);
  }
}

}  // namespace wire
}  // namespace io
}  // namespace cartographer

#endif  // CARTOGRAPHER_IO_PROTO_WIRE_H_
