// A reader (and, further down, a writer) for the protocol-buffers wire format, enough to take serialized proto::Grid2D and
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

// The writing side, canonical as libprotobuf and upb emit it: the caller adds fields in field-number
// order; a proto3 scalar that is zero is left out (floating point: all bits zero), a repeated scalar
// travels packed and is left out when empty, a sub-message that is present is always written, even
// when empty.  int32 / int64 / enum values are sign-extended to 64 bits before the varint.
class Writer {
 public:
  const std::string& str() const { return out_; }
  void Varint(uint64_t v) {
    while (v >= 0x80) {
      out_.push_back(static_cast<char>((v & 0x7F) | 0x80));
      v >>= 7;
    }
    out_.push_back(static_cast<char>(v));
  }
  void Tag(const int field, const int wire_type) { Varint((static_cast<uint64_t>(field) << 3) | wire_type); }
  void Int64(const int field, const int64_t v) {
    if (v == 0) return;
    Tag(field, 0);
    Varint(static_cast<uint64_t>(v));
  }
  void Int32(const int field, const int32_t v) { Int64(field, static_cast<int64_t>(v)); }
  void UInt32(const int field, const uint32_t v) { Int64(field, static_cast<int64_t>(v)); }
  void Bool(const int field, const bool v) { Int64(field, v ? 1 : 0); }
  void Float(const int field, const float v) {
    uint32_t bits;
    std::memcpy(&bits, &v, 4);
    if (bits == 0) return;
    Tag(field, 5);
    out_.append(reinterpret_cast<const char*>(&bits), 4);  // little endian on the wire and on this machine
  }
  void Double(const int field, const double v) {
    uint64_t bits;
    std::memcpy(&bits, &v, 8);
    if (bits == 0) return;
    Tag(field, 1);
    out_.append(reinterpret_cast<const char*>(&bits), 8);
  }
  void Message(const int field, const std::string& body) {
    Tag(field, 2);
    Varint(body.size());
    out_.append(body);
  }
  // repeated int32 / sint32 / float, packed; `get(i)` returns element i
  template <typename F>
  void PackedInt32(const int field, const size_t n, const F& get) {
    if (n == 0) return;
    Writer payload;
    for (size_t i = 0; i != n; ++i) payload.Varint(static_cast<uint64_t>(static_cast<int64_t>(static_cast<int32_t>(get(i)))));
    Message(field, payload.str());
  }
  template <typename F>
  void PackedSInt32(const int field, const size_t n, const F& get) {
    if (n == 0) return;
    Writer payload;
    for (size_t i = 0; i != n; ++i) {
      const int32_t v = static_cast<int32_t>(get(i));
      payload.Varint(static_cast<uint32_t>((static_cast<uint32_t>(v) << 1) ^ static_cast<uint32_t>(v >> 31)));
    }
    Message(field, payload.str());
  }
  template <typename F>
  void PackedFloat(const int field, const size_t n, const F& get) {
    if (n == 0) return;
    std::string payload(4 * n, '\0');
    for (size_t i = 0; i != n; ++i) {
      const float v = get(i);
      std::memcpy(&payload[4 * i], &v, 4);
    }
    Message(field, payload);
  }

 private:
  std::string out_;
};

}  // namespace wire
}  // namespace io
}  // namespace cartographer

#endif  // CARTOGRAPHER_IO_PROTO_WIRE_H_
