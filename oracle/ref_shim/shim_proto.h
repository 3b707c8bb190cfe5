// TEST INFRASTRUCTURE (oracle/_ref build only): protoc and libprotobuf are not in this image, so
// the generated *.pb.h the reference's sources include are replaced by plain structs with the
// accessor names protoc would generate (field(), set_field(), mutable_field(), has_field(),
// repeated field_size()/add_field()).  No serialisation: the scan-matcher path only carries
// option values and grid limits through them.
#ifndef ORACLE_REF_SHIM_PROTO_H_
#define ORACLE_REF_SHIM_PROTO_H_

#include <cstdint>
#include <string>
#include <vector>

namespace google {
namespace protobuf {
template <typename T>
class RepeatedField : public std::vector<T> {
 public:
  using std::vector<T>::vector;
  RepeatedField() = default;
  void Add(const T& v) { this->push_back(v); }
  void Reserve(int n) { this->reserve(n); }
};
template <typename T>
using RepeatedPtrField = RepeatedField<T>;
}  // namespace protobuf
}  // namespace google

#define SHIM_SCALAR(type, name)                \
 public:                                       \
  type name() const { return name##_; }        \
  void set_##name(type v) { name##_ = v; }     \
  void clear_##name() { name##_ = type(); }    \
                                               \
 private:                                      \
  type name##_ = type();

#define SHIM_STRING(name)                                  \
 public:                                                   \
  const std::string& name() const { return name##_; }      \
  void set_##name(const std::string& v) { name##_ = v; }   \
  std::string* mutable_##name() { return &name##_; }       \
                                                           \
 private:                                                  \
  std::string name##_;

#define SHIM_MESSAGE(Type, name)                      \
 public:                                              \
  const Type& name() const { return name##_; }        \
  Type* mutable_##name() {                            \
    has_##name##_ = true;                             \
    return &name##_;                                  \
  }                                                   \
  bool has_##name() const { return has_##name##_; }   \
  void clear_##name() {                               \
    name##_ = Type();                                 \
    has_##name##_ = false;                            \
  }                                                   \
                                                      \
 private:                                             \
  Type name##_;                                       \
  bool has_##name##_ = false;

#define SHIM_REPEATED(Type, name)                                                       \
 public:                                                                                \
  const ::google::protobuf::RepeatedField<Type>& name() const { return name##_; }       \
  ::google::protobuf::RepeatedField<Type>* mutable_##name() { return &name##_; }        \
  const Type& name(int i) const { return name##_[i]; }                                  \
  int name##_size() const { return static_cast<int>(name##_.size()); }                  \
  void add_##name(const Type& v) { name##_.push_back(v); }                              \
  Type* add_##name() {                                                                  \
    name##_.emplace_back();                                                             \
    return &name##_.back();                                                             \
  }                                                                                     \
  void clear_##name() { name##_.clear(); }                                              \
                                                                                        \
 private:                                                                               \
  ::google::protobuf::RepeatedField<Type> name##_;

#endif  // ORACLE_REF_SHIM_PROTO_H_
