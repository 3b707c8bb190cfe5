// TEST INFRASTRUCTURE (oracle/_ref build only): abseil is not in this image.  Substitute is used
// by the DebugString() members only.
#ifndef ORACLE_REF_SHIM_ABSL_SUBSTITUTE_H_
#define ORACLE_REF_SHIM_ABSL_SUBSTITUTE_H_
#include <sstream>
#include <string>
#include <vector>
namespace absl {
template <typename... Args>
std::string Substitute(const std::string& format, const Args&... args) {
  std::vector<std::string> text;
  auto add = [&text](const auto& a) {
    std::ostringstream os;
    os << a;
    text.push_back(os.str());
  };
  (add(args), ...);
  std::string out;
  for (size_t i = 0; i < format.size(); ++i) {
    if (format[i] == '$' && i + 1 < format.size() && format[i + 1] >= '0' && format[i + 1] <= '9') {
      const size_t k = format[i + 1] - '0';
      if (k < text.size()) out += text[k];
      ++i;
    } else {
      out += format[i];
    }
  }
  return out;
}
}  // namespace absl
#endif
