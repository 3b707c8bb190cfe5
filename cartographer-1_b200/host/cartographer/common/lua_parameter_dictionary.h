// Stand-in for common::LuaParameterDictionary (reference: cartographer/common/
// lua_parameter_dictionary.{h,cc}), which needs a Lua interpreter this image does not have.  It
// reads what the Create*Options functions of the scan matchers are given: one flat table
// "return { key = number | true | false, ... }".  Nested tables and expressions are not
// supported; in a full Cartographer tree the real header replaces this file.
#ifndef CARTOGRAPHER_COMMON_LUA_PARAMETER_DICTIONARY_H_
#define CARTOGRAPHER_COMMON_LUA_PARAMETER_DICTIONARY_H_

#include <cctype>
#include <cstdlib>
#include <map>
#include <memory>
#include <string>

#include "cartographer_compat.h"

namespace cartographer {
namespace common {

class LuaParameterDictionary {
 public:
  explicit LuaParameterDictionary(const std::string& code) {
    size_t at = code.find('{');
    CSM_CHECK(at != std::string::npos);
    ++at;
    while (at < code.size()) {
      while (at < code.size() && !(std::isalpha(static_cast<unsigned char>(code[at])) || code[at] == '_')) {
        if (code[at] == '}') return;
        ++at;
      }
      size_t end = at;
      while (end < code.size() && (std::isalnum(static_cast<unsigned char>(code[end])) || code[end] == '_')) ++end;
      const std::string key = code.substr(at, end - at);
      at = code.find('=', end);
      CSM_CHECK(at != std::string::npos);
      ++at;
      while (at < code.size() && std::isspace(static_cast<unsigned char>(code[at]))) ++at;
      end = at;
      while (end < code.size() && code[end] != ',' && code[end] != '}') ++end;
      std::string value = code.substr(at, end - at);
      while (!value.empty() && std::isspace(static_cast<unsigned char>(value.back()))) value.pop_back();
      CSM_CHECK(!value.empty() && value[0] != '{');  // nested tables are not supported
      values_[key] = value;
      at = end;
    }
  }
  bool HasKey(const std::string& key) const { return values_.count(key) != 0; }
  double GetDouble(const std::string& key) { return std::atof(Take(key).c_str()); }
  int GetInt(const std::string& key) { return std::atoi(Take(key).c_str()); }
  int GetNonNegativeInt(const std::string& key) {  // lua_parameter_dictionary.cc: CHECK_GE(value, 0)
    const int value = GetInt(key);
    CSM_CHECK(value >= 0);
    return value;
  }
  bool GetBool(const std::string& key) { return Take(key) == "true"; }

 private:
  const std::string& Take(const std::string& key) const {
    auto it = values_.find(key);
    if (it == values_.end()) {
      std::fprintf(stderr, "Key '%s' not in dictionary.\n", key.c_str());
      std::abort();
    }
    return it->second;
  }
  std::map<std::string, std::string> values_;
};

// common/lua_parameter_dictionary_test_helpers.h:29-33
inline std::unique_ptr<LuaParameterDictionary> MakeDictionary(const std::string& code) {
  return std::unique_ptr<LuaParameterDictionary>(new LuaParameterDictionary(code));
}

}  // namespace common
}  // namespace cartographer

#endif  // CARTOGRAPHER_COMMON_LUA_PARAMETER_DICTIONARY_H_
