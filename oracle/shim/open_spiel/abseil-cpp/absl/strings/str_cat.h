// Stand-in for absl/strings/str_cat.h (Abseil is not vendored in the reference tree).
// TEST INFRASTRUCTURE ONLY: lets the unmodified reference game sources compile for the oracle.
#ifndef BSG_SHIM_ABSL_STR_CAT_H_
#define BSG_SHIM_ABSL_STR_CAT_H_
#include <sstream>
#include <string>
namespace absl {
namespace shim_detail {
inline void Put(std::ostringstream&) {}
template <typename T, typename... Rest>
void Put(std::ostringstream& os, const T& v, const Rest&... rest) {
  os << v;
  Put(os, rest...);
}
}  // namespace shim_detail
template <typename... Args>
std::string StrCat(const Args&... args) {
  std::ostringstream os;
  shim_detail::Put(os, args...);
  return os.str();
}
template <typename... Args>
void StrAppend(std::string* dst, const Args&... args) {
  std::ostringstream os;
  shim_detail::Put(os, args...);
  dst->append(os.str());
}
}  // namespace absl
#endif
