// Stand-in for absl/strings/str_format.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_STR_FORMAT_H_
#define BSG_SHIM_ABSL_STR_FORMAT_H_
#include <cstdio>
#include <string>
namespace absl {
namespace shim_detail {
template <typename T>
const T& FmtArg(const T& v) { return v; }
inline const char* FmtArg(const std::string& s) { return s.c_str(); }
}  // namespace shim_detail
template <typename... Args>
std::string StrFormat(const char* fmt, const Args&... args) {
  int n = std::snprintf(nullptr, 0, fmt, shim_detail::FmtArg(args)...);
  std::string out(n > 0 ? n : 0, '\0');
  if (n > 0) std::snprintf(&out[0], n + 1, fmt, shim_detail::FmtArg(args)...);
  return out;
}
}  // namespace absl
#endif
