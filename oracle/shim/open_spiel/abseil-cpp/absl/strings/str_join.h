// Stand-in for absl/strings/str_join.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_STR_JOIN_H_
#define BSG_SHIM_ABSL_STR_JOIN_H_
#include <sstream>
#include <string>
namespace absl {
template <typename Container, typename Sep>
std::string StrJoin(const Container& c, const Sep& sep) {
  std::ostringstream os;
  bool first = true;
  for (const auto& v : c) {
    if (!first) os << sep;
    os << v;
    first = false;
  }
  return os.str();
}
}  // namespace absl
#endif
