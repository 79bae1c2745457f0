// Stand-in for absl/strings/str_split.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_STR_SPLIT_H_
#define BSG_SHIM_ABSL_STR_SPLIT_H_
#include <string>
#include <vector>
namespace absl {
inline std::vector<std::string> StrSplit(const std::string& s, char delim) {
  std::vector<std::string> out;
  std::string cur;
  for (char ch : s) {
    if (ch == delim) {
      out.push_back(cur);
      cur.clear();
    } else {
      cur.push_back(ch);
    }
  }
  out.push_back(cur);
  return out;
}
}  // namespace absl
#endif
