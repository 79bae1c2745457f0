// Stand-in for absl/algorithm/container.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_CONTAINER_H_
#define BSG_SHIM_ABSL_CONTAINER_H_
#include <algorithm>
#include <numeric>
namespace absl {
template <typename C, typename T>
T c_accumulate(const C& c, T init) { return std::accumulate(c.begin(), c.end(), init); }
template <typename C, typename P>
auto c_find_if(C& c, P p) -> decltype(c.begin()) { return std::find_if(c.begin(), c.end(), p); }
template <typename C, typename F>
void c_for_each(C&& c, F f) { std::for_each(c.begin(), c.end(), f); }
template <typename C, typename T>
auto c_count(const C& c, const T& v) -> decltype(std::count(c.begin(), c.end(), v)) {
  return std::count(c.begin(), c.end(), v);
}
template <typename C, typename T>
void c_fill(C& c, const T& v) { std::fill(c.begin(), c.end(), v); }
}  // namespace absl
#endif
