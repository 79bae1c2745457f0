// Stand-in for absl/random/bit_gen_ref.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_BIT_GEN_REF_H_
#define BSG_SHIM_ABSL_BIT_GEN_REF_H_
#include <cstddef>
#include <cstdint>
#include <random>
namespace absl {
// Type-erased reference to any URBG producing at least 32 bits per call.
class BitGenRef {
 public:
  using result_type = uint64_t;
  template <typename G>
  BitGenRef(G& g) : obj_(&g), call_([](void* p) -> uint64_t {
    G& gen = *static_cast<G*>(p);
    uint64_t hi = static_cast<uint32_t>(gen());
    uint64_t lo = static_cast<uint32_t>(gen());
    return (hi << 32) | lo;
  }) {}
  static constexpr result_type min() { return 0; }
  static constexpr result_type max() { return ~uint64_t{0}; }
  result_type operator()() { return call_(obj_); }
 private:
  void* obj_;
  uint64_t (*call_)(void*);
};
// Real Abseil makes absl::Uniform visible through this header's includes (spiel.cc:339 relies
// on that), so the stand-in defines it here; distributions.h just includes this file.
template <typename G>
double Uniform(G&& g, double lo, double hi) {
  return std::uniform_real_distribution<double>(lo, hi)(g);
}
template <typename G>
size_t Uniform(G&& g, unsigned lo, size_t hi) {
  return std::uniform_int_distribution<size_t>(lo, hi - 1)(g);
}
}  // namespace absl
#endif
