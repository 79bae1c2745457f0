// Host-side helpers of the C ABI: game dispatch, packing histories, single-state accessors.
// They reuse the same __host__ __device__ game code as the kernels (bsg_games.cuh), so the
// CPU test-suite exercises the exact rules the GPU runs — but there is deliberately no host
// playout loop here: rollouts exist only as CUDA kernels.
#pragma once
#include <string.h>

#include "../../include/bsg.h"
#include "bsg_games.cuh"

namespace bsg {

typedef HexT<5> Hex5;

template <class G> struct TablesOf {
  static const typename G::Tables& get() {
    static typename G::Tables* t = [] {
      auto* p = new typename G::Tables();
      G::fill_tables(*p);
      return p;
    }();
    return *t;
  }
};

// apply() takes the tables only for hex; unify.
template <class G> inline bool apply_any(typename G::State& s, int a) { return G::apply(s, a); }
template <> inline bool apply_any<Hex>(Hex::State& s, int a) { return Hex::apply(s, a, TablesOf<Hex>::get()); }
template <> inline bool apply_any<Hex5>(Hex5::State& s, int a) { return Hex5::apply(s, a, TablesOf<Hex5>::get()); }

// Calls f.template operator()<G>() for the game id; returns BSG_ERR_BAD_GAME otherwise.
template <class F> inline int dispatch_game(int game, F&& f) {
  switch (game) {
    case BSG_TIC_TAC_TOE: return f.template operator()<TicTacToe>();
    case BSG_CONNECT_FOUR: return f.template operator()<ConnectFour>();
    case BSG_BREAKTHROUGH: return f.template operator()<Breakthrough>();
    case BSG_HEX: return f.template operator()<Hex>();
    case BSG_HEX5_HOST_ONLY: return f.template operator()<Hex5>();
    default: break;
  }
  return BSG_ERR_BAD_GAME;
}
// The four games that have device kernels.
template <class F> inline int dispatch_device_game(int game, F&& f) {
  switch (game) {
    case BSG_TIC_TAC_TOE: return f.template operator()<TicTacToe>();
    case BSG_CONNECT_FOUR: return f.template operator()<ConnectFour>();
    case BSG_BREAKTHROUGH: return f.template operator()<Breakthrough>();
    case BSG_HEX: return f.template operator()<Hex>();
    default: break;
  }
  return BSG_ERR_BAD_GAME;
}

struct PackOp {
  const int64_t* flat; const int64_t* off; int64_t n; void* out; int64_t cap;
  template <class G> int operator()() const {
    for (int64_t i = 0; i < n; ++i) {
      typename G::State s;
      G::init(s);
      for (int64_t j = off[i]; j < off[i + 1]; ++j)
        if (!apply_any<G>(s, (int)flat[j])) return BSG_ERR_ILLEGAL_ACTION;
      G::store(s, out, cap, i);
    }
    return BSG_OK;
  }
};
struct InfoOp {
  const void* p; int64_t cap, i; int* cur; int* term; int* ret0;
  template <class G> int operator()() const {
    typename G::State s;
    G::load(s, p, cap, i);
    bool t = G::terminal(s);
    if (term) *term = t ? 1 : 0;
    if (cur) *cur = t ? BSG_TERMINAL_PLAYER : G::mover(s);
    if (ret0) *ret0 = outcome_ret0(G::outcome(s));
    return BSG_OK;
  }
};
struct LegalOp {
  const void* p; int64_t cap, i; int64_t* out; int capacity;
  template <class G> int operator()() const {
    typename G::State s;
    G::load(s, p, cap, i);
    int tmp[128];
    int n = G::legal(s, tmp);
    if (n > capacity) return BSG_ERR_BAD_ARG;
    for (int k = 0; k < n; ++k) out[k] = tmp[k];
    return n;
  }
};
struct ApplyOp {
  void* p; int64_t cap, i; int64_t a;
  template <class G> int operator()() const {
    typename G::State s;
    G::load(s, p, cap, i);
    if (a < 0 || a >= G::kNumActions || !apply_any<G>(s, (int)a)) return BSG_ERR_ILLEGAL_ACTION;
    G::store(s, p, cap, i);
    return BSG_OK;
  }
};
// One replay step through the SAME step() the kernels run (table look-ups, fast modulo,
// rank-select).  Debug/test hook: lets the CPU suite check the device selection logic
// without a GPU.  Not used by any product path.
struct StepOp {
  void* p; int64_t cap, i; uint32_t r; int64_t* action;
  template <class G> int operator()() const {
    typename G::State s;
    G::load(s, p, cap, i);
    if (G::terminal(s)) return BSG_ERR_ILLEGAL_ACTION;
    int tmp[128];
    if (G::legal(s, tmp) == 0) return BSG_ERR_STUCK_LANE;
    int a = -1;
    G::step(s, r, TablesOf<G>::get(), a);
    if (action) *action = a;
    G::store(s, p, cap, i);
    return BSG_OK;
  }
};
struct ObsOp {
  const void* p; int64_t cap, i; int player; float* out;
  template <class G> int operator()() const {
    typename G::State s;
    G::load(s, p, cap, i);
    // through the bit string the batched device writer expands, cross-checked per element
    u32 words[G::kObsWords];
    G::obs_words(s, player, words);
    for (int k = 0; k < G::kObsSize; ++k) {
      u32 b = (words[k >> 5] >> (k & 31)) & 1u;
      if (b != G::obs_bit(s, player, k)) return BSG_ERR_BAD_ARG;  // internal inconsistency
      out[k] = (float)b;
    }
    return BSG_OK;
  }
};
struct MaskOp {
  const void* p; int64_t cap, i; uint8_t* out;
  template <class G> int operator()() const {
    typename G::State s;
    G::load(s, p, cap, i);
    u32 words[G::kMaskWords];
    G::mask_words(s, words);
    for (int a = 0; a < G::kNumActions; ++a) {
      u32 b = (words[a >> 5] >> (a & 31)) & 1u;
      if (b != G::legal_bit(s, a)) return BSG_ERR_BAD_ARG;  // internal inconsistency
      out[a] = (uint8_t)b;
    }
    return BSG_OK;
  }
};

}  // namespace bsg
