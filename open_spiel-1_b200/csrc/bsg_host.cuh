// Host-side helpers of the C ABI: game dispatch, packing histories, single-state accessors.
// They reuse the same __host__ __device__ game code as the kernels (bsg_games.cuh), so the
// CPU test-suite exercises the exact rules the GPU runs — but there is deliberately no host
// playout loop here: rollouts exist only as CUDA kernels.
#pragma once
#include <string.h>

#include <string>
#include <type_traits>

#include "../../include/bsg.h"
#include "bsg_games.cuh"

namespace bsg {

typedef HexT<5> Hex5;
typedef YT<9> Y9;
typedef HavannahT<4> Havannah4;
typedef GoT<7, 9> Go7;      // the playthrough's go(board_size=7,komi=4.5)
constexpr int kMaxLegal = 320;   // pentago: 36 cells x 8 rotations

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
template <> inline bool apply_any<Pentago>(Pentago::State& s, int a) { return Pentago::apply(s, a, TablesOf<Pentago>::get()); }
template <> inline bool apply_any<Y>(Y::State& s, int a) { return Y::apply(s, a, TablesOf<Y>::get()); }
template <> inline bool apply_any<Y9>(Y9::State& s, int a) { return Y9::apply(s, a, TablesOf<Y9>::get()); }
template <> inline bool apply_any<Havannah>(Havannah::State& s, int a) { return Havannah::apply(s, a, TablesOf<Havannah>::get()); }
template <> inline bool apply_any<Havannah4>(Havannah4::State& s, int a) { return Havannah4::apply(s, a, TablesOf<Havannah4>::get()); }

// Calls f.template operator()<G>() for the game id; returns BSG_ERR_BAD_GAME otherwise.
template <class F> inline int dispatch_game(int game, F&& f) {
  switch (game) {
    case BSG_TIC_TAC_TOE: return f.template operator()<TicTacToe>();
    case BSG_CONNECT_FOUR: return f.template operator()<ConnectFour>();
    case BSG_BREAKTHROUGH: return f.template operator()<Breakthrough>();
    case BSG_HEX: return f.template operator()<Hex>();
    case BSG_HEX5_HOST_ONLY: return f.template operator()<Hex5>();
    case BSG_PENTAGO: return f.template operator()<Pentago>();
    case BSG_OTHELLO: return f.template operator()<Othello>();
    case BSG_Y: return f.template operator()<Y>();
    case BSG_Y9_HOST_ONLY: return f.template operator()<Y9>();
    case BSG_HAVANNAH: return f.template operator()<Havannah>();
    case BSG_HAVANNAH4_HOST_ONLY: return f.template operator()<Havannah4>();
    case BSG_GO: return f.template operator()<Go>();
    case BSG_GO7_HOST_ONLY: return f.template operator()<Go7>();
    default: break;
  }
  return BSG_ERR_BAD_GAME;
}
// The games that have device kernels.
template <class F> inline int dispatch_device_game(int game, F&& f) {
  switch (game) {
    case BSG_TIC_TAC_TOE: return f.template operator()<TicTacToe>();
    case BSG_CONNECT_FOUR: return f.template operator()<ConnectFour>();
    case BSG_BREAKTHROUGH: return f.template operator()<Breakthrough>();
    case BSG_HEX: return f.template operator()<Hex>();
    case BSG_PENTAGO: return f.template operator()<Pentago>();
    case BSG_OTHELLO: return f.template operator()<Othello>();
    case BSG_Y: return f.template operator()<Y>();
    case BSG_HAVANNAH: return f.template operator()<Havannah>();
    case BSG_GO: return f.template operator()<Go>();
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
    int tmp[kMaxLegal];
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
    int tmp[kMaxLegal];
    if (G::legal(s, tmp) == 0) return BSG_ERR_STUCK_LANE;
    int a = -1;
    if constexpr (G::kLazyEnd) {
      // the lean kernel's step_lazy must arrive where step() does: the same action and state,
      // a terminal state one action-less iteration later
      typename G::State z = s;
      int az = -1; u32 made = 0;
      G::step_lazy(z, r, TablesOf<G>::get(), az, made);
      G::step(s, r, TablesOf<G>::get(), a);
      if (made != 1u || az != a || G::terminal(z)) return BSG_ERR_BAD_ARG;  // internal inconsistency
      if (G::terminal(s)) {
        G::step_lazy(z, r ^ 0x9E3779B9u, TablesOf<G>::get(), az, made);
        if (made != 0u || !G::terminal(z)) return BSG_ERR_BAD_ARG;
      }
      if (!G::same_state(s, z)) return BSG_ERR_BAD_ARG;
    } else {
      G::step(s, r, TablesOf<G>::get(), a);
    }
    if (action) *action = a;
    G::store(s, p, cap, i);
    return BSG_OK;
  }
};
template <class G, class = void> struct host_fillable : std::false_type {};
template <class G> struct host_fillable<G, std::enable_if_t<G::kFillable>> : std::true_type {};
struct FillOp {
  const void* p; int64_t cap, i; uint64_t seed, playout; int* ret0; int* moves;
  template <class G> int operator()() const {
    if constexpr (host_fillable<G>::value) {
      typename G::State s;
      G::load(s, p, cap, i);
      PhiloxKeys ks;
      philox_schedule(seed, ks);
      int t = 0;
      while (!G::fill_done(s)) {
        u32 rr[4];
        philox4x32_10((u32)playout, (u32)(playout >> 32), (u32)t >> 2, 0u, ks, rr);
        int a = -1;
        G::step_fill(s, rr[t & 3], TablesOf<G>::get(), a);
        ++t;
      }
      u32 board[G::kFillWords];
      G::fill_store(s, board);
      if (ret0) *ret0 = G::fill_winner(board) == 0 ? 1 : -1;
      if (moves) *moves = t;
      return BSG_OK;
    } else {
      return BSG_ERR_BAD_ARG;
    }
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

// ---------------------------------------------------------------------------------------
// Text formats of the reference: State::ToString and the board-string serialisation
// (State::Serialize / Game::DeserializeState).  tic_tac_toe and hex serialise as their action
// history (spiel.cc:375-384), which a packed state does not carry — the Python mirror keeps
// the history and handles those two itself; connect_four and breakthrough serialise the board.
// ---------------------------------------------------------------------------------------
template <class G> struct Text {
  static std::string to_string(const typename G::State&) { return std::string(); }
  static bool has_board_format() { return false; }
  static std::string serialize(const typename G::State&) { return std::string(); }
  static int deserialize(typename G::State&, const char*) { return BSG_ERR_BAD_ARG; }
};
template <> struct Text<TicTacToe> {   // tic_tac_toe.cc:126-137
  static std::string to_string(const TicTacToe::State& s) {
    u32 x = TicTacToe::mover(s) ? s.opp : s.cur, o = TicTacToe::mover(s) ? s.cur : s.opp;
    std::string str;
    for (int r = 0; r < 3; ++r) {
      for (int c = 0; c < 3; ++c) str += ((x >> (3 * r + c)) & 1) ? 'x' : (((o >> (3 * r + c)) & 1) ? 'o' : '.');
      if (r < 2) str += '\n';
    }
    return str;
  }
  static bool has_board_format() { return false; }
  static std::string serialize(const TicTacToe::State&) { return std::string(); }
  static int deserialize(TicTacToe::State&, const char*) { return BSG_ERR_BAD_ARG; }
};
template <> struct Text<ConnectFour> {  // connect_four.cc:166-175 (rows top-down), :227-279
  static std::string to_string(const ConnectFour::State& s) {
    u64 x = ConnectFour::mover(s) ? s.opp : s.cur, o = ConnectFour::mover(s) ? s.cur : s.opp;
    std::string str;
    for (int row = 5; row >= 0; --row) {
      for (int col = 0; col < 7; ++col) {
        int b = row * 8 + col;
        str += ((x >> b) & 1) ? 'x' : (((o >> b) & 1) ? 'o' : '.');
      }
      str += '\n';
    }
    return str;
  }
  static bool has_board_format() { return true; }
  static std::string serialize(const ConnectFour::State& s) { return to_string(s); }
  static int deserialize(ConnectFour::State& s, const char* str) {
    u64 x = 0, o = 0;
    int xs = 0, os = 0, r = 5, c = 0;
    for (const char* p = str; *p; ++p) {
      char ch = *p;
      if (ch != '.' && ch != 'x' && ch != 'o') continue;
      if (r < 0) return BSG_ERR_BAD_ARG;
      if (ch == 'x') { x |= 1ull << (r * 8 + c); ++xs; }
      if (ch == 'o') { o |= 1ull << (r * 8 + c); ++os; }
      if (++c >= 7) { --r; c = 0; }
    }
    if (!(xs == os || xs == os + 1) || r != -1 || c != 0) return BSG_ERR_BAD_ARG;
    // gravity: an empty cell below a stone cannot be packed (the marker encoding keeps only the
    // height of a column); the reference keeps such a board as written, no playout starts there
    const u64 occ = x | o;
    if ((occ >> 8) & ~occ & 0x00007F7F7F7F7F7Full) return BSG_ERR_BAD_ARG;
    int mv = (xs == os) ? 0 : 1;                    // connect_four.cc:270
    bool line_x = ConnectFour::has_line(x), line_o = ConnectFour::has_line(o);
    // a win is always the last mover's in a reachable position; the reference would also
    // accept boards where the side to move already has a line — no playout can start there
    if ((line_x && mv == 0) || (line_o && mv == 1) || (line_x && line_o)) return BSG_ERR_BAD_ARG;
    s.cur = mv ? o : x;
    s.opp = mv ? x : o;
    s.left = (u32)(42 - xs - os);
    s.win = (line_x || line_o) ? 1u : 0u;
    return BSG_OK;
  }
};
template <> struct Text<Breakthrough> {  // breakthrough.cc:262-282, :389-435
  static void boards(const Breakthrough::State& s, u64& black, u64& white) {
    black = s.mv ? s.opp : s.own; white = s.mv ? s.own : s.opp;
  }
  static std::string to_string(const Breakthrough::State& s) {
    u64 b, w;
    boards(s, b, w);
    std::string str;
    for (int r = 0; r < 8; ++r) {
      str += (char)('1' + (7 - r));
      for (int c = 0; c < 8; ++c) str += ((b >> (8 * r + c)) & 1) ? 'b' : (((w >> (8 * r + c)) & 1) ? 'w' : '.');
      str += '\n';
    }
    str += ' ';
    for (int c = 0; c < 8; ++c) str += (char)('a' + c);
    str += '\n';
    return str;
  }
  static bool has_board_format() { return true; }
  static std::string serialize(const Breakthrough::State& s) {
    u64 b, w;
    boards(s, b, w);
    std::string str;
    for (int i = 0; i < 64; ++i) str += ((b >> i) & 1) ? 'b' : (((w >> i) & 1) ? 'w' : '.');
    return str;
  }
  // The reference keeps the side to move of a fresh state (black) and sets no winner
  // (breakthrough.cc:399-434): the state is terminal only when a side has no pieces left.
  static int deserialize(Breakthrough::State& s, const char* str) {
    if (strlen(str) != 64) return BSG_ERR_BAD_ARG;
    u64 b = 0, w = 0;
    for (int i = 0; i < 64; ++i) {
      if (str[i] == 'b') b |= 1ull << i;
      else if (str[i] == 'w') w |= 1ull << i;
      else if (str[i] != '.') return BSG_ERR_BAD_ARG;
    }
    // A piece on its goal row means the game is over (breakthrough.cc:187-191).  The reference's
    // reader never restores winner_ (breakthrough.cc:399-435), so it reports such a board as
    // unfinished with that piece immobile; here the winner is restored instead — the move
    // generator relies on "no mover piece on the goal row in an unfinished game".
    bool b_home = (b >> 56) != 0, w_home = (w & 0xFFull) != 0;
    if (b_home && w_home) return BSG_ERR_BAD_ARG;
    s.own = b; s.opp = w; s.mv = 0;
    bool p0 = w == 0 || b_home, p1 = b == 0 || w_home;
    if (p0 && p1) return BSG_ERR_BAD_ARG;
    s.term = (p0 || p1) ? 1u : 0u;
    s.outc = p0 ? (u32)kOutcomeP0 : (p1 ? (u32)kOutcomeP1 : (u32)kOutcomeNone);
    return BSG_OK;
  }
};
template <int N> struct Text<HexT<N>> {  // hex.cc:229-246, cell characters hex.cc:123-146
  typedef HexT<N> H;
  static std::string to_string(const typename H::State& s0) {
    typename H::State s = s0;
    H::finish(s);
    const char names[9] = {'O', 'p', 'q', 'o', '.', 'x', 'z', 'y', 'X'};  // plane order of the observation
    std::string str;
    int line = 0;
    for (int cell = 0; cell < N * N; ++cell) {
      if (cell && !(cell % N)) { str += '\n'; ++line; str += std::string(line, ' '); }
      char ch = '?';
      for (int p = 0; p < 9; ++p)
        if ((H::plane_word(s, p, cell >> 5) >> (cell & 31)) & 1u) ch = names[p];
      str += ch;
      str += ' ';
    }
    return str;
  }
  static bool has_board_format() { return false; }
  static std::string serialize(const typename H::State&) { return std::string(); }
  static int deserialize(typename H::State&, const char*) { return BSG_ERR_BAD_ARG; }
};

template <> struct Text<Pentago> {  // pentago.cc:168-217 (plain output, no ANSI colours)
  static std::string to_string(const Pentago::State& s) {
    const u64 p0 = Pentago::mover(s) ? s.opp : s.cur, p1 = Pentago::mover(s) ? s.cur : s.opp;
    const char* left[6] = {"v", "s", " ", " ", "z", "^"};
    const char* right[6] = {"v", "v", " ", " ", "w", "^"};
    std::string out = "    > t     u <\n    a b c d e f\n";
    for (int y = 0; y < 6; ++y) {
      out += left[y];
      out += ' ';
      out += (char)('1' + y);
      out += ' ';
      for (int x = 0; x < 6; ++x) {
        const int pos = y * 8 + x;
        out += ((p0 >> pos) & 1ull) ? 'O' : (((p1 >> pos) & 1ull) ? '@' : '.');
        out += ' ';
      }
      out += right[y];
      out += '\n';
    }
    out += "    > y     x <\n";
    return out;
  }
  static bool has_board_format() { return false; }
  static std::string serialize(const Pentago::State&) { return std::string(); }
  static int deserialize(Pentago::State&, const char*) { return BSG_ERR_BAD_ARG; }
};

template <> struct Text<Othello> {  // othello.cc:271-282; labels :68-92
  static std::string to_string(const Othello::State& s) {
    const u64 black = s.mv ? s.opp : s.own, white = s.mv ? s.own : s.opp;
    // CurrentPlayer() is kTerminalPlayerId (-4) at the end, printed as a number (othello.cc:82-92)
    std::string str = s.term ? "-4" : (s.mv ? "White (o)" : "Black (x)");
    str += " to play:\n  a b c d e f g h  \n";
    for (int r = 0; r < 8; ++r) {
      str += (char)('1' + r);
      str += ' ';
      for (int c = 0; c < 8; ++c) {
        const int cell = r * 8 + c;
        str += ((black >> cell) & 1ull) ? 'x' : (((white >> cell) & 1ull) ? 'o' : '-');
        str += ' ';
      }
      str += (char)('1' + r);
      str += '\n';
    }
    str += "  a b c d e f g h  ";
    return str;
  }
  static bool has_board_format() { return false; }
  static std::string serialize(const Othello::State&) { return std::string(); }
  static int deserialize(Othello::State&, const char*) { return BSG_ERR_BAD_ARG; }
};

template <int N> struct Text<YT<N>> {  // y.cc:155-225 (plain output: ansi_color_output = false)
  typedef YT<N> G;
  static std::string to_string(const typename G::State& s) {
    std::string out = " ";
    for (int x = 0; x < N; ++x) { out += ' '; out += (char)('a' + x); }
    out += '\n';
    const int last = (int)s.last - 1;
    for (int y = 0; y < N; ++y) {
      out += std::string(y + ((y + 1) < 10 ? 1 : 0), ' ');   // leading space
      out += std::to_string(y + 1);
      bool found_last = false;
      for (int x = 0; x < N - y; ++x) {
        const int cell = x + y * N;
        // spacing and last-move highlight
        if (found_last) { out += ']'; found_last = false; }
        else if (cell == last) { out += '['; found_last = true; }
        else out += ' ';
        const bool stone = (s.A[cell >> 5] >> (cell & 31)) & 1u, p1 = (s.P[cell >> 5] >> (cell & 31)) & 1u;
        out += !stone ? '.' : (p1 ? '@' : 'O');
      }
      if (found_last) out += ']';
      out += '\n';
    }
    return out;
  }
  static bool has_board_format() { return false; }
  static std::string serialize(const typename G::State&) { return std::string(); }
  static int deserialize(typename G::State&, const char*) { return BSG_ERR_BAD_ARG; }
};

template <int S> struct Text<HavannahT<S>> {  // havannah.cc:197-267 (plain output: ansi_color_output = false)
  typedef HavannahT<S> G;
  static std::string to_string(const typename G::State& s) {
    const int D = G::D;
    std::string out(S + 3, ' ');
    for (int x = 0; x < S; ++x) { out += ' '; out += (char)('a' + x); }
    out += '\n';
    const int last = (int)s.last - 1;
    for (int y = 0; y < D; ++y) {
      const int indent = (S - 1 - y < 0 ? y - (S - 1) : S - 1 - y) + 1 + ((y + 1) < 10 ? 1 : 0);
      out += std::string(indent, ' ');
      out += std::to_string(y + 1);                          // leading y coordinate
      bool found_last = false;
      const int start_x = y < S ? 0 : y - S + 1, end_x = y < S ? S + y : D;
      for (int x = start_x; x < end_x; ++x) {
        const int cell = x + y * D;
        if (found_last) { out += ']'; found_last = false; }
        else if (cell == last) { out += '['; found_last = true; }
        else out += ' ';
        out += G::has(s.st[0], cell) ? 'O' : (G::has(s.st[1], cell) ? '@' : '.');
      }
      if (found_last) out += ']';
      if (y < S - 1) { out += ' '; out += (char)('a' + S + y); }   // trailing x coordinate
      out += '\n';
    }
    return out;
  }
  static bool has_board_format() { return false; }
  static std::string serialize(const typename G::State&) { return std::string(); }
  static int deserialize(typename G::State&, const char*) { return BSG_ERR_BAD_ARG; }
};

template <int N, int K2> struct Text<GoT<N, K2>> {   // go.cc:150-156 + go_board.cc:478-493
  typedef GoT<N, K2> G;
  static std::string to_string(const typename G::State& s) {
    std::string out = "GoState(komi=";
    out += std::to_string(K2 / 2);
    if (K2 & 1) out += ".5";
    out += ", to_play=";
    out += s.mv ? "W" : "B";
    out += ", history.size()=" + std::to_string(s.moves) + ")\n\n";
    for (int row = N - 1; row >= 0; --row) {
      if (row + 1 < 10) out += ' ';
      out += std::to_string(row + 1);
      out += ' ';
      for (int col = 0; col < N; ++col) {
        const int c = row * N + col;
        out += G::has(s.st[0], c) ? 'X' : (G::has(s.st[1], c) ? 'O' : '+');
      }
      out += '\n';
    }
    out += "   ";
    out += std::string("ABCDEFGHJKLMNOPQRST").substr(0, N);
    out += '\n';
    return out;
  }
  static bool has_board_format() { return false; }
  static std::string serialize(const typename G::State&) { return std::string(); }
  static int deserialize(typename G::State&, const char*) { return BSG_ERR_BAD_ARG; }
};

struct ToStringOp {
  const void* p; int64_t cap, i; int serialize; char* buf; int capacity;
  template <class G> int operator()() const {
    typename G::State s;
    G::load(s, p, cap, i);
    if (serialize && !Text<G>::has_board_format()) return BSG_ERR_BAD_ARG;
    std::string str = serialize ? Text<G>::serialize(s) : Text<G>::to_string(s);
    if ((int)str.size() + 1 > capacity) return BSG_ERR_BAD_ARG;
    memcpy(buf, str.c_str(), str.size() + 1);
    return (int)str.size();
  }
};
struct DeserializeOp {
  const char* str; void* p; int64_t cap, i;
  template <class G> int operator()() const {
    typename G::State s;
    G::init(s);
    int rc = Text<G>::deserialize(s, str);
    if (rc != BSG_OK) return rc;
    G::store(s, p, cap, i);
    return BSG_OK;
  }
};

}  // namespace bsg
