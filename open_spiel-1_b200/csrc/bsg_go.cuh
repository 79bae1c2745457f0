// go — reference: open_spiel/games/go.{h,cc}, open_spiel/games/go/go_board.{h,cc}
// 
//
//   N x N board, action id = row * N + col, pass = N * N (go.h:66-70); black = player 0 moves
//   first; no handicap.  A move is legal on an empty point that is not the ko point and that
//   has an empty neighbour, or joins a friendly chain that keeps a liberty, or captures
//   (GoBoard::IsLegalMove, go_board.cc:416-441); it captures every enemy chain it leaves without
//   liberties (:257-289, :336-352); a stone played into an enemy eye that captures exactly one
//   stone makes that point the ko point for one move (:262-283).  The game ends after two
//   passes, at 2 N^2 moves, or when a non-pass move recreates an earlier position (superko, a
//   draw: go.cc:166-172, 204-211); otherwise Tromp-Taylor area scoring with komi decides
//   (go_board.cc:525-598).
//
//   The reference keeps chains as linked lists with pseudo-liberty sums.  Here a position is two
//   N^2-bit sets; a chain is a 4-neighbourhood flood of a set, its liberties one more dilation.
//   Legality is bit-parallel for every point with an empty neighbour; for fully surrounded
//   points one flood finds every chain with a liberty that is not such a point, and only the
//   chains that are left are flooded one by one, each once, up to the liberty count that decides.
//   Two things have to be reproduced, not just re-derived:
//    * superko needs every earlier position of the game: the state carries an open-addressed set
//      of their 64-bit hashes (any collision-free hash gives the reference's answers; its own is
//      a Zobrist table of another library), and a PACKED state carries the move history and is
//      replayed at load — by the playout kernel in lockstep with the other lanes' steps
//      (kReplayLoad, load_lazy);
//    * the reference's territory search stops exploring a region as soon as it has seen both
//      colours and leaves the rest unmarked, so a later start inside the same region may see only
//      one colour and count that part (go_board.cc:525-553): the order-dependent search is
//      emulated frame by frame with an explicit stack.
//   `in_atari` of the reference is exact ("one distinct liberty") as long as its 16 / 32-bit
//   pseudo-liberty sums cannot wrap, which holds up to 9 x 9 (vertex ids <= 198); on 19 x 19 its
//   uint16 sum wraps for large chains.  Together with the uint8 move counts of the playout ABI
//   (a 19 x 19 game has up to 722 moves) that is why the device game is go(board_size=9).
//   packed: SoA, plane 0 / 1: black / white stones (4 words each), plane 2: u32 meta = mover |
//   last action was a pass << 1 | two passes << 2 | superko << 3 | (ko point + 1) << 8 | moves <<
//   16 | outcome << 24, plane 3: the move history, one byte per action (kHistBytes).
#pragma once
#include "bsg_common.cuh"

#ifndef BSG_GO_ROLLED
#define BSG_GO_ROLLED 1
#endif
#ifndef BSG_GO_MINBLOCKS
#define BSG_GO_MINBLOCKS 1
#endif

namespace bsg {

template <int N, int K2>   // K2 = 2 * komi
struct GoT {
  static constexpr int kId = kGameGo;
  static constexpr int kCells = N * N;
  static constexpr int kPass = N * N;
  static constexpr int W = (kCells + 31) / 32;
  static_assert(W <= 4 && kCells <= 120, "written for boards up to 10 x 10");
  static constexpr int kNumActions = kCells + 1, kMaxLen = 2 * kCells, kMaxDepth = kMaxLen, kObsSize = 4 * kCells;
  static constexpr int kHistBytes = (kMaxLen + 15) / 16 * 16;
  static constexpr int kSeenSlots = kMaxLen <= 85 ? 128 : (kMaxLen <= 170 ? 256 : 512);   // load factor <= 2/3
  static constexpr int kStateBytes = 16 + 16 + 4 + kHistBytes;
  static constexpr int kObsWords = (kObsSize + 31) / 32, kMaskWords = (kNumActions + 31) / 32;
  static constexpr int kPlanes = 4;
  BSG_HD static int plane_bytes(int p) { return p < 2 ? 16 : (p == 2 ? 4 : kHistBytes); }
  static constexpr bool kCanStick = false;      // a pass is always legal
  static constexpr bool kAsync = false;
  static constexpr bool kMoverCanLose = true;   // the winner is decided by counting
  static constexpr bool kLazyEnd = false;
  static constexpr bool kLocalHeavy = true;   // playout kernel: L1 over shared memory (the state is local memory)
  static constexpr bool kRolledGroup = BSG_GO_ROLLED;   // the step loop of the playout kernel is not unrolled
  static constexpr int kMinBlocks = BSG_GO_MINBLOCKS;   // lean kernel: resident blocks per SM asked of the compiler
  struct State {
    u32 st[2][W];                  // black / white stones
    u32 mv;                        // player to move (0 black)
    u32 ko;                        // ko point + 1 (0: none)
    u32 moves;                     // actions made (passes included)
    u32 last_pass, two_pass, superko;
    u32 replay, replay_on;         // playout kernel only: history moves still to replay / this group replays (load_lazy)
    u32 outc;                      // kOutcomeNone until the game is over
    u64 hash;                      // of the stones on the board
    u64 seen[kSeenSlots];          // open-addressed set of the hashes of every earlier non-empty position (0: free)
    __align__(4) unsigned char acts[kHistBytes];
  };
  // The playout kernel does not replay a root's history inside the claim (a few lanes at work,
  // the rest of the warp waiting): load_lazy() only copies the history and the lane then replays
  // one move per iteration of the step loop, beside the other lanes' ordinary steps, in whole
  // groups so that its first real step starts a Philox block (see playout_kernel).
  static constexpr bool kReplayLoad = true;
  struct Tables {
    u32 modM[128];
    unsigned char sel8[256][8];
  };
  static void fill_tables(Tables& t) {
    for (u32 n = 0; n < 128; ++n) t.modM[n] = modM_entry(n);
    for (int b = 0; b < 256; ++b) {
      int k = 0;
      for (int i = 0; i < 8; ++i) t.sel8[b][i] = 0;
      for (int i = 0; i < 8; ++i) if ((b >> i) & 1) t.sel8[b][k++] = (unsigned char)i;
    }
  }
  // Compile-time masks of word w: board cells (kind 0), cells not in column 0 (1), not in column N-1 (2).
  BSG_HDC static u32 word_mask(int w, int kind) {
    u32 m = 0;
    for (int b = 0; b < 32; ++b) {
      const int cell = 32 * w + b;
      if (cell >= kCells) break;
      const int col = cell % N;
      if (kind == 0 || (kind == 1 ? col != 0 : col != N - 1)) m |= 1u << b;
    }
    return m;
  }
  BSG_HD static u32 board_word(int w) { constexpr u32 m[4] = {word_mask(0, 0), word_mask(1, 0), word_mask(2, 0), word_mask(3, 0)}; return m[w]; }
  BSG_HD static u32 not_col0(int w) { constexpr u32 m[4] = {word_mask(0, 1), word_mask(1, 1), word_mask(2, 1), word_mask(3, 1)}; return m[w]; }
  BSG_HD static u32 not_col_last(int w) { constexpr u32 m[4] = {word_mask(0, 2), word_mask(1, 2), word_mask(2, 2), word_mask(3, 2)}; return m[w]; }
  // Word / bit access by selects: a set indexed with a run-time word number would be forced out
  // of registers into local memory (every lane of the playout kernel carries several sets).
  BSG_HD static u32 word_of(const u32* set, int w) {
    u32 x = set[0];
#pragma unroll
    for (int j = 1; j < W; ++j) x = (w == j) ? set[j] : x;
    return x;
  }
  BSG_HD static bool has(const u32* set, int cell) { return (word_of(set, cell >> 5) >> (cell & 31)) & 1u; }
  BSG_HD static void set_bit(u32* set, int cell) {
    const u32 b = 1u << (cell & 31);
    const int w = cell >> 5;
#pragma unroll
    for (int j = 0; j < W; ++j) set[j] |= (w == j) ? b : 0u;
  }
  BSG_HD static void clear_bit(u32* set, int cell) {
    const u32 b = 1u << (cell & 31);
    const int w = cell >> 5;
#pragma unroll
    for (int j = 0; j < W; ++j) set[j] &= (w == j) ? ~b : ~0u;
  }
  // Lowest cell of a non-empty set.
  BSG_HD static int first_cell(const u32* set) {
    int c = 0;
    bool found = false;
#pragma unroll
    for (int j = 0; j < W; ++j) {
      const bool here = !found && set[j] != 0u;
      c = here ? 32 * j + ctz32(set[j]) : c;
      found = found || here;
    }
    return c;
  }
  BSG_HD static void stones_of(const State& s, u32 player, u32* out) {
#pragma unroll
    for (int w = 0; w < W; ++w) out[w] = player ? s.st[1][w] : s.st[0][w];
  }
  BSG_HD static int count(const u32* set) { int n = 0; for (int w = 0; w < W; ++w) n += popc32(set[w]); return n; }
  BSG_HD static bool any(const u32* set) { u32 a = 0; for (int w = 0; w < W; ++w) a |= set[w]; return a != 0; }

  // The four neighbours of every cell of f (go_board.cc:28-34), on the board.
  BSG_HD static void nb4(const u32* f, u32* out) {
    u32 e[W], x[W], up[W], dn[W];
    shl_words<1, W>(f, e);
    shr_words<1, W>(f, x);
    shl_words<N, W>(f, up);
    shr_words<N, W>(f, dn);
    for (int w = 0; w < W; ++w)
      out[w] = ((e[w] & not_col0(w)) | (x[w] & not_col_last(w)) | up[w] | dn[w]) & board_word(w);
  }
  // The chain of `set` that contains the cell — a 4-neighbourhood flood — stopped as soon as the
  // chain is known to have `enough` liberties: returns min(liberties, enough).  `out` is the whole chain only when the result is below `enough`;
  // otherwise it is the part flooded so far (still stones of that one chain).
  BSG_HD static int chain_liberties_capped(const u32* set, int cell, const u32* empty, int enough, u32* out) {
    for (int w = 0; w < W; ++w) out[w] = 0;
    set_bit(out, cell);
    for (;;) {
      u32 d[W];
      nb4(out, d);
      int libs = 0;
      u32 grew = 0;
      for (int w = 0; w < W; ++w) {
        libs += popc32(d[w] & empty[w]);
        const u32 g = (out[w] | d[w]) & set[w];
        grew |= g ^ out[w];
        out[w] = g;
      }
      if (libs >= enough) return enough;
      if (!grew) return libs;
    }
  }
  // Neighbour of a cell in the reference's order (row + 1, col + 1, col - 1, row - 1), -1 off the board.
  BSG_HD static int neighbour(int cell, int d) {
    const int r = cell / N, c = cell - r * N;
    const int r2 = r + (d == 0 ? 1 : (d == 3 ? -1 : 0)), c2 = c + (d == 1 ? 1 : (d == 2 ? -1 : 0));
    return (r2 >= 0 && r2 < N && c2 >= 0 && c2 < N) ? r2 * N + c2 : -1;
  }
  BSG_HD static u64 zobrist(int cell, u32 colour) {     // splitmix64 of (cell, colour): collision-free in practice
    u64 z = (u64)(2 * cell + (int)colour + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }

  BSG_HD static void empties(const State& s, u32* e) {
    for (int w = 0; w < W; ++w) e[w] = ~(s.st[0][w] | s.st[1][w]) & board_word(w);
  }
  // GoBoard::IsLegalMove for every point at once (go_board.cc:416-441), for the player to move.
  BSG_HD static void legal_points(const State& s, u32* out) {
    u32 e[W], ne[W];
    empties(s, e);
    nb4(e, ne);
    u32 mine[W], theirs[W];
    stones_of(s, s.mv, mine);
    stones_of(s, s.mv ^ 1u, theirs);
    u32 rest[W];
    for (int w = 0; w < W; ++w) { out[w] = e[w] & ne[w]; rest[w] = e[w] & ~ne[w]; }
    // points without an empty neighbour: legal next to a friendly chain that is not in atari or
    // next to an enemy chain that is (its one liberty is then this point).  Neither property
    // depends on the point, so every chain that touches such a point is flooded ONCE and the
    // verdicts are applied to all of them together.
    if (any(rest)) {
      u32 cand[W], known[W], good[W];
      nb4(rest, cand);
      for (int w = 0; w < W; ++w) { cand[w] &= mine[w] | theirs[w]; known[w] = 0; good[w] = 0; }
      // First every chain with a liberty that is NOT a surrounded point, both colours in one flood
      // that all lanes of a warp run together: such a friendly chain has that liberty and the
      // point (>= 2), such an enemy chain is not captured there.  What is left for the per-chain
      // loop are chains whose liberties are all surrounded points.
      {
        u32 fa[W], fb[W], nbo[W];
        nb4(out, nbo);                                   // out: the empty points with an empty neighbour
        for (int w = 0; w < W; ++w) { fa[w] = mine[w] & nbo[w]; fb[w] = theirs[w] & nbo[w]; }
        for (;;) {
          u32 da[W], db[W];
          nb4(fa, da);
          nb4(fb, db);
          u32 grew = 0;
          for (int w = 0; w < W; ++w) {
            const u32 na = (fa[w] | da[w]) & mine[w], nb = (fb[w] | db[w]) & theirs[w];
            grew |= (na ^ fa[w]) | (nb ^ fb[w]);
            fa[w] = na;
            fb[w] = nb;
          }
          if (!grew) break;
        }
        for (int w = 0; w < W; ++w) { known[w] = fa[w] | fb[w]; good[w] = fa[w]; }
      }
      for (;;) {
        u32 todo[W];
        for (int w = 0; w < W; ++w) todo[w] = cand[w] & ~known[w];
        if (!any(todo)) break;
        const int q = first_cell(todo);
        const bool friendly = has(mine, q);
        u32 own[W], ch[W];
        for (int w = 0; w < W; ++w) own[w] = friendly ? mine[w] : theirs[w];
        const int libs = chain_liberties_capped(own, q, e, 2, ch);   // a partial chain is fine: the
        const bool ok = friendly ? libs >= 2 : libs == 1;            // rest of it is looked at again
        for (int j = 0; j < W; ++j) { known[j] |= ch[j]; good[j] |= ok ? ch[j] : 0u; }
      }
      u32 ng[W];
      nb4(good, ng);
      for (int w = 0; w < W; ++w) out[w] |= rest[w] & ng[w];
    }
    if (s.ko) clear_bit(out, (int)s.ko - 1);
  }

  // NumSurroundedPoints / TrompTaylorScore (go_board.cc:525-598): black's area minus white's,
  // with the reference's order-dependent early stop; returns 2 * score (komi included).
  BSG_HD static int score2(const State& s) {
    // Regions of empty points that touch one colour only (or none) are searched to the end by the
    // reference whatever the order, so they are counted with two floods; only the regions that
    // touch BOTH colours go through the order-dependent search below.
    u32 e[W], eb[W], ew[W], mark[W];
    empties(s, e);
    nb4(s.st[0], eb);
    nb4(s.st[1], ew);
    for (int w = 0; w < W; ++w) { eb[w] &= e[w]; ew[w] &= e[w]; }
    for (;;) {
      u32 db[W], dw[W];
      nb4(eb, db);
      nb4(ew, dw);
      u32 grew = 0;
      for (int w = 0; w < W; ++w) {
        const u32 nb = (eb[w] | db[w]) & e[w], nw = (ew[w] | dw[w]) & e[w];
        grew |= (nb ^ eb[w]) | (nw ^ ew[w]);
        eb[w] = nb;
        ew[w] = nw;
      }
      if (!grew) break;
    }
    int delta = count(s.st[0]) - count(s.st[1]);
    u32 both[W];
    for (int w = 0; w < W; ++w) {
      delta += popc32(eb[w] & ~ew[w]) - popc32(ew[w] & ~eb[w]);
      both[w] = eb[w] & ew[w];
      mark[w] = 0;
    }
    unsigned char cell[kCells + 1], dir[kCells + 1];
    for (;;) {
      u32 todo[W];
      for (int w = 0; w < W; ++w) todo[w] = both[w] & ~mark[w];
      if (!any(todo)) break;
      const int p = first_cell(todo);                // the reference's scan order: ascending points
      bool rb = false, rw = false;
      int n = 1, top = 0;
      set_bit(mark, p);
      cell[0] = (unsigned char)p; dir[0] = 0;
      while (top >= 0) {
        if (dir[top] == 4) { --top; continue; }
        const int q = neighbour(cell[top], dir[top]++);
        if (q < 0) continue;                       // the border is ignored
        if (has(s.st[0], q)) rb = true;
        else if (has(s.st[1], q)) rw = true;
        else if (!has(mark, q) && !(rb && rw)) {   // the recursion's entry test
          set_bit(mark, q);
          ++n;
          ++top;
          cell[top] = (unsigned char)q; dir[top] = 0;
        }
      }
      if (rb && !rw) delta += n;
      else if (!rb && rw) delta -= n;
    }
    return 2 * delta - K2;
  }

  BSG_HD static bool terminal(const State& s) {          // go.cc:166-172
    return s.moves >= 2u && (s.moves >= (u32)kMaxLen || s.superko || s.two_pass);
  }
  BSG_HD static void settle(State& s) {                  // go.cc:174-199
    if (!terminal(s)) { s.outc = kOutcomeNone; return; }
    if (s.superko) { s.outc = kOutcomeDraw; return; }
    const int sc = score2(s);
    s.outc = sc > 0 ? kOutcomeP0 : (sc < 0 ? kOutcomeP1 : kOutcomeDraw);
  }
  // GoState::DoApplyAction + GoBoard::PlayMove (go.cc:201-211, go_board.cc:257-289); the action
  // must be legal.
  BSG_HD static void play(State& s, int a) {
    const u32 me = s.mv;
    if (s.moves < (u32)kHistBytes) s.acts[s.moves] = (unsigned char)a;
    if (a == kPass) {
      s.ko = 0;
      s.two_pass = s.last_pass;
      s.last_pass = 1;
    } else {
      u32 mine[W], theirs[W], e[W];
      stones_of(s, me, mine);
      stones_of(s, me ^ 1u, theirs);
      empties(s, e);
      u32 at[W], cand[W], ne[W];
      for (int w = 0; w < W; ++w) at[w] = 0;
      set_bit(at, a);
      nb4(at, cand);                                      // the neighbours of the new stone
      // played in an enemy eye: no neighbour is friendly or empty (the border counts as enemy)
      u32 open = 0;
      for (int w = 0; w < W; ++w) open |= cand[w] & (mine[w] | e[w]);
      const bool eye = open == 0u;
      for (int w = 0; w < W; ++w) { mine[w] |= at[w]; e[w] &= ~at[w]; }
      s.hash ^= zobrist(a, me);
      int captured = 0, cap_cell = 0;
      // enemy stones next to the new one that have no liberty of their own: only their chains can
      // have been left without liberties (the order of the captures does not matter: the hash is
      // a XOR and the ko point is only set when exactly one stone was taken)
      nb4(e, ne);
      for (int w = 0; w < W; ++w) cand[w] &= theirs[w] & ~ne[w];
      while (any(cand)) {
        const int q = first_cell(cand);
        u32 ch[W];
        const int libs = chain_liberties_capped(theirs, q, e, 1, ch);
        for (int w = 0; w < W; ++w) cand[w] &= ~ch[w];
        if (libs != 0) continue;
        for (int w = 0; w < W; ++w) { theirs[w] &= ~ch[w]; e[w] |= ch[w]; }   // RemoveChain
        while (any(ch)) {
          cap_cell = first_cell(ch);
          clear_bit(ch, cap_cell);
          s.hash ^= zobrist(cap_cell, me ^ 1u);
          ++captured;
        }
      }
      for (int w = 0; w < W; ++w) {
        s.st[0][w] = me ? theirs[w] : mine[w];
        s.st[1][w] = me ? mine[w] : theirs[w];
      }
      s.ko = (eye && captured == 1) ? (u32)cap_cell + 1u : 0u;
      s.last_pass = 0;
      // superko: a non-pass move that recreates an earlier position (never the empty board: a
      // stone was just played).  One probe sequence both looks the new position up and files it;
      // at most kMaxLen positions are ever filed, so a free slot always ends the sequence.
      const u64 key = s.hash | (u64)(s.hash == 0);             // 0 marks a free slot
      for (u32 slot = (u32)(key >> 40);; ++slot) {
        const u64 k = s.seen[slot & (u32)(kSeenSlots - 1)];
        if (k == key) { s.superko = 1; break; }
        if (k == 0) { s.seen[slot & (u32)(kSeenSlots - 1)] = key; break; }
      }
    }
    s.moves += 1u;
    s.mv = me ^ 1u;
    settle(s);
  }

  BSG_HD static void init(State& s) {
    for (int w = 0; w < W; ++w) s.st[0][w] = s.st[1][w] = 0;
    s.mv = 0; s.ko = 0; s.moves = 0; s.last_pass = 0; s.two_pass = 0; s.superko = 0;
    s.replay = 0; s.replay_on = 0;
    s.outc = kOutcomeNone;
    s.hash = 0;
    for (int i = 0; i < kSeenSlots; ++i) s.seen[i] = 0;
    for (int i = 0; i < kHistBytes; ++i) s.acts[i] = 0;
  }
  // A packed state is its move history (plane 3): replaying it rebuilds the position AND the set
  // of earlier positions superko needs.  Planes 0-2 are what the observation writer and the
  // host-side accessors read without replaying; load() cross-checks nothing, it trusts plane 3.
  BSG_HD static void load(State& s, const void* base, int64_t n, int64_t i) {
    const u32 meta = ((const u32*)((const char*)base + (int64_t)32 * n))[i];
    const unsigned char* h = (const unsigned char*)base + (int64_t)36 * n + i * kHistBytes;
    const u32 moves = (meta >> 16) & 0xFFu;
    init(s);
    for (u32 k = 0; k < moves && k < (u32)kMaxLen; ++k) play(s, h[k]);
  }
  // The lazy load: an empty board that still has `moves` history moves to replay; what the kernel
  // needs of the ROOT at once (its mover and outcome) comes from the meta word.
  BSG_HD static void load_lazy(State& s, const void* base, int64_t n, int64_t i, u32& root_mover, u32& root_outcome) {
    const u32 meta = ((const u32*)((const char*)base + (int64_t)32 * n))[i];
    const u32* h = (const u32*)((const char*)base + (int64_t)36 * n + i * kHistBytes);
    u32 moves = (meta >> 16) & 0xFFu;
    moves = moves < (u32)kMaxLen ? moves : (u32)kMaxLen;
    init(s);
    u32* dst = (u32*)s.acts;
    for (u32 k = 0; 4u * k < moves; ++k) dst[k] = h[k];
    s.replay = moves;
    root_mover = meta & 1u;
    root_outcome = meta >> 24;
  }
  // Called at the start of every group: does this lane spend the group replaying?
  BSG_HD static u32 group_begin(State& s) { s.replay_on = s.replay ? 1u : 0u; return s.replay_on; }
  BSG_HD static void store(const State& s, void* base, int64_t n, int64_t i) {
    u32* p0 = (u32*)base + i * 4;
    u32* p1 = (u32*)((char*)base + (int64_t)16 * n) + i * 4;
    for (int w = 0; w < 4; ++w) { p0[w] = w < W ? s.st[0][w] : 0u; p1[w] = w < W ? s.st[1][w] : 0u; }
    ((u32*)((char*)base + (int64_t)32 * n))[i] = s.mv | (s.last_pass << 1) | (s.two_pass << 2) | (s.superko << 3) |
                                                 (s.ko << 8) | (s.moves << 16) | (s.outc << 24);
    unsigned char* h = (unsigned char*)base + (int64_t)36 * n + i * kHistBytes;
    for (int k = 0; k < kHistBytes; ++k) h[k] = k < (int)s.moves ? s.acts[k] : (unsigned char)0;
  }
  BSG_HD static bool won(const State& s) {               // the LAST mover won
    return terminal(s) && s.outc == (s.mv ^ 1u);
  }
  BSG_HD static bool lost(const State& s) { return terminal(s) && s.outc == s.mv; }
  BSG_HD static int mover(const State& s) { return (int)s.mv; }
  BSG_HD static int outcome(const State& s) { return terminal(s) ? (int)s.outc : kOutcomeNone; }

  // k-th (ascending) set bit of a W-word set, k < count
  BSG_HD static int select_bit(const u32* set, u32 k, const Tables& tab) {
    u32 x = set[0], below = 0, pos = 0, cum = 0;
    for (int w = 0; w + 1 < W; ++w) {
      cum += (u32)popc32(set[w]);
      const bool g = k >= cum;
      x = g ? set[w + 1] : x;
      below = g ? cum : below;
      pos = g ? 32u * (u32)(w + 1) : pos;
    }
    u32 kk = k - below;
    u32 h = (u32)popc32(x & 0xFFFFu);
    if (kk >= h) { kk -= h; x >>= 16; pos += 16; }
    h = (u32)popc32(x & 0xFFu);
    if (kk >= h) { kk -= h; x >>= 8; pos += 8; }
    const u32 cell = pos + tab.sel8[x & 0xFFu][kk & 7u];
    return (int)(cell < (u32)kCells ? cell : 0u);
  }
  BSG_HD static u32 step(State& s, u32 r, const Tables& tab, int& action) {
    // One call of play() serves the lanes that step and the lanes that replay (load_lazy): the
    // replaying ones sit out the legality search and join the others for the move itself.
    int a = -1;                                          // -1: nothing to play in this iteration
    u32 n = 1u;
    if (s.replay_on) {                                   // a replaying group: one history move
      if (s.replay) { a = s.acts[s.moves]; --s.replay; }
    } else if (!terminal(s)) {                           // (a finished lane runs dead iterations)
      u32 L[W];
      legal_points(s, L);
      const u32 np = (u32)count(L);
      n = np + 1u;                                       // go.cc:127-137: the legal points, then the pass
      const u32 k = fastmod(r, n, tab.modM[n & 127u]);
      a = k == np ? kPass : select_bit(L, k, tab);
    }
    if (a >= 0) play(s, a);
    action = a >= 0 ? a : kPass;
    return n;
  }
  BSG_HD static int legal(const State& s, int* out) {
    if (terminal(s)) return 0;
    u32 L[W];
    legal_points(s, L);
    int n = 0;
    for (int c = 0; c < kCells; ++c) if (has(L, c)) out[n++] = c;
    out[n++] = kPass;
    return n;
  }
  BSG_HD static bool apply(State& s, int a) {
    if (terminal(s) || a < 0 || a > kPass) return false;
    if (a != kPass) {
      u32 L[W];
      legal_points(s, L);
      if (!has(L, a)) return false;
    }
    play(s, a);
    return true;
  }
  BSG_HD static u32 legal_bit(const State& s, int a) {
    if (terminal(s)) return 0u;
    if (a == kPass) return 1u;
    u32 L[W];
    legal_points(s, L);
    return has(L, a) ? 1u : 0u;
  }
  // go.cc:104-125: planes black, white, empty, then a plane of ones when white is to play; the
  // observer does not matter.
  BSG_HD static u32 plane_word(const State& s, int plane, int w) {
    if (plane < 2) return s.st[plane][w];
    if (plane == 2) return ~(s.st[0][w] | s.st[1][w]) & board_word(w);
    return s.mv ? board_word(w) : 0u;
  }
  BSG_HD static u32 obs_bit(const State& s, int player, int idx) {
    (void)player;
    const int plane = idx / kCells, cell = idx - plane * kCells;
    return (plane_word(s, plane, cell >> 5) >> (cell & 31)) & 1u;
  }
  BSG_HD static void obs_words(const State& s, int player, u32* out) {
    (void)player;
    for (int w = 0; w < kObsWords; ++w) out[w] = 0;
    for (int p = 0; p < 4; ++p) {
      const int off = p * kCells, wi = off >> 5, sh = off & 31;
      for (int j = 0; j < W; ++j) {
        const u32 m = plane_word(s, p, j);
        if (wi + j < kObsWords) out[wi + j] |= m << sh;
        if (sh != 0 && wi + j + 1 < kObsWords) out[wi + j + 1] |= m >> ((32 - sh) & 31);
      }
    }
  }
  BSG_HD static void mask_words(const State& s, u32* out) {
    for (int w = 0; w < kMaskWords; ++w) out[w] = 0;
    if (terminal(s)) return;
    u32 L[W];
    legal_points(s, L);
    for (int w = 0; w < W; ++w) out[w] = L[w];
    out[kPass >> 5] |= 1u << (kPass & 31);
  }
};
typedef GoT<9, 15> Go;       // go(board_size=9), komi 7.5 (the reference's default komi, go.h:34)

}  // namespace bsg
